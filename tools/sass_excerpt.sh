#!/bin/bash
# The memory instructions of the shipped kernels, from the library that is actually loaded (cuobjdump -sass).
# Usage: bash tools/sass_excerpt.sh > profiles/r02_sass_excerpt.txt      (needs no GPU)
LIB=${1:-gpumounter_b200/libgpumounter_probe.so}
echo "# cuobjdump -sass $LIB : per kernel, the count of each memory / bulk-copy / barrier mnemonic and the first line of each"
echo "# architectures in the image: $(cuobjdump -lelf $LIB | sed 's/.*\.sm_/sm_/' | sort -u | tr '\n' ' ')"
cuobjdump -sass "$LIB" | awk '
  /Function :/ { fn=$3; next }
  /(LDG|STG|LDS|STS|UBLKCP|SYNCS|ATOMG|RED|MEMBAR|FENCE|DEPBAR|BAR\.SYNC|CCTL)/ {
      match($0, /(LDG|STG|LDS|STS|UBLKCP|SYNCS|ATOMG|RED|MEMBAR|FENCE|DEPBAR|BAR\.SYNC|CCTL)[A-Za-z0-9_.]*/);
      m=substr($0, RSTART, RLENGTH); key=fn SUBSEP m; cnt[key]++;
      if (!(key in first)) { line=$0; sub(/^[ \t]*\/\*[0-9a-f]+\*\/[ \t]*/, "", line); sub(/[ \t]*\/\*.*$/, "", line); first[key]=line }
  }
  END { for (k in cnt) { split(k, a, SUBSEP); printf "%s\t%s\t%d\t%s\n", a[1], a[2], cnt[k], first[k] } }' | sort | c++filt -p 2>/dev/null | awk -F'\t' '
  { if ($1 != last) { print ""; print $1; last=$1 } printf "    %-28s x%-4d %s\n", $2, $3, $4 }'
