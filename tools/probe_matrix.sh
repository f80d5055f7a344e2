#!/bin/bash
# The HBM probe per pattern and per buffer size (SURVEY 8d metric 2): 1 warm-up + 10 timed iterations, best and median.
P=gpumounter_b200/gpumounter-probe
echo "["
first=1
for gib in 1 2 4 8; do
  for pat in 0 1 2 3 4; do
    [ $first -eq 0 ] && echo ","
    first=0
    $P --bytes ${gib}G --iters 10 --patterns $((1 << pat)) --dma-bytes 0 --skip-p2p --skip-nvml | python -c "
import sys,json; j=json.loads(sys.stdin.read()); h=j['hbm'][0]
print(json.dumps({'gib': $gib, 'pattern': ['splitmix','zero','ones','checker','address'][$pat], 'status': j['status'], **{k: h[k] for k in ('write_gbs','write_gbs_median','read_gbs','read_gbs_median','copy_gbs','copy_gbs_median','mismatch_lanes')}}), end='')"
  done
done
echo "]"
