#!/usr/bin/env python
"""Generate tests/golden/api_filedescriptor.pb from the reference (run in the authoring container only).

protoc-gen-go embeds the gzipped FileDescriptorProto of api.proto in pkg/api/gpu-mount/api.pb.go
(`var fileDescriptor_00212fb1f9d3bf1c = []byte{...}`, ref: api.pb.go near line 300).  That blob is the one
machine-checkable statement of the wire contract the reference itself holds, so it is extracted once, committed as a
fixture, and tests/test_api_wire.py compares the hand-built descriptor in gpumounter_b200/api.py against it.
"""
import gzip
import os
import re
import sys

ref = sys.argv[1] if len(sys.argv) > 1 else "/root/reference"
src = open(os.path.join(ref, "pkg/api/gpu-mount/api.pb.go")).read()
m = re.search(r"var fileDescriptor_[0-9a-f]+ = \[\]byte\{(.*?)\n\}", src, re.S)
body = re.sub(r"//.*", "", m.group(1))
raw = bytes(int(x, 16) for x in re.findall(r"0x([0-9a-fA-F]{2})", body))
out = gzip.decompress(raw)
dst = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden", "api_filedescriptor.pb")
open(dst, "wb").write(out)
print(f"{len(raw)} gzipped bytes -> {len(out)} bytes -> {dst}")
