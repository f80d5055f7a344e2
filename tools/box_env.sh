#!/bin/bash
# What the GPU box offers (toolchains, cgroups, namespaces, topology). Output is a plain text report.
echo "== nproc: $(nproc)"; echo "== uname: $(uname -r)"
for t in go javac node kind kubectl docker nsenter unshare mknod; do printf "%s: " $t; command -v $t || echo absent; done
go version 2>&1 | head -1
echo "== id: $(id)"
echo "== cgroup mounts"; grep cgroup /proc/mounts | head -20
echo "== /sys/fs/cgroup"; ls /sys/fs/cgroup | head -30
echo "== self cgroup"; cat /proc/self/cgroup | head -20
echo "== devices cgroup writable?"; ls -la /sys/fs/cgroup/devices 2>&1 | head -8
(mkdir /sys/fs/cgroup/devices/gmp_probe_test 2>&1 && echo "mkdir devices cgroup OK" && rmdir /sys/fs/cgroup/devices/gmp_probe_test) || echo "mkdir devices cgroup FAILED"
echo "== unshare -m"; unshare -m true 2>&1 && echo "unshare -m OK"
echo "== mknod test"; (mknod -m 666 /tmp/gmp_nvtest c 195 254 2>&1 && ls -la /tmp/gmp_nvtest && rm /tmp/gmp_nvtest) || echo "mknod FAILED"
echo "== capabilities"; grep Cap /proc/self/status
echo "== /dev/nvidia*"; ls -la /dev/nvidia* 2>&1 | head -30
echo "== nvidia-smi"; nvidia-smi -L; nvidia-smi --query-gpu=index,uuid,pci.bus_id,persistence_mode,clocks.max.sm,clocks.max.mem --format=csv
nvidia-smi topo -m 2>&1 | head -30
echo "== libs"; ldconfig -p | grep -E "libcuda.so|libnvidia-ml" 
echo "== mem"; free -g | head -2
