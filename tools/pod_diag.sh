#!/bin/bash
# What the stand-in pod (tools/standin_pod.py, pod_like) will meet on this box: how the driver creates device nodes,
# which majors it registered, what the box's own device cgroup allows.  Plain text report.
echo "== /proc/driver/nvidia/params"; grep -E "ModifyDeviceFiles|DeviceFile" /proc/driver/nvidia/params 2>&1
echo "== /proc/devices (nvidia)"; grep -i nvidia /proc/devices
echo "== /dev/nvidia*"; ls -la /dev/nvidia* /dev/nvidia-caps 2>&1 | head -40
echo "== /proc/driver/nvidia/gpus"; ls /proc/driver/nvidia/gpus 2>&1; for f in /proc/driver/nvidia/gpus/*/information; do echo "-- $f"; cat "$f"; done 2>&1 | head -60
echo "== self cgroup (devices)"; grep devices /proc/self/cgroup; cat /sys/fs/cgroup/devices/devices.list 2>&1 | head -40
echo "== pid namespace"; readlink /proc/self/ns/pid /proc/1/ns/pid; echo "pid of this shell: $$"
echo "== nvidia-smi"; nvidia-smi -L; nvidia-smi --query-gpu=index,uuid,pci.bus_id,persistence_mode --format=csv,noheader
echo "== which"; for t in nvidia-smi nvidia-modprobe strace go; do printf "%s: " $t; command -v $t || echo absent; done
echo "== env"; env | grep -E "CUDA|NVIDIA_VIS"
