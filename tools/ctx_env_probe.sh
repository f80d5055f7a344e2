#!/bin/bash
# How long does primary-context creation take under different driver environment settings?
P=gpumounter_b200/gpumounter-probe
run() { echo "== $1"; for i in 1 2 3; do env $1 $P --bytes 256M --iters 1 --dma-bytes 0 --skip-copy --skip-nvml | python -c "
import sys,json; j=json.loads(sys.stdin.read()); h=j['hbm'][0]; print('  cuda_init', j['ms_cuda_init'], 'context', h['ms_context'], 'alloc', h['ms_alloc'], 'first', h['ms_first_kernel'], 'to_usable', j['ms_to_usable'])"; done; }
run "X=1"
run "CUDA_DEVICE_MAX_CONNECTIONS=1"
run "CUDA_DEVICE_MAX_CONNECTIONS=32"
run "CUDA_MODULE_LOADING=EAGER"
run "CUDA_DEVICE_MAX_COPY_CONNECTIONS=1 CUDA_DEVICE_MAX_CONNECTIONS=1"
run "CUDA_SCALE_LAUNCH_QUEUES=0.25x"
run "CUDA_DISABLE_PERF_BOOST=1"
