#!/bin/bash
# One call on an N-GPU box: topology, P2P tests, cold probe of all GPUs, the scaling bench at 1/2/4/N, mount-to-usable.
N=${1:-8}
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/topo_${N}gpu.txt 2>&1
python -m pytest tests/test_probe_gpu.py -m gpu -x -q 2>&1 | tail -4
python tools/time_probe.py --bytes 1G --iters 1 --dma-bytes 16M --p2p-bytes 1G --p2p-iters 2 --lib-baselines > gpurun_out/probe_${N}gpu.json
python -c "
import json; d=json.load(open('gpurun_out/probe_${N}gpu.json')); print({k:d[k] for k in ('wall_s','status','n_gpus','ms')}); p=d.get('p2p',{});
print('match',p.get('matrix_match'),'alltoall',p.get('alltoall_gbs'),'bidir',p.get('bidir_gbs'),'links',p.get('nvlink_active'))
for row in p.get('write_gbs',[]): print('W',[round(x) for x in row])
for row in p.get('read_gbs',[]): print('R',[round(x) for x in row])
for row in p.get('memcpy_peer_gbs',[]): print('M',[round(x) for x in row])
"
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 tools/nccl_crosscheck.py 2>/dev/null | tail -1 | tee gpurun_out/nccl_${N}gpu.json
for g in 1 2 4 $N; do
  [ $g -gt $N ] && continue
  if [ $g -eq 1 ]; then
    python bench.py --steps 10 --warmup 3 --mount-iters 12 2>gpurun_out/scale_${g}.err | tail -1 > gpurun_out/scale_${g}.json
  else
    python -m torch.distributed.run --nnodes=1 --nproc-per-node $g --master-addr 127.0.0.1 --master-port 2951$g bench.py --gpus $g --steps 10 --warmup 3 --mount-iters 12 2>gpurun_out/scale_${g}.err | tail -1 > gpurun_out/scale_${g}.json
  fi
  python -c "
import json; d=json.load(open('gpurun_out/scale_${g}.json')); m=d.get('mount',{}); u=m.get('mount_to_usable',{})
print('N=${g}', 'value', round(d['value']), 'e2e', round(d['e2e']['value']), 'ms/call', round(d['e2e']['ms_per_call'],1), 'mount ref/native p50 ms', m.get('reference_fork_exec_mount_ms',{}).get('p50'), m.get('native_mount_ms',{}).get('p50'), 'to-usable p50/p99 s', u.get('p50_s'), u.get('p99_s'), 'p2p', json.dumps(u.get('p2p'))[:400], m.get('unavailable'))
"
done
