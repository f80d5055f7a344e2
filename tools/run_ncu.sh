#!/bin/bash
# ncu evidence for the round (1 GPU): launch list of the bench command, one full capture of every shipped kernel.
#   gpurun --timeout 900 -- 'bash tools/run_ncu.sh'
mkdir -p gpurun_out
BENCH="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-mount --e2e-steps 2 --no-clocks"
timeout 200 $BENCH > gpurun_out/ncu_bench_plain.json 2> gpurun_out/ncu_bench_plain.err || { echo "bench failed without ncu"; tail -5 gpurun_out/ncu_bench_plain.err; exit 1; }
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_launches.csv $BENCH > gpurun_out/ncu_bench.log 2>&1; echo "launch list rc=$?"
timeout 120 python tools/ncu_case.py || { echo "ncu_case failed without ncu"; exit 1; }
# the second round of launches: skip the warm-up round (7 launches: write, verify, copy_bulk, copy remainder, segments, fold, bulk write, + final verify)
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"k_hbm|k_noalloc" -s 8 -c 8 -o gpurun_out/r02_full -f python tools/ncu_case.py > gpurun_out/ncu_full.log 2>&1; echo "full capture rc=$?"
ls -la gpurun_out/*.ncu-rep
