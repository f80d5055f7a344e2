#!/bin/bash
# NVLink TX/RX byte counters of one peer store (vector), one bulk-engine push and one peer load, against algorithmic bytes.
#   gpurun --gpus 2 --timeout 600 -- 'bash tools/run_2gpu_ncu.sh'
mkdir -p gpurun_out
timeout 120 python tools/ncu_p2p_case.py || { echo "case failed without ncu"; exit 1; }
M=gpu__time_duration.sum,nvltx__bytes.sum,nvltx__bytes_data_user.sum,nvltx__bytes_data_protocol.sum,nvlrx__bytes.sum,nvlrx__bytes_data_user.sum,nvlrx__bytes_data_protocol.sum,nvltx__bytes_packet_request.sum,nvltx__bytes_packet_response.sum,nvlrx__bytes_packet_request.sum,nvlrx__bytes_packet_response.sum,lts__t_sectors_srcunit_tex_aperture_peer.sum
timeout 400 ncu --metrics $M --clock-control none -k regex:k_hbm -s 3 -c 3 --csv --log-file gpurun_out/r02_p2p_ncu.csv python tools/ncu_p2p_case.py > gpurun_out/ncu_p2p.log 2>&1; echo "ncu rc=$?"
tail -5 gpurun_out/ncu_p2p.log; grep -c k_hbm gpurun_out/r02_p2p_ncu.csv
