#!/bin/bash
# GPU-box job (round 2, first slot): parity suite, bench lines, kernel variants,
# ncu captures. Everything lands in gpurun_out/.
O=gpurun_out
mkdir -p $O
nproc > $O/box.txt; free -g | head -2 >> $O/box.txt; nvidia-smi --query-gpu=name,memory.total --format=csv >> $O/box.txt
( time timeout 1500 python -m pytest tests -m gpu -q --durations=15 ) > $O/pytest_gpu.log 2>&1
tail -40 $O/pytest_gpu.log
timeout 600 python bench.py --workload C2 --steps 5 --warmup 3 > $O/r02_b_c2.json 2> $O/r02_b_c2.err; cut -c1-300 $O/r02_b_c2.json; tail -3 $O/r02_b_c2.err
( time timeout 1200 python bench.py ) > $O/r02_b_c4.json 2> $O/r02_b_c4.err; cut -c1-300 $O/r02_b_c4.json; tail -5 $O/r02_b_c4.err
# C4 line-FFT variants (device times per stage from the library's own events)
SKA_SDP_FUNC_ES_DEBUG=1 timeout 300 python profiles/run_path.py C4 grid 2 > $O/c4_default.txt 2>&1; tail -3 $O/c4_default.txt | cut -c1-400
SKA_SDP_FUNC_ES_DEBUG=1 SKA_SDP_FUNC_ES_FFT_RADICES=16,14,11,10 SKA_SDP_FUNC_ES_FFT_CFG=1024,1024,0,0 timeout 300 python profiles/run_path.py C4 grid 2 > $O/c4_wide.txt 2>&1; tail -3 $O/c4_wide.txt | cut -c1-400
SKA_SDP_FUNC_ES_DEBUG=1 SKA_SDP_FUNC_ES_FFT_RADICES=16,14,11,10 SKA_SDP_FUNC_ES_FFT_CFG=512,512,0,0 timeout 300 python profiles/run_path.py C4 grid 2 > $O/c4_r16_512.txt 2>&1; tail -3 $O/c4_r16_512.txt | cut -c1-400
timeout 600 ncu --set full --clock-control none --import-source on -k regex:es_fft_grid -s 6 -c 2 -o $O/r02_ncu_c2_grid_a -f python profiles/run_path.py C2 grid 3 > $O/ncu_c2.log 2>&1; tail -2 $O/ncu_c2.log
timeout 900 ncu --set full --clock-control none --import-source on -k regex:es_fft_grid -s 6 -c 2 -o $O/r02_ncu_c4_grid_a -f python profiles/run_path.py C4 grid 1 > $O/ncu_c4.log 2>&1; tail -2 $O/ncu_c4.log
