#!/bin/bash
cd /root/repo
timeout 900 python -m pytest tests -q -m gpu > gpurun_out/r2_gpu_tests_final4.log 2>&1; tail -2 gpurun_out/r2_gpu_tests_final4.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -1
bash tools/profile_step.sh r2final4 > gpurun_out/profile_step_r2final4.log 2>&1; tail -c 200 gpurun_out/profile_step_r2final4.log
timeout 300 python tools/decode_ab.py --workloads slab,config3,noise --kernels persistent,classic --fill-mode 2 --iters 30 --check --probes > gpurun_out/decode_ab_final4.jsonl 2> gpurun_out/decode_ab_final4.err
timeout 400 ncu --set full --clock-control none --import-source on -k regex:kd_decode6 -s 5 -c 1 -f -o gpurun_out/dec_c3_final4 python tools/decode_ab.py --workloads config3 --kernels persistent --fill-mode 2 --iters 4 > gpurun_out/ncu_c3_final4.log 2>&1
python tools/ncu_summary.py gpurun_out/dec_c3_final4.ncu-rep > gpurun_out/dec_c3_final4.ncu.txt 2>&1
ls gpurun_out | grep final4 | wc -l
