ncu --set full --clock-control none --import-source on -k regex:march_kernel -s 4 -c 1 -f -o gpurun_out/march_dist python bench.py --steps 3 --warmup 3 --no-cpu --no-graph --no-extras > gpurun_out/ncu_mar_dist.log 2>&1
python tools/ncu_summary.py gpurun_out/march_dist.ncu-rep > gpurun_out/march_dist.ncu.txt 2>&1
grep -E "time_duration|inst_executed.sum|issue_active|thread_inst" gpurun_out/march_dist.ncu.txt
