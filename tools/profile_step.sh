# usage: bash tools/profile_step.sh <tag>   (on the GPU box; writes gpurun_out/*_<tag>.*)
# 1. the bench line as the driver runs it (never under a profiler); 2. ncu launch list of the same command (per-launch times
# are cold-cache and serialised: the kernels' SHARES of the step must agree with the CUDA-event stages, not the absolutes);
# 3. one `--set full` capture each of the decode and the march kernel (DRAM traffic, stall reasons, source view).
tag=$1
mkdir -p gpurun_out
timeout 600 python bench.py > gpurun_out/bench_$tag.json 2> gpurun_out/bench_$tag.err || exit 1
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_$tag.csv \
  python bench.py --steps 12 --warmup 3 --no-cpu --no-graph --no-extras > gpurun_out/ncu_list_$tag.log 2>&1
timeout 400 ncu --set full --clock-control none --import-source on -k regex:kd_decode6 -s 4 -c 1 -f -o gpurun_out/decode_$tag \
  python bench.py --steps 3 --warmup 3 --no-cpu --no-graph --no-extras > gpurun_out/ncu_dec_$tag.log 2>&1
timeout 400 ncu --set full --clock-control none --import-source on -k regex:march_kernel -s 4 -c 1 -f -o gpurun_out/march_$tag \
  python bench.py --steps 3 --warmup 3 --no-cpu --no-graph --no-extras > gpurun_out/ncu_mar_$tag.log 2>&1
python tools/ncu_summary.py gpurun_out/decode_$tag.ncu-rep > gpurun_out/decode_$tag.ncu.txt 2>&1
python tools/ncu_summary.py gpurun_out/march_$tag.ncu-rep > gpurun_out/march_$tag.ncu.txt 2>&1
tail -c 400 gpurun_out/bench_$tag.json
