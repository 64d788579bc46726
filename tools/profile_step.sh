# usage: bash tools/profile_step.sh <tag>   (on the GPU box; writes gpurun_out/*_<tag>.*)
tag=$1
mkdir -p gpurun_out
timeout 120 ./tools/fill_probe > gpurun_out/fill_probe_$tag.txt 2>&1
timeout 400 python bench.py --steps 50 --warmup 3 > gpurun_out/bench_$tag.json 2> gpurun_out/bench_$tag.err || exit 1
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_$tag.csv \
  python bench.py --steps 12 --warmup 3 --no-cpu --no-graph > gpurun_out/ncu_list_$tag.log 2>&1
timeout 400 ncu --set full --clock-control none --import-source on -k regex:kd_decode_tiles -s 4 -c 1 -f -o gpurun_out/decode_$tag \
  python bench.py --steps 3 --warmup 3 --no-cpu --no-graph > gpurun_out/ncu_dec_$tag.log 2>&1
timeout 400 ncu --set full --clock-control none --import-source on -k regex:march_kernel -s 4 -c 1 -f -o gpurun_out/march_$tag \
  python bench.py --steps 3 --warmup 3 --no-cpu --no-graph > gpurun_out/ncu_mar_$tag.log 2>&1
tail -c 600 gpurun_out/bench_$tag.json
