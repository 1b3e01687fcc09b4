set -x
mkdir -p gpurun_out/v7
python bench.py > gpurun_out/v7/bench_config2.json 2> gpurun_out/v7/bench_config2.err || exit 1
tail -c 600 gpurun_out/v7/bench_config2.json
python bench.py --steps 2 --warmup 3 > gpurun_out/v7/pre.json 2>/dev/null || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/v7/launches.csv python bench.py --steps 2 --warmup 3 > gpurun_out/v7/ncu_list.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:hnsw_search_kernel -s 18 -c 1 -f -o gpurun_out/v7/search_v7 python bench.py --steps 2 --warmup 3 > gpurun_out/v7/ncu_full.log 2>&1
ls -la gpurun_out/v7
