# scatter kernel with an unpredicated body for whole tiles: aggregate tests, launch list at 2.5e8 rows, C4 bench at 5e8
set -x
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_kmeans_ops.py -q -k "aggregate or segment or kmeans or full_size_agg" --timeout 300 2>&1 | tail -3
timeout 300 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,smsp__inst_executed.sum --clock-control none -k regex:"agg_|radix_" -c 400 --csv --log-file gpurun_out/r2k_c4.csv python bench.py --workload c4 --rows 250000000 --steps 1 --warmup 3 --no-e2e --no-cpu > gpurun_out/r2k.log 2>&1
timeout 300 python bench.py --workload c4 --steps 5 --warmup 3 --no-e2e --no-cpu > gpurun_out/r2k_bench.json 2>&1
echo done
