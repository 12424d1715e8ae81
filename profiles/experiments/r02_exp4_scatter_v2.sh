# scatter kernel v2 (loads before stores, key prefetch): aggregate tests, then C4 at 5e8 rows with 11- and 8-bit passes
set -x
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_kmeans_ops.py -q -k "aggregate or segment or kmeans or full_size_agg" --timeout 300 2>&1 | tail -3
for bits in 11 8; do
  TFS_RADIX_BITS=$bits timeout 300 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:"agg_|radix_" -c 400 --csv --log-file gpurun_out/r2f_c4_bits${bits}.csv python bench.py --workload c4 --rows 250000000 --steps 1 --warmup 3 --no-e2e --no-cpu > gpurun_out/r2f_bits${bits}.log 2>&1
  TFS_RADIX_BITS=$bits timeout 300 python bench.py --workload c4 --steps 5 --warmup 3 --no-e2e --no-cpu > gpurun_out/r2f_bench_bits${bits}.json 2>&1
done
echo done
