# C4 at the full 5e8 rows: per-kernel times with 11-bit passes (default) and with 8-bit passes; plus the new chunk test
set -x
timeout 300 python -m pytest tests/test_gpu_parity.py -q -k "cut_into_row_ranges or far_more_keys or config4" --timeout 200 2>&1 | tail -3
for bits in 11 8; do
  TFS_RADIX_BITS=$bits timeout 300 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:"agg_|radix_" -c 400 --csv --log-file gpurun_out/r2d_c4_bits${bits}.csv python bench.py --workload c4 --steps 1 --warmup 3 --no-e2e --no-cpu > gpurun_out/r2d_bits${bits}.log 2>&1
done
for bits in 11 8; do
  TFS_RADIX_BITS=$bits timeout 300 python bench.py --workload c4 --steps 5 --warmup 3 --no-e2e --no-cpu > gpurun_out/r2d_bench_bits${bits}.json 2>&1
done
echo done
