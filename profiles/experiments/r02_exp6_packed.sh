# packed two-pass row partition (default) against the generic 8-bit passes (TFS_AGG_PACKED=off)
set -x
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_kmeans_ops.py -q -k "aggregate or segment or kmeans or full_size_agg" --timeout 300 2>&1 | tail -3
for mode in on off; do
  TFS_AGG_PACKED=$mode timeout 300 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:"agg_|radix_" -c 400 --csv --log-file gpurun_out/r2h_c4_packed_${mode}.csv python bench.py --workload c4 --rows 250000000 --steps 1 --warmup 3 --no-e2e --no-cpu > gpurun_out/r2h_${mode}.log 2>&1
  TFS_AGG_PACKED=$mode timeout 300 python bench.py --workload c4 --steps 5 --warmup 3 --no-e2e --no-cpu > gpurun_out/r2h_bench_${mode}.json 2>&1
done
echo done
