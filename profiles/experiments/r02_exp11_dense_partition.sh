# rows partitioned by dense group number (20 bits for 1e6 groups: two 10-bit packed passes) against by table slot (23 bits: three passes)
set -x
timeout 1200 python -m pytest tests -x -q -m gpu -k "aggregate or agg or kmeans or c4 or segment" > gpurun_out/r2n_tests.log 2>&1; tail -3 gpurun_out/r2n_tests.log
for mode in insert bounds; do
  TFS_AGG_COUNT=$mode timeout 300 python bench.py --workload c4 --steps 10 --warmup 3 --no-e2e --no-cpu > gpurun_out/r2n_$mode.json 2> gpurun_out/r2n_$mode.err; tail -c 400 gpurun_out/r2n_$mode.json
done
for pb in 9 10 11; do
  TFS_RADIX_PACKED_BITS=$pb timeout 300 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:"agg_insert|radix_scatter|agg_gather|agg_mark|agg_counts|radix_hist|radix_map" --launch-skip 40 -c 40 --csv --log-file gpurun_out/r2n_pb$pb.csv python bench.py --workload c4 --rows 250000000 --steps 1 --warmup 3 --no-e2e --no-cpu > gpurun_out/r2n_ncu_$pb.log 2>&1
done
echo done
