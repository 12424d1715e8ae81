# rows per slot: counted at insert (one scattered atomic per row) vs read off the last partition pass (segment bounds)
set -x
timeout 900 python -m pytest tests -x -q -m gpu -k "aggregate or agg or kmeans or c4 or segment" > gpurun_out/r2m_tests.log 2>&1; tail -3 gpurun_out/r2m_tests.log
for mode in insert bounds; do
  TFS_AGG_COUNT=$mode timeout 300 python bench.py --workload c4 --steps 10 --warmup 3 --no-e2e --no-cpu > gpurun_out/r2m_$mode.json 2> gpurun_out/r2m_$mode.err; tail -c 600 gpurun_out/r2m_$mode.json
  TFS_AGG_COUNT=$mode timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"agg_insert|radix_scatter|agg_gather|agg_mark|agg_counts|radix_hist" --launch-skip 40 -c 40 --csv --log-file gpurun_out/r2m_$mode.csv python bench.py --workload c4 --rows 250000000 --steps 1 --warmup 3 --no-e2e --no-cpu > gpurun_out/r2m_ncu_$mode.log 2>&1
done
echo done
