# hash-table load factor of the insert pass: table slots = factor x max_groups (rounded up to a power of two)
set -x
for f in 2 4 8; do
  TFS_AGG_TABLE_FACTOR=$f timeout 300 ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum --clock-control none -k regex:"agg_insert|radix_scatter|agg_gather" --launch-skip 30 -c 30 --csv --log-file gpurun_out/r2l_f$f.csv python bench.py --workload c4 --rows 250000000 --steps 1 --warmup 3 --no-e2e --no-cpu > gpurun_out/r2l_f$f.log 2>&1
done
echo done
