# insert pass variants (kInsU keys per thread; probe scheme): launch list of one C4 step at 2.5e8 rows + aggregate tests
set -x
timeout 900 python -m pytest tests -x -q -m gpu -k "aggregate or agg or kmeans or segment" > gpurun_out/r2t_tests.log 2>&1; tail -2 gpurun_out/r2t_tests.log
timeout 300 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sectors_srcunit_tex_op_read.sum,smsp__inst_executed.sum --clock-control none -k regex:"agg_insert" --launch-skip 4 -c 2 --csv --log-file gpurun_out/r2t.csv python bench.py --workload c4 --rows 250000000 --steps 1 --warmup 3 --no-e2e --no-cpu > gpurun_out/r2t_ncu.log 2>&1
grep -o '"agg_insert[^"]*","[^"]*","[^"]*","[a-z_.]*","[^"]*","[^"]*"' gpurun_out/r2t.csv | tail -5
