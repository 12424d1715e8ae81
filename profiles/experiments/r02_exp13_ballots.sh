# ranks inside a tile from one ballot per digit bit instead of MATCH.ANY
set -x
timeout 1200 python -m pytest tests -x -q -m gpu -k "aggregate or agg or kmeans or c4 or segment or radix or multi" > gpurun_out/r2p_tests.log 2>&1; tail -3 gpurun_out/r2p_tests.log
timeout 300 python bench.py --workload c4 --steps 10 --warmup 3 --no-e2e --no-cpu > gpurun_out/r2p_c4.json 2> gpurun_out/r2p_c4.err; tail -c 400 gpurun_out/r2p_c4.json
timeout 300 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:"agg_insert|radix_scatter|agg_gather|agg_mark|agg_counts|radix_hist" --launch-skip 40 -c 40 --csv --log-file gpurun_out/r2p.csv python bench.py --workload c4 --rows 250000000 --steps 1 --warmup 3 --no-e2e --no-cpu > gpurun_out/r2p_ncu.log 2>&1
echo done
