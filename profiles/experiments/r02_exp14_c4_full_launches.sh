# every launch of our kernels in one C4 step at the configured size (5e8 rows: two row ranges and the merge)
set -x
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:"agg_|radix_|scan_|fill_u64|fold_parts" --launch-skip 300 -c 150 --csv --log-file gpurun_out/r2q_c4_full.csv python bench.py --workload c4 --steps 1 --warmup 3 --no-e2e --no-cpu > gpurun_out/r2q.log 2>&1
echo done
