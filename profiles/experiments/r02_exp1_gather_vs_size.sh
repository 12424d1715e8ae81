set -x
for rows in 250000000 350000000; do
  timeout 300 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:"agg_gather|radix_scatter_kernel<unsigned int|agg_insert" -c 40 --csv --log-file gpurun_out/r2c_c4_${rows}.csv python bench.py --workload c4 --rows $rows --steps 1 --warmup 3 --no-e2e --no-cpu > gpurun_out/r2c_${rows}.log 2>&1
done
TFS_RADIX_BITS=8 timeout 300 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:"agg_gather|radix_scatter_kernel<unsigned int|agg_insert" -c 40 --csv --log-file gpurun_out/r2c_c4_bits8.csv python bench.py --workload c4 --rows 250000000 --steps 1 --warmup 3 --no-e2e --no-cpu > gpurun_out/r2c_bits8.log 2>&1
echo done
