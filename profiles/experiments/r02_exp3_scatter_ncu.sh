# ncu --set full on the radix scatter launches of one 2.5e8-row aggregate block (6 small group-sort passes, 2 row passes)
set -x
for bits in 11 8; do
  n=8; [ $bits = 8 ] && n=11
  TFS_RADIX_BITS=$bits timeout 600 ncu --set full --import-source on --clock-control none -k regex:radix_scatter_kernel -c $n -o gpurun_out/r2e_scatter_bits${bits} -f python bench.py --workload c4 --rows 250000000 --steps 1 --warmup 3 --no-e2e --no-cpu > gpurun_out/r2e_bits${bits}.log 2>&1
done
ls -la gpurun_out/
echo done
