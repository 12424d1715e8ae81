# pass width sweep of the row partition (packed pairs between passes): 5, 6, 7, 8 bits per pass
set -x
for b in 5 6 7 8; do
  TFS_RADIX_BITS=$b timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"radix_scatter|radix_hist" --launch-skip 60 -c 60 --csv --log-file gpurun_out/r2m_b$b.csv python bench.py --workload c4 --rows 250000000 --steps 1 --warmup 3 --no-e2e --no-cpu > gpurun_out/r2m_b$b.log 2>&1
done
echo done
