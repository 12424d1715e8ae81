# source-level profile of the middle (packed in, packed out) pass of the row partition
set -x
timeout 600 ncu --set full --import-source on --clock-control none -k regex:radix_scatter --launch-skip 7 -c 1 -o gpurun_out/r2o_scatter -f python bench.py --workload c4 --rows 250000000 --steps 1 --warmup 1 --no-e2e --no-cpu > gpurun_out/r2o.log 2>&1
ls -la gpurun_out/
echo done
