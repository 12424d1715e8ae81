set -x
TFS_RADIX_BITS=8 timeout 600 ncu --set full --import-source on --clock-control none -k regex:radix_scatter_kernel --launch-skip 8 -c 3 -o gpurun_out/r2g_scatter8 -f python bench.py --workload c4 --rows 250000000 --steps 1 --warmup 3 --no-e2e --no-cpu > gpurun_out/r2g.log 2>&1
echo done
