# ncu --set full of the keyed-reduction kernels of one 2.5e8-row block (insert, 8 + 3 partition passes, gather)
set -x
mkdir -p gpurun_out/ev
timeout 900 ncu --set full --import-source on --clock-control none -k 'regex:agg_insert|agg_gather|radix_scatter|radix_hist' --launch-skip 72 -c 24 -o gpurun_out/ev/full_c4 -f python bench.py --workload c4 --rows 250000000 --steps 1 --warmup 3 --no-e2e --no-cpu > gpurun_out/ev/ncu_full_c4.log 2>&1; echo "full c4 rc=$?"
