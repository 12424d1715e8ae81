# Round-2 evidence: the default bench line, the reference arm, a launch list of the library's kernels over the default
# bench, and one ncu --set full capture per workload's dominant kernels.  One GPU.
set -x
mkdir -p gpurun_out/ev
python bench.py --steps 20 --warmup 5 > gpurun_out/ev/bench_n1.json 2> gpurun_out/ev/bench_n1.err; echo "bench rc=$?"
python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/ev/bench_reference.json 2> gpurun_out/ev/bench_reference.err; echo "ref rc=$?"
K='regex:reduce_axis0|chain_kernel|agg_|radix_|matmul_tc|scan_tile|scan_apply|fill_u64|split_weights|fold_|eltwise'
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k "$K" -c 2000 --csv --log-file gpurun_out/ev/launches_default_bench.csv python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu > gpurun_out/ev/ncu_launches.log 2>&1; echo "launches rc=$?"
for wl in c2 c3 c5; do
  case $wl in c2) kk='regex:reduce_axis0_tma';; c3) kk='regex:chain_kernel';; c5) kk='regex:matmul_tc_pair';; esac
  timeout 600 ncu --set full --import-source on --clock-control none -k "$kk" --launch-skip 3 -c 1 -o gpurun_out/ev/full_$wl -f python bench.py --workload $wl --steps 2 --warmup 3 --no-e2e --no-cpu > gpurun_out/ev/ncu_full_$wl.log 2>&1; echo "full $wl rc=$?"
done
timeout 900 ncu --set full --import-source on --clock-control none -k 'regex:agg_insert|agg_gather|radix_scatter|radix_hist' --launch-skip 30 -c 12 -o gpurun_out/ev/full_c4 -f python bench.py --workload c4 --rows 250000000 --steps 1 --warmup 3 --no-e2e --no-cpu > gpurun_out/ev/ncu_full_c4.log 2>&1; echo "full c4 rc=$?"
ls -la gpurun_out/ev
