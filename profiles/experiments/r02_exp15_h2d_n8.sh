# 8-GPU box: topology, copy-only H2D at N = 1, 2, 4, 8 (default placement of the pinned block and GPU-local), then the default bench at N = 8
set -x
nvidia-smi topo -m > gpurun_out/r2s_topo.txt 2>&1
(lscpu | grep -i "numa\|socket\|model name\|^CPU(s)"; numactl -H 2>/dev/null; cat /sys/devices/system/node/node*/meminfo 2>/dev/null | grep MemTotal) >> gpurun_out/r2s_topo.txt 2>&1
S=profiles/experiments/r02_h2d_copy_only.py
python $S > gpurun_out/r2s_h2d.jsonl 2>/dev/null
python $S numa_local >> gpurun_out/r2s_h2d.jsonl 2>/dev/null
for n in 2 4 8; do
  for mode in "" "numa_local"; do
    timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 $S $mode >> gpurun_out/r2s_h2d.jsonl 2>gpurun_out/r2s_h2d_err.txt
  done
done
cat gpurun_out/r2s_h2d.jsonl
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 8 --steps 20 --warmup 5 > gpurun_out/r2s_bench_n8.json 2> gpurun_out/r2s_bench_n8.err; echo "bench rc=$?"
echo done
