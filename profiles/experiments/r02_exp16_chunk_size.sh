# row-range size of the chunked aggregate: 64 GiB (default), 32 GiB, 16 GiB of payload per range at 5e8 rows
set -x
for mb in 65536 32768 16384; do
  TFS_AGG_CHUNK_MB=$mb timeout 300 python bench.py --workload c4 --steps 10 --warmup 3 --no-e2e --no-cpu 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('chunk_mb', $mb, 'ms', d['ms_per_step'], d['clocks'])"
done
