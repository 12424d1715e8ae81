# hash-table slots per group, again, now that the insert pass no longer counts rows (no second table of atomics in L2)
for f in 4 8; do
  TFS_AGG_TABLE_FACTOR=$f timeout 200 python bench.py --workload c4 --rows 250000000 --steps 10 --warmup 3 --no-e2e --no-cpu 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('factor', $f, 'ms', d['ms_per_step'], d['clocks']['sm_mhz'])"
done
