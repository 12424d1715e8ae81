# two keyed aggregates at once: do their phases overlap?  (gather CTAs per SM uncapped / capped by unused shared memory)
set -x
for pad in 0 40 46; do
  TFS_AGG_GATHER_PAD_KB=$pad timeout 600 python profiles/experiments/r02_two_sessions.py 2>&1 | tail -2
done
