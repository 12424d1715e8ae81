#!/bin/bash
# Counts of the SASS mnemonics that show what the kernels are made of (TMA / mbarrier / tcgen05 / shared-memory atomics /
# warp collectives), per object file of the last build:   bash profiles/sass_mnemonics.sh > profiles/r02_sass.txt
for f in kernels_reduce kernels_matmul_tc kernels_radix kernels_aggregate kernels_eltwise kernels_chain_f64 kernels_noalloc; do
  o=build/obj/$f.cu.o
  [ -f "$o" ] || continue
  echo "== $f.cu.o (arch sm_100a)"
  cuobjdump -sass "$o" 2>/dev/null | grep -oE "\b(UBLKCP[.A-Z0-9]*|SYNCS[.A-Z0-9]*|UTCHMMA[.A-Z0-9]*|UTMALDG[.A-Z0-9]*|UTMASTG[.A-Z0-9]*|UTCBAR[.A-Z0-9]*|LDTM[.a-zA-Z0-9]*|MATCH[.A-Z]*|VOTE[.A-Z]*|ATOMS[.A-Z0-9]*|REDUX[.A-Z]*|LDS\.128|STS\.128|STS\.64|LDS\.64|LDG\.E\.128[.A-Z]*|REDG[.A-Z0-9]*|ATOMG[.A-Z0-9]*|BRX|JMX[.A-Z]*)\b" | sort | uniq -c | sort -rn | head -14
done
