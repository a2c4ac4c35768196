#!/bin/bash
# Per-kernel SASS opcode histogram of libhonk_b200.so: the mnemonics that prove which hardware path a kernel uses
# (UTC*MMA = tcgen05.mma, LDTM/STTM = tcgen05.ld/st, UBLKCP = cp.async.bulk, UTMALDG = TMA tensor load, DMMA = FP64 tensor
# core mma.sync, HMMA = legacy f16 mma.sync).  usage: scripts/sass_histogram.sh > profiles/r04_sass_histogram.txt
LIB=${1:-honk_b200/libhonk_b200.so}
cuobjdump -sass "$LIB" | awk '
  /Function :/ { fn=$3; next }
  { for (i=1;i<=NF;i++) if ($i ~ /^(UTC[A-Z]*MMA(\.2CTA)?|UTCBAR(\.2CTA)?(\.MULTICAST)?|LDTM(\.[x0-9]+)?|STTM(\.[x0-9]+)?|UBLKCP(\.[A-Z.]+)?|UTMALDG(\.[0-9A-Z.]+)?|DMMA\.[0-9x]+|HMMA\.[0-9]+(\.[A-Z0-9]+)*|IMMA\.[0-9]+(\.[A-Z0-9]+)*|SYNCS(\.[A-Z0-9.]+)?|UTCATOMSWS(\.[A-Z0-9.]+)?|DFMA|DADD|DMUL)$/) c[fn" "$i]++ }
  END { for (k in c) print k, c[k] }' | sort | while read fn op n; do printf "%-120s %-28s %6d\n" "$(echo $fn | c++filt | cut -c1-118)" "$op" "$n"; done
