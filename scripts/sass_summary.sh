#!/usr/bin/env bash
# Counts the Blackwell-only SASS mnemonics (UTCHMMA = tcgen05.mma, LDTM = tcgen05.ld, UTMALDG /
# UTMASTG = TMA load / store, UTCBAR = tcgen05.commit, SYNCS = mbarrier) per kernel of the built
# library; profiles/rNN_sass_summary.txt is its output.
cuobjdump -sass "$(dirname "$0")/../dcvc_b200/libdcvc_b200.so" | awk '/Function :/{fn=$3}
  {for (i = 1; i <= NF; i++) if ($i ~ /^(UTCHMMA|LDTM|STTM|UTMALDG|UTMASTG|UTCBAR|SYNCS|UTMACCTL)/) {split($i, b, ";"); c[fn" "b[1]]++}}
  END {for (k in c) print c[k], k}' | sort -k2,2 -k1,1nr
