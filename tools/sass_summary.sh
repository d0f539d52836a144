#!/bin/bash
# Per-kernel counts of the SASS mnemonics that prove tcgen05 / TMEM / TMA / cluster use (B200_PROFILING.md), from the
# built library.  Runs without a GPU:  tools/sass_summary.sh > profiles/r02_sass_summary.txt
SO=convex-optimization_b200/glprox/libglprox.so
echo "# cuobjdump -sass $SO : mnemonic counts per kernel (sm_100a)"
echo "# UTCHMMA = tcgen05.mma (.2CTA = cta_group::2), LDTM = tcgen05.ld, UTMALDG = cp.async.bulk.tensor (TMA),"
echo "# UBLKCP = cp.async.bulk, UTCBAR = tcgen05.commit, SYNCS = mbarrier, UCGABAR = barrier.cluster, ACQBULK / UTCATOMSWS = TMEM alloc"
cuobjdump -sass $SO 2>/dev/null | awk '
  /Function :/ { fn=$3 }
  { for (i=1;i<=NF;i++) { t=$i; sub(/;$/,"",t);
      if (t ~ /^(UTCHMMA|UTCQMMA|UTCOMMA|LDTM|STTM|UTMALDG|UTMASTG|UBLKCP|UTCBAR|UTCATOMSWS|UCGABAR_ARV|UCGABAR_WAIT|ACQBULK|HMMA|DFMA|DADD|FFMA)/) { split(t,a,"."); key=a[1]; if (t ~ /2CTA/) key=key".2CTA"; if (t ~ /MULTICAST/) key=key".MULTICAST"; c[fn" "key]++ } } }
  END { for (k in c) print k, c[k] }' | sort | c++filt 2>/dev/null | awk '{ n=$NF; $NF=""; printf "%6d  %s\n", n, $0 }' | sort -k2 | grep -E "UTC|LDTM|UTMA|UBLKCP|UCGABAR|ACQBULK" 
