#!/bin/sh
# Blackwell-native instruction evidence: per object file and kernel, the count of tcgen05 / TMEM / TMA / bulk-copy / DSMEM SASS
# mnemonics (UTCHMMA = tcgen05.mma kind::f16, .2CTA = cta_group::2; LDTM = tcgen05.ld; UTMALDG = cp.async.bulk.tensor;
# UBLKCP = cp.async.bulk; STAS = st.async; SYNCS = mbarrier; UTCBAR = tcgen05.commit).  Usage: scripts/sass_summary.sh > profiles/r02_sass_summary.txt
cd "$(dirname "$0")/.." || exit 1
for o in pointcloudprocessing_b200/csrc/_obj/*.o; do
  cuobjdump -sass "$o" 2>/dev/null | awk -v obj="$(basename "$o")" '
    /Function : /{fn=$3}
    /UTCHMMA|UTCQMMA|UTCBAR|LDTM|STTM|UTMALDG|UTMASTG|UBLKCP|STAS|SYNCS|UTCCP|UTCATOM|REDUX|FADD2|FFMA2/{
      m=$0; sub(/^ *\/\*[0-9a-f]+\*\/ */,"",m); split(m,a," "); op=a[1]; if (op ~ /^@/) op=a[2];
      sub(/\..*$/,"",op); if ($0 ~ /2CTA/) op=op".2CTA"; c[fn" "op]++ }
    END{for(k in c) print obj, k, c[k]}' | sort
done
