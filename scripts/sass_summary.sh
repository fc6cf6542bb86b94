#!/bin/bash
# Per-kernel count of the SASS mnemonics that prove the Blackwell paths (tcgen05.mma = UTCHMMA, tcgen05.ld/st =
# LDTM/STTM, TMA = UTMALDG / UTMAPF (L2 prefetch), tcgen05.commit = UTCBAR, warp-level MMA = HMMA, ldmatrix = LDSM).
# Usage: scripts/sass_summary.sh > profiles/sass_blackwell_mnemonics.txt
cd "$(dirname "$0")/.."
cuobjdump -sass jax_llama_b200/ops/_C/libjllama_b200.so | awk '
  /Function :/ { fn = $3 }
  { for (i = 1; i <= NF; ++i) if ($i ~ /^(UTCHMMA|LDTM|STTM|UTMALDG|UTMAPF|UTCBAR|HMMA|LDSM|UTCATOMSWS)/) { split($i, a, ";"); c[fn " :: " a[1]]++ } }
  END { for (k in c) print c[k], k }' | sort -k2
