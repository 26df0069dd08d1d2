#!/usr/bin/env bash
# Per-kernel SASS instruction counts of the in-tree library: the evidence behind DESIGN.md's statements about bulk
# copies (UBLKCP), cp.async (LDGSTS), vector reductions (RED ... F32x4), mbarrier waits (SYNCS), shared-memory atomics
# (ATOMS) and packed FP32 (FFMA2 / FMUL2 / FADD2). Writes a table to stdout; run after 3dgs-mcmc_b200/build.sh.
set -euo pipefail
cd "$(dirname "$0")/.."
LIB=3dgs-mcmc_b200/diff_gaussian_rasterization/lib3dgs_mcmc_b200.so
echo "# $(basename $LIB) sha256 $(sha256sum $LIB | cut -c1-16), git $(git rev-parse --short HEAD 2>/dev/null || echo '?'), $(nvcc --version | tail -2 | head -1)"
cuobjdump -sass "$LIB" | awk '
  /Function :/ { fn=$3; order[++n]=fn; next }
  fn=="" { next }
  { total[fn]++ }
  /UBLKCP/ { c[fn,"UBLKCP"]++ } /LDGSTS/ { c[fn,"LDGSTS"]++ } /RED\.E\.ADD\.F32x4|REDG\.E\.ADD\.F32x4/ { c[fn,"RED.F32x4"]++ }
  /SYNCS/ { c[fn,"SYNCS"]++ } /ATOMS/ { c[fn,"ATOMS"]++ } /FFMA2/ { c[fn,"FFMA2"]++ } /FMUL2/ { c[fn,"FMUL2"]++ } /FADD2/ { c[fn,"FADD2"]++ }
  /MUFU\.EX2/ { c[fn,"MUFU.EX2"]++ } /FFMA / { c[fn,"FFMA"]++ } /MATCH/ { c[fn,"MATCH"]++ } /REDUX/ { c[fn,"REDUX"]++ }
  END {
    printf "%-72s %7s %6s %6s %9s %5s %5s %5s %5s %5s %6s %8s %5s %5s\n", "kernel (mangled, truncated)", "instrs", "UBLKCP", "LDGSTS", "RED.F32x4", "SYNCS", "ATOMS", "FFMA2", "FMUL2", "FADD2", "FFMA", "MUFU.EX2", "MATCH", "REDUX"
    for (i = 1; i <= n; i++) { f = order[i];
      printf "%-72s %7d %6d %6d %9d %5d %5d %5d %5d %5d %6d %8d %5d %5d\n", substr(f, 1, 72), total[f], c[f,"UBLKCP"], c[f,"LDGSTS"], c[f,"RED.F32x4"], c[f,"SYNCS"], c[f,"ATOMS"], c[f,"FFMA2"], c[f,"FMUL2"], c[f,"FADD2"], c[f,"FFMA"], c[f,"MUFU.EX2"], c[f,"MATCH"], c[f,"REDUX"] }
  }'
