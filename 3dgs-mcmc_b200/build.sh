#!/usr/bin/env bash
# Build lib3dgs_mcmc_b200.so (C ABI in include/gs_b200.h) for sm_100a, in-tree.
# nvcc defaults on purpose: -fmad=true and NO -use_fast_math, like the reference's build
# (submodules/diff-gaussian-rasterization/setup.py:17-35), because bit-exact parity of the
# binning keys depends on identical FP32 contraction.
set -euo pipefail
cd "$(dirname "$0")"
NVCC=${NVCC:-nvcc}
OUT=diff_gaussian_rasterization/lib3dgs_mcmc_b200.so
SRCS="csrc/swin.cu csrc/knn.cu csrc/activations.cu csrc/loss.cu csrc/optim.cu csrc/dp_exchange.cu csrc/api.cu csrc/preprocess.cu csrc/binning.cu csrc/render_fwd.cu csrc/render_bwd.cu csrc/preprocess_bwd.cu csrc/mcmc.cu"
$NVCC -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo \
  -Xcompiler -fPIC -shared ${GS_EXTRA_NVCC_FLAGS:-} -o "$OUT" $SRCS -ldl
echo "built $OUT"
