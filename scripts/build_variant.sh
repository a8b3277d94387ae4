#!/bin/bash
# A/B tuning helper: compiles csrc/fs2_cabi.cu with extra -D flags into
# final-stage-path-tracer-2-0_b200/_exp/<name>/libfs2cuda.so (git-ignored; travels to the GPU box).
#   scripts/build_variant.sh base
#   scripts/build_variant.sh rep4 -DFS2_WF_REP_ENTER=4
set -e
cd "$(dirname "$0")/.."
name=$1; shift
pkg=final-stage-path-tracer-2-0_b200
mkdir -p $pkg/_exp/$name
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -fmad=false \
  -Xcompiler -fPIC,-ffp-contract=off --expt-relaxed-constexpr --extended-lambda -shared \
  -Xptxas=-v "$@" -Iinclude -o $pkg/_exp/$name/libfs2cuda.so $pkg/csrc/fs2_cabi.cu 2>&1 \
  | grep -A2 "k_wf_traceILb0ELb[01]" | grep -E "spill|registers" | tr '\n' ' '
echo
echo "$name: $*" > $pkg/_exp/$name/flags.txt
