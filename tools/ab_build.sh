#!/bin/bash
# Builds two more variants of the library for same-box A/B runs (select one with LK_LIBRARY=<path>, llama.kotlin_b200/_binding.py):
#   tools/_ab/lib_r1.so     the round-1 library (commit faf9642)
#   tools/_ab/lib_nofix.so  the current sources with -DLK_GEMV_NO_RANGE_FIX (K1 without the range detection / fix-up)
# tools/_ab/ is git-ignored; the .so files travel to the GPU box.
set -e
cd "$(dirname "$0")/.."
ROOT=$PWD
mkdir -p tools/_ab/r1 tools/_ab/nofix tools/_ab/include_r1
cd tools/_ab/r1
for f in Makefile lk_attention.cu lk_comm.cu lk_common.cuh lk_elementwise.cu lk_gemm_tc.cu lk_gemv.cu lk_kernels.h lk_kquants.cu lk_shim.cu lk_shim.h; do
  git -C "$ROOT" show faf9642:llama.kotlin_b200/csrc/$f > $f
done
git -C "$ROOT" show faf9642:include/llamakotlin_cuda.h > ../include_r1/llamakotlin_cuda.h
sed -i 's#../../include/llamakotlin_cuda.h#../include_r1/llamakotlin_cuda.h#g; s#OUT       := ../libllamakotlin_cuda.so#OUT       := ../lib_r1.so#' Makefile lk_shim.h
make -j4 > build.log 2>&1
cd "$ROOT"
# variant <name> <extra nvcc flags>: the current sources with the flags -> tools/_ab/lib_<name>.so
variant() {
  mkdir -p "$ROOT/tools/_ab/$1"
  cd "$ROOT/tools/_ab/$1"
  cp "$ROOT"/llama.kotlin_b200/csrc/{Makefile,*.cu,*.cuh,*.h} .
  sed -i "s#OUT       := ../libllamakotlin_cuda.so#OUT       := ../lib_$1.so#; s#NVCCFLAGS := #NVCCFLAGS := $2 #; s#../../include/llamakotlin_cuda.h#../../../include/llamakotlin_cuda.h#g" Makefile lk_shim.h
  make -j4 > build.log 2>&1
  cd "$ROOT"
}
variant nofix -DLK_GEMV_NO_RANGE_FIX
for v in "$@"; do      # further variants as name=flags, e.g.  chain_nodetect=-DLK_CHAIN_DETECT=false
  variant "${v%%=*}" "${v#*=}"
done
ls -la "$ROOT"/tools/_ab/*.so
