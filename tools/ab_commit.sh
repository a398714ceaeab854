#!/bin/bash
# Builds the library of an earlier commit for same-box A/B runs:  tools/ab_commit.sh <commit> <name>  ->  tools/_ab/lib_<name>.so
# (select it with LK_LIBRARY=<path>, llama.kotlin_b200/_binding.py; tools/_ab/ is git-ignored, the .so files travel to the GPU box)
set -e
cd "$(dirname "$0")/.."
ROOT=$PWD
mkdir -p tools/_ab/$2
cd tools/_ab/$2
for f in $(git -C "$ROOT" ls-tree --name-only $1 llama.kotlin_b200/csrc/ | xargs -n1 basename); do
  git -C "$ROOT" show $1:llama.kotlin_b200/csrc/$f > $f
done
sed -i "s#OUT       := ../libllamakotlin_cuda.so#OUT       := ../lib_$2.so#; s#../../include/llamakotlin_cuda.h#../../../include/llamakotlin_cuda.h#g" Makefile lk_shim.h
make -j4 > build.log 2>&1 || (tail -5 build.log; exit 1)
ls -la "$ROOT"/tools/_ab/lib_$2.so
