#!/bin/bash
# Build a variant of libswgpu.so with extra -D flags for in-box A/B runs (scripts/ab_lib.py):
#   scripts/build_variant.sh <name> [-DK1S_...=...]...   ->  seq-align-gpu_b200/libswgpu_v<name>.so
# (a copy of the sources is built out of tree with K1S_DEFS="<flags>", so the in-tree objects stay those of the default build)
set -e
name=$1; shift
root="$(cd "$(dirname "$0")/.." && pwd)"
tmp=$(mktemp -d)
mkdir -p $tmp/seq-align-gpu_b200 && cp -r $root/seq-align-gpu_b200/csrc $root/seq-align-gpu_b200/Makefile $tmp/seq-align-gpu_b200/ && cp -r $root/include $tmp/
rm -f $tmp/seq-align-gpu_b200/csrc/*.o
make -s -j8 -C $tmp/seq-align-gpu_b200 $tmp/seq-align-gpu_b200/libswgpu.so K1S_DEFS="$*"
cp $tmp/seq-align-gpu_b200/libswgpu.so $root/seq-align-gpu_b200/libswgpu_v$name.so
rm -rf $tmp
ls -la $root/seq-align-gpu_b200/libswgpu_v$name.so
