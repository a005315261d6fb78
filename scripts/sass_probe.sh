#!/bin/bash
# scripts/sass_probe.sh <name> [flags]: nvcc -cubin of one k1s<PR, PG, PM, PS> instantiation (defaults 32, 32, false, 1), registers /
# spills from ptxas -v, SASS to /tmp/probe/<name>.sass, then the loop statistics (scripts/sass_loop_moves.py).  No GPU needed.
name=$1; shift
root="$(cd "$(dirname "$0")/.." && pwd)"
mkdir -p /tmp/probe
cat > /tmp/probe/probe.cu <<'EOF2'
#include "sw_k1s.cuh"
#ifndef PR
#define PR 32
#endif
#ifndef PG
#define PG 32
#endif
#ifndef PM
#define PM false
#endif
#ifndef PS
#define PS 1
#endif
template __global__ void swk::k1s<PR, PG, PM, PS>(const swk::K1SParams);
EOF2
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -I $root/seq-align-gpu_b200/csrc -Xptxas -v "$@" -cubin -o /tmp/probe/$name.cubin /tmp/probe/probe.cu 2>&1 | grep -E "registers|spill|error" | head -3
cuobjdump -sass /tmp/probe/$name.cubin | grep -E "^\s+/\*[0-9a-f]{4}\*/" | sed -E 's/\s+\/\* 0x[0-9a-f]+ \*\///; s/^\s+//' > /tmp/probe/$name.sass
python $root/scripts/sass_loop_moves.py $name 80
