#!/bin/bash
# Host C layer (libalign.so: CLI parser, matrix loader, FASTA readers, mmap parser, .swdb cache) and the packer under
# AddressSanitizer + UBSan: builds seq-align-gpu_b200/asan/ and runs the CPU host-layer tests against it.
#   scripts/host_asan.sh [extra pytest args]      -> profiles/r2_sanitize_host_asan.txt
cd "$(dirname "$0")/.."
make -s -C seq-align-gpu_b200 asan || exit 1
ASAN_SO=$(gcc -print-file-name=libasan.so)
export SWGPU_HOST_LIB=$PWD/seq-align-gpu_b200/asan/libalign.so SWGPU_HOST_BIN=$PWD/seq-align-gpu_b200/asan/smith_waterman
# python itself is not instrumented: preload the runtime, do not report the interpreter's own (intentional) leaks
export LD_PRELOAD=$ASAN_SO ASAN_OPTIONS=detect_leaks=0:abort_on_error=1 UBSAN_OPTIONS=print_stacktrace=1:halt_on_error=1
{ echo "# $(date -u +%FT%TZ) host layer under -fsanitize=address,undefined ($(gcc --version | head -1))";
  python -m pytest tests/test_host_layer.py tests/test_db_loader.py -q -m "not gpu" -p no:cacheprovider "$@" 2>&1 | tail -n 15;
  echo "# db_pack.c (packer invariants, layout from lengths) through the same sanitised library";
  SWGPU_LIB=$SWGPU_HOST_LIB python -m pytest tests/test_synth_layout.py -q -m "not gpu" -p no:cacheprovider -k "packer or layout_from" 2>&1 | tail -n 15; } | tee profiles/r2_sanitize_host_asan.txt
