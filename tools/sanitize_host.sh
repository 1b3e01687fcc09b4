#!/bin/sh
# Host-side C under AddressSanitizer + UndefinedBehaviorSanitizer, as the reference's CI does for its C unit tests
# (/root/reference/Makefile:161-163, .github/workflows/ci.yml:352-372): the graph builder (csrc/hnsw_build.c) and
# the CPU oracle (oracle/hnsw_oracle.c) are rebuilt instrumented into build/sanitize/ and the CPU tests that drive
# them run against those copies. The CUDA side has its own tool: tools/sanitize_device.sh (needs a GPU).
set -e
ROOT=$(cd "$(dirname "$0")/.." && pwd)
OUT="$ROOT/build/sanitize"
CC=/usr/bin/gcc
mkdir -p "$OUT"
SAN="-g -O1 -fno-omit-frame-pointer -fsanitize=address,undefined -fno-sanitize-recover=all"
$CC $SAN -std=c11 -fPIC -fopenmp -Wall -Wextra -I"$ROOT/include" -shared -o "$OUT/libmuninn_host.so" "$ROOT/sqlite_muninn_b200/csrc/hnsw_build.c" -lm
$CC $SAN -std=c11 -fPIC -Wall -Wextra -shared -o "$OUT/liboracle.so" "$ROOT/oracle/hnsw_oracle.c" -lm -lpthread
cd "$ROOT"
# python itself is not instrumented: preload the runtime; CPython's own allocations make leak reports useless
LD_PRELOAD="$($CC -print-file-name=libasan.so) $($CC -print-file-name=libubsan.so)" \
ASAN_OPTIONS=detect_leaks=0:abort_on_error=1 UBSAN_OPTIONS=print_stacktrace=1:halt_on_error=1 \
MUNINN_HOST_SO="$OUT/libmuninn_host.so" MUNINN_ORACLE_SO="$OUT/liboracle.so" \
python -m pytest -q -x -m "not gpu" -p no:cacheprovider \
    tests/test_device_build.py tests/test_host.py tests/test_oracle_golden.py tests/test_oracle_vs_reference.py \
    tests/test_touched_set.py tests/test_multi.py "$@"
