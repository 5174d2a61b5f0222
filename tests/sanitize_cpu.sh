#!/bin/bash
# The CPU test suite against a copy of the library whose host code is built with AddressSanitizer and UBSan
# (loader, blob validator, shard builder, TSV writer and reader, side-car, the host side of the C-ABI; no kernel runs
# without a GPU).  Builds in a scratch copy so that the in-tree library stays the product build.
#   tests/sanitize_cpu.sh [scratch_dir]              -> prints the pytest summary and every sanitizer report
#   SAN=thread tests/sanitize_cpu.sh [scratch_dir]   -> the same under ThreadSanitizer (the loader, the writers and the table
#                                                       reader are multithreaded); reports inside libtorch's gloo are not ours,
#                                                       and the tests that run the reference binary as a child fail under the preload
#   FUZZ=1 tests/sanitize_cpu.sh [scratch_dir]      -> after the suite, tests/fuzz_host.py (damaged blobs, index files, tables) on the same build
set -euo pipefail
ROOT="$(cd "$(dirname "$0")/.." && pwd)"
DST="${1:-$(mktemp -d /tmp/ksb_asan.XXXXXX)}"
if [ "${SAN:-address}" = thread ]; then
    FLAGS="-fsanitize=thread"; PRE="$(gcc -print-file-name=libtsan.so)"
else
    FLAGS="-fsanitize=address,-fsanitize=undefined"; PRE="$(gcc -print-file-name=libasan.so) $(gcc -print-file-name=libubsan.so)"
fi
mkdir -p "$DST"
tar -C "$ROOT" --exclude=.git --exclude=gpurun_out --exclude='kspider_b200/csrc/build' --exclude='libkspider_b200.so' \
    --exclude='kspider_b200/bin' -cf - . | tar -xf - -C "$DST"
# (the `pairwise` binary of the Makefile's `all` does not link against a sanitized library: only the library is made)
make -s -j"$(nproc)" -C "$DST/kspider_b200/csrc" "$DST/kspider_b200/libkspider_b200.so" \
    OUT="$DST/kspider_b200/libkspider_b200.so" \
    CXXFLAGS="-O1 -g -std=c++17 -Xcompiler -fPIC,-pthread,$FLAGS,-fno-omit-frame-pointer" \
    > "$DST/sanitize_build.log" 2>&1 || { tail -n 30 "$DST/sanitize_build.log"; exit 1; }
cd "$DST"
LOG="$DST/sanitize_cpu.log"
LD_PRELOAD="$PRE" TSAN_OPTIONS="halt_on_error=0 report_signal_unsafe=0" \
ASAN_OPTIONS=detect_leaks=0:halt_on_error=1 UBSAN_OPTIONS=print_stacktrace=1:halt_on_error=0 \
    python -m pytest tests/ -q -s -m "not gpu" --deselect tests/test_dropin_binary.py > "$LOG" 2>&1 || true
tail -n 3 "$LOG"
echo "sanitizer reports: $(grep -c -E 'runtime error|AddressSanitizer|WARNING: ThreadSanitizer' "$LOG" || true)," \
     "lines naming libkspider_b200: $(grep -c libkspider_b200 "$LOG" || true) (log: $LOG)"
grep -E 'runtime error|AddressSanitizer' "$LOG" | sort | uniq -c | sort -rn | head -n 40 || true
if [ "${FUZZ:-0}" != 0 ]; then
    LD_PRELOAD="$PRE" TSAN_OPTIONS="halt_on_error=0 report_signal_unsafe=0" \
    ASAN_OPTIONS=detect_leaks=0:halt_on_error=1 UBSAN_OPTIONS=print_stacktrace=1:halt_on_error=0 \
        python tests/fuzz_host.py "${FUZZ_SEED:-0}" > "$DST/fuzz_host.log" 2>&1 || echo "fuzz_host.py did not finish: see $DST/fuzz_host.log"
    grep -E 'accepted|return code|BLOB REFUSED|runtime error|AddressSanitizer|ThreadSanitizer' "$DST/fuzz_host.log" | head -n 20 || true
fi
