#!/bin/bash
# libpymasc_io.so under the sanitizers (host code only, no GPU):
#   1. AddressSanitizer + UBSan build, tests/test_io_native.py through it (fuzzed DEFLATE streams,
#      hostile BGZF / BAM / BigWig files, records cut by blocks, shares and batches);
#   2. ThreadSanitizer build, a BAM load in many small batches and a BigWig prefetch on 6 threads.
# Usage: tools/sanitize_io.sh [file.bam file.bigwig]   (step 2 needs the two files, e.g. what
# tools/bam_e2e.py leaves in /tmp/pmb_e2e).  The release library is rebuilt at the end.
set -e
ROOT="$(cd "$(dirname "$0")/.." && pwd)"
CSRC="$ROOT/pymasc_b200/csrc"
TMP="$(mktemp -d)"
trap 'cd "$ROOT" && python -c "from pymasc_b200 import build; build.build_io(True)"; rm -rf "$TMP"' EXIT
cd "$CSRC"
g++ -O1 -g -std=c++17 -shared -fPIC -Wall -fsanitize=address,undefined -fno-omit-frame-pointer \
    -o "$ROOT/pymasc_b200/libpymasc_io.so" pmb_io.cpp -lz -lpthread
cd "$ROOT"
LD_PRELOAD="$(g++ -print-file-name=libasan.so)" ASAN_OPTIONS=detect_leaks=0 UBSAN_OPTIONS=print_stacktrace=1 \
    python -m pytest tests/test_io_native.py -x -q -s -p no:cacheprovider > "$TMP/asan.log" 2>&1 || { tail -n 40 "$TMP/asan.log"; exit 1; }
if grep -q "runtime error\|AddressSanitizer" "$TMP/asan.log"; then grep -m 20 "runtime error\|AddressSanitizer" "$TMP/asan.log"; exit 1; fi
echo "ASan + UBSan: $(tail -n 1 "$TMP/asan.log")"
if [ -n "$2" ]; then
    cd "$CSRC"
    g++ -O1 -g -std=c++17 -shared -fPIC -fsanitize=thread -o "$TMP/libpymasc_io.so" pmb_io.cpp -lz -lpthread
    cat > "$TMP/main.cpp" <<CPP
#include "$ROOT/include/pymasc_io.h"
#include <stdio.h>
int main(int argc, char** argv) {
    for (int rep = 0; rep < 2; rep++) {
        pmio_bam* b = nullptr;
        if (pmio_bam_load(argv[1], 1, 6, &b)) { printf("bam: %s\n", pmio_last_error()); return 1; }
        pmio_bam_stats st; pmio_bam_stats_get(b, &st); pmio_bam_free(b);
        pmio_bw* w = nullptr;
        if (pmio_bw_open(argv[2], &w) || pmio_bw_prefetch(w, 1.0f, 6)) { printf("bigwig: %s\n", pmio_last_error()); return 1; }
        pmio_bw_close(w);
        printf("accepted %llu\n", (unsigned long long)st.n_accepted);
    }
}
CPP
    g++ -O1 -g -fsanitize=thread -o "$TMP/main" "$TMP/main.cpp" -L"$TMP" -lpymasc_io -Wl,-rpath,"$TMP" -pthread
    PMIO_BATCH_BLOCKS=64 "$TMP/main" "$1" "$2" > "$TMP/tsan.log" 2>&1 || { tail -n 40 "$TMP/tsan.log"; exit 1; }
    if grep -q "ThreadSanitizer" "$TMP/tsan.log"; then grep -m 20 -A 12 "ThreadSanitizer" "$TMP/tsan.log"; exit 1; fi
    echo "TSan: clean ($(tail -n 1 "$TMP/tsan.log"))"
fi
