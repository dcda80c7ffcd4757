#!/bin/bash
# The CUDA kernels with their own device-side bounds checks (-DPMB_CHECKED, pmb_window.cuh):
# compute-sanitizer is closed on the B200 pool (profiles/r02_sanitizer_closed_on_this_pool.log),
# so this is the memcheck of this repo.  Builds pymasc_b200/libpymasc_b200_checked.so and runs the
# whole -m gpu suite against it; a failed check turns into an error of the call, so tests fail.
#     tools/checked_build.sh build      (here, no GPU)
#     gpurun -- 'tools/checked_build.sh run'
set -eu
cd "$(dirname "$0")/.."
out=pymasc_b200/libpymasc_b200_checked.so
case "${1:-run}" in
build)
    (cd pymasc_b200/csrc && nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 --shared \
        -Xcompiler -fPIC -DPMB_CHECKED -o ../libpymasc_b200_checked.so pmb_api.cu)
    echo "built $out"
    ;;
run)
    mkdir -p gpurun_out
    [ -f "$out" ] || { echo "$out missing: run '$0 build' first"; exit 1; }
    export PYMASC_B200_LIB="$PWD/$out"
    # 1. the chain is live: with every TMA range check forced to fail, the call must report it
    python - <<'PY' | tee gpurun_out/checked_selftest.log
import numpy as np
from pymasc_b200 import _shim
d = _shim.Device(0)
d.set_option("path", 2)
d.set_option("check_selftest", 1)
try:
    d.run_batch([_shim.Job(50000, 36, 200, 1, np.arange(10, 40000, 7, dtype=np.uint32), np.arange(50, 40000, 9, dtype=np.uint32))])
    print("SELFTEST FAILED: a forced check failure went unreported")
except _shim.DeviceError as e:
    print("selftest ok:", e)
PY
    grep -q "selftest ok" gpurun_out/checked_selftest.log
    # 2. the GPU suite against the checked build
    python -m pytest tests -x -q -m gpu -p no:cacheprovider 2>&1 | tail -6 | tee gpurun_out/checked_pytest.log
    ;;
esac
