#!/bin/bash
# The CUDA kernels under compute-sanitizer (the reference's counterpart is the
# always-on BitArray asserts, bit_array.c:663, :670, :825).  ONE tool per call (B200_PROFILING.md):
#     gpurun -- 'tools/sanitize_gpu.sh memcheck'      (or racecheck | synccheck | initcheck)
# Runs a subset of the -m gpu tests that covers all three device algorithms, grouped asynchronous
# H2D (two streams), shared-track followers, the unsorted fallbacks, max_shift 9000 and the
# allocation-free entry point -- at sizes that finish under the tools' 10-100x slow-down.
# Log: gpurun_out/sanitizer_<tool>.log (copy the summary into profiles/).
set -u
tool=${1:-memcheck}
mkdir -p gpurun_out
sel='random_chromosomes_bit_exact or edge_inputs or mixed_jobs or large_max_shift or mappability_gather or packed_read_arrays or pipelined_input_copy or sharing_one_track or shared_track_position or unsorted_input or no_alloc or inputs_packed or position_shards_with_halos'
extra=""
[ "$tool" = "racecheck" ] && extra="--racecheck-report all"
[ "$tool" = "memcheck" ] && extra="--leak-check no"
timeout 3000 compute-sanitizer --tool "$tool" $extra --error-exitcode 0 --print-limit 40 \
    --log-file "gpurun_out/sanitizer_${tool}.log" \
    python -m pytest tests/test_gpu_parity.py tests/test_zz_no_alloc.py -x -q -m gpu -k "$sel" -p no:cacheprovider \
    > "gpurun_out/sanitizer_${tool}_pytest.log" 2>&1
echo "pytest exit $?" >> "gpurun_out/sanitizer_${tool}_pytest.log"
tail -4 "gpurun_out/sanitizer_${tool}_pytest.log"
grep -E "ERROR SUMMARY|RACECHECK SUMMARY|Error:|Race reported|Hazard" "gpurun_out/sanitizer_${tool}.log" | sort | uniq -c | sort -rn | head -20
