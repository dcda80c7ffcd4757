#!/bin/bash
# One B200: the GPU suite, the allocation-free test alone, the -DPMB_CHECKED suite, an ncu capture of
# every kernel of one list-path step (-> the measured-traffic file bench.py reads), the bench line and
# the ncu launch list of the bench command.  Everything lands in gpurun_out/ under the given tag.
#     gpurun --timeout 900 -- 'tools/round_measure.sh e'
set -u
tag=${1:-x}
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q -p no:cacheprovider > gpurun_out/r02_pytest_gpu_$tag.log 2>&1; tail -1 gpurun_out/r02_pytest_gpu_$tag.log
python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q -p no:cacheprovider > gpurun_out/r02_pytest_noalloc_alone_$tag.log 2>&1; tail -1 gpurun_out/r02_pytest_noalloc_alone_$tag.log
tools/checked_build.sh run > gpurun_out/r02_checked_$tag.log 2>&1; tail -2 gpurun_out/r02_checked_$tag.log
python tools/list_prof.py --execs 3 > gpurun_out/list_prof_$tag.json 2>&1 &&
ncu --set full --import-source on --clock-control none -s 8 -c 4 -f -o gpurun_out/r02_list_step_$tag \
    python tools/list_prof.py --execs 3 > gpurun_out/ncu_step_$tag.log 2>&1
ncu -i gpurun_out/r02_list_step_$tag.ncu-rep --page raw --csv > gpurun_out/r02_list_step_${tag}_raw.csv 2>/dev/null &&
python tools/ncu_traffic.py gpurun_out/r02_list_step_${tag}_raw.csv profiles/r02_traffic_lists.json 9.8e10 > gpurun_out/ncu_traffic_$tag.log 2>&1 &&
cp profiles/r02_traffic_lists.json gpurun_out/r02_traffic_lists_$tag.json
python bench.py --steps 20 --warmup 3 > gpurun_out/r02_bench_1gpu_$tag.json 2> gpurun_out/r02_bench_1gpu_$tag.err; tail -c 600 gpurun_out/r02_bench_1gpu_$tag.json
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_launches_lists_$tag.csv \
    python bench.py --steps 5 --warmup 3 --no-configs --no-bam --no-cpu-baseline --single-path > gpurun_out/ncu_bench_$tag.log 2>&1
echo finished
