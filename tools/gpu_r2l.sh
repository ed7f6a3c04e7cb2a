#!/bin/bash
# round-2 call L: full parity suite, ncu captures of the bench kernels (chain pass at N=1, box persistent), short bench
set -u
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r2l_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r2l_pytest.log
python tools/prof_case.py cfg4_open 3840 2160 64 auto 3 > gpurun_out/r2l_plain_open.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'rt(SpecPrimary|Chain)Kernel' -s 2 -c 2 -f -o gpurun_out/r2_chain_open \
   python tools/prof_case.py cfg4_open 3840 2160 64 auto 2 > gpurun_out/r2l_ncu_open.log 2>&1
echo "ncu open rc=$?"; cat gpurun_out/r2l_plain_open.log
python tools/prof_case.py cfg4_box 3840 2160 8 persistent 3 > gpurun_out/r2l_plain_box.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'rtVoteKernel' -s 1 -c 1 -f -o gpurun_out/r2_vote_box8 \
   python tools/prof_case.py cfg4_box 3840 2160 8 persistent 2 > gpurun_out/r2l_ncu_box.log 2>&1
echo "ncu box rc=$?"; cat gpurun_out/r2l_plain_box.log
python bench.py --steps 5 --warmup 3 > gpurun_out/r2l_bench.json 2> gpurun_out/r2l_bench.err; echo "bench rc=$?"; cut -c1-1500 gpurun_out/r2l_bench.json; tail -3 gpurun_out/r2l_bench.err
