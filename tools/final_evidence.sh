#!/bin/bash
# One-GPU evidence run of a round (under gpurun): tests, bench lines, latency, launch list, ncu captures.
# usage: tools/final_evidence.sh <tag>      writes gpurun_out/<tag>_*
tag=${1:-r03}
out=gpurun_out
mkdir -p $out
python -m pytest tests -m gpu -q > $out/${tag}_pytest_gpu.txt 2>&1
tail -1 $out/${tag}_pytest_gpu.txt
make -s -C tests/cpp > /dev/null 2>&1
for r in 1 2; do tests/cpp/test_group.bin $r 2000 24; done > $out/${tag}_cpp_group_test.txt 2>&1; tail -1 $out/${tag}_cpp_group_test.txt
python bench.py --steps 20 --warmup 5 > $out/${tag}_bench.json 2> $out/${tag}_bench.err; echo "bench rc=$?"
python tools/latency.py > $out/${tag}_latency.json 2> $out/${tag}_latency.err; echo "latency rc=$?"
# launch list of the same command (cold, serialised: shares only)
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $out/${tag}_launches.csv \
    python bench.py --steps 2 --warmup 3 --no-ref > $out/${tag}_ncu_launches.log 2>&1; echo "launch list rc=$?"
# the three kernels of a C5 step, full set with source
ncu --set full --clock-control none --import-source on -k regex:'k_tc_shortlist|k_rerank|k_resolve_bow' --launch-skip 9 \
    --launch-count 3 -f -o $out/${tag}_c5 python bench.py --steps 2 --warmup 3 --no-ref > $out/${tag}_ncu_c5.log 2>&1; echo "ncu c5 rc=$?"
# the tracking kernels (C2 / C3)
ncu --set full --clock-control none --import-source on -k regex:'k_proj_fused|k_stereo_match' --launch-skip 40 \
    --launch-count 3 -f -o $out/${tag}_tracking python tools/latency.py --only C2,C3 > $out/${tag}_ncu_tracking.log 2>&1; echo "ncu tracking rc=$?"
python bench.py --impl reference --steps 20 --warmup 5 > $out/${tag}_bench_reference_arm.json 2> $out/${tag}_ref.err; echo "reference arm rc=$?"
