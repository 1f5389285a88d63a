#!/bin/bash
# First GPU call of round 2: tests, A/B of the guarded kernel, full bench.
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/a_smi.txt 2>&1
timeout 900 python -m pytest tests/test_gpu_guarded.py -x -q -m gpu > gpurun_out/a_pytest_guarded.log 2>&1
echo "guarded rc=$?" >> gpurun_out/a_pytest_guarded.log
L=mlcomp_b200/lib
( timeout 300 python tools/bench_guarded.py tta sharp c5 rle
  MLC_LIB=$L/libmlcomp_b200_dev.so timeout 300 python tools/bench_guarded.py tta sharp c5
  MLC_LIB=$L/libmlcomp_b200_dev.so MLC_HIST_VARIANT=6 timeout 300 python tools/bench_guarded.py tta c5
  for v in g384x2 g256x3 g512x1 g256x4 g1024x1 g768x1; do
    MLC_LIB=$L/libmlcomp_b200_$v.so timeout 200 python tools/bench_guarded.py tta c5
  done ) > gpurun_out/a_bench_guarded.jsonl 2> gpurun_out/a_bench_guarded.err
timeout 900 python bench.py > gpurun_out/a_bench.json 2> gpurun_out/a_bench.err
echo "bench rc=$?" >> gpurun_out/a_bench.err
timeout 1200 python -m pytest tests -x -q -m gpu > gpurun_out/a_pytest_all.log 2>&1
echo "all rc=$?" >> gpurun_out/a_pytest_all.log
tail -3 gpurun_out/a_pytest_guarded.log gpurun_out/a_pytest_all.log
cat gpurun_out/a_bench_guarded.jsonl
tail -c 3000 gpurun_out/a_bench.json
tail -5 gpurun_out/a_bench.err
