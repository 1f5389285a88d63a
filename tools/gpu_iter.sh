#!/bin/bash
# Iteration run: guarded tests, kernel timings (default + dev), optional ncu.
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
TAG=${1:-it}
timeout 600 python -m pytest tests/test_gpu_guarded.py tests/test_gpu_argmax.py -x -q -m gpu > gpurun_out/${TAG}_pytest.log 2>&1
echo "pytest rc=$?"; tail -2 gpurun_out/${TAG}_pytest.log
L=mlcomp_b200/lib
( timeout 300 python tools/bench_guarded.py tta c5
  [ -f $L/libmlcomp_b200_dev.so ] && MLC_LIB=$L/libmlcomp_b200_dev.so timeout 300 python tools/bench_guarded.py tta c5
  for v in $VARIANTS; do
    MLC_LIB=$L/libmlcomp_b200_$v.so timeout 200 python tools/bench_guarded.py tta c5
  done ) > gpurun_out/${TAG}_bench.jsonl 2> gpurun_out/${TAG}_bench.err
cat gpurun_out/${TAG}_bench.jsonl; tail -3 gpurun_out/${TAG}_bench.err
if [ -n "$NCU" ]; then
python tools/bench_guarded.py tta > gpurun_out/${TAG}_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:seg_hist_guarded -s ${NCU_SKIP:-3} -c 1 -f -o gpurun_out/${TAG}_guarded python tools/bench_guarded.py tta > gpurun_out/${TAG}_ncu.log 2>&1
echo "ncu rc=$?"
fi
