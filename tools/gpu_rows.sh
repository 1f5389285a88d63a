#!/bin/bash
# Iteration run for the TMA-rows kernel: its tests, C3 timing, optional ncu.
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
TAG=${1:-rows}
timeout ${PYTEST_TIMEOUT:-400} python -m pytest tests/test_gpu_guarded.py -x -q -m gpu -k "${KEXPR:-unaligned or tma}" > gpurun_out/${TAG}_pytest.log 2>&1
echo "pytest rc=$?"; tail -15 gpurun_out/${TAG}_pytest.log
L=mlcomp_b200/lib
( timeout 300 python tools/bench_configs.py ${CONFIGS:-c3:1024 c3}
  for v in $VARIANTS; do
    MLC_LIB=$L/libmlcomp_b200_$v.so timeout 300 python tools/bench_configs.py ${CONFIGS:-c3:1024 c3}
  done ) > gpurun_out/${TAG}_bench.jsonl 2> gpurun_out/${TAG}_bench.err
cat gpurun_out/${TAG}_bench.jsonl; tail -3 gpurun_out/${TAG}_bench.err
if [ -n "$NCU" ]; then
ncu --set full --clock-control none --import-source on -k regex:seg_hist_guarded -s ${NCU_SKIP:-3} -c 1 -f -o gpurun_out/${TAG}_rows python tools/bench_configs.py c3:1024 > gpurun_out/${TAG}_ncu.log 2>&1
echo "ncu rc=$?"
fi
