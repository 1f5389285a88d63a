#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
python tools/bench_guarded.py tta > gpurun_out/n_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:seg_hist_guarded -s 3 -c 2 -f -o gpurun_out/r02_guarded_v1 python tools/bench_guarded.py tta > gpurun_out/n_ncu.log 2>&1
echo "rc=$?"; tail -3 gpurun_out/n_ncu.log
