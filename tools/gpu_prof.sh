#!/bin/bash
# Launch list of the bench's timed region + full capture of its top kernel,
# at the bench's full size (IMAGES=0) unless IMAGES says otherwise.
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
IM=${IMAGES:-0}
CMD="python bench.py --steps 2 --warmup 3 --no-cpu --no-side --images $IM --e2e-images 64 --e2e-steps 1"
$CMD > gpurun_out/p_plain.json 2> gpurun_out/p_plain.err &&
ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/r02_launches_pipeline.csv $CMD > gpurun_out/p_ncu1.log 2>&1
echo "launch list rc=$?"
$CMD > gpurun_out/p_plain2.json 2> gpurun_out/p_plain2.err &&
ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:seg_hist_guarded -c 1 -f -o gpurun_out/r02_seg_hist_guarded_pipeline $CMD > gpurun_out/p_ncu2.log 2>&1
echo "full capture rc=$?"
tail -2 gpurun_out/p_ncu2.log
