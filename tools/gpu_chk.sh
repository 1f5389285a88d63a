#!/bin/bash
# The GPU suite on the bounds-checked library (every load of the
# segmentation kernels checked against the registered tensor ranges).
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
MLC_LIB=$PWD/mlcomp_b200/lib/libmlcomp_b200_chk.so timeout 2400 python -m pytest tests -x -q -m gpu \
  --deselect tests/test_zz_no_alloc.py > gpurun_out/chk_pytest.log 2>&1
echo "bounds-checked suite rc=$?"; tail -4 gpurun_out/chk_pytest.log
grep -c "MLC_BOUNDS_CHECK" gpurun_out/chk_pytest.log || true
