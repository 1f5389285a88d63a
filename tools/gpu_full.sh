#!/bin/bash
# Full check: all GPU tests, smoke, the bench (N=1) and its reference arm.
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
TAG=${1:-full}
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/${TAG}_pytest.log 2>&1
echo "pytest rc=$?"; tail -3 gpurun_out/${TAG}_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${TAG}_smoke.log 2>&1
echo "smoke rc=$?"; tail -2 gpurun_out/${TAG}_smoke.log
timeout 900 python bench.py > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err
echo "bench rc=$?"; tail -c 600 gpurun_out/${TAG}_bench.err
python - $TAG <<'PY'
import json,sys
try:
    d=json.loads(open('gpurun_out/'+sys.argv[1]+'_bench.json').read().strip().splitlines()[-1])
    print({k:d[k] for k in ('value','ms_per_step','validated')})
    print('roofline', {k:d['roofline'][k] for k in ('achieved','frac','avg_launch_ms')}, d['roofline']['whole_step']['frac'])
    print('e2e', d['e2e']['value'], d['e2e'].get('h2d_gbs_all_gpus'))
    print('cpu', d['cpu_baseline']['value'], d['cpu_baseline']['kind'])
    print({k:d[k] for k in d if k in ('pipeline','pipeline_rereading_sources','c2_single_source')})
    print('parity', {k:v for k,v in d['parity'].items() if k not in ('best','straddlers_listed','sigmoid')})
except Exception as e:
    print('no bench json', e)
PY
