# usage: bash tools/gpu_exp5.sh -- dataset-path tests + the default bench (all e2e feeds)
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_overlap_input.py -m gpu -x -q 2>&1 | tail -15
timeout 900 python bench.py --no-cpu-baseline > gpurun_out/bench_e2e.json 2> gpurun_out/bench_e2e.err; python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_e2e.json').read().strip().splitlines()[-1])
print({k:d.get(k) for k in ('value','ms_per_step','gpu_launches','clocks')})
print(json.dumps(d['e2e'], indent=1))
PY
tail -3 gpurun_out/bench_e2e.err
