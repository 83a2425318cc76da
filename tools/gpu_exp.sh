# usage: bash tools/gpu_exp.sh  -- selftest, GPU tests, probe, bench
mkdir -p gpurun_out
timeout 300 magic_b200/csrc/build/selftest_gemm bench > gpurun_out/selftest.log 2>&1; echo "selftest rc=$?"; grep -i -E "fail|error|PASSED|bench (hidden|dec hidden)" gpurun_out/selftest.log | grep -v cluster | head -30
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
timeout 300 python tests/perf_probe.py 8d 4096 bf16 2>&1 | tee gpurun_out/probe_8d_bf16.txt | grep -E "8d B=|ms/step"
timeout 600 python bench.py --steps 100 --warmup 5 --no-cpu-baseline --no-e2e > gpurun_out/bench_try.json 2> gpurun_out/bench_try.err; python - <<'PY'
import json,sys
d=json.loads(open('gpurun_out/bench_try.json').read().strip().splitlines()[-1])
print({k:d[k] for k in ('value','ms_per_step','with_generator','ms_per_step_with_per_launch_events','clocks')})
print(d['kernel_time_shares'])
PY
tail -3 gpurun_out/bench_try.err
