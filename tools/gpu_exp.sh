# usage: bash tools/gpu_exp.sh  -- selftest, GPU tests, probe, bench (early decoder Adam on/off)
mkdir -p gpurun_out
timeout 300 magic_b200/csrc/build/selftest_gemm > gpurun_out/selftest.log 2>&1; echo "selftest rc=$?"; grep -i -E "fail|error|PASSED" gpurun_out/selftest.log | head
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
timeout 300 python tests/perf_probe.py 8d 4096 bf16 2>&1 | tee gpurun_out/probe_8d_bf16.txt | grep -E "8d B=|ms/step"
for ea in 1 0; do
MAGIC_EARLY_ADAM=$ea timeout 600 python bench.py --steps 100 --warmup 5 --no-cpu-baseline --no-e2e > gpurun_out/bench_try$ea.json 2> gpurun_out/bench_try.err; python - $ea <<'PY'
import json,sys
d=json.loads(open('gpurun_out/bench_try%s.json'%sys.argv[1]).read().strip().splitlines()[-1])
print("EARLY_ADAM", sys.argv[1], {k:d[k] for k in ('value','ms_per_step','with_generator','ms_per_step_with_per_launch_events','clocks')})
print(d['kernel_time_shares'])
PY
done
tail -3 gpurun_out/bench_try.err
