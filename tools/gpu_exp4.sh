# usage: bash tools/gpu_exp4.sh -- selftest, GPU tests, fused-vs-separate small wgrads
mkdir -p gpurun_out
timeout 300 magic_b200/csrc/build/selftest_gemm > gpurun_out/selftest.log 2>&1; echo "selftest rc=$?"; grep -i -E "fail|error|PASSED|appended" gpurun_out/selftest.log | head -30
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
for f in 1 0; do
MAGIC_FUSE_WGRAD=$f timeout 600 python bench.py --steps 100 --warmup 5 --no-cpu-baseline --no-e2e > gpurun_out/bench_fw$f.json 2> gpurun_out/bench_fw$f.err; F=$f python - <<'PY'
import json,sys,os
d=json.loads(open('gpurun_out/bench_fw%s.json'%os.environ['F']).read().strip().splitlines()[-1])
print('FUSE', os.environ['F'], {k:d.get(k) for k in ('value','ms_per_step','with_generator','ms_per_step_with_per_launch_events','gpu_launches','clocks')})
print(d['kernel_time_shares']); print(d['roofline'])
PY
tail -3 gpurun_out/bench_fw$f.err
done
