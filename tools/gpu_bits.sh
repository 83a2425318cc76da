# usage: bash tools/gpu_bits.sh -- packed first-layer operand: native self-test, GPU tests, bench in the operand modes
mkdir -p gpurun_out
for es in 1; do
  timeout 600 magic_b200/csrc/build/selftest_gemm bench > gpurun_out/selftest_bits$es.log 2>&1; rc=$?; echo "selftest split=$es rc=$rc"
  grep -E "bits|FAIL|SELFTEST|enc L0" gpurun_out/selftest_bits$es.log | grep -v "OK$" | cut -c1-150
  if [ $rc -ne 0 ]; then exit 1; fi
done
if [ "$1" != "quick" ]; then timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -8; fi
timeout 600 python bench.py --steps 60 --warmup 5 --no-cpu-baseline > gpurun_out/bench_bits1.json 2> gpurun_out/bench_bits1.err; echo "bench rc=$?"
MAGIC_SKIP_ZERO=0 timeout 600 python bench.py --steps 60 --warmup 5 --no-cpu-baseline > gpurun_out/bench_planes.json 2> gpurun_out/bench_planes.err; echo "bench(dense) rc=$?"
python - <<'PY'
import json
for f in ("bits1", "planes"):
    try:
        d = json.load(open("gpurun_out/bench_%s.json" % f))
        print(f, "value %.0f ms %.4f gen %.4f e2e %.4f" % (d["value"], d["ms_per_step"], d["with_generator"]["ms_per_step"], d["e2e"]["ms_per_step"]),
              {k: v for k, v in d["kernel_time_shares"].items() if "enc0" in k}, d["roofline"]["avg_launch_ms"], d["clocks"]["sm_mhz"])
    except Exception as e:
        print(f, "failed", e)
PY
