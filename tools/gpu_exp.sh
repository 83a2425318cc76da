# usage: bash tools/gpu_exp.sh  -- 3-D TMA boxes for MN-major operands: selftest + mainloop-only micro-benchmarks + tests + probe
mkdir -p gpurun_out
for d in 4 0; do
MAGIC_EPI_DBG=$d timeout 300 magic_b200/csrc/build/selftest_gemm bench > gpurun_out/selftest_dbg$d.log 2>&1; echo "dbg=$d selftest rc=$?"; grep -i -E "fail|error|PASSED|bench" gpurun_out/selftest_dbg$d.log | grep -v "cluster" | head -40
done
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
timeout 300 python tests/perf_probe.py 8d 4096 bf16 2>&1 | tee gpurun_out/probe_8d_bf16.txt | grep -E "8d B=|ms/step"
timeout 600 python bench.py --steps 100 --warmup 5 --no-cpu-baseline --e2e-steps 3 > gpurun_out/bench_try.json 2> gpurun_out/bench_try.err; tail -c 3000 gpurun_out/bench_try.json; tail -3 gpurun_out/bench_try.err
