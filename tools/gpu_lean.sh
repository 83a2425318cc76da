# usage: bash tools/gpu_lean.sh -- after the lean producer / MMA-issuer loops: native checks + per-shape timings, the chain tests,
# the GPU suite, and the bench line at the three chain levels
mkdir -p gpurun_out
timeout 600 magic_b200/csrc/build/selftest_gemm bench > gpurun_out/selftest_bench.log 2>&1; echo "selftest rc=$?"; grep -c " OK" gpurun_out/selftest_bench.log; grep "FAIL\|SELFTEST" gpurun_out/selftest_bench.log | head; grep "^bench" gpurun_out/selftest_bench.log | grep -v "CTA pairs" | cut -c1-150
MAGIC_PARITY_TABLE=gpurun_out/parity_table.txt timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/pytest_gpu.log | cut -c1-250
for lvl in 2 1 0; do
MAGIC_CHAIN=$lvl timeout 600 python bench.py --no-cpu-baseline --no-extra-legs --no-e2e --steps 100 > gpurun_out/bench_chain$lvl.json 2> gpurun_out/bench_chain$lvl.err; echo "bench MAGIC_CHAIN=$lvl rc=$?"; cut -c1-160 gpurun_out/bench_chain$lvl.json
done
