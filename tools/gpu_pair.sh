# usage: bash tools/gpu_pair.sh -- the CTA-pair (tcgen05 cta_group::2) kernels: native checks + timings against the single-CTA
# kernels, then the GPU suite and the bench line with and without them
mkdir -p gpurun_out
timeout 300 magic_b200/csrc/build/selftest_gemm pair > gpurun_out/selftest_pair.log 2>&1; echo "selftest pair rc=$?"; grep -c " OK" gpurun_out/selftest_pair.log; grep "FAIL\|error\|bench" gpurun_out/selftest_pair.log | head -40
timeout 300 magic_b200/csrc/build/selftest_gemm > gpurun_out/selftest.log 2>&1; echo "selftest rc=$?"; tail -1 gpurun_out/selftest.log
MAGIC_PARITY_TABLE=gpurun_out/parity_table.txt timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --no-cpu-baseline --no-extra-legs --steps 100 > gpurun_out/bench_pair.json 2> gpurun_out/bench_pair.err; echo "bench rc=$?"; cut -c1-200 gpurun_out/bench_pair.json; tail -3 gpurun_out/bench_pair.err
MAGIC_CG2=0 timeout 600 python bench.py --no-cpu-baseline --no-extra-legs --no-e2e --steps 100 > gpurun_out/bench_nopair.json 2> gpurun_out/bench_nopair.err; echo "bench(no pairs) rc=$?"; cut -c1-200 gpurun_out/bench_nopair.json
