# usage: bash tools/gpu_check.sh   (under gpurun): native self-test, GPU tests, per-kernel probe of the bench workload
mkdir -p gpurun_out
timeout 300 magic_b200/csrc/build/selftest_gemm > gpurun_out/selftest.log 2>&1; echo "selftest rc=$?"; grep -c -i fail gpurun_out/selftest.log; grep -i -E "fail|error" gpurun_out/selftest.log | head
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -15
timeout 300 python tests/perf_probe.py 8d 4096 bf16 2>&1 | tee gpurun_out/probe_8d_bf16.txt
