# usage: bash tools/gpu_selftest.sh  -- GEMM selftest only
mkdir -p gpurun_out
timeout 300 magic_b200/csrc/build/selftest_gemm > gpurun_out/selftest.log 2>&1; echo "selftest rc=$?"; grep -i -E "fail|error|PASSED|extras" gpurun_out/selftest.log | head -40; tail -5 gpurun_out/selftest.log
