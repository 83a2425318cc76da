# usage: bash tools/gpu_trace.sh -- per-CTA phase stamps of the GEMM micro-benchmarks
mkdir -p gpurun_out
MAGIC_TRACE=1 timeout 300 magic_b200/csrc/build/selftest_gemm bench > gpurun_out/selftest_trace.log 2>&1; echo "rc=$?"
grep -A18 "trace hidden fwd, tanh epilogue" gpurun_out/selftest_trace.log | head -22
grep -A18 "trace hidden dgrad, tanh' epilogue" gpurun_out/selftest_trace.log | head -22
