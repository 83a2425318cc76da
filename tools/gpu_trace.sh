# usage: bash tools/gpu_trace.sh -- per-CTA phase stamps of the GEMM micro-benchmarks
mkdir -p gpurun_out
MAGIC_TRACE=1 timeout 300 magic_b200/csrc/build/selftest_gemm bench > gpurun_out/selftest_trace.log 2>&1; echo "rc=$?"
grep -v "^ok\|^test" gpurun_out/selftest_trace.log | grep -A11 "trace hidden\|trace dec out fwd  \|trace enc L0 wgrad  \|trace dec-out shape, K=256" | head -80
