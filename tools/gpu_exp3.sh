mkdir -p gpurun_out
timeout 300 magic_b200/csrc/build/selftest_gemm bench > gpurun_out/selftest_s3.log 2>&1; echo "rc=$?"; grep -E "PASSED|FAILED|^bench" gpurun_out/selftest_s3.log | grep -v cluster
PROBE_STEPS=30 timeout 300 python tests/perf_probe.py 8d 4096 bf16 2>&1 | grep -E "8d B=|wgrad_enc0|fwd_enc0|dec_out"
