# usage: bash tools/gpu_check2.sh  (under gpurun): self-test with GEMM micro-benchmarks, GPU tests, probe, and the ncu launch list of 2 steps
mkdir -p gpurun_out
timeout 300 magic_b200/csrc/build/selftest_gemm bench > gpurun_out/selftest.log 2>&1; echo "selftest rc=$?"; grep -i -E "fail|error|bench|PASSED" gpurun_out/selftest.log | head -40
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -15
MAGIC_VERBOSE=1 timeout 300 python tests/perf_probe.py 8d 4096 bf16 2>&1 | tee gpurun_out/probe_8d_bf16.txt
MAGIC_CLUSTER_OUT=1 timeout 300 python tests/perf_probe.py 8d 4096 bf16 2>&1 | grep -E "ms/step" | head -8
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e"
timeout 600 $CMD > gpurun_out/launch_plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --kernel-name-base demangled -c 400 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/launch_ncu.log 2>&1
tail -2 gpurun_out/launch_ncu.log
