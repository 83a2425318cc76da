# usage: bash tools/gpu_final.sh [ncu] -- native self-test, GPU tests (per-tensor parity table -> gpurun_out/parity_table.txt),
# smoke, default bench line, reference arm, ncu launch list of the bench command; with "ncu": also one full-set capture
# each of the two first-layer GEMMs
mkdir -p gpurun_out
rm -f gpurun_out/parity_table.txt
timeout 300 magic_b200/csrc/build/selftest_gemm > gpurun_out/selftest.log 2>&1; echo "selftest rc=$?"; tail -1 gpurun_out/selftest.log
MAGIC_PARITY_TABLE=gpurun_out/parity_table.txt timeout 1500 python -m pytest tests -m gpu -q ${PYTEST_ARGS} > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/pytest_gpu.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -3
timeout 900 python bench.py ${BENCH_ARGS} > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "bench rc=$?"; cut -c1-600 gpurun_out/bench_default.json; tail -5 gpurun_out/bench_default.err
if [ "$1" != "quick" ]; then
timeout 600 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err; echo "ref rc=$?"; cut -c1-300 gpurun_out/bench_reference.json
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-extra-legs"
timeout 600 $CMD > gpurun_out/launch_plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --kernel-name-base demangled -c 800 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/launch_ncu.log 2>&1
tail -2 gpurun_out/launch_ncu.log | cut -c1-200
fi
if [ "$1" = "ncu" ]; then
  timeout 900 bash tools/gpu_ncu_one.sh wgrad_enc0_bits 'gemm_tcgen05_kernel<.bool.1, .bool.1, .int.0, .int.256, .bool.1>' 3
  timeout 900 bash tools/gpu_ncu_one.sh fwd_enc0_bits 'gemm_tcgen05_kernel<.bool.0, .bool.1, .int.0, .int.256, .bool.1>' 3
  ls -la gpurun_out/*.ncu-rep
fi
