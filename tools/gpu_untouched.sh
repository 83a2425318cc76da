# usage: bash tools/gpu_untouched.sh -- validates and times the experimental option skip_untouched_rows on one GPU:
# native kernel check, the engine-level bit-equality test, then the bench loop with the option off and on (same box)
mkdir -p gpurun_out
timeout 60 magic_b200/csrc/build/selftest_gemm adam | tail -6
MAGIC_TEST_EXPERIMENTAL=1 timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -q -k untouched 2>&1 | tail -4
for v in 0 1; do
  MAGIC_SKIP_UNTOUCHED=$v timeout 300 python bench.py --steps 100 --warmup 10 --no-cpu-baseline --no-e2e \
    > gpurun_out/bench_untouched_$v.json 2> gpurun_out/bench_untouched_$v.err
  echo "skip_untouched_rows=$v rc=$?"; cut -c1-260 gpurun_out/bench_untouched_$v.json; echo
done
