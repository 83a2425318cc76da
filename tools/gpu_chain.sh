# usage: bash tools/gpu_chain.sh -- the hidden-layer chain kernels: on/off equality tests, then the bench line with and without
mkdir -p gpurun_out
MAGIC_VERBOSE=1 timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "chains" > gpurun_out/chain_tests.log 2>&1; echo "chain tests rc=$?"; grep "chain:" gpurun_out/chain_tests.log | sort | uniq -c | head; tail -12 gpurun_out/chain_tests.log | cut -c1-250
timeout 600 python bench.py --no-cpu-baseline --no-extra-legs --no-e2e --steps 100 > gpurun_out/bench_chain.json 2> gpurun_out/bench_chain.err; echo "bench rc=$?"; cut -c1-200 gpurun_out/bench_chain.json; tail -3 gpurun_out/bench_chain.err
MAGIC_CHAIN=0 timeout 600 python bench.py --no-cpu-baseline --no-extra-legs --no-e2e --steps 100 > gpurun_out/bench_nochain.json 2> gpurun_out/bench_nochain.err; echo "bench(no chain) rc=$?"; cut -c1-200 gpurun_out/bench_nochain.json
