# usage: bash tools/gpu_adam.sh -- the first layer's Adam in the wgrad epilogue: parity tests, then the bench line with the
# option on / off and the generator gate / dgrad chain variants
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -q -x -k "adam_in_the_wgrad or ten_steps or update or hidden_layer_chains" > gpurun_out/pytest_adam.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_adam.log | cut -c1-250
run() { # name, env...
  name=$1; shift
  env "$@" timeout 600 python bench.py --no-cpu-baseline --no-extra-legs --no-e2e --steps 100 > gpurun_out/bench_$name.json 2> gpurun_out/bench_$name.err
  echo "bench $name ($*) rc=$?: $(python -c "import json,sys; d=json.load(open('gpurun_out/bench_$name.json')); print(round(d['value']), 'pairs/s', round(d['ms_per_step'],4), 'ms')" 2>&1 | tail -1)"
}
run fuse1 MAGIC_FUSE_ADAM=1
run fuse0 MAGIC_FUSE_ADAM=0



MAGIC_VERBOSE=1 timeout 600 python bench.py --no-cpu-baseline --steps 50 > gpurun_out/bench_full.json 2> gpurun_out/bench_full.err; echo "full bench rc=$?"; grep "magic\]" gpurun_out/bench_full.err | sort | uniq -c | head; cut -c1-1500 gpurun_out/bench_full.json
