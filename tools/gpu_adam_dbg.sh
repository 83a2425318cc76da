# usage: bash tools/gpu_adam_dbg.sh -- where does the fused first-layer Adam lose its time: timing experiments (wrong results)
mkdir -p gpurun_out
run() { name=$1; shift
  env "$@" timeout 600 python bench.py --no-cpu-baseline --no-extra-legs --no-e2e --steps 60 > gpurun_out/bench_$name.json 2> gpurun_out/bench_$name.err
  echo "bench $name ($*) rc=$?: $(python -c "import json,sys; d=json.load(open('gpurun_out/bench_$name.json')); print(round(d['value']), 'pairs/s', round(d['ms_per_step'],4), 'ms; wgrad_enc0 launch', d['roofline']['avg_launch_ms'])" 2>&1 | tail -1)"
}
run plain MAGIC_FUSE_ADAM=0
run fused MAGIC_FUSE_ADAM=1
run adamkernel_plainorder_allstored MAGIC_FUSE_ADAM=1 MAGIC_ADAM_DBG=24
run fused_nomem MAGIC_FUSE_ADAM=1 MAGIC_ADAM_DBG=1
run fused_nothing MAGIC_FUSE_ADAM=1 MAGIC_ADAM_DBG=4
