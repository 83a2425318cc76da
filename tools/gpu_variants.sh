run() { name=$1; shift
  env "$@" timeout 300 python bench.py --no-cpu-baseline --no-extra-legs --no-e2e --steps 100 > gpurun_out/bench_v_$name.json 2> gpurun_out/bench_v_$name.err
  echo "bench $name ($*): $(python -c "import json; d=json.load(open('gpurun_out/bench_v_$name.json')); print(round(d['value']), 'pairs/s', round(d['ms_per_step'],4), 'ms')" 2>&1 | tail -1)"
}
run default MAGIC_GEN_GATE=1
run gate0 MAGIC_GEN_GATE=0
run chain2 MAGIC_CHAIN=2
run early_adam0 MAGIC_EARLY_ADAM=0
run default_again MAGIC_GEN_GATE=1
