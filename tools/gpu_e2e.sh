# usage: bash tools/gpu_e2e.sh -- API tests around magic_read_losses, then the bench line with the end-to-end leg, plus the
# generator-gate / dgrad-chain variants of the headline loop
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_tango_api.py tests/test_abi.py -q -x > gpurun_out/pytest_api.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_api.log | cut -c1-250
run() { name=$1; shift
  env "$@" timeout 600 python bench.py --no-cpu-baseline --no-extra-legs --no-e2e --steps 60 > gpurun_out/bench_$name.json 2> gpurun_out/bench_$name.err
  echo "bench $name ($*) rc=$?: $(python -c "import json,sys; d=json.load(open('gpurun_out/bench_$name.json')); print(round(d['value']), 'pairs/s', round(d['ms_per_step'],4), 'ms; wgrad_enc0 launch', d['roofline']['avg_launch_ms'])" 2>&1 | tail -1)"
}
run gate1 MAGIC_GEN_GATE=1
run gate2 MAGIC_GEN_GATE=2
run gate2_chain2 MAGIC_GEN_GATE=2 MAGIC_CHAIN=2
timeout 900 python bench.py --no-cpu-baseline --steps 100 > gpurun_out/bench_full.json 2> gpurun_out/bench_full.err; echo "full bench rc=$?"
python - <<'PY'
import json
d=json.load(open('gpurun_out/bench_full.json')); r=d['roofline']; e=d['e2e']
print('value', round(d['value']), d['ms_per_step'], 'resident', r.get('resident_value'), 'e2e', round(e['value']), e['ms_per_step'], 'f32 feed', round(e['float32_feed']['value']))
PY
