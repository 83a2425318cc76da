# usage (gpurun --gpus N): bash tools/gpu_scale.sh N -- the bench line at N GPUs (headline loop + end-to-end leg, no CPU leg), fused
# reduce-scatter path; at N = 2 the data-parallel parity tests first
N=$1
mkdir -p gpurun_out
if [ "$N" = "2" ]; then
timeout 600 python -m pytest tests/test_dp_gpu.py -m gpu -q > gpurun_out/dp_tests.log 2>&1; echo "dp tests (fused) rc=$?"; tail -3 gpurun_out/dp_tests.log
MAGIC_DP_MODE=allreduce timeout 600 python -m pytest tests/test_dp_gpu.py -m gpu -q > gpurun_out/dp_tests_allreduce.log 2>&1; echo "dp tests (NCCL all-reduce baseline) rc=$?"; tail -3 gpurun_out/dp_tests_allreduce.log
MAGIC_DP_FUSE_SMALL=0 MAGIC_DP_XCHG=0 timeout 600 python -m pytest tests/test_dp_gpu.py -m gpu -q > gpurun_out/dp_tests_variants.log 2>&1; echo "dp tests (small layers on their own, NCCL exchanges) rc=$?"; tail -3 gpurun_out/dp_tests_variants.log
fi
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 100 --warmup 10 --no-cpu-baseline ${SCALE_ARGS} > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; echo "bench n$N rc=$?"
python - <<PY
import json
d=json.load(open('gpurun_out/bench_n$N.json')); r=d['roofline']; e=d.get('e2e') or {}
print('N=$N value', round(d['value']), d['ms_per_step'], 'e2e', e.get('value'), e.get('ms_per_step'), 'dp_parity', r.get('dp_parity'))
PY
tail -3 gpurun_out/bench_n$N.err
