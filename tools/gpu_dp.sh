# usage (gpurun --gpus 2): bash tools/gpu_dp.sh -- 2-GPU parity tests (fused reduce-scatter path and the NCCL all-reduce baseline),
# then the bench line at N = 2 (carries dp_parity and the other BASELINE configs at this N)
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_dp_gpu.py -m gpu -q > gpurun_out/dp_tests.log 2>&1; echo "dp tests (fused) rc=$?"; tail -4 gpurun_out/dp_tests.log
MAGIC_DP_MODE=allreduce timeout 600 python -m pytest tests/test_dp_gpu.py -m gpu -q > gpurun_out/dp_tests_allreduce.log 2>&1; echo "dp tests (allreduce) rc=$?"; tail -3 gpurun_out/dp_tests_allreduce.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 100 --warmup 5 > gpurun_out/bench_n2.json 2> gpurun_out/bench_n2.err; echo "bench n2 rc=$?"; cut -c1-300 gpurun_out/bench_n2.json; tail -5 gpurun_out/bench_n2.err
