# under gpurun --gpus 2: data-parallel parity test + 1- and 2-GPU bench lines
mkdir -p gpurun_out
nvidia-smi -L
timeout 900 python -m pytest tests/test_dp_gpu.py -m gpu -x -q 2>&1 | tail -15
MAGIC_DP_MODE=allreduce timeout 900 python -m pytest tests/test_dp_gpu.py -m gpu -x -q 2>&1 | tail -5
timeout 600 python bench.py --steps 100 --warmup 10 --no-cpu-baseline --no-e2e > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; tail -3 gpurun_out/bench_n1.err; cat gpurun_out/bench_n1.json
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 100 --warmup 10 --no-e2e > gpurun_out/bench_n2.json 2> gpurun_out/bench_n2.err; tail -5 gpurun_out/bench_n2.err; cat gpurun_out/bench_n2.json
MAGIC_DP_MODE=allreduce timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 100 --warmup 10 --no-e2e > gpurun_out/bench_n2_allreduce.json 2> gpurun_out/bench_n2_ar.err; tail -5 gpurun_out/bench_n2_ar.err; cat gpurun_out/bench_n2_allreduce.json
