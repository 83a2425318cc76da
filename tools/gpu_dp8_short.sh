# under gpurun --gpus 8: N=1 and N=8 bench lines of the default workload on the same box (fused DP, then the NCCL baseline)
mkdir -p gpurun_out
timeout 300 python bench.py --steps 100 --warmup 10 --no-cpu-baseline --no-e2e > gpurun_out/bench8_n1.json 2> gpurun_out/bench8_n1.err; cut -c1-200 gpurun_out/bench8_n1.json
run() { local name=$1; shift; local envs=$1; shift
  env $envs timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29520 bench.py --gpus 8 --steps 100 --warmup 10 --no-e2e "$@" > gpurun_out/$name.json 2> gpurun_out/$name.err
  tail -2 gpurun_out/$name.err | cut -c1-300; cut -c1-200 gpurun_out/$name.json; }
run bench8_n8_fused "MAGIC_X=1"
run bench8_n8_allreduce "MAGIC_DP_MODE=allreduce"
