# under gpurun --gpus 8: scaling bench lines (fused peer-memory DP vs NCCL all-reduce baseline), three configs
mkdir -p gpurun_out
run() { # name, extra env, args...
  local name=$1; shift
  local envs=$1; shift
  env $envs timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29520 bench.py --gpus 8 --steps 100 --warmup 10 --no-e2e "$@" > gpurun_out/$name.json 2> gpurun_out/$name.err
  tail -2 gpurun_out/$name.err | cut -c1-300; cut -c1-700 gpurun_out/$name.json
}
nvidia-smi -L | wc -l
run bench_n8_8d_fused "MAGIC_X=1"
run bench_n8_8d_allreduce "MAGIC_DP_MODE=allreduce"
run bench_n8_8e_fused "MAGIC_X=1" --exp 8e
run bench_n8_9a_fused "MAGIC_X=1" --exp 9a
run bench_n8_11a_fused "MAGIC_X=1" --exp 11a
