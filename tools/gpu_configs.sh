# usage: bash tools/gpu_configs.sh -- one short bench line per BASELINE.json config on one GPU (per-GPU batch of the config)
mkdir -p gpurun_out
for spec in "8c 100" "8e 2048" "9a 8192" "11a 4096"; do
  set -- $spec
  timeout 200 python bench.py --exp $1 --batch $2 --steps 60 --warmup 5 --e2e-steps 6 --no-cpu-baseline \
    > gpurun_out/bench_$1_n1.json 2> gpurun_out/bench_$1_n1.err
  echo "$1 rc=$?"; cut -c1-330 gpurun_out/bench_$1_n1.json; echo
done
