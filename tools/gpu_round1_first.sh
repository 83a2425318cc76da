set -x
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.limit --format=csv
nproc
python -c "import tensorflow" 2>&1 | tail -1
python -m pytest tests -m gpu -x -q 2>&1 | tail -15
python __graft_entry__.py smoke 2>&1 | tail -8
python bench.py --steps 100 --warmup 10 > gpurun_out/bench_r1_a.json 2> gpurun_out/bench_r1_a.err; echo "bench exit $?"; cat gpurun_out/bench_r1_a.json; tail -5 gpurun_out/bench_r1_a.err
python tests/perf_probe.py 8d 4096 bf16 > gpurun_out/probe_8d_bf16.txt 2>&1; cat gpurun_out/probe_8d_bf16.txt
python tests/perf_probe.py 8d 4096 fp32 > gpurun_out/probe_8d_fp32.txt 2>&1; tail -22 gpurun_out/probe_8d_fp32.txt
python tests/perf_probe.py 8c 100 bf16 > gpurun_out/probe_8c_bf16.txt 2>&1; tail -22 gpurun_out/probe_8c_bf16.txt
