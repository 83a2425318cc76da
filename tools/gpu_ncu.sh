# usage: bash tools/gpu_ncu.sh -- ncu launch list of the bench command + one full-set capture each of the dgrad chain kernel and of
# the two packed first-layer GEMMs (reports under gpurun_out/*.ncu-rep, summaries by tools/ncu_summary.py)
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-extra-legs"
timeout 600 $CMD > gpurun_out/launch_plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --kernel-name-base demangled -c 800 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/launch_ncu.log 2>&1
tail -1 gpurun_out/launch_ncu.log | cut -c1-150
timeout 900 bash tools/gpu_ncu_one.sh chain_dgrad 'chain_tcgen05_kernel<.bool.1>' 3
timeout 900 bash tools/gpu_ncu_one.sh chain_fwd 'chain_tcgen05_kernel<.bool.0>' 3
timeout 900 bash tools/gpu_ncu_one.sh wgrad_enc0_bits 'gemm_tcgen05_kernel<.bool.1, .bool.1, .int.0, .int.256, .bool.1, .bool.0>' 3
timeout 900 bash tools/gpu_ncu_one.sh fwd_enc0_bits 'gemm_tcgen05_kernel<.bool.0, .bool.1, .int.0, .int.256, .bool.1, .bool.0>' 3
ls -la gpurun_out/*.ncu-rep
