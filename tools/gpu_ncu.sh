# usage: bash tools/gpu_ncu.sh <tag>      (run under gpurun; writes gpurun_out/<tag>_*)
set -x
TAG=${1:-r01}
CMD="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e"
mkdir -p gpurun_out
$CMD > gpurun_out/${TAG}_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/${TAG}_launches.csv $CMD > gpurun_out/${TAG}_ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:gemm_tcgen05 -s 66 -c 22 -o gpurun_out/${TAG}_gemm $CMD > gpurun_out/${TAG}_ncu2.log 2>&1
ncu --set full --clock-control none --import-source on -k 'regex:generate_pairs|adam_dual|latent_' -s 12 -c 5 -o gpurun_out/${TAG}_misc $CMD > gpurun_out/${TAG}_ncu3.log 2>&1
tail -3 gpurun_out/${TAG}_ncu1.log gpurun_out/${TAG}_ncu2.log gpurun_out/${TAG}_ncu3.log
ls -la gpurun_out
