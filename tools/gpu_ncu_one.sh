# usage: bash tools/gpu_ncu_one.sh <tag> <demangled-name regex> [skip]   -- one full-set capture of one kernel of the bench step
TAG=$1; PAT=$2; SKIP=${3:-3}
CMD="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e --no-extra-legs"
mkdir -p gpurun_out
$CMD > gpurun_out/${TAG}_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k "regex:$PAT" -s $SKIP -c 1 -o gpurun_out/${TAG} $CMD > gpurun_out/${TAG}_ncu.log 2>&1
tail -3 gpurun_out/${TAG}_ncu.log
