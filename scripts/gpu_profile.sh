# usage: bash scripts/gpu_profile.sh <tag>   -- launch list + full ncu capture of the synthesis kernel (1 GPU)
TAG=${1:-r01}
ARGS="--n-det 8192 --steps 2 --warmup 3 --no-e2e --no-cpu"
mkdir -p gpurun_out
python bench.py $ARGS > gpurun_out/plain_$TAG.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_$TAG.csv python bench.py $ARGS > gpurun_out/ncu_launch_$TAG.log 2>&1
python bench.py $ARGS > gpurun_out/plain2_$TAG.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:synth_kernel -s 3 -c 1 -o gpurun_out/prof_synth_$TAG python bench.py $ARGS > gpurun_out/ncu_full_$TAG.log 2>&1
tail -2 gpurun_out/plain_$TAG.log
ls -la gpurun_out
