# usage: bash scripts/gpu_profile_full.sh <tag> -- launch list + full capture of the synthesis kernel for the DEFAULT bench workload (C4)
TAG=${1:-r01}
ARGS="--steps 2 --warmup 3 --no-e2e --no-cpu"
mkdir -p gpurun_out
python bench.py $ARGS > gpurun_out/plain_$TAG.json 2>gpurun_out/plain_$TAG.err || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 800 --csv --log-file gpurun_out/launches_$TAG.csv python bench.py $ARGS > gpurun_out/ncu_launch_$TAG.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:synth_f32_kernel -s 3 -c 1 -o gpurun_out/prof_c4_$TAG python bench.py $ARGS > gpurun_out/ncu_full_$TAG.log 2>&1
tail -1 gpurun_out/plain_$TAG.json | cut -c1-200; ls -la gpurun_out | tail -5
