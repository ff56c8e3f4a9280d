# usage: bash scripts/gpu_final.sh <tag> -- sequential / depth sweep (device only) + launch list of the default bench workload (C4)
TAG=${1:-x}
B="--no-e2e --no-cpu"
mkdir -p gpurun_out
timeout 600 python bench.py $B --no-overlap > gpurun_out/seq_$TAG.json 2> gpurun_out/seq_$TAG.err
for d in 2 3 4; do timeout 600 python bench.py $B --depth $d > gpurun_out/depth${d}_$TAG.json 2> gpurun_out/depth${d}_$TAG.err; done
ARGS="--steps 2 --warmup 3 $B"
timeout 600 python bench.py $ARGS > gpurun_out/plain_$TAG.json 2>gpurun_out/plain_$TAG.err || exit 1
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 800 --csv --log-file gpurun_out/launches_$TAG.csv python bench.py $ARGS > gpurun_out/ncu_launch_$TAG.log 2>&1
for f in seq depth2 depth3 depth4; do cut -c1-220 gpurun_out/${f}_$TAG.json; done
