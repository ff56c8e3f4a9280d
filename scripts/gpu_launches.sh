TAG=${1:-x}
ARGS="--n-det 8192 --steps 2 --warmup 3 --no-e2e --no-cpu"
mkdir -p gpurun_out
python bench.py $ARGS > gpurun_out/plain_$TAG.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_$TAG.csv python bench.py $ARGS > gpurun_out/ncu_launch_$TAG.log 2>&1
tail -1 gpurun_out/plain_$TAG.log | cut -c1-400
