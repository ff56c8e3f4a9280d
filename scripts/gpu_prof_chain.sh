TAG=${1:-x}
ARGS="--n-det 2048 --steps 1 --warmup 3 --no-e2e --no-cpu --no-overlap"
mkdir -p gpurun_out
python bench.py $ARGS > gpurun_out/plainc_$TAG.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:ar_chain_v4 -s 31 -c 1 -o gpurun_out/prof_chain_$TAG python bench.py $ARGS > gpurun_out/ncu_chain_$TAG.log 2>&1
tail -3 gpurun_out/ncu_chain_$TAG.log
