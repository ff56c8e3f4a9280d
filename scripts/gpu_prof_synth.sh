# usage: bash scripts/gpu_prof_synth.sh <tag> <kernel regex>  -- parity, bench, then one full ncu capture of the synthesis kernel
TAG=${1:-x}
K=${2:-synth_f32_kernel}
ARGS="${PROF_ARGS:---n-det 8192} --steps 2 --warmup 3 --no-e2e --no-cpu"
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_$TAG.log 2>&1; tail -4 gpurun_out/pytest_$TAG.log
python bench.py $ARGS > gpurun_out/bench8k_$TAG.log 2>&1 && tail -1 gpurun_out/bench8k_$TAG.log | grep -o '"roofline.*kernel_ms[^,]*,[^,]*'
ncu --set full --clock-control none --import-source on -k regex:$K -s 3 -c 1 -o gpurun_out/prof_$TAG python bench.py $ARGS > gpurun_out/ncu_$TAG.log 2>&1
tail -2 gpurun_out/ncu_$TAG.log
