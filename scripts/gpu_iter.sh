# usage: bash scripts/gpu_iter.sh <tag>  -- parity tests + mid-size bench (+ launch list) of the current build
TAG=${1:-x}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_$TAG.log 2>&1; tail -15 gpurun_out/pytest_$TAG.log
timeout 600 python bench.py --n-det 8192 --steps 3 --warmup 3 --no-e2e --no-cpu > gpurun_out/bench8k_$TAG.log 2>&1; tail -1 gpurun_out/bench8k_$TAG.log | cut -c1-300; tail -1 gpurun_out/bench8k_$TAG.log | grep -o '"roofline.*kernel_ms[^,]*,[^,]*'
