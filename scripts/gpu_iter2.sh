# usage: bash scripts/gpu_iter2.sh <tag>  -- parity tests + 100k-detector / 600 s bench (full detector density)
TAG=${1:-x}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_$TAG.log 2>&1; tail -6 gpurun_out/pytest_$TAG.log
timeout 600 python bench.py --n-det 100000 --duration 600 --steps 3 --warmup 3 --no-e2e --no-cpu --no-overlap > gpurun_out/bench100k_$TAG.log 2>&1; tail -1 gpurun_out/bench100k_$TAG.log | grep -o '"roofline.*step_ms_outside_kernel[^}]*' | cut -c1-420
