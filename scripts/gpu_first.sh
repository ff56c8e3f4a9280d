set -x
mkdir -p gpurun_out
python __graft_entry__.py > gpurun_out/smoke.log 2>&1; tail -5 gpurun_out/smoke.log
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; tail -8 gpurun_out/pytest_gpu.log
timeout 900 python bench.py --n-det 4096 --duration 600 --steps 3 --warmup 3 --cpu-dets 16 > gpurun_out/bench_small.log 2>&1; tail -3 gpurun_out/bench_small.log | cut -c1-1500
timeout 1200 python bench.py --steps 3 --warmup 3 > gpurun_out/bench_default.log 2>&1; tail -3 gpurun_out/bench_default.log | cut -c1-1500
