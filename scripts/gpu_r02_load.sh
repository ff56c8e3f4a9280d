TAG=r02load
mkdir -p gpurun_out
timeout 300 python scripts/chain_bench.py --reps 2 > gpurun_out/chain_idle_$TAG.txt 2>&1; grep process gpurun_out/chain_idle_$TAG.txt
timeout 600 python scripts/chain_bench.py --reps 2 --load > gpurun_out/chain_load_$TAG.txt 2>&1; grep process gpurun_out/chain_load_$TAG.txt
for n in 1 2; do
  timeout 300 python -m pytest tests/test_gpu_configs.py -m gpu -q -x -k "pipelined or memory_system" > gpurun_out/pipe_${n}_$TAG.log 2>&1
  echo "run $n: $(tail -1 gpurun_out/pipe_${n}_$TAG.log)"
done
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/pytest_$TAG.log 2>&1; tail -4 gpurun_out/pytest_$TAG.log
