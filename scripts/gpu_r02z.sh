# dedicated far warp
TAG=r02z
mkdir -p gpurun_out
timeout 300 python scripts/chain_bench.py > gpurun_out/chain_c4_$TAG.txt 2>&1; grep process gpurun_out/chain_c4_$TAG.txt
for n in 1 2 3; do
  timeout 300 python -m pytest tests/test_gpu_configs.py -m gpu -q -x -k "pipelined or memory_system" > gpurun_out/pipe_${n}_$TAG.log 2>&1
  echo "run $n: $(tail -1 gpurun_out/pipe_${n}_$TAG.log) $(grep -o 'AssertionError: ([0-9], [0-9])' gpurun_out/pipe_${n}_$TAG.log | head -1)"
done
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/pytest_$TAG.log 2>&1; tail -5 gpurun_out/pytest_$TAG.log
timeout 600 python scripts/chain_bench.py --config c5 > gpurun_out/chain_c5_$TAG.txt 2>&1; grep -E "process [567]" gpurun_out/chain_c5_$TAG.txt
