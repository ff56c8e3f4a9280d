# far samples through delay lines in shared memory: is the pipelined test stable now, what does it cost, parity
TAG=r02v
mkdir -p gpurun_out
for n in 1 2 3 4 5 6 7 8; do
  timeout 300 python -m pytest tests/test_gpu_configs.py -m gpu -q -x -k "pipelined" > gpurun_out/pipe_${n}_$TAG.log 2>&1
  echo "run $n: $(tail -1 gpurun_out/pipe_${n}_$TAG.log) $(grep -o 'AssertionError: ([0-9], [0-9])' gpurun_out/pipe_${n}_$TAG.log | head -1)"
done
timeout 300 python scripts/chain_bench.py > gpurun_out/chain_c4_$TAG.txt 2>&1; grep process gpurun_out/chain_c4_$TAG.txt
timeout 600 python scripts/chain_bench.py --config c5 > gpurun_out/chain_c5_$TAG.txt 2>&1; grep -E "process [67]" gpurun_out/chain_c5_$TAG.txt
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/pytest_$TAG.log 2>&1; tail -5 gpurun_out/pytest_$TAG.log
nvcc -O3 -gencode arch=compute_100a,code=sm_100a scripts/microbench/dfma.cu -o /tmp/dfma 2>/dev/null && timeout 60 /tmp/dfma > gpurun_out/dfma_$TAG.txt 2>&1; cat gpurun_out/dfma_$TAG.txt | head -20
