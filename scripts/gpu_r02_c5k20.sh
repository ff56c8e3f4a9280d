# C5 at 8 GPUs with 20 timed steps: the drain of the pipeline at the end of the timed region (the chains of the last realisations
# started, ~85 ms) is amortised over 20 steps instead of 5
TAG=r02c5k
mkdir -p gpurun_out
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29521"
timeout 150 $T bench.py --gpus 8 --config c5 --steps 20 --warmup 3 --no-e2e --no-extras > gpurun_out/bench_c5_n8_k20_$TAG.json 2> gpurun_out/bench_c5_n8_k20_$TAG.err
grep -o '"ms_per_step": [0-9.]*' gpurun_out/bench_c5_n8_k20_$TAG.json | head -1; grep -o '"value": [0-9.e+]*' gpurun_out/bench_c5_n8_k20_$TAG.json | head -1
