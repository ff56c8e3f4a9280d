# 8 GPUs, final kernels of round 2: C5 as BASELINE.json states it (pipeline depth from the chain length), then C4 (weak + strong block)
TAG=r02f9
mkdir -p gpurun_out
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29521"
timeout 170 $T bench.py --gpus 8 --config c5 --steps 5 --warmup 3 --no-e2e > gpurun_out/bench_c5_n8_$TAG.json 2> gpurun_out/bench_c5_n8_$TAG.err
tail -2 gpurun_out/bench_c5_n8_$TAG.err | cut -c1-300; grep -o '"value": [0-9.e+]*' gpurun_out/bench_c5_n8_$TAG.json | head -1
timeout 200 $T bench.py --gpus 8 --steps 10 --warmup 3 --e2e-steps 1 > gpurun_out/bench_c4_n8_$TAG.json 2> gpurun_out/bench_c4_n8_$TAG.err
tail -2 gpurun_out/bench_c4_n8_$TAG.err | cut -c1-300; grep -o '"value": [0-9.e+]*' gpurun_out/bench_c4_n8_$TAG.json | head -1
