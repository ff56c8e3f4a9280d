# what bounds the pipelined C5 step at 8 GPUs: pipeline depth or the exchange?
TAG=r02c5d
mkdir -p gpurun_out
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29521"
timeout 120 $T bench.py --gpus 8 --config c5 --steps 5 --warmup 3 --no-e2e --no-extras --depth 16 > gpurun_out/bench_c5_n8_depth16_$TAG.json 2> gpurun_out/bench_c5_n8_depth16_$TAG.err
grep -o '"ms_per_step": [0-9.]*' gpurun_out/bench_c5_n8_depth16_$TAG.json | head -1
timeout 120 $T bench.py --gpus 8 --config c5 --steps 5 --warmup 3 --no-e2e --no-extras --nccl-max-ctas 16 > gpurun_out/bench_c5_n8_ctas16_$TAG.json 2> gpurun_out/bench_c5_n8_ctas16_$TAG.err
grep -o '"ms_per_step": [0-9.]*' gpurun_out/bench_c5_n8_ctas16_$TAG.json | head -1
