TAG=r02n4
mkdir -p gpurun_out
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29521"
timeout 70 $T bench.py --gpus 4 --steps 10 --warmup 3 --e2e-steps 1 > gpurun_out/bench_c4_n4_$TAG.json 2> gpurun_out/bench_c4_n4_$TAG.err
tail -2 gpurun_out/bench_c4_n4_$TAG.err | cut -c1-300; grep -o '"value": [0-9.e+]*' gpurun_out/bench_c4_n4_$TAG.json | head -1
