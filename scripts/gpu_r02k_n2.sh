TAG=r02n
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_multirank.py -q > gpurun_out/pytest_multirank_n2_$TAG.log 2>&1; tail -3 gpurun_out/pytest_multirank_n2_$TAG.log
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --steps 6 --warmup 3 > gpurun_out/bench_n2_$TAG.json 2> gpurun_out/bench_n2_$TAG.err
tail -5 gpurun_out/bench_n2_$TAG.err; cut -c1-400 gpurun_out/bench_n2_$TAG.json
