# final build of the chain kernel (256 threads, up to 48 samples per lane): C5 chain latency, parity, bench line, launch list, ncu
TAG=r02s
mkdir -p gpurun_out
timeout 600 python scripts/chain_bench.py --config c5 > gpurun_out/chain_c5_$TAG.txt 2>&1; grep process gpurun_out/chain_c5_$TAG.txt
timeout 900 python -m pytest tests/test_gpu_configs.py tests/test_gpu_parity.py -m gpu -q -x -k "recurrence or golden or stages or c5 or pipelined or 3d" > gpurun_out/pytest_chain_$TAG.log 2>&1; tail -3 gpurun_out/pytest_chain_$TAG.log
timeout 900 python bench.py > gpurun_out/bench_c4_$TAG.json 2> gpurun_out/bench_c4_$TAG.err; tail -2 gpurun_out/bench_c4_$TAG.err; cut -c1-400 gpurun_out/bench_c4_$TAG.json
timeout 900 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_c4_ref_$TAG.json 2> gpurun_out/bench_c4_ref_$TAG.err; cut -c1-200 gpurun_out/bench_c4_ref_$TAG.json
ARGS="--steps 2 --warmup 3 --no-e2e --no-cpu --no-extras"
timeout 600 python bench.py $ARGS > gpurun_out/plain_$TAG.json 2> gpurun_out/plain_$TAG.err
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 2500 --csv --log-file gpurun_out/launches_$TAG.csv python bench.py $ARGS > gpurun_out/ncu_launch_$TAG.log 2>&1
tail -1 gpurun_out/ncu_launch_$TAG.log | cut -c1-200
timeout 900 ncu --set full --clock-control none --import-source on -k regex:ar_chain_v4_kernelILi8ELi40ELi5ELi5 -s 2 -c 1 -f -o gpurun_out/prof_ar_chain_$TAG python bench.py $ARGS --no-overlap > gpurun_out/ncu_chain_$TAG.log 2>&1
tail -1 gpurun_out/ncu_chain_$TAG.log
