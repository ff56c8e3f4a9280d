# final N = 1 evidence of round 2: smoke, bench line, reference arm, launch list, full ncu capture of the chain kernels
TAG=r02f2
mkdir -p gpurun_out
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_$TAG.log 2>&1; tail -3 gpurun_out/smoke_$TAG.log
timeout 900 python bench.py > gpurun_out/bench_c4_$TAG.json 2> gpurun_out/bench_c4_$TAG.err; tail -2 gpurun_out/bench_c4_$TAG.err; cut -c1-300 gpurun_out/bench_c4_$TAG.json
timeout 900 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_c4_ref_$TAG.json 2> gpurun_out/bench_c4_ref_$TAG.err; cut -c1-200 gpurun_out/bench_c4_ref_$TAG.json
ARGS="--steps 2 --warmup 3 --no-e2e --no-cpu --no-extras"
timeout 600 python bench.py $ARGS > gpurun_out/plain_$TAG.json 2> gpurun_out/plain_$TAG.err
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 2500 --csv --log-file gpurun_out/launches_$TAG.csv python bench.py $ARGS > gpurun_out/ncu_launch_$TAG.log 2>&1
tail -1 gpurun_out/ncu_launch_$TAG.log | cut -c1-200



