# new bench.py (sequential block, per-kernel times, d2h ceiling, honest CPU arm) + tests + launch list + full captures
TAG=r02i
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -q > gpurun_out/pytest_$TAG.log 2>&1; tail -8 gpurun_out/pytest_$TAG.log
timeout 1500 python bench.py > gpurun_out/bench_$TAG.json 2> gpurun_out/bench_$TAG.err; tail -3 gpurun_out/bench_$TAG.err; cut -c1-600 gpurun_out/bench_$TAG.json
timeout 900 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref_$TAG.json 2> gpurun_out/bench_ref_$TAG.err; cut -c1-300 gpurun_out/bench_ref_$TAG.json
ARGS="--steps 2 --warmup 3 --no-e2e --no-cpu --no-extras"
timeout 600 python bench.py $ARGS > gpurun_out/plain_$TAG.json 2> gpurun_out/plain_$TAG.err
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 2500 --csv --log-file gpurun_out/launches_$TAG.csv python bench.py $ARGS > gpurun_out/ncu_launch_$TAG.log 2>&1
tail -1 gpurun_out/ncu_launch_$TAG.log | cut -c1-200
S="--steps 2 --warmup 3 --no-e2e --no-cpu --no-extras --no-overlap --chunk-mib 60000 --chunk-serial"
for K in synth_eval_f32 loading_f32; do
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:$K -s 3 -c 1 -f -o gpurun_out/prof_${K}_$TAG python bench.py $S > gpurun_out/ncu_${K}_$TAG.log 2>&1
  tail -1 gpurun_out/ncu_${K}_$TAG.log
done
