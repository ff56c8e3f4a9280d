# round 2, first GPU call: parity of the two-kernel synthesis, then chunk / overlap sweep at full C4 size
TAG=r02a
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_$TAG.log 2>&1; tail -5 gpurun_out/pytest_$TAG.log
B="--no-e2e --no-cpu --steps 5 --warmup 3"
for c in 8 16 32 64; do
  timeout 300 python bench.py $B --chunk-mib $c > gpurun_out/bench_c${c}_$TAG.json 2> gpurun_out/bench_c${c}_$TAG.err
  echo "chunk $c overlap: $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/bench_c${c}_$TAG.json) $(grep -o '"kernel_ms": [0-9.]*' gpurun_out/bench_c${c}_$TAG.json)"
done
for c in 32 60000; do
  timeout 300 python bench.py $B --chunk-mib $c --chunk-serial > gpurun_out/bench_s${c}_$TAG.json 2> gpurun_out/bench_s${c}_$TAG.err
  echo "chunk $c serial: $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/bench_s${c}_$TAG.json) $(grep -o '"kernel_ms": [0-9.]*' gpurun_out/bench_s${c}_$TAG.json)"
done
timeout 300 python bench.py $B --no-overlap --chunk-mib 32 > gpurun_out/bench_seq_$TAG.json 2> gpurun_out/bench_seq_$TAG.err
echo "sequential: $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/bench_seq_$TAG.json) $(grep -o '"kernel_ms": [0-9.]*' gpurun_out/bench_seq_$TAG.json)"
# per-kernel times: single chunk, serial, under ncu (launch list)
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'loading_f32|synth_eval' -c 12 --csv --log-file gpurun_out/launches_ab_$TAG.csv python bench.py --no-e2e --no-cpu --steps 2 --warmup 3 --no-overlap --chunk-mib 60000 --chunk-serial > gpurun_out/ncu_ab_$TAG.log 2>&1
grep -o 'loading_f32_kernel.*\|synth_eval_f32_kernel.*' gpurun_out/launches_ab_$TAG.csv | awk -F'","' '{print $1, $NF}' | tail -6
