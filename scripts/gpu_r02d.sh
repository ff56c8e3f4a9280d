# eval kernel v3 (4 consecutive samples per thread, packed over sample pairs, st.v4): parity, variants, ncu
TAG=r02d
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_$TAG.log 2>&1; tail -5 gpurun_out/pytest_$TAG.log
B="--no-e2e --no-cpu --steps 5 --warmup 3"
cp maria_b200/libmaria_atm.so /tmp/lib_default.so
run() {  # name, extra args
  timeout 300 python bench.py $B $2 > gpurun_out/bench_$1_$TAG.json 2> gpurun_out/bench_$1_$TAG.err
  echo "$1: $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/bench_$1_$TAG.json) $(grep -o '"kernel_ms": [0-9.]*' gpurun_out/bench_$1_$TAG.json)"
}
run default ""
run default_seq "--no-overlap"
run default_seq_single "--no-overlap --chunk-mib 60000 --chunk-serial"
for v in u2m3 u2m4 u4m2 u4m3kt96; do
  cp maria_b200/variants/libmaria_atm_$v.so maria_b200/libmaria_atm.so
  run ${v}_seq "--no-overlap"
done
cp /tmp/lib_default.so maria_b200/libmaria_atm.so
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'loading_f32|synth_eval' -c 12 --csv --log-file gpurun_out/launches_ab_$TAG.csv python bench.py --no-e2e --no-cpu --steps 2 --warmup 3 --no-overlap --chunk-mib 60000 --chunk-serial > gpurun_out/ncu_ab_$TAG.log 2>&1
grep -o 'loading_f32_kernel.*\|synth_eval_f32_kernel.*' gpurun_out/launches_ab_$TAG.csv | awk -F'","' '{print $1, $NF}' | tail -4
timeout 900 ncu --set full --clock-control none --import-source on -k regex:synth_eval_f32 -s 3 -c 1 -f -o gpurun_out/prof_synth_eval_f32_$TAG python bench.py --no-e2e --no-cpu --steps 2 --warmup 3 --no-overlap --chunk-mib 60000 --chunk-serial > gpurun_out/ncu_eval_$TAG.log 2>&1
tail -1 gpurun_out/ncu_eval_$TAG.log
