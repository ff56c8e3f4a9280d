TAG=r02f
mkdir -p gpurun_out
timeout 600 python scripts/diag_c2.py > gpurun_out/diag_c2_$TAG.txt 2>&1; cat gpurun_out/diag_c2_$TAG.txt | tail -8
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/pytest_$TAG.log 2>&1; tail -5 gpurun_out/pytest_$TAG.log
B="--no-e2e --no-cpu --steps 5 --warmup 3"
run() {  # name, extra args
  timeout 300 python bench.py $B $2 > gpurun_out/bench_$1_$TAG.json 2> gpurun_out/bench_$1_$TAG.err
  echo "$1: $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/bench_$1_$TAG.json) $(grep -o '"kernel_ms": [0-9.]*' gpurun_out/bench_$1_$TAG.json)"
}
run default ""
run default_seq "--no-overlap"
run default_seq_single "--no-overlap --chunk-mib 60000 --chunk-serial"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'loading_f32|synth_eval' -c 12 --csv --log-file gpurun_out/launches_ab_$TAG.csv python bench.py --no-e2e --no-cpu --steps 2 --warmup 3 --no-overlap --chunk-mib 60000 --chunk-serial > gpurun_out/ncu_ab_$TAG.log 2>&1
grep -o 'loading_f32_kernel.*\|synth_eval_f32_kernel.*' gpurun_out/launches_ab_$TAG.csv | awk -F'","' '{print $1, $NF}' | tail -2
timeout 900 ncu --set full --clock-control none --import-source on -k regex:loading_f32 -s 3 -c 1 -f -o gpurun_out/prof_loading_f32_$TAG python bench.py --no-e2e --no-cpu --steps 2 --warmup 3 --no-overlap --chunk-mib 60000 --chunk-serial > gpurun_out/ncu_eval_$TAG.log 2>&1
tail -1 gpurun_out/ncu_eval_$TAG.log
