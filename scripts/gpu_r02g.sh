# loading kernel variants: knots per thread x layers in flight x occupancy (per-kernel times from the library's events)
TAG=r02g
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_$TAG.log 2>&1; tail -3 gpurun_out/pytest_$TAG.log
B="--no-e2e --no-cpu --steps 5 --warmup 3 --kernel-times"
cp maria_b200/libmaria_atm.so /tmp/lib_default.so
run() {  # name, extra args
  timeout 300 python bench.py $B $2 > gpurun_out/bench_$1_$TAG.json 2> gpurun_out/bench_$1_$TAG.err
  echo "$1: $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/bench_$1_$TAG.json) $(grep -o '"kernel_ms": [0-9.]*' gpurun_out/bench_$1_$TAG.json) $(grep -o '"per_kernel": {[^}]*}' gpurun_out/bench_$1_$TAG.json)"
}
run default ""
run default_seq_single "--no-overlap --chunk-mib 60000 --chunk-serial"
for v in g4lb2 g1lb8 g2lb4m4 g4lb4m2 g4lb1 g2lb2m4; do
  cp maria_b200/variants/libmaria_atm_$v.so maria_b200/libmaria_atm.so
  run ${v}_seq_single "--no-overlap --chunk-mib 60000 --chunk-serial"
done
cp /tmp/lib_default.so maria_b200/libmaria_atm.so
timeout 900 ncu --set full --clock-control none --import-source on -k regex:loading_f32 -s 3 -c 1 -f -o gpurun_out/prof_loading_f32_$TAG python bench.py --no-e2e --no-cpu --steps 2 --warmup 3 --no-overlap --chunk-mib 60000 --chunk-serial > gpurun_out/ncu_eval_$TAG.log 2>&1
tail -1 gpurun_out/ncu_eval_$TAG.log
