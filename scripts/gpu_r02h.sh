# new parity tests (2 processes, C5 fp64, 3d) + second sweep of the loading kernel
TAG=r02h
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -q > gpurun_out/pytest_$TAG.log 2>&1; tail -15 gpurun_out/pytest_$TAG.log
B="--no-e2e --no-cpu --steps 5 --warmup 3 --kernel-times"
cp maria_b200/libmaria_atm.so /tmp/lib_default.so
run() {  # name, extra args
  timeout 300 python bench.py $B $2 > gpurun_out/bench_$1_$TAG.json 2> gpurun_out/bench_$1_$TAG.err
  echo "$1: $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/bench_$1_$TAG.json) $(grep -o '"kernel_ms": [0-9.]*' gpurun_out/bench_$1_$TAG.json) $(grep -o '"per_kernel": {[^}]*}' gpurun_out/bench_$1_$TAG.json)"
}
run default ""
run default_seq_single "--no-overlap --chunk-mib 60000 --chunk-serial"
for v in g4lb1m4 g2lb1m4 g2lb2m5 g2lb2m4ka64 g2lb2m4ka256; do
  cp maria_b200/variants/libmaria_atm_$v.so maria_b200/libmaria_atm.so
  run ${v}_seq_single "--no-overlap --chunk-mib 60000 --chunk-serial"
done
cp /tmp/lib_default.so maria_b200/libmaria_atm.so
