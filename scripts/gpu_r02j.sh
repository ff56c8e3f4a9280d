TAG=r02j
mkdir -p gpurun_out
B="--no-e2e --no-cpu --no-extras --steps 6 --warmup 3"
for c in 16 32 64 128 256 1024 60000; do
  timeout 300 python bench.py $B --chunk-mib $c > gpurun_out/bench_c${c}_$TAG.json 2> gpurun_out/bench_c${c}_$TAG.err
  echo "chunk $c: $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/bench_c${c}_$TAG.json | head -1) $(grep -o '"kernel_ms": [0-9.]*' gpurun_out/bench_c${c}_$TAG.json)"
done
for d in 2 4; do
  timeout 300 python bench.py $B --depth $d > gpurun_out/bench_d${d}_$TAG.json 2> gpurun_out/bench_d${d}_$TAG.err
  echo "depth $d: $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/bench_d${d}_$TAG.json | head -1) $(grep -o '"kernel_ms": [0-9.]*' gpurun_out/bench_d${d}_$TAG.json)"
done
