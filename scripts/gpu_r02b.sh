# store-pattern microbenchmark + full ncu captures of the two synthesis kernels (single chunk, no chains underneath)
TAG=r02b
mkdir -p gpurun_out
./scripts/microbench/store_pattern 32768 180000 3072 > gpurun_out/store_pattern_$TAG.txt 2>&1; cat gpurun_out/store_pattern_$TAG.txt
ARGS="--no-e2e --no-cpu --steps 2 --warmup 3 --no-overlap --chunk-mib 60000 --chunk-serial"
for K in synth_eval_f32 loading_f32; do
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:$K -s 3 -c 1 -f -o gpurun_out/prof_${K}_$TAG python bench.py $ARGS > gpurun_out/ncu_${K}_$TAG.log 2>&1
  tail -2 gpurun_out/ncu_${K}_$TAG.log
done
ls -la gpurun_out/*.ncu-rep
