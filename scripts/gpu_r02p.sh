TAG=r02p
mkdir -p gpurun_out
# the allocation-free entry point: alone in a fresh process, then after every other GPU test
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > gpurun_out/pytest_noalloc_alone_$TAG.log 2>&1; tail -15 gpurun_out/pytest_noalloc_alone_$TAG.log
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/pytest_$TAG.log 2>&1; tail -5 gpurun_out/pytest_$TAG.log
# C1-C3 at full size: the CUDA arm and the CPU port on the same configuration
for c in c1 c3 c2; do
  timeout 900 python bench.py --config $c --steps 10 --warmup 3 > gpurun_out/bench_${c}_$TAG.json 2> gpurun_out/bench_${c}_$TAG.err
  echo "$c: $(grep -o '"value": [0-9.e+]*' gpurun_out/bench_${c}_$TAG.json | head -1) $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/bench_${c}_$TAG.json | head -1)"
  timeout 900 python bench.py --impl reference --config $c --steps 2 --warmup 1 > gpurun_out/bench_ref_${c}_$TAG.json 2> gpurun_out/bench_ref_${c}_$TAG.err
  echo "ref $c: $(grep -o '"value": [0-9.e+]*' gpurun_out/bench_ref_${c}_$TAG.json | head -1)"
done
