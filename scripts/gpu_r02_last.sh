# last check of the committed state: GPU test-suite, smoke, a short bench line
TAG=r02last
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_$TAG.log 2>&1; tail -3 gpurun_out/pytest_$TAG.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_$TAG.log 2>&1; tail -2 gpurun_out/smoke_$TAG.log
timeout 300 python bench.py --steps 10 --warmup 3 --no-e2e --no-cpu --no-extras > gpurun_out/bench_c4_short_$TAG.json 2> gpurun_out/bench_c4_short_$TAG.err; cut -c1-260 gpurun_out/bench_c4_short_$TAG.json
