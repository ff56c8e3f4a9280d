# full ncu capture of the final chain kernels (one realisation, sequential mode) + GPU test-suite on the committed state
TAG=r02f3
mkdir -p gpurun_out
ARGS="--steps 2 --warmup 3 --no-e2e --no-cpu --no-extras"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:ar_chain_v4_kernel -c 8 -f -o gpurun_out/prof_ar_chain_$TAG python bench.py $ARGS --no-overlap > gpurun_out/ncu_chain_$TAG.log 2>&1
tail -1 gpurun_out/ncu_chain_$TAG.log | cut -c1-200
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_$TAG.log 2>&1; tail -3 gpurun_out/pytest_$TAG.log
