# usage: bash scripts/gpu_lds_probe.sh <tag> -- source-level ncu counters of the synthesis kernel on a small workload
TAG=${1:-x}
ARGS="--n-det 8192 --duration 600 --steps 1 --warmup 1 --no-e2e --no-cpu --no-overlap"
mkdir -p gpurun_out
ncu --section SourceCounters --section WarpStateStats --metrics gpu__time_duration.sum --clock-control none --import-source on -k regex:synth_f32_kernel -s 1 -c 1 -o gpurun_out/probe_$TAG python bench.py $ARGS > gpurun_out/probe_$TAG.log 2>&1
tail -2 gpurun_out/probe_$TAG.log
