# usage: bash scripts/gpu_ar_v4.sh <tag> -- parity tests with the signal-driven chain (v4), then chain timing (sequential bench)
TAG=${1:-x}
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_$TAG.log 2>&1; tail -5 gpurun_out/pytest_$TAG.log
for CFG in "1 0" "0 0"; do
  set -- $CFG
  MATM_AR_V3=$1 MATM_AR_MINL=$2 timeout 300 python bench.py --n-det 2048 --steps 3 --warmup 3 --no-e2e --no-cpu --no-overlap > gpurun_out/ar_${1}_${2}_$TAG.log 2>&1
  echo "MATM_AR_V3=$1 MINL=$2 $(tail -1 gpurun_out/ar_${1}_${2}_$TAG.log | grep -o '"step_ms_outside_kernel[^}]*')"
done
