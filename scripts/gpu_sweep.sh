# usage: bash scripts/gpu_sweep.sh "<flags1>" "<flags2>" ...   -- rebuild with each flag set, run the 8k-detector bench
mkdir -p gpurun_out
i=0
for F in "$@"; do
  i=$((i+1))
  MATM_NVCC_FLAGS="$F" python -m maria_b200._build > gpurun_out/build_$i.log 2>&1 || { echo "build failed: $F"; tail -5 gpurun_out/build_$i.log; continue; }
  grep -A2 "synth_f32" maria_b200/csrc/ptxas.log | grep -E "Used" | head -1
  python bench.py --n-det 8192 --steps 3 --warmup 3 --no-e2e --no-cpu > gpurun_out/sweep_$i.log 2>&1
  echo "[$F] $(tail -1 gpurun_out/sweep_$i.log | grep -o '"kernel_ms[^,]*')"
done
