mkdir -p gpurun_out
for L in 0 8 16 32; do
  MATM_AR_MINL=$L python bench.py --n-det 2048 --steps 3 --warmup 3 --no-e2e --no-cpu > gpurun_out/ar_$L.log 2>&1
  echo "MINL=$L $(tail -1 gpurun_out/ar_$L.log | grep -o '"step_ms_outside_kernel[^}]*')"
done
