# usage: bash scripts/gpu_ab.sh  -- parity + C4 bench with and without the vec4 path
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
for e in "" "MATM_SYNTH_NOVEC4=1"; do
  env $e python bench.py --steps 8 --warmup 3 --no-e2e --no-cpu 2>&1 | tail -1 | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('$e', d['value'], d['ms_per_step'], d['roofline']['kernel_ms'], d['roofline']['frac'])"
done
