# chain kernel with 256 threads / 16-byte loads / unrolled by four as the production path: latency, parity, headline
TAG=r02r
mkdir -p gpurun_out
timeout 300 python scripts/chain_bench.py > gpurun_out/chain_c4_$TAG.txt 2>&1; grep process gpurun_out/chain_c4_$TAG.txt
timeout 600 python scripts/chain_bench.py --config c5 > gpurun_out/chain_c5_$TAG.txt 2>&1; grep process gpurun_out/chain_c5_$TAG.txt
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/pytest_$TAG.log 2>&1; tail -5 gpurun_out/pytest_$TAG.log
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_c4_$TAG.json 2> gpurun_out/bench_c4_$TAG.err
python - <<'PY'
import json
j=json.loads([l for l in open("gpurun_out/bench_c4_r02r.json") if l.startswith("{")][0])
print("c4", j["value"], j["ms_per_step"], j["roofline"]["frac"], {k:v for k,v in j["sequential"].items() if isinstance(v,float)})
PY
