set -x
python __graft_entry__.py 2>&1 | tail -5
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
timeout 900 python bench.py --n-det 4096 --duration 600 --steps 3 --warmup 3 --cpu-dets 16 2>&1 | tail -5
timeout 1200 python bench.py --steps 3 --warmup 3 --no-e2e --no-cpu 2>&1 | tail -5
