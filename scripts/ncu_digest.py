"""Digest of an .ncu-rep: key raw metrics + instruction / stall-sample share per SASS region.
usage: python scripts/ncu_digest.py report.ncu-rep [n_output_samples] [bucket] [kernel index in the report]"""
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
n_out = float(sys.argv[2]) if len(sys.argv) > 2 else None
bucket = int(sys.argv[3]) if len(sys.argv) > 3 else 100
kid = int(sys.argv[4]) if len(sys.argv) > 4 else 0
sel = ["--launch-skip", str(kid), "--launch-count", "1"] if kid else []
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv", *sel], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, vals = rows[0], rows[1], rows[2]
want = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__occupancy_limit_registers",
        "launch__occupancy_limit_shared_mem", "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "lts__t_sectors.sum", "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_xu.sum", "sm__inst_executed_pipe_lsu.sum", "launch__shared_mem_per_block_dynamic", "launch__grid_size",
        "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct", "sm__inst_executed_pipe_fma.sum", "sm__inst_executed_pipe_alu.sum",
        "smsp__thread_inst_executed.sum"]
d = {}
for w in want:
    if w in hdr:
        i = hdr.index(w)
        d[w] = vals[i]
        print(f"{w:70s} {vals[i]:>22s} {units[i]}")
if n_out and "smsp__thread_inst_executed.sum" in d:
    print(f"thread instructions per output sample: {float(d['smsp__thread_inst_executed.sum'].replace(',', '')) / n_out:.1f}")
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", *sel], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
h = rows[1]
ia, ie, isamp = h.index("Source"), h.index("Instructions Executed"), h.index("# Samples")
data = [r for r in rows[2:] if len(r) > max(ie, isamp) and r[ie].isdigit()]
tot = sum(int(r[ie]) for r in data)
ts = sum(int(r[isamp]) for r in data) or 1
print(f"SASS instructions: {len(data)}, warp-instructions executed {tot}")
for s in range(0, len(data), bucket):
    e = sum(int(r[ie]) for r in data[s:s + bucket])
    sm = sum(int(r[isamp]) for r in data[s:s + bucket])
    if e / tot > 0.005:
        print(f"  sass[{s:5d}:{s + bucket:5d}] {e / tot * 100:6.2f}% of instructions  {sm / ts * 100:6.2f}% of stall samples")
with open(rep + ".sass.txt", "w") as f:
    for k, r in enumerate(data):
        f.write(f"{k:5d} {int(r[ie]) / tot * 100:6.3f}% {int(r[isamp]) / ts * 100:6.3f}%  {r[ia][:120]}\n")
