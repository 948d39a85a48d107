"""Digest of one ncu report: key metrics of the raw page and the stall-reason shares of the source page.
usage: python tools/ncu_digest.py gpurun_out/x.ncu-rep [top_n]"""
import csv, io, subprocess, sys

rep = sys.argv[1]
top_n = int(sys.argv[2]) if len(sys.argv) > 2 else 12
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, vals = rows[0], rows[-1]
want = ["gpu__time_duration.sum", "launch__registers_per_thread", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__inst_executed.sum", "smsp__inst_executed.sum",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
        "sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_active", "sm__pipe_fmalite_cycles_active.avg.pct_of_peak_sustained_active",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "smsp__warps_eligible.avg.per_cycle_active",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__average_warp_latency_per_inst_issued.ratio",
        "l1tex__data_bank_conflicts_pipe_lsu.sum", "sm__cycles_elapsed.max"]
for h, v in zip(hdr, vals):
    if h in want:
        print(f"{h},{rows[1][hdr.index(h)]},{v}")
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
hdr, data = rows[1], rows[2:]
ix = {h: i for i, h in enumerate(hdr)}
def f(r, k):
    try:
        return float(r[ix[k]])
    except Exception:
        return 0.0
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
n = sum(f(r, "# Samples") for r in data)
print("samples", int(n), "warp_instructions", int(sum(f(r, "Instructions Executed") for r in data)))
for s_, v in sorted(((s_, sum(f(r, s_) for r in data)) for s_ in stalls), key=lambda x: -x[1]):
    if v:
        print(f"{s_},{100 * v / n:.1f}%")
for r in sorted(data, key=lambda r: -f(r, "# Samples"))[:top_n]:
    print(r[0][-5:], r[1][:64].strip(), int(f(r, "# Samples")), {s_[6:]: int(f(r, s_)) for s_ in stalls if f(r, s_) > 0.2 * f(r, "# Samples")})
