"""Digest of several `ncu --set full` captures (one kernel launch each) + refresh of profiles/ncu_traffic.json.

  python tools/ncu_kernels_digest.py OUT.txt label=report.ncu-rep [label=report.ncu-rep ...]

label is the key bench.py looks up in profiles/ncu_traffic.json (e.g. env_step_pair_f32_1048576); the traffic figure is
dram__bytes_read.sum + dram__bytes_write.sum of that launch."""
import csv, io, json, os, subprocess, sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
WANT = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "launch__registers_per_thread",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "smsp__inst_executed.sum",
        "smsp__sass_inst_executed_op_local_ld.sum", "smsp__sass_inst_executed_op_local_st.sum", "launch__grid_size",
        "launch__shared_mem_per_block_dynamic"]
UNIT = {"Mbyte": 1e6, "Gbyte": 1e9, "Kbyte": 1e3, "byte": 1.0}

out, pairs = sys.argv[1], [a.split("=", 1) for a in sys.argv[2:]]
traffic_path = os.path.join(ROOT, "profiles", "ncu_traffic.json")
traffic = json.load(open(traffic_path)) if os.path.exists(traffic_path) else {}
lines = []
for label, rep in pairs:
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units, vals = rows[0], rows[1], rows[-1]
    d = {h: (u, v) for h, u, v in zip(hdr, units, vals)}
    lines.append(f"== {label}  ({d['Kernel Name'][1][:96]})")
    for k in WANT:
        if k in d:
            lines.append(f"   {k} = {d[k][1]} {d[k][0]}")
    byt = sum(float(d[k][1]) * UNIT[d[k][0]] for k in ("dram__bytes_read.sum", "dram__bytes_write.sum"))
    lines.append(f"   dram bytes per launch = {byt / 1e6:.1f} MB")
    st = {k[len("smsp__pcsamp_warps_issue_stalled_"):]: float(v[1]) for k, v in d.items()
          if k.startswith("smsp__pcsamp_warps_issue_stalled_") and not k.endswith("_not_issued") and v[1] not in ("", "0")}
    tot = sum(st.values()) or 1.0
    lines.append("   stall shares: " + ", ".join(f"{k} {100 * v / tot:.0f}%" for k, v in sorted(st.items(), key=lambda kv: -kv[1])[:7]))
    traffic[label] = int(byt)
open(out, "w").write("\n".join(lines) + "\n")
traffic["_source"] = (f"ncu --set full --clock-control none, dram__bytes_read.sum + dram__bytes_write.sum per launch; captures of the final "
                      f"round-2 build ({os.path.relpath(out, ROOT)}); env_step_f32 / env_step_f64 (one-aircraft kernels): r01c captures")
json.dump(traffic, open(traffic_path, "w"), indent=1)
print("\n".join(lines))
