#!/usr/bin/env python
"""Summarise an .ncu-rep (read here, no GPU): headline metrics, opcode mix, hottest source lines.

    python tools/ncu_summary.py gpurun_out/prof.ncu-rep [voxels_per_launch]
"""
import collections, csv, io, re, subprocess, sys

rep = sys.argv[1]
nvox = float(sys.argv[2]) if len(sys.argv) > 2 else None


def ncu(*args):
    """ncu's CSV for the FIRST captured launch only (a report with several launches repeats the tables)."""
    out = subprocess.run(["ncu", "-i", rep, *args], capture_output=True, text=True).stdout
    if "--page" in args and args[args.index("--page") + 1] == "source":
        parts = out.split('"Kernel Name"')
        if len(parts) > 2:
            out = '"Kernel Name"' + parts[1]
    return out


rows = list(csv.reader(io.StringIO(ncu("--page", "raw", "--csv"))))
h, u, v = rows[0], rows[1], rows[2]
want = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "launch__registers_per_thread",
        "launch__grid_size", "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "smsp__inst_executed.sum", "smsp__thread_inst_executed_per_inst_executed.ratio",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct",
        "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__lsu_writeback_active.avg.pct_of_peak_sustained_elapsed",
        "SM_A.TriageCompute.l1tex__data_pipe_lsu_wavefronts.avg", "SM_A.TriageCompute.l1tex__data_pipe_lsu_wavefronts_mem_shared.avg",
        "SM_A.TriageCompute.l1tex__data_pipe_lsu_wavefronts_mem_lgds.avg", "sm__cycles_elapsed.avg",
        "l1tex__t_output_wavefronts_pipe_lsu_mem_local_op_ld.sum", "l1tex__t_output_wavefronts_pipe_lsu_mem_local_op_st.sum",
        "l1tex__t_output_wavefronts_pipe_lsu_mem_global_op_ld.sum", "l1tex__t_output_wavefronts_pipe_lsu_mem_global_op_st.sum"]
print("kernel:", v[h.index("Kernel Name")] if "Kernel Name" in h else "")
for k in want:
    if k in h:
        i = h.index(k)
        print(f"{k:70s} {v[i]:>16s} {u[i]}")

rows = list(csv.reader(io.StringIO(ncu("--page", "source", "--csv"))))
hh = rows[1]
iS, iE, iSm = hh.index("Source"), hh.index("Instructions Executed"), hh.index("# Samples")
data = [(r[iS].strip(), int(r[iE]), int(r[iSm])) for r in rows[2:] if len(r) > iSm]
tot = sum(d[1] for d in data); ts = sum(d[2] for d in data)
print(f"\nwarp instructions {tot}  samples {ts}" + (f"  thread-instr/voxel {tot * 32 / nvox:.0f}" if nvox else ""))
c, cs = collections.Counter(), collections.Counter()
for s, e, sm in data:
    m = re.match(r"(@!?U?P\d+\s+)?([A-Z0-9_]+)", s)
    op = m.group(2) if m else s
    c[op] += e; cs[op] += sm
for op, e in c.most_common(28):
    per = f"{e * 32 / nvox:7.1f}/vox" if nvox else ""
    print(f"  {op:10s} {e / tot * 100:5.1f}% inst {per}  {cs[op] / ts * 100:5.1f}% samples")

rows = list(csv.reader(io.StringIO(ncu("--page", "source", "--print-source", "cuda,sass", "--csv"))))
cur, agg = None, []
for r in rows:
    if r and r[0] == "File Path":
        cur = r[1].split("/")[-1]; continue
    if len(r) > 8 and r[0].strip().isdigit() and r[2] == "-":
        try:
            agg.append((int(r[7]), int(r[4]), cur, int(r[0]), r[1].strip()[:90]))
        except ValueError:
            pass
agg.sort(reverse=True)
print("\nhottest source lines (inlined callers repeat their callees' instructions)")
for a in agg[:45]:
    per = f"{a[0] * 32 / nvox:6.1f}/vox" if nvox else ""
    print(f"  {a[0] / tot * 100:5.1f}% {per} {a[1]:6d} smp  {a[2]}:{a[3]}  {a[4]}")

# warp-state sampling: why warps that could not issue were stalled (all instructions summed)
rows = list(csv.reader(io.StringIO(ncu("--page", "source", "--csv"))))
hh = rows[1]
cols = [i for i, n in enumerate(hh) if n.startswith("stall_") and "Not Issued" not in n]
tot_st = collections.Counter()
for r in rows[2:]:
    for i in cols:
        if i < len(r) and r[i].isdigit():
            tot_st[hh[i]] += int(r[i])
ts2 = sum(tot_st.values()) or 1
print("\nstall reasons (all samples): " + ", ".join(f"{k[6:]} {v / ts2 * 100:.1f}%" for k, v in tot_st.most_common(10)))
