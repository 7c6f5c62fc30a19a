"""Quick look at a .ncu-rep: headline metrics, stall reasons and instruction mix of each captured launch.
Usage: python scripts/ncu_quick.py file.ncu-rep [--src]"""
import csv, subprocess, sys, io
from collections import Counter
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
idx = {h: i for i, h in enumerate(hdr)}
KEYS = ["gpu__time_duration.sum", "sm__cycles_elapsed.avg.per_second", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "l1tex__m_xbar2l1tex_read_bytes.sum", "l1tex__m_xbar2l1tex_read_bytes.sum.per_second",
        "sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__warps_active.avg.per_cycle_active",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "launch__registers_per_thread"]
for r in rows[2:]:
    print("==", r[idx["Kernel Name"]][:70])
    for k in KEYS:
        if k in idx:
            print("   %-85s %s %s" % (k, r[idx[k]], units[idx[k]]))
    st = [(h.split("stalled_")[1], float(r[idx[h]].replace(",", "") or 0)) for h in hdr
          if "pcsamp_warps_issue_stalled" in h and "not_issued" not in h]
    tot = sum(v for _, v in st) or 1
    print("   stalls:", ", ".join("%s %.0f%%" % (k, 100 * v / tot) for k, v in sorted(st, key=lambda x: -x[1])[:9]))
if "--src" in sys.argv:
    src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(src)))
    blocks, cur = [], None
    for r in rows:
        if r and r[0] == "Kernel Name":
            cur = {"name": r[1], "rows": []}; blocks.append(cur); continue
        if r and r[0] == "Address":
            cur["hdr"] = r; continue
        if cur is not None and r:
            cur["rows"].append(r)
    b = blocks[0]; h = b["hdr"]; ia = h.index("Warp Stall Sampling (All Samples)"); ie = h.index("Instructions Executed")
    ci, cs = Counter(), Counter()
    for r in b["rows"]:
        t = r[1].split()
        op = (t[1] if t[0].startswith("@") else t[0]).split(".")[0]
        ci[op] += int(r[ie]); cs[op] += int(r[ia])
    print("instr (M):", [(k, round(v / 1e6, 1)) for k, v in ci.most_common(18)], "total", round(sum(ci.values()) / 1e6, 1))
    print("samples:", cs.most_common(10), "total", sum(cs.values()))
    for r in sorted(b["rows"], key=lambda r: -int(r[ia]))[:14]:
        print(r[ia].rjust(6), r[ie].rjust(9), r[0][-5:], r[1][:90])
