"""Summarise an .ncu-rep: per-kernel headline metrics, stall mix, hottest SASS lines.
usage: python scripts/ncu_summary.py report.ncu-rep [--sass N]"""
import csv, subprocess, sys, io

rep = sys.argv[1]
nsass = int(sys.argv[sys.argv.index("--sass") + 1]) if "--sass" in sys.argv else 12
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr = rows[0]
want = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "dram__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__grid_size",
        "launch__block_size", "launch__shared_mem_per_block_dynamic", "smsp__inst_executed.sum",
        "sm__inst_executed_pipe_tensor.sum", "l1tex__data_bank_conflicts_pipe_lsu.sum",
        "smsp__issue_active.avg.pct_of_peak_sustained_active"]
for r in rows[2:]:
    print("==", r[hdr.index("Kernel Name")][:90], "id", r[hdr.index("ID")])
    for w in want:
        if w in hdr:
            print(f"   {w:62s} {r[hdr.index(w)]} {rows[1][hdr.index(w)]}")
    st = [(float(r[i]), h.split("issue_stalled_")[1].split("_per")[0]) for i, h in enumerate(hdr)
          if "smsp__average_warps_issue_stalled" in h and "not_issued" not in h and r[i]]
    print("   stalls/issue:", ", ".join(f"{n}={v:.2f}" for v, n in sorted(st, reverse=True)[:7]))
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
cur = None
table = []
for row in csv.reader(io.StringIO(src)):
    if row and row[0] == "Kernel Name":
        if cur: table.append(cur)
        cur = {"name": row[1], "hdr": None, "rows": []}
    elif cur is not None and row and row[0] == "Address":
        cur["hdr"] = row
    elif cur is not None and cur["hdr"] and len(row) == len(cur["hdr"]):
        cur["rows"].append(row)
if cur: table.append(cur)
seen = set()
for k in table:
    if k["name"] in seen: continue
    seen.add(k["name"])
    h = k["hdr"]; si = h.index("# Samples"); so = h.index("Source")
    tot = sum(int(r[si]) for r in k["rows"] if r[si].isdigit())
    print("== SASS hot spots", k["name"][:80], "samples", tot)
    top = sorted(((int(r[si]), i) for i, r in enumerate(k["rows"]) if r[si].isdigit()), reverse=True)[:nsass]
    for n, i in top:
        print(f"   {n:6d} {100*n/max(tot,1):5.1f}%  [{i}] {k['rows'][i][so]}")
