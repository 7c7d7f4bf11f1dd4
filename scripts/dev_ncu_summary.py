"""dev tool: per-kernel summary of an `ncu --set full` capture exported with
    ncu -i capture.ncu-rep --page raw --csv > raw.csv
    python scripts/dev_ncu_summary.py raw.csv out.json
averages per kernel name: duration, DRAM bytes per launch (read + write), issue-slot utilisation,
the largest warp-stall reasons (cycles per issued instruction)."""
import csv, json, sys
from collections import defaultdict

raw, out = sys.argv[1], sys.argv[2]
rows = list(csv.reader(open(raw)))
hdr = next(r for r in rows if "Kernel Name" in r)
i0 = rows.index(hdr)
units = rows[i0 + 1]
col = {h: i for i, h in enumerate(hdr)}


def find(sub):
    c = [h for h in hdr if sub in h]
    return c[0] if c else None


def val(r, name):
    if name is None:
        return None
    x = r[col[name]].replace(",", "")
    try:
        v = float(x)
    except ValueError:
        return None
    u = units[col[name]]
    scale = {"Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "byte": 1.0, "ns": 1e-3, "us": 1.0, "ms": 1e3, "usecond": 1.0, "nsecond": 1e-3, "msecond": 1e3}
    return v * scale.get(u, 1.0)


dur, rd, wr = find("gpu__time_duration.sum"), find("dram__bytes_read.sum"), find("dram__bytes_write.sum")
issue = find("sm__inst_issued.avg.pct_of_peak_sustained_active") or find("smsp__issue_active.avg.pct")
stalls = [h for h in hdr if "smsp__average_warp" in h and "issue_stalled" in h and h.endswith("_per_warp_active.pct")] or \
         [h for h in hdr if "smsp__average_warps_issue_stalled" in h and "per_issue_active" in h]
acc = defaultdict(lambda: defaultdict(list))
for r in rows[i0 + 2:]:
    if len(r) != len(hdr):
        continue
    k = r[col["Kernel Name"]].split("(")[0]
    a = acc[k]
    a["us"].append(val(r, dur))
    a["dram_read"].append(val(r, rd))
    a["dram_write"].append(val(r, wr))
    a["issue"].append(val(r, issue))
    for s in stalls:
        a["stall:" + s].append(val(r, s))
res = {"source": " ".join(sys.argv[3:]) or raw, "kernels": {}}
mean = lambda xs: (sum(x for x in xs if x is not None) / max(1, sum(1 for x in xs if x is not None))) if xs else None
for k, a in acc.items():
    st = sorted(((mean(v), n[6:]) for n, v in a.items() if n.startswith("stall:") and mean(v)), reverse=True)[:4]
    e = {"launches": len(a["us"]), "avg_us": mean(a["us"]), "dram_read_bytes": mean(a["dram_read"]), "dram_write_bytes": mean(a["dram_write"]),
         "dram_bytes_per_launch": (mean(a["dram_read"]) or 0) + (mean(a["dram_write"]) or 0), "issue_active_pct": mean(a["issue"]),
         "top_stalls": [{"metric": n, "value": v} for v, n in st]}
    res["kernels"][k] = e
    res[k] = {"dram_bytes_per_launch": e["dram_bytes_per_launch"]}
json.dump(res, open(out, "w"), indent=1)
for k, e in res["kernels"].items():
    print(f"{k:28s} n={e['launches']:2d} {e['avg_us'] or 0:8.1f} us  dram {e['dram_bytes_per_launch'] / 1e6:6.2f} MB  issue {e['issue_active_pct'] or 0:5.1f}%  " +
          "  ".join(f"{s['metric'].split('issue_stalled_')[-1].split('_per')[0]}={s['value']:.1f}" for s in e["top_stalls"]))
