"""Aggregate an ncu launch list (--metrics gpu__time_duration.sum --csv) by kernel: python tools/launch_summary.py CSV [top]"""
import collections, csv, json, sys


def summarize(path, top=25):
    rows = list(csv.reader(open(path)))
    h = [i for i, r in enumerate(rows) if "Kernel Name" in r][0]
    hdr = rows[h]
    ki, vi, ui = (hdr.index(k) for k in ("Kernel Name", "Metric Value", "Metric Unit"))
    agg = collections.defaultdict(lambda: [0, 0.0])
    total = 0.0
    for r in rows[h + 1:]:
        if len(r) < len(hdr):
            continue
        v = float(r[vi].replace(",", ""))
        v = v / 1e6 if r[ui] in ("ns", "nsecond") else v / 1e3 if r[ui].startswith("us") else v
        short = r[ki].split("(")[0].replace("void ", "")[:90]
        agg[short][0] += 1
        agg[short][1] += v
        total += v
    items = sorted(agg.items(), key=lambda kv: -kv[1][1])
    return {"launches": sum(v[0] for v in agg.values()), "serialised_ms": round(total, 3),
            "ours_share": round(sum(v[1] for k, v in agg.items() if k.startswith("cgs::")) / total, 4),
            "kernels": [{"name": k, "n": v[0], "ms": round(v[1], 3), "share": round(v[1] / total, 4),
                         "us_per_launch": round(v[1] / v[0] * 1e3, 2)} for k, v in items[:top]]}


if __name__ == "__main__":
    s = summarize(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 25)
    print(json.dumps({k: v for k, v in s.items() if k != "kernels"}))
    for k in s["kernels"]:
        print(f'{k["ms"]:8.3f} ms {k["share"]*100:5.1f}%  n={k["n"]:4d}  {k["us_per_launch"]:8.2f} us  {k["name"]}')
