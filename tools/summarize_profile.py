"""Summarise the ncu outputs of tools/profile_step.sh into profiles/ (run here, where ncu can read the reports).
usage: python tools/summarize_profile.py TAG      (reads gpurun_out/launches.csv and gpurun_out/prof_top.ncu-rep)"""
import collections, csv, io, json, os, subprocess, sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
tag = sys.argv[1]


# the first launches of the timed step (level-1 SpatialTransformer of SDXL at 1024^2, B2 = 16): what each capture is
SHAPES = [
    ("gemm_kernel<1, 0, 0, 0>", "GEGLU GEMM (one CTA per tile) M=65536 K=640 N=2560 (2N=5120)", 425984000, 429496729600.0),
    ("gemm_kernel<1, 1, 0, 0>", "GEGLU GEMM (CTA pairs) M=16384 K=1280 N=5120 (2N=10240)", 236060672, 429496729600.0),
    ("gemm_kernel<0, 0, 0, 0>", "linear + bias + residual (one CTA per tile) to_out M=65536 K=640 N=640", 252477440, 53687091200.0),
    ("gemm_kernel<0, 1, 0, 0>", "linear + bias + residual (CTA pairs): first captured = ff.net.2 level 1 M=65536 K=2560 N=640", 506593280, 214748364800.0),
    ("gemm_kernel<0, 1, 1, 1>", "3x3 convolution, 320-column tiles: first captured = 16 x 128^2, 320 -> 320", 337387520, 483183820800.0),
    ("gemm_kernel<0, 1, 1, 0>", "3x3 convolution, 256-column tiles: first captured = 16 x 32^2, 640 -> 1280", 77660160, 241591910400.0),
    ("attn_fwd_kernel<1, 3, 4", "self-attention B2=16 H=10 N=4096 d=64", 335544320, 687194767360.0),
    ("attn_fwd_kernel<1, 1, 1, 96", "cross-attention B2=16 H=10 Nq=4096 Nk=77 d=64 (96-key tile)", 170926080, 12918456320.0),
    ("layernorm_stream_kernel<3, 16>", "LayerNorm 65536 x 640", 167772160, 0.0),
]


def launches():
    rows = list(csv.reader(open(os.path.join(ROOT, "gpurun_out", "launches.csv"))))
    h = [i for i, r in enumerate(rows) if "Kernel Name" in r][0]
    hdr = rows[h]
    ki, gi, vi, ui = (hdr.index(k) for k in ("Kernel Name", "Grid Size", "Metric Value", "Metric Unit"))
    agg = collections.defaultdict(lambda: [0, 0.0])
    total, n = 0.0, 0
    for r in rows[h + 1:]:
        if len(r) < len(hdr):
            continue
        v = float(r[vi].replace(",", ""))
        v = v / 1e6 if r[ui] in ("ns", "nsecond") else v / 1e3 if r[ui].startswith("us") else v
        name = r[ki]
        short = name.split("(")[0].replace("void ", "")[:80]
        agg[short][0] += 1
        agg[short][1] += v
        total += v
        n += 1
    top = sorted(agg.items(), key=lambda kv: -kv[1][1])
    ours = sum(v[1] for k, v in agg.items() if k.startswith("cgs::"))
    out = {"source": "ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off "
                     "python bench.py --steps 1 --warmup 3 --profile (one timed denoise step, SDXL 1024^2, batch 8)",
           "note": "per-launch times are cold-cache and serialised; shares, not absolutes, are comparable with the step",
           "launches": n, "serialised_ms": round(total, 3), "share_of_our_kernels": round(ours / total, 4),
           "kernels": [{"name": k, "launches": v[0], "ms": round(v[1], 3), "share": round(v[1] / total, 4)} for k, v in top[:40]]}
    json.dump(out, open(os.path.join(ROOT, "profiles", f"{tag}_launches_summary.json"), "w"), indent=1)
    with open(os.path.join(ROOT, "profiles", f"{tag}_launches.csv"), "w") as f:
        w = csv.writer(f)
        w.writerow(["id", "kernel", "grid", "block", "ms"])
        bi = hdr.index("Block Size")
        for r in rows[h + 1:]:
            if len(r) < len(hdr):
                continue
            v = float(r[vi].replace(",", ""))
            v = v / 1e6 if r[ui] in ("ns", "nsecond") else v / 1e3 if r[ui].startswith("us") else v
            w.writerow([r[0], r[ki][:100], r[gi], r[bi], round(v, 5)])
    print("launches:", n, "serialised ms:", round(total, 2), "ours:", round(ours / total, 3))


def full():
    rep = os.path.join(ROOT, "gpurun_out", "prof_top.ncu-rep")
    raw = os.path.join(ROOT, "gpurun_out", "prof_top_raw.csv")        # exported on the GPU box by tools/profile_step.sh
    if os.path.exists(raw):
        txt = open(raw).read()
    else:
        txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(txt)))
    hdr = rows[0]
    want = {"gpu__time_duration.sum": "ms_under_ncu", "dram__bytes_read.sum": "dram_read_bytes",
            "dram__bytes_write.sum": "dram_write_bytes",
            "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active": "tensor_pipe_active_pct",
            "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active": "xu_pipe_pct",
            "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active": "fma_pipe_pct",
            "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active": "alu_pipe_pct",
            "smsp__issue_active.avg.pct_of_peak_sustained_active": "issue_active_pct",
            "launch__registers_per_thread": "registers", "lts__t_sector_hit_rate.pct": "l2_hit_pct",
            "sm__cycles_elapsed.avg.per_second": "sm_ghz", "launch__grid_size": "grid", "launch__block_size": "block"}
    idx = {h: i for i, h in enumerate(hdr)}
    units = rows[1]
    seen, kernels = set(), []
    for r in rows[2:]:
        name = r[idx["Kernel Name"]]
        key = (name, r[idx["launch__grid_size"]] if "launch__grid_size" in idx else "")
        if key in seen:
            continue
        seen.add(key)
        k = {"name": name[:110]}
        for m, label in want.items():
            if m in idx:
                try:
                    v = float(r[idx[m]].replace(",", ""))
                except ValueError:
                    v = r[idx[m]]
                if m == "gpu__time_duration.sum" and isinstance(v, float):
                    u = units[idx[m]]
                    v = v / 1e6 if u.startswith("ns") else v / 1e3 if u.startswith("us") else v
                if m.startswith("dram__bytes") and isinstance(v, float):
                    u = units[idx[m]]
                    v = v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(u, 1)
                k[label] = v
        if "dram_read_bytes" in k:
            k["dram_bytes_per_launch"] = k["dram_read_bytes"] + k["dram_write_bytes"]
        for sub, shape, abytes, aflops in SHAPES:
            if sub in name:
                k.update({"shape": shape, "algorithmic_bytes": abytes, "algorithmic_flops": aflops})
        kernels.append(k)
    out = {"source": "ncu --set full --clock-control none --import-source on --profile-from-start off -k "
                     "regex:'gemm_kernel|attn_fwd_kernel|layernorm_stream' -c 48 python bench.py --steps 1 --warmup 3 --profile "
                     "(gpurun, B200; report gpurun_out/prof_top.ncu-rep, not committed)",
           "note": "first launches of the timed step = SDXL level-1 SpatialTransformer (B2 = 16, 4096 tokens, C = 640); one "
                   "entry per distinct kernel / grid", "kernels": kernels}
    json.dump(out, open(os.path.join(ROOT, "profiles", f"{tag}_ncu_top_kernels.json"), "w"), indent=1)
    for k in kernels:
        print(k["name"][:60], k.get("grid"), k.get("ms_under_ncu"), k.get("tensor_pipe_active_pct"), k.get("dram_bytes_per_launch"))


launches()
full()
