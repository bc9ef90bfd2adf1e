"""Turn one round of ncu output into the two small JSON files bench.py quotes (profiles/r2_traffic.json, profiles/r2_launch_summary.json).

    python tools/make_profile_json.py --stamp <hx_kernel_stamp> --launches gpurun_out/X_launches.csv --rep gpurun_out/X_ks.ncu-rep

* --launches: the CSV of `ncu --metrics gpu__time_duration.sum --clock-control none --csv` over one bench step. Launches are grouped into
  the engine's kernel families (the names hx_prof_name() reports) by kernel name and grid: the 32-rotation batches of the hoisted key
  switch are the launches whose grid spans >= 31 rotations.
* --rep: one `ncu --set full` capture holding the four kernels of a 32-rotation batch; traffic = dram__bytes_read.sum +
  dram__bytes_write.sum summed over the group (K1 + K2 + K3 on the FP64 limbs + K3 on the 60-bit limb), averaged over the batches captured.

Both files carry the kernel stamp of the library they were measured on; bench.py ignores them when its library's stamp differs.
Needs `ncu` on PATH to read the .ncu-rep (no GPU).
"""
import argparse, collections, csv, io, json, re, subprocess, sys


def grid(s):
    return tuple(int(x) for x in re.findall(r"\d+", s))


def family(name, g):
    """Kernel family of one launch. `g` is the launch grid (x, y, z)."""
    if name.startswith("k_ks_p_inner") or name.startswith("k_ks_q_finish"):
        return "ks_hoisted" if g[2] >= 31 else "ks_fused"
    if name.startswith("k_moddown_w"):
        return "ks_hoisted" if g[1] >= 62 else "ks_fused_or_modup_middle"
    if name.startswith("k_conv_fwd_col"):
        return "ks_hoisted" if g[2] >= 31 else "ks_fused"
    if name.startswith("k_inv_cols") or name.startswith("k_inv_rows"):
        return "inverse_ntt (front of modup / moddown / rescale)"
    if name.startswith("k_modup_rows"):
        return "modup"
    if name.startswith("k_rescale"):
        return "rescale"
    if name.startswith("k_diag_mac"):
        return "diag_mac"
    if name.startswith("k_tensor"):
        return "tensor"
    if name in ("k_ew", "k_wsum", "k_add_scalar", "k_fold_mod"):
        return "elementwise"
    return name


def launch_summary(path):
    rows = [r for r in csv.DictReader(l for l in open(path) if l.startswith('"'))]
    fam = collections.defaultdict(lambda: {"launches": 0, "us": 0.0, "kernels": collections.Counter()})
    for r in rows:
        if r.get("Metric Name") != "gpu__time_duration.sum":
            continue
        name = re.sub(r"\(.*", "", r["Kernel Name"]).replace("void ", "")
        f = fam[family(name, grid(r["Grid Size"]))]
        f["launches"] += 1
        f["us"] += float(r["Metric Value"].replace(",", "")) / 1e3
        f["kernels"][name] += 1
    total = sum(f["us"] for f in fam.values())
    out = {}
    for k, f in sorted(fam.items(), key=lambda kv: -kv[1]["us"]):
        out[k] = {"launches": f["launches"], "total_us": round(f["us"], 1), "share": round(f["us"] / total, 4), "kernels": dict(f["kernels"])}
    return out, total, sum(f["launches"] for f in fam.values())


def traffic(rep):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units = rows[0], rows[1]
    ix = {h: i for i, h in enumerate(hdr)}
    scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    per = collections.defaultdict(lambda: {"n": 0, "bytes": 0.0, "us": 0.0})
    for r in rows[2:]:
        name = re.sub(r"\(.*", "", r[ix["Kernel Name"]]).replace("void ", "")
        g = grid(r[ix["Grid Size"]])
        if family(name, g) != "ks_hoisted" or (name.startswith("k_moddown_w") and g[1] != 64) or (not name.startswith("k_moddown_w") and g[2] != 32):
            continue
        b = sum(float(r[ix[m]].replace(",", "")) * scale[units[ix[m]]] for m in ("dram__bytes_read.sum", "dram__bytes_write.sum"))
        p = per[name]
        p["n"] += 1
        p["bytes"] += b
        p["us"] += float(r[ix["gpu__time_duration.sum"]].replace(",", "")) * {"us": 1.0, "ns": 1e-3, "ms": 1e3}[units[ix["gpu__time_duration.sum"]]]
    return {k: {"captured": v["n"], "dram_bytes_per_launch": round(v["bytes"] / v["n"]), "us_per_launch_under_ncu": round(v["us"] / v["n"], 2)}
            for k, v in per.items()}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--stamp", required=True)
    ap.add_argument("--launches")
    ap.add_argument("--rep")
    ap.add_argument("--alg-bytes", type=float, default=None, help="algorithmic bytes per 32-rotation batch (bench.py's roofline.algorithmic_bytes_per_launch)")
    ap.add_argument("--out-prefix", default="profiles/r2")
    a = ap.parse_args()
    if a.launches:
        fam, total, n = launch_summary(a.launches)
        json.dump({"kernel_stamp": a.stamp, "source": a.launches.replace("gpurun_out/", "profiles/"),
                   "_comment": "one bench step (membership + index, configs[1]) under `ncu --metrics gpu__time_duration.sum --clock-control none`: "
                               "serialized, cold-cache launch times; shares, not absolutes, are comparable with the live run",
                   "launches": n, "total_us": round(total, 1), "families": fam}, open(a.out_prefix + "_launch_summary.json", "w"), indent=1)
        print(json.dumps({k: (v["launches"], v["total_us"], v["share"]) for k, v in fam.items()}, indent=1))
    if a.rep:
        per = traffic(a.rep)
        need = [k for k in per]
        tot = sum(v["dram_bytes_per_launch"] for v in per.values())
        d = {"kernel_stamp": a.stamp, "source": a.rep.replace("gpurun_out/", "gpurun_out/ (not committed) "),
             "_comment": "dram__bytes_read.sum + dram__bytes_write.sum per launch from one `ncu --set full --clock-control none` capture, B200, "
                         "configs[1]; ks_hoisted = one group of 32 hoisted rotations = the kernels listed under `kernels`",
             "ks_hoisted": {"traffic_bytes_per_launch": tot, "algorithmic_bytes_per_launch": a.alg_bytes, "kernels": per,
                            "launch": "32 hoisted rotations, N=2^15, 12+4 limbs, dnum 3"}}
        json.dump(d, open(a.out_prefix + "_traffic.json", "w"), indent=1)
        print(json.dumps(d["ks_hoisted"], indent=1), need)


if __name__ == "__main__":
    sys.exit(main())
