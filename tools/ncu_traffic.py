"""profiles/traffic.json from an ncu CSV: DRAM bytes (read + write) per launch of every kernel of one bench step.

    python tools/ncu_traffic.py profiles/r02_ncu_step_metrics.csv > profiles/traffic.json

The CSV is `ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none --csv`
over tools/one_step.py (the third step: -s 64 -c 32).  Kernel names are mapped to bench.py's launch tags; launches with
the same tag are averaged."""
import csv, json, re, sys

def tag_of(name):
    m = re.search(r"edgeconv_tc_kernel<(\d+), (\d+), \d+, (\d+)>", name)
    if m: return f"edgeconv_tc_g{m.group(1)}_c{m.group(2)}_k{m.group(3)}"
    m = re.search(r"linear_split_kernel<(\d+),", name)
    if m:
        K = int(m.group(1)) * 8
        return f"linear_split_k{K}_n{256 if K == 32 else 512}"
    m = re.search(r"cond_chain_tc_kernel<(\d+)>", name)
    if m: return f"cond_chain_tc_c{m.group(1)}"
    m = re.search(r"flow_tc_kernel<(\w+)>", name)
    if m: return "flow_inv_tc" if m.group(1) in ("1", "true") else "flow_fwd_tc"
    if "interp_tc_kernel" in name: return "interp_tc"
    m = re.search(r"knn_kernel<(\d+)>", name)
    if m: return "chamfer_nn" if m.group(1) == "1" else f"knn_topk_k{m.group(1)}"
    if "nn_grid_kernel" in name: return "chamfer_nn"
    if "knn_patch_kernel" in name: return "knn_patch"
    if any(k in name for k in ("fps_bucket_kernel", "fps_queue_kernel", "fps_cqueue_kernel", "fps_cluster_kernel")): return "fps_n40960"
    if re.search(r"\bfps_kernel", name) or "fps_cta_kernel" in name: return "fps_n2048"
    if "linear_kernel" in name: return "linear_k3_n128"
    m = re.search(r"normalize_kernel<(\w+)>", name)
    if m: return "gather_normalize" if m.group(1) in ("1", "true") else "normalize_pc"
    for k in ("gather_points", "denorm_concat", "remove_outliers"):
        if k in name: return k
    return None

def main():
    rows = list(csv.reader(l for l in open(sys.argv[1]) if not l.startswith("==")))
    hdr = rows[0]
    iname, imet, ival, iid = hdr.index("Kernel Name"), hdr.index("Metric Name"), hdr.index("Metric Value"), hdr.index("ID")
    acc = {}
    for r in rows[1:]:
        if len(r) <= ival or not r[imet].startswith("dram__bytes"):
            continue
        t = tag_of(r[iname])
        if t is None:
            continue
        acc.setdefault(t, {}).setdefault(r[iid], 0.0)
        acc[t][r[iid]] += float(r[ival].replace(",", ""))
    print(json.dumps({t: round(sum(v.values()) / len(v)) for t, v in sorted(acc.items())}, indent=1))


if __name__ == "__main__":
    main()
