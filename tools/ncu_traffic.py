"""DRAM traffic per launch of each product kernel from an `ncu --set full` report -> profiles/<name>.json
(bench.py copies the dominant kernel's figure into roofline.traffic): python tools/ncu_traffic.py rep out.json"""
import csv, io, json, subprocess, sys, collections
rep, out = sys.argv[1], sys.argv[2]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
H, U = rows[0], rows[1]
ki, ri, wi, ti = H.index("Kernel Name"), H.index("dram__bytes_read.sum"), H.index("dram__bytes_write.sum"), H.index("gpu__time_duration.sum")
scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
acc = collections.defaultdict(list)
STAGE = {"blend_bwd_kernel": "blend_bwd", "blend_fwd_kernel": "blend_fwd", "preprocess_fwd_kernel": "preprocess_fwd", "preprocess_fwd_views_kernel": "preprocess_fwd_views",
         "preprocess_bwd_kernel": "preprocess_bwd", "preprocess_bwd_views_kernel": "preprocess_bwd_views", "duplicate_tiles_kernel": "duplicate_keys", "tile_ranges_kernel": "tile_ranges",
         "onesweep_pass_kernel": "sort_pass", "scan_tiles_kernel": "scan"}
for r in rows[2:]:
    for k, st in STAGE.items():
        if k in r[ki]:
            f = lambda i: float(r[i].replace(",", "")) * scale.get(U[i], 1.0)
            g = lambda name: float(r[H.index(name)].replace(",", "")) if name in H and r[H.index(name)] else None
            acc[st].append((f(ri) + f(wi), float(r[ti].replace(",", "")), g("smsp__inst_executed.sum"),
                            g("smsp__issue_active.avg.pct_of_peak_sustained_active"),
                            g("l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed")))
avg = lambda v, i: (sum(x[i] for x in v if x[i] is not None) / max(1, sum(1 for x in v if x[i] is not None)))
res = {st: {"dram_bytes_per_launch": avg(v, 0), "launches": len(v), "avg_us_under_ncu": avg(v, 1),
            "warp_inst_executed_per_launch": avg(v, 2), "issue_active_pct": round(avg(v, 3), 1),
            "l1tex_data_pipe_wavefronts_pct": round(avg(v, 4), 1)} for st, v in acc.items()}
res["_source"] = rep.split("/")[-1] + " (ncu --set full --clock-control none, bench.py --steps 1 --views 2, C2)"
json.dump(res, open(out, "w"), indent=1)
print(json.dumps(res, indent=1))
