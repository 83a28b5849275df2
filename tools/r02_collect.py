"""Copies one evidence run (gpurun_out/<tag>_*) into profiles/: raw ncu exports, launch list, traffic, bench lines; derives
profiles/search_roofline.json and the plain keys of profiles/traffic.json that bench.py reads."""
import csv, collections, json, os, shutil, subprocess, sys
tag = sys.argv[1]
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
G, P = os.path.join(ROOT, "gpurun_out"), os.path.join(ROOT, "profiles")
for f in os.listdir(G):
    if f.startswith(tag + "_") and f.endswith((".json", ".csv")) and "b_" not in f:
        shutil.copy(os.path.join(G, f), os.path.join(P, f))
for n in ("chain_batch", "search", "edge_plane"):
    rep = os.path.join(G, f"{tag}_{n}.ncu-rep")
    if os.path.exists(rep):
        open(os.path.join(P, f"{tag}_{n}_ncu_raw.csv"), "w").write(subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout)
have_search = os.path.exists(os.path.join(P, f"{tag}_search_ncu_raw.csv"))  # a chain-only run (SKIP_SEARCH=1) leaves the search files alone
rows = list(csv.reader(open(os.path.join(P, f"{tag}_search_ncu_raw.csv")))) if have_search else [[], []]
h = rows[0]
for r in rows[2:]:
    d = dict(zip(h, r))
    if "cdef_search3" in d["Kernel Name"]:
        out = {"source": f"profiles/{tag}_search_ncu_raw.csv (ncu --set full --clock-control none, 4K 10-bit, 64 strengths, one launch)",
               "warp_instructions_per_launch": int(float(d["smsp__inst_executed.sum"])),
               "issue_active_pct": round(float(d["smsp__issue_active.avg.pct_of_peak_sustained_active"]), 1),
               "alu_pipe_pct": round(float(d["sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active"]), 1),
               "fma_pipe_pct": round(float(d["sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active"]), 1),
               "lsu_pipe_pct": round(float(d["sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active"]), 1),
               "gpu_time_us_under_ncu": float(d["gpu__time_duration.sum"]), "sm_mhz_assumed": 1965}
        json.dump(out, open(os.path.join(P, "search_roofline.json"), "w"), indent=1)
        print(out)
        break
agg = collections.defaultdict(lambda: collections.defaultdict(list))
for f in (f"{tag}_traffic_batch.csv",) + ((f"{tag}_traffic_search.csv",) if have_search else ()):
    rows = [r for r in csv.reader(open(os.path.join(P, f))) if len(r) > 10]
    h = rows[0]
    for r in rows[1:]:
        d = dict(zip(h, r))
        agg[d["Kernel Name"].replace("void ", "").split("<")[0].split("(")[0]][d["Metric Name"]].append(float(d["Metric Value"].replace(",", "")))
m = {k: {a: sum(b) / len(b) for a, b in v.items()} for k, v in agg.items()}
t = json.load(open(os.path.join(P, "traffic.json")))
for k, name in (("deblock_kernel", "deblock_kernel"), ("cdef5_kernel", "cdef5_kernel+ccso")) + ((("cdef_search3_kernel", "cdef_search3_kernel"),) if have_search else ()):
    t[name] = int(m[k]["dram__bytes_read.sum"] + m[k]["dram__bytes_write.sum"])
    t[f"{tag}_{name}"] = {"read": int(m[k]["dram__bytes_read.sum"]), "write": int(m[k]["dram__bytes_write.sum"]), "gpu_time_ns": int(m[k]["gpu__time_duration.sum"])}
t["_note_" + tag] = t.get("_note_r02f", "").replace("r02f", tag)
json.dump(t, open(os.path.join(P, "traffic.json"), "w"), indent=1)
print({k: t[k] for k in t if k.startswith(tag)})
