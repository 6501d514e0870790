"""Fold an `ncu --set full` capture of one bench step into profiles/ncu_summary.json (read here, no GPU).

  python scripts/ncu_extract.py gpurun_out/x.ncu-rep <workload> <src_hash> [profiles/<round>_ncu_<kernel>_<workload>.txt]

Per kernel family (wv_extend, wv_shade, wv_resolve, rt_trace_shade, rt_path_trace) over all captured launches:
DRAM bytes (read + write, summed = per step when the capture covers one step), time-weighted L1 / L2 sector hit
rates, issue-slot utilisation, lanes per instruction, ALU / FMA pipe utilisation.  bench.py reads the file and
uses an entry only when its src_hash equals the hash of the sources it runs from (bench.py source_hash())."""
import csv, io, json, os, re, subprocess, sys

rep, workload, src_hash = sys.argv[1], sys.argv[2], sys.argv[3]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
col = {k: i for i, k in enumerate(hdr)}


def num(r, k):
    if k not in col:
        return None
    try:
        v = float(r[col[k]].replace(",", ""))
    except ValueError:
        return None
    u = units[col[k]].lower()
    scale = {"gbyte": 1e9, "mbyte": 1e6, "kbyte": 1e3, "byte": 1.0, "ms": 1e-3, "us": 1e-6, "ns": 1e-9, "s": 1.0,
             "msecond": 1e-3, "usecond": 1e-6, "nsecond": 1e-9, "second": 1.0}.get(u)
    return v * scale if scale and ("byte" in u or "second" in u or u in ("ms", "us", "ns", "s")) else v


fam = {}
for r in rows[2:]:
    name = re.sub(r"^void\s+", "", r[col["Kernel Name"]]).split("<")[0].split("(")[0].strip()
    t = num(r, "gpu__time_duration.sum") or 0.0
    f = fam.setdefault(name, dict(launches=0, time_s=0.0, dram=0.0, w={}))
    f["launches"] += 1
    f["time_s"] += t
    f["dram"] += (num(r, "dram__bytes_read.sum") or 0.0) + (num(r, "dram__bytes_write.sum") or 0.0)
    for key, metric in (("l1_hit", "l1tex__t_sector_hit_rate.pct"), ("l2_hit", "lts__t_sector_hit_rate.pct"),
                        ("issue", "smsp__issue_active.avg.pct_of_peak_sustained_active"),
                        ("lanes", "smsp__thread_inst_executed_per_inst_executed.ratio"),
                        ("alu", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active"),
                        ("fma", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active"),
                        ("lsu", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active"),
                        ("l1wf", "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed"),
                        ("warps", "smsp__warps_active.avg.per_cycle_active"),
                        ("eligible", "smsp__warps_eligible.avg.per_cycle_active")):
        v = num(r, metric)
        if v is not None:
            a = f["w"].setdefault(key, [0.0, 0.0])
            a[0] += v * t
            a[1] += t
kernels = {}
for name, f in fam.items():
    g = lambda k: (f["w"][k][0] / f["w"][k][1]) if k in f["w"] and f["w"][k][1] > 0 else None
    kernels[name] = {"launches": f["launches"], "time_ms_under_ncu": f["time_s"] * 1e3, "dram_bytes_per_step": f["dram"],
                     "l1_hit_rate": (g("l1_hit") / 100.0) if g("l1_hit") is not None else None,
                     "l2_hit_rate": (g("l2_hit") / 100.0) if g("l2_hit") is not None else None,
                     "issue_active_pct": g("issue"), "lanes_per_inst": g("lanes"), "alu_pipe_pct": g("alu"),
                     "fma_pipe_pct": g("fma"), "lsu_pipe_pct": g("lsu"), "l1_wavefronts_pct": g("l1wf"),
                     "warps_active_per_smsp": g("warps"), "warps_eligible_per_smsp": g("eligible")}
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
path = os.environ.get("NCU_SUMMARY_OUT") or os.path.join(root, "profiles", "ncu_summary.json")
try:
    allrec = json.load(open(path))
except Exception:
    allrec = {}
allrec[workload] = {"src_hash": src_hash, "file": sys.argv[4] if len(sys.argv) > 4 else os.path.basename(rep),
                    "note": "time-weighted means over the captured launches of each kernel family; dram bytes summed (one step)",
                    "kernels": kernels}
json.dump(allrec, open(path, "w"), indent=1, sort_keys=True)
print(json.dumps(allrec[workload], indent=1))
