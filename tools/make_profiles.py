"""Turn the scratch outputs of tools/gpu_full.sh (gpurun_out/) into the tracked evidence
under profiles/ for round RR:  python tools/make_profiles.py r01"""
import csv, json, os, shutil, subprocess, sys

RR = sys.argv[1] if len(sys.argv) > 1 else "r01"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GO, PR = os.path.join(ROOT, "gpurun_out"), os.path.join(ROOT, "profiles")
os.makedirs(PR, exist_ok=True)

WANT = ["Kernel Name", "Grid Size", "Block Size", "gpu__time_duration.sum", "dram__bytes_read.sum",
        "dram__bytes_write.sum", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "launch__registers_per_thread",
        "launch__shared_mem_per_block_dynamic", "launch__shared_mem_per_block_static",
        "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
        "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_tensor_subpipe_dmma_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "lts__t_sector_hit_rate.pct"] + [
        f"smsp__average_warps_issue_stalled_{k}_per_issue_active.ratio" for k in
        ("barrier", "wait", "short_scoreboard", "long_scoreboard", "math_pipe_throttle", "mio_throttle",
         "no_instruction", "not_selected", "dispatch_stall", "branch_resolving")]


def ncu_full(wl):
    rep = os.path.join(GO, f"full_{wl}.ncu-rep")
    rawcsv = os.path.join(GO, f"full_{wl}.raw.csv")
    if os.path.exists(rawcsv) and os.path.getsize(rawcsv):
        raw = open(rawcsv).read()
    elif os.path.exists(rep):
        raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    else:
        return None
    rows = list(csv.reader(raw.splitlines()))
    hdr, units, data = rows[0], rows[1], rows[2:]
    out = os.path.join(PR, f"{RR}_ncu_full_{wl}.csv")
    traffic = {}
    with open(out, "w") as f:
        f.write("metric,unit," + ",".join(f"launch{i}" for i in range(len(data))) + "\n")
        for h in WANT:
            if h not in hdr:
                continue
            i = hdr.index(h)
            vals = [d[i].split("(")[0].replace("void ", "").strip() if h == "Kernel Name"
                    else d[i].replace(",", "") if h not in ("Grid Size", "Block Size") else d[i].replace(",", " ")
                    for d in data]
            f.write(f"{h},{units[i]}," + ",".join(vals) + "\n")
    ir, iw, ik, it = (hdr.index(x) for x in ("dram__bytes_read.sum", "dram__bytes_write.sum", "Kernel Name",
                                              "gpu__time_duration.sum"))
    scale = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0}
    for d in data:
        name = d[ik].split("(")[0].replace("void ", "").split("<")[0].split("::")[-1]
        traffic[name] = {"dram_bytes": (float(d[ir]) * scale[units[ir]] + float(d[iw]) * scale[units[iw]]),
                         "ms_under_ncu": float(d[it]) * {"ms": 1.0, "us": 1e-3, "ns": 1e-6, "msecond": 1.0,
                                                         "usecond": 1e-3, "nsecond": 1e-6}[units[it]]}
    return traffic


def launch_list(wl):
    src = os.path.join(GO, f"launches_{wl}.csv")
    if not os.path.exists(src):
        return None
    dst = os.path.join(PR, f"{RR}_launches_{wl}.csv")
    lines = [l for l in open(src) if not l.startswith("==")]
    open(dst, "w").writelines(lines)
    agg = {}
    for r in csv.DictReader(lines):
        k = r["Kernel Name"].split("(")[0].replace("void ", "").split("<")[0].split("::")[-1]
        a = agg.setdefault(k, [0, 0.0])
        a[0] += 1; a[1] += float(r["Metric Value"]) * 1e-6
    return agg


summary = {"round": RR, "bench": {}, "ncu": {}}
for name in ("default", "reference", "c2", "c1", "c3", "c4", "c5", "horizons"):
    p = os.path.join(GO, f"bench_{name}.json")
    if os.path.exists(p) and os.path.getsize(p):
        line = open(p).read().strip().splitlines()[-1]
        open(os.path.join(PR, f"{RR}_bench_{name}.json"), "w").write(line + "\n")
        summary["bench"][name] = json.loads(line)
for wl in ("c1b", "c2", "c3", "c5", "c1bx", "c2x"):
    summary["ncu"][wl] = {"full": ncu_full(wl), "launches": launch_list(wl)}
for f in ("pytest_gpu.log", "smoke.log"):
    if os.path.exists(os.path.join(GO, f)):
        shutil.copy(os.path.join(GO, f), os.path.join(PR, f"{RR}_{f}"))
json.dump(summary["ncu"], open(os.path.join(PR, f"{RR}_ncu_digest.json"), "w"), indent=1)
for wl, d in summary["ncu"].items():
    print(wl, json.dumps(d, indent=1))
subprocess.run([sys.executable, os.path.join(ROOT, "tools", "sass_summary.py"), RR], stdout=subprocess.DEVNULL)
for k, b in summary["bench"].items():
    if "value" not in b:
        continue
    r = b.get("roofline") or {}
    print(k, "%.0f %s" % (b["value"], b["unit"]), "ms/step %.3f" % b["ms_per_step"], r.get("bound"), r.get("frac"),
          "e2e", (b.get("e2e") or {}).get("value"))
