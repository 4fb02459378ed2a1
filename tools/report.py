"""Turn the benchmark JSON lines under profiles/bench/ into a roofline table (markdown on stdout).

Model FLOPs per optimizer step = 12 * tokens * D * F * L (six GEMMs of 2*T*D*F per block; the reference executes a
seventh).  Denominator: the driver's MEASURED sustained cuBLAS bf16 rate (MEASURED_PEAKS.json, fallback 1394 TF/s)."""
import glob, json, os, re, sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
peak = 1394.0
try:
    peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["bf16_tflops_sustained"])
except Exception:
    pass
rows = []
records = []
for f in sorted(glob.glob(os.path.join(ROOT, "profiles", "bench", "bench_*_n*.json")) + glob.glob(os.path.join(ROOT, "profiles", "bench", "r2", "bench_*_n*.json"))):
    try:
        d = json.load(open(f))
    except Exception:
        continue
    tag = "r2 " + os.path.basename(f)[6:-5] if os.sep + "r2" + os.sep in f else "r1"
    records.append((d, tag))
    for sub in (d.get("also") or {}).values():          # round 2: DDP headline + also.fsdp / also.mp in one line
        if isinstance(sub, dict) and "metric" in sub:
            records.append((sub, tag))
for d, tag in records:
    m = re.search(r"d(\d+)_f(\d+)_l(\d+)", d["config"]["model"])
    D, F, L = map(int, m.groups())
    n = d["n_gpus"]
    tokens_global = d["config"]["global_batch"]
    par = d["config"]["parallelism"]
    flops = 12.0 * tokens_global * D * F * L                      # whole job, per step
    gpus_sharing = n                                              # dp: n x the work on n GPUs; tp: the work split over n
    tf_per_gpu = flops / (d["ms_per_step"] * 1e-3) / 1e12 / gpus_sharing
    rows.append((d["metric"].split("_")[2], n, d.get("impl", "ours"), d["value"], d["ms_per_step"], tf_per_gpu, tf_per_gpu / peak,
                 d["clocks"]["sm_mhz"] if d.get("clocks") else None, par, tag))
print(f"| method | GPUs | arm | tokens/s | ms/step | model TFLOP/s per GPU | of measured sustained cuBLAS ({peak:.0f}) | SM MHz | run |")
print("|---|---|---|---|---|---|---|---|---|")
for r in sorted(rows, key=lambda r: (r[0], r[1], r[2], r[9])):
    print(f"| {r[0]} | {r[1]} | {r[2]} | {r[3]:,.0f} | {r[4]:.1f} | {r[5]:.0f} | {r[6]:.2f} | {r[7]} | {r[9]} |")
