"""Turn `ncu --page raw --csv` exports (tools/run_r2_ncu.sh) into the per-kernel summaries kept under profiles/ncu/.

    python tools/ncu_summary.py gpurun_out/r2ncu profiles/ncu/r2
Writes <case>.summary.csv (metric,unit,value) per capture and SUMMARY.md (one row per kernel family)."""
import csv, glob, os, sys

KEYS = ["gpu__time_duration.sum", "gpc__cycles_elapsed.avg.per_second", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__inst_executed.sum", "launch__registers_per_thread", "launch__block_size",
        "launch__grid_size", "launch__cluster_dim_x", "launch__shared_mem_per_block_dynamic", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__m_xbar2l1tex_read_bytes.sum", "l1tex__m_l1tex2xbar_write_bytes.sum",
        "lts__t_sector_hit_rate.pct", "sm__warps_active.avg.pct_of_peak_sustained_active", "sm__cycles_active.avg"]
UNIT = {"Gbyte": 1.0, "Mbyte": 1e-3, "Kbyte": 1e-6, "byte": 1e-9}
src, dst = sys.argv[1], sys.argv[2]
os.makedirs(dst, exist_ok=True)
rows_md = []
for f in sorted(glob.glob(os.path.join(src, "prof_*.raw.csv"))):
    rows = list(csv.reader(open(f)))
    if len(rows) < 3:
        continue
    h, u, r = rows[0], rows[1], rows[2]
    d, un = dict(zip(h, r)), dict(zip(h, u))
    case = os.path.basename(f)[5:-8]
    with open(os.path.join(dst, f"{case}.summary.csv"), "w") as o:
        o.write(f"Kernel Name,,{d.get('Kernel Name', '')}\n")
        for k in KEYS:
            if k in d:
                o.write(f"{k},{un.get(k, '')},{d[k]}\n")

    def g(k, scale=1.0, nd=1):
        try:
            return f"{float(d[k]) * scale:.{nd}f}"
        except (KeyError, ValueError):
            return "-"

    def gb(k):
        try:
            return float(d[k]) * UNIT.get(un.get(k, ""), 1.0)
        except (KeyError, ValueError):
            return 0.0

    dur = d.get("gpu__time_duration.sum", "-") + " " + un.get("gpu__time_duration.sum", "")
    rows_md.append(f"| `{case}` | `{d.get('Kernel Name', '')[:70]}` | {dur} | {g('sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active')} | "
                   f"{g('smsp__inst_executed.sum', 1e-6)} | {d.get('launch__registers_per_thread', '-')} x {d.get('launch__block_size', '-')} | "
                   f"{gb('dram__bytes_read.sum'):.2f} / {gb('dram__bytes_write.sum'):.2f} | {g('gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed')} | "
                   f"{gb('l1tex__m_xbar2l1tex_read_bytes.sum'):.2f} / {gb('l1tex__m_l1tex2xbar_write_bytes.sum'):.2f} | {g('gpc__cycles_elapsed.avg.per_second', 1.0, 2)} |")
with open(os.path.join(dst, "SUMMARY.md"), "w") as o:
    o.write("| case (`tools/ncu_gemm.py`) | kernel | duration | tensor pipe % | M warp-instr | regs x threads | DRAM rd / wr GB | DRAM % of peak | L2->SM / SM->L2 GB | GHz |\n")
    o.write("|---|---|---|---|---|---|---|---|---|---|\n")
    o.write("\n".join(rows_md) + "\n")
print(open(os.path.join(dst, "SUMMARY.md")).read())
