"""Regenerate the static evidence under profiles/sass and profiles/ptxas from the built objects (no GPU needed).

    python tools/sass_report.py

* profiles/sass/mnemonic_counts.txt   every kernel of libb200ffn.so: instruction count, registers, spills and the counts of
                                      the mnemonics that prove what it is built on - UTCHMMA (tcgen05.mma), UTMALDG (TMA
                                      load), LDTM (tcgen05.ld), UTCBAR (tcgen05.commit), SYNCS (mbarrier), *.STRONG.SYS
                                      (peer-memory flags / volatile peer loads), LDGMC / multimem (NVLS), RED/ATOM
* profiles/sass/<kernel>.key_lines.sass  the tensor / TMA / barrier / peer-memory instructions of the flagship kernels
* profiles/ptxas/*.ptxas.log          nvcc -Xptxas -v output per translation unit
"""
from __future__ import annotations

import re
import shutil
import subprocess
import sys
from collections import Counter
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
NATIVE = ROOT / "distributed_llm_code_samples_b200" / "_native"
OUT = ROOT / "profiles"

PROOF = re.compile(r"^(UTCHMMA|UTCQMMA|UTMALDG|UTMASTG|UTMACCTL|LDTM|STTM|UTCBAR|UTCATOMSWS|UTCCP|SYNCS|LDGMC|REDG|RED\b|ATOMG|ATOMS|MEMBAR|FENCE|ERRBAR|CGAERRBAR|"
                   r"LDG\.E\.128|LDG\.E\.STRONG|LDG\.E\.128\.STRONG|STG\.E\.128|STG\.E\.STRONG|NANOSLEEP|CCTL|HMMA|MUFU|ELECT|UCGABAR|ACQBULK|BAR\b)")
KEY = re.compile(r"UTCHMMA|UTMALDG|LDTM|UTCBAR|UTCATOMSWS|SYNCS|STRONG\.SYS|STRONG\.GPU|MEMBAR|FENCE|REDG|LDGMC|UCGABAR|NANOSLEEP|ERRBAR")
FLAGSHIP = {
    "gemm_pair_relu_mask": ("gemm", "gemm_tcgen05_pair_kernelINS_8LocalEpiILi8E"),
    "gemm_pair_fused_kernel": ("fused", "gemm_tcgen05_pair_kernelINS_9FusedEpiTILi4ELi32ELi0ELi0E"),
    "gemm_1cta_fused_kernel": ("fused", "gemm_tcgen05_kernelINS_9FusedEpiTILi4ELi32ELi0ELi0E"),
    "gemm_1cta_fused_nvls_kernel": ("fused", "gemm_tcgen05_kernelINS_9FusedEpiTILi4ELi32ELi0ELi1E"),
    "gemm_pair_fused_adam_kernel": ("fused", "gemm_tcgen05_pair_kernelINS_9FusedEpiTILi4ELi32ELi1ELi0E"),
    "gemm_pair_gather_kernel": ("gather_gemm", "gemm_tcgen05_pair_kernelINS_9GatherEpiILi8E"),
    "gemm_pair_f32_split_kernel": ("gemm_f32", "gemm_tcgen05_pair_kernelINS_6F32EpiILi1E"),
    "nvls_all_reduce_kernel": ("vmm", "nvls_coll_kernelILi0ELb0E"),
}


def demangle(names: list[str]) -> dict[str, str]:
    if not names or shutil.which("c++filt") is None:
        return {n: n for n in names}
    out = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True).stdout.splitlines()
    return dict(zip(names, out))


def kernels(obj: Path):
    text = subprocess.run(["cuobjdump", "-sass", str(obj)], capture_output=True, text=True, check=True).stdout
    cur, out = None, {}
    for line in text.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = m.group(1)
            out[cur] = []
            continue
        m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?)\s*;", line)
        if cur and m:
            out[cur].append(m.group(2))
    return out


def resources(log: Path) -> dict[str, str]:
    res, cur = {}, None
    if not log.exists():
        return res
    for line in log.read_text().splitlines():
        m = re.search(r"Compiling entry function '(\S+)'", line)
        if m:
            cur = m.group(1)
        m = re.search(r"(\d+) bytes stack frame, (\d+) bytes spill stores, (\d+) bytes spill loads", line)
        if m and cur:
            res[cur] = f"stack {m.group(1)} B, spill st/ld {m.group(2)}/{m.group(3)} B"
        m = re.search(r"Used (\d+) registers", line)
        if m and cur:
            res[cur] = f"{m.group(1)} regs, " + res.get(cur, "")
    return res


def main() -> int:
    sys.path.insert(0, str(ROOT))
    from distributed_llm_code_samples_b200.build_native import build
    build()
    (OUT / "sass").mkdir(parents=True, exist_ok=True)
    (OUT / "ptxas").mkdir(parents=True, exist_ok=True)
    lines = ["# cuobjdump -sass of every object of libb200ffn.so (sm_100a): per kernel, the instruction count, ptxas resources and",
             "# the counts of the mnemonics that show what it is built on.  UTCHMMA = tcgen05.mma, UTMALDG = TMA tensor load,",
             "# LDTM = tcgen05.ld, UTCBAR = tcgen05.commit, SYNCS = mbarrier, *.STRONG.SYS = system-scope flag / peer accesses,",
             "# REDG...GPU = device-scope release counter, FENCE.VIEW.ASYNC = generic->async proxy fence, LDGMC = multimem (NVLS).",
             "# Regenerate with: python tools/sass_report.py", ""]
    total = 0
    for obj in sorted(NATIVE.glob("*.o")):
        ks = kernels(obj)
        log = NATIVE / (obj.stem + ".ptxas.log")
        if log.exists():
            shutil.copyfile(log, OUT / "ptxas" / log.name)
        res = resources(log)
        names = demangle(list(ks))
        lines.append(f"## {obj.stem}.cu  ({len(ks)} kernels)")
        lines.append("")
        for name, body in ks.items():
            total += 1
            cnt = Counter()
            for ins in body:
                op = re.sub(r"^@!?U?P\w+\s+", "", ins).split()[0]
                if PROOF.match(op):
                    cnt[op] += 1
            pretty = re.sub(r"\(.*", "", names[name])
            lines.append(f"{pretty}")
            lines.append(f"  {len(body)} instructions; {res.get(name, 'resources: see ptxas log')}")
            lines.append("  " + (", ".join(f"{k} x{v}" for k, v in sorted(cnt.items())) or "(none of the tracked mnemonics)"))
            lines.append("")
        for tag, (stem, needle) in FLAGSHIP.items():
            if stem != obj.stem:
                continue
            for name, body in ks.items():
                if needle in name:
                    key = [f"{i:5d}  {ins}" for i, ins in enumerate(body) if KEY.search(ins)]
                    (OUT / "sass" / f"{tag}.key_lines.sass").write_text(
                        f"// {names[name]}\n// {len(body)} instructions; tensor / TMA / barrier / peer-memory lines only (index = instruction number)\n"
                        + "\n".join(key) + "\n")
    (OUT / "sass" / "mnemonic_counts.txt").write_text("\n".join(lines))
    print(f"{total} kernels -> profiles/sass/mnemonic_counts.txt")
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
