"""Build the in-tree native library ``csrc/*.cu -> distributed_llm_code_samples_b200/_native/libb200ffn.so``.

nvcc cross-compiles sm_100a without a GPU, so this runs on the CPU dev box; the resulting ``.so`` is
git-ignored but travels with ``gpurun`` snapshots.  Everything is rebuilt whenever the SHA-256 of the sources,
headers and flags differs from the one recorded with the objects (``_native/fingerprint.txt``).

    python -m distributed_llm_code_samples_b200.build_native [--force] [--verbose]
"""
from __future__ import annotations

import argparse
import concurrent.futures as cf
import os
import shutil
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
CSRC = ROOT / "csrc"
INCLUDE = ROOT / "include"
OUT_DIR = Path(__file__).resolve().parent / "_native"
LIB_PATH = OUT_DIR / "libb200ffn.so"

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-std=c++17", "-lineinfo",
    "-Xcompiler", "-fPIC",
    "--expt-relaxed-constexpr",
    "-Xptxas", "-v",
    "-I", str(INCLUDE), "-I", str(CSRC),
]


def find_nvcc() -> str:
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(nvcc):
        raise RuntimeError("nvcc not found (needed to build libb200ffn.so)")
    return nvcc


def source_fingerprint(defines: tuple = ()) -> str:
    """SHA-256 over every source / header the library is built from plus the compiler flags.  The build is up to date
    iff the fingerprint recorded next to the objects matches - modification times say nothing about objects that arrived
    with a snapshot of the tree (``build()`` "reused" stale objects)."""
    import hashlib
    h = hashlib.sha256()
    files = sorted(CSRC.glob("*.cu")) + sorted(CSRC.glob("*.cuh")) + sorted(CSRC.glob("*.h")) + sorted(INCLUDE.glob("*.h"))
    for f in files:
        h.update(f.name.encode())
        h.update(f.read_bytes())
    h.update(" ".join(NVCC_FLAGS[:-4] + list(defines)).encode())        # flags without the absolute -I paths
    return h.hexdigest()


def _compile_one(nvcc: str, src: Path, obj: Path, verbose: bool, defines: tuple = ()) -> str:
    cmd = [nvcc, *NVCC_FLAGS, *[f"-D{d}" for d in defines], "-c", str(src), "-o", str(obj)]
    p = subprocess.run(cmd, capture_output=True, text=True)
    log = p.stdout + p.stderr
    (obj.with_suffix(".ptxas.log")).write_text(log)
    if p.returncode != 0:
        raise RuntimeError(f"nvcc failed for {src.name}:\n{log}")
    if verbose:
        print(log)
    return log


def build(force: bool = False, verbose: bool = False) -> Path:
    """Compile every ``csrc/*.cu`` for sm_100a and link ``libb200ffn.so``. Returns the library path."""
    nvcc = find_nvcc()
    OUT_DIR.mkdir(parents=True, exist_ok=True)
    srcs = sorted(CSRC.glob("*.cu"))
    if not srcs:
        raise RuntimeError(f"no CUDA sources under {CSRC}")
    stamp = OUT_DIR / "fingerprint.txt"
    fp = source_fingerprint()
    fresh = stamp.exists() and stamp.read_text().strip() == fp and LIB_PATH.exists()
    jobs = []
    objs = []
    for src in srcs:
        obj = OUT_DIR / (src.stem + ".o")
        objs.append(obj)
        if force or not fresh or not obj.exists():      # any source, header or flag changed: rebuild everything (seconds)
            jobs.append((src, obj))
    if jobs:
        with cf.ThreadPoolExecutor(max_workers=min(len(jobs), os.cpu_count() or 4)) as ex:
            futs = [ex.submit(_compile_one, nvcc, s, o, verbose) for s, o in jobs]
            for f in futs:
                f.result()
    if jobs or not LIB_PATH.exists():
        cmd = [nvcc, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", str(LIB_PATH), *map(str, objs)]
        p = subprocess.run(cmd, capture_output=True, text=True)
        if p.returncode != 0:
            raise RuntimeError(f"link failed:\n{p.stdout}{p.stderr}")
        stamp.write_text(fp + "\n")
    build.last_compiled = len(jobs)
    return LIB_PATH


build.last_compiled = 0


def build_variant(tag: str, defines: tuple) -> Path:
    """A/B builds: the same sources with extra ``-D`` defines -> ``_native/variants/<tag>/libb200ffn_<tag>.so``
    (select it at run time with ``B200_NATIVE_LIB=<path>``)."""
    nvcc = find_nvcc()
    out = OUT_DIR / "variants" / tag
    out.mkdir(parents=True, exist_ok=True)
    srcs = sorted(CSRC.glob("*.cu"))
    objs = [out / (s.stem + ".o") for s in srcs]
    with cf.ThreadPoolExecutor(max_workers=min(len(srcs), os.cpu_count() or 4)) as ex:
        for f in [ex.submit(_compile_one, nvcc, s, o, False, tuple(defines)) for s, o in zip(srcs, objs)]:
            f.result()
    lib = out / f"libb200ffn_{tag}.so"
    p = subprocess.run([nvcc, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", str(lib), *map(str, objs)], capture_output=True, text=True)
    if p.returncode != 0:
        raise RuntimeError(f"link failed:\n{p.stdout}{p.stderr}")
    return lib


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--force", action="store_true")
    ap.add_argument("--verbose", action="store_true")
    ap.add_argument("--variant", default=None, help="TAG:DEFINE[=V][,DEFINE...]  e.g. hint20k:B200_MBAR_HINT_NS=20000")
    a = ap.parse_args()
    if a.variant:
        tag, _, defs = a.variant.partition(":")
        print(build_variant(tag, tuple(d for d in defs.split(",") if d)))
        return 0
    path = build(force=a.force, verbose=a.verbose)
    print(path)
    return 0


if __name__ == "__main__":
    sys.exit(main())
