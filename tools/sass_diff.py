"""Check that a refactor left the machine code of existing kernels unchanged.

    python tools/sass_diff.py snapshot            # store the SASS of every kernel of every object under gpurun_out/sass_snapshot/
    python tools/sass_diff.py compare             # rebuild, then compare kernel by kernel (by mangled name)

Instruction text is compared with the encoding comments stripped; kernels that only exist on one side are listed.
Used while no GPU is available: a change whose kernels compare equal cannot have changed validated behaviour.
"""
from __future__ import annotations

import json
import re
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
NATIVE = ROOT / "distributed_llm_code_samples_b200" / "_native"
SNAP = ROOT / "gpurun_out" / "sass_snapshot"


def kernels(obj: Path) -> dict[str, list[str]]:
    text = subprocess.run(["cuobjdump", "-sass", str(obj)], capture_output=True, text=True, check=True).stdout
    out: dict[str, list[str]] = {}
    cur = None
    for line in text.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = m.group(1)
            out[cur] = []
        elif cur and re.match(r"\s+/\*[0-9a-f]{4}\*/", line):
            out[cur].append(re.sub(r"/\* 0x[0-9a-f]+ \*/", "", line).strip())
    return out


def main() -> int:
    mode = sys.argv[1] if len(sys.argv) > 1 else "compare"
    sys.path.insert(0, str(ROOT))
    from distributed_llm_code_samples_b200.build_native import build
    build()
    cur = {o.stem: kernels(o) for o in sorted(NATIVE.glob("*.o"))}
    if mode == "snapshot":
        SNAP.mkdir(parents=True, exist_ok=True)
        (SNAP / "kernels.json").write_text(json.dumps(cur))
        print(f"stored {sum(len(v) for v in cur.values())} kernels")
        return 0
    old = json.loads((SNAP / "kernels.json").read_text())
    bad = 0
    for obj, ks in old.items():
        for name, body in ks.items():
            now = cur.get(obj, {}).get(name)
            if now is None:
                # a template that gained a (defaulted) parameter changes the mangled name: accept an identical body
                twin = [n for n, b in cur.get(obj, {}).items() if b == body and n not in ks]
                if twin:
                    print(f"renamed {obj}: {name[:70]} -> {twin[0][:70]} (identical code)")
                else:
                    print(f"MISSING {obj}: {name[:100]}")
                    bad += 1
            elif now != body:
                print(f"CHANGED {obj}: {name[:100]} ({len(body)} -> {len(now)} instructions)")
                bad += 1
    new = [(o, n) for o, ks in cur.items() for n in ks if n not in old.get(o, {})]
    for o, n in new:
        print(f"new     {o}: {n[:100]}")
    print(f"{bad} changed/missing, {len(new)} new, {sum(len(v) for v in old.values()) - bad} identical")
    return 1 if bad else 0


if __name__ == "__main__":
    raise SystemExit(main())
