"""bench.py's contract pieces that can be checked without a GPU: it always prints ONE JSON line and exits 0; both arms
describe the run with IDENTICAL `config` blocks; the extra strategies are chosen and launched as documented."""
import argparse
import json
import os
import subprocess
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU answer")
@pytest.mark.parametrize("impl", ["ours", "reference"])
def test_bench_without_a_gpu_prints_one_json_line_and_exits_zero(impl):
    p = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", impl], capture_output=True, text=True, timeout=300)
    assert p.returncode == 0, p.stderr[-500:]
    lines = [l for l in p.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1
    rec = json.loads(lines[0])
    assert rec["impl"] == impl and "unavailable" in rec


def _args(**kw):
    base = dict(method="ddp", also="auto", seq_len=2048, batch_size=8, steps=5, warmup=3, model_size=None, ffn_dim=None, layers=None)
    base.update(kw)
    return argparse.Namespace(**base)


def test_both_arms_emit_the_same_config_block():
    """The driver compares the two arms' `config`; everything arm-specific lives in `impl_detail` (same_config)."""
    a = _args()
    for method, world in (("ddp", 8), ("fsdp", 4), ("mp", 8), ("single", 1)):
        d, f, l = bench.shape_defaults(a, method if method != "single" else "ddp")
        ours = bench.record("ours", method, world, a, 1.0, 10.0, None, None, 7, bench.base_config(method, world, d, f, l, a, 123), {"comm": "x"})
        ref = bench.record("reference", method, world, a, 1.0, 10.0, None, None, 0, bench.base_config(method, world, d, f, l, a, 123), {"comm": "y"})
        assert ours["config"] == ref["config"]
        assert ours["metric"] == ref["metric"] and ours["scaling"] == ("weak" if method in ("ddp", "fsdp") else "strong")
        assert set(ours) == set(ref)
    assert bench.shape_defaults(a, "mp") == (8192, 28672, 16) and bench.shape_defaults(a, "ddp") == (4096, 16384, 32)


def test_also_strategies_selection():
    assert bench.also_methods(_args()) == ["fsdp", "mp"]
    assert bench.also_methods(_args(method="mp")) == []
    assert bench.also_methods(_args(also="none")) == []
    assert bench.also_methods(_args(also="mp,ddp_recompute")) == ["mp", "ddp_recompute"]


def test_child_environment_drops_the_elastic_agent_store(monkeypatch):
    """A child bench on another port must host its own rendezvous store: with torchrun's TORCHELASTIC_* variables it would
    wait for the agent's store on a port nobody serves (measured: a 15-minute hang)."""
    seen = {}

    class P:
        stdout, stderr, returncode = '{"value": 1.0, "metric": "m"}\n', "", 0

    def fake_run(cmd, env=None, **kw):
        seen["cmd"], seen["env"] = cmd, env
        return P()

    monkeypatch.setattr(subprocess, "run", fake_run)
    monkeypatch.setenv("TORCHELASTIC_USE_AGENT_STORE", "True")
    monkeypatch.setenv("TORCHELASTIC_RUN_ID", "x")
    monkeypatch.setenv("MASTER_PORT", "29500")
    monkeypatch.setenv("RANK", "0")
    monkeypatch.setenv("WORLD_SIZE", "1")
    a = _args(gpus=1, comm="auto", ops="native", mp_mode="ar", fsdp_gather="fused", mp_input="sharded", nvls="auto", pool=3, no_e2e=False, recompute=False)
    rec = bench.also_in_child(a, "ours", "mp", 0)
    assert rec == {"value": 1.0, "metric": "m"}
    assert not any(k.startswith("TORCHELASTIC_") for k in seen["env"])
    assert "--also" in seen["cmd"] and seen["cmd"][seen["cmd"].index("--also") + 1] == "none"
    assert seen["cmd"][seen["cmd"].index("--method") + 1] == "mp"
