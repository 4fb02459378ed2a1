"""CLI parity on CPU (gloo, world 2): the reference's flags, prints and method-0 DDP-vs-FSDP check
(train_ffns.py:343-391), plus checkpoint save / cross-strategy resume and the --profile trace."""
import os

import pytest
import torch

import train_ffns

BASE = ["-s", "4", "-bs", "2", "-n", "8", "-l", "2", "-d", "16", "-r", "3", "--device", "cpu", "--backend", "gloo",
        "--world_size", "2"]


def test_parser_has_reference_flags_and_defaults():
    a = train_ffns.build_parser().parse_args([])
    assert (a.num_steps, a.batch_size, a.seq_len, a.layers, a.model_size, a.method, a.random_seed) == (1, 8, 1024, 1, 4, 0, 0)
    b = train_ffns.build_parser().parse_args(["-s", "2", "-bs", "3", "-n", "5", "-l", "7", "-d", "11", "-m", "4", "-r", "13"])
    assert (b.num_steps, b.batch_size, b.seq_len, b.layers, b.model_size, b.method, b.random_seed) == (2, 3, 5, 7, 11, 4, 13)


@pytest.mark.timeout(600)
def test_method_0_runs_all_four_and_ddp_equals_fsdp(capsys):
    assert train_ffns.main(BASE + ["-m", "0"]) == 0
    out = capsys.readouterr().out
    assert "ARGS:\n num_steps: 4\n BS: 2\n N: 8\n D: 16\n FFN: 64" in out                 # :353
    assert "PARAMS: 4_096 (size" in out                                                  # :366
    assert "initial layers_params[0] torch.Size([64, 16]) torch.Size([16, 64])" in out   # :370
    for name in ("train_1gpu", "train_ddp", "train_fsdp", "train_tp"):                   # :382-384
        assert f"{name} takes" in out and f"final {name} layers_params[0] torch.Size([64, 16])" in out
    assert "SoftAssertionError" not in out                                               # :386-391
    assert "tokens/s" in out


@pytest.mark.timeout(600)
def test_checkpoint_written_under_ddp_resumes_under_fsdp_and_mp(tmp_path, capsys):
    ck = str(tmp_path / "ck.pt")
    assert train_ffns.main(BASE + ["-m", "2", "--save_path", ck]) == 0
    saved = torch.load(ck, weights_only=False)
    assert set(saved) >= {"model", "optimizer", "step", "args", "seeds"} and saved["step"] == 4
    assert saved["model"]["layers.1.ffn2.weight"].shape == (16, 64)
    for m in ("3", "4"):                                                                 # reshard on load
        assert train_ffns.main(BASE + ["-m", m, "--load_path", ck, "-s", "6"]) == 0
    out = capsys.readouterr().out
    assert out.count("resumed from") == 2


@pytest.mark.timeout(600)
def test_profile_flag_exports_chrome_trace(tmp_path, monkeypatch):
    monkeypatch.chdir(tmp_path)
    assert train_ffns.main(["-s", "2", "-bs", "2", "-n", "8", "-l", "1", "-d", "16", "-m", "1", "--device", "cpu", "--profile"]) == 0
    assert os.path.getsize(tmp_path / "trace_profiler_trace.json") > 0                   # :139


@pytest.mark.timeout(600)
def test_save_every_writes_intermediate_full_checkpoints(tmp_path):
    ck = str(tmp_path / "ck.pt")
    assert train_ffns.main(BASE + ["-m", "3", "--save_path", ck, "--save_every", "1"]) == 0      # FSDP: shards gathered
    a = torch.load(ck + ".step2", weights_only=False)
    b = torch.load(ck + ".step4", weights_only=False)
    end = torch.load(ck, weights_only=False)
    assert a["step"] == 2 and b["step"] == 4 and a["model"]["layers.0.ffn1.weight"].shape == (64, 16)
    assert not torch.equal(a["model"]["layers.0.ffn1.weight"], b["model"]["layers.0.ffn1.weight"])
    assert torch.equal(b["model"]["layers.0.ffn1.weight"], end["model"]["layers.0.ffn1.weight"])


def test_bench_contract_without_a_gpu(capsys, monkeypatch):
    """On a box with no CUDA device both arms print one JSON line saying so and exit 0."""
    import json
    import sys
    import bench
    for impl in ("ours", "reference"):
        monkeypatch.setattr(sys, "argv", ["bench.py", "--impl", impl])
        if torch.cuda.is_available():
            pytest.skip("GPU present")
        assert bench.main() == 0
        line = capsys.readouterr().out.strip().splitlines()[-1]
        assert "unavailable" in json.loads(line)
