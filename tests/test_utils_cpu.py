"""Small host-side utilities the benchmark contract depends on."""
from distributed_llm_code_samples_b200.utils.clocks import ClockSampler
from distributed_llm_code_samples_b200.utils.timing import tokens_per_second


def test_clock_summary_reports_median_clock_and_active_throttle_reasons():
    s = ClockSampler()
    s.rows = [
        ["0", "1965", "1965", "410.2", "0x0000000000000000", "Not Active", "Not Active", "Not Active", "Not Active"],
        ["0", "1410", "1965", "998.7", "0x0000000000000004", "Not Active", "Not Active", "Not Active", "Active"],
        ["0", "1380", "1965", "1001.3", "0x0000000000000004", "Not Active", "Not Active", "Not Active", "Active"],
        ["0", "[N/A]", "1965", "1.0", "x", "Not Active", "Not Active", "Not Active", "Not Active"],      # unparsable row is skipped
    ]
    out = s.summary()
    assert out["sm_mhz"] == 1410 and out["sm_max_mhz"] == 1965 and out["samples"] == 3
    assert out["reasons"] == ["sw_power_cap"] and out["power_w_max"] == 1001.3


def test_clock_summary_without_samples_is_explicit():
    assert ClockSampler().summary() == {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}


def test_sampler_survives_a_box_without_nvidia_smi():
    s = ClockSampler().start()          # no GPU / no nvidia-smi here: must not raise, must stop cleanly
    assert s.stop()["samples"] >= 0


def test_tokens_per_second():
    assert tokens_per_second(1000, 4, 2000.0) == 2000.0


def test_benchmark_shapes_take_the_fused_paths_at_every_world_size():
    """The driver runs bench.py at 1, 2, 4 and 8 GPUs: at every multi-GPU size the headline configurations must satisfy
    the fused trainers' shape rules (so the run measures the fused kernels, not a fallback), with sane buffer sizes."""
    import ctypes as C
    from distributed_llm_code_samples_b200.config import TrainConfig
    from distributed_llm_code_samples_b200.ops import native as nv
    from distributed_llm_code_samples_b200.parallel import api
    lib = nv.load()

    class Ops:
        name = "native"

    class Comm:
        def __init__(self, world):
            self.rank, self.world = 0, world

    shapes = {"ddp": (4096, 16384, 32), "fsdp": (4096, 16384, 32), "mp": (8192, 28672, 16)}
    for world in (2, 4, 8):
        for method, (d, f, l) in shapes.items():
            cfg = TrainConfig(batch_size=8, seq_len=2048, layers=l, model_size=d, ffn_dim=f, dtype="bf16", comm="fused", impl="native")
            assert api._fused_available(Ops(), Comm(world), cfg, method), (method, world)
            T = cfg.tokens
            outs = [(f, d), (d, f)] if method != "mp" else [(T, d)]
            mode = {"ddp": 0, "fsdp": 1, "mp": 0}[method]
            cap = max(lib.b200_fused_owned_tiles(M, N, world, mode) for M, N in outs)
            inbox = lib.b200_fused_inbox_bytes(cap, world)
            assert 0 < inbox <= 2 << 30, (method, world, inbox)
            grid = 148
            assert (cap + grid - 1) // grid <= 64, "owned tiles per CTA fit the done-counter scratch"


def _crashing_rank(rank, world, dev):
    import time
    if rank == 1:
        raise ValueError("boom from rank 1")
    time.sleep(120)          # rank 0 'hangs in a collective'


def test_launcher_fails_fast_when_a_peer_crashes():
    """A crashed rank must fail the launch within seconds even though its peer is still blocked (the old
    in-order join waited on rank 0 for up to an hour)."""
    import time
    import pytest
    from distributed_llm_code_samples_b200.parallel.launcher import launch
    t0 = time.monotonic()
    with pytest.raises(RuntimeError) as ei:
        launch(2, _crashing_rank, (), backend="gloo", device="cpu")
    assert time.monotonic() - t0 < 60
    assert "boom from rank 1" in str(ei.value) and "(1, 'exit code 1')" in str(ei.value)
