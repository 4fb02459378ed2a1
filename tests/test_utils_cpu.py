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
