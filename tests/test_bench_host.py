"""bench.py's host-side pieces on a CPU-only box: the module imports, its byte accounting matches SURVEY.md §8d,
the synthetic inputs are deterministic, and the CLI keeps the driver's contract (flags, defaults)."""

import importlib
import sys

import numpy as np


def _bench():
    sys.argv = ["bench.py"]
    return importlib.import_module("bench")


def test_byte_accounting_matches_the_survey():
    b = _bench()
    assert b.lookup_bytes_per_query(4, 4) == 2904      # 1600 + 8 + 1296
    assert b.lookup_bytes_per_query(3, 4) == 1816      # 1024 + 8 + 784
    own = b.f16_accounting(1000, 1.0, 6545.9)
    assert own["own_algorithmic_bytes"] == 1000 * (800 + 8 + 1296)
    assert abs(own["own_frac_of_hbm_peak"] - own["own_hbm_gbs"] / 6545.9) < 1e-12
    assert (b.H, b.W, b.ITERS) == (440, 1024, 32) and b.UNIT == "pairs/s"
    assert b.SWEEP_SIZES == ((46, 128), (55, 128), (68, 120), (90, 160), (136, 240))


def test_synthetic_pairs_are_deterministic_and_in_range():
    b = _bench()
    a1, a2 = b.synthetic_pairs(1, seed=1)
    c1, c2 = b.synthetic_pairs(1, seed=1)
    assert a1.shape == (1, 440, 1024, 3) and a1.dtype == np.float32
    assert np.array_equal(a1, c1) and np.array_equal(a2, c2)
    assert a1.min() >= -1 and a1.max() <= 1 and a2.min() >= -1 and a2.max() <= 1
    d1, _ = b.synthetic_pairs(1, seed=2)
    assert not np.array_equal(a1, d1)


def test_cli_contract(monkeypatch):
    b = _bench()
    seen = {}
    monkeypatch.setattr(b, "run_b200", lambda args: seen.update(vars(args)))
    monkeypatch.setattr(b, "run_reference", lambda args: seen.update(vars(args), ref=True))
    monkeypatch.setattr(sys, "argv", ["bench.py"])
    b.main()
    assert seen["gpus"] == 1 and seen["steps"] == 10 and seen["warmup"] >= 3 and seen["impl"] == "b200" and seen["batch"] == 1
    monkeypatch.setattr(sys, "argv", ["bench.py", "--gpus", "8", "--steps", "5", "--warmup", "1"])
    b.main()
    assert seen["gpus"] == 8 and seen["steps"] == 5 and seen["warmup"] == 3  # at least three warm-up steps
    monkeypatch.setattr(sys, "argv", ["bench.py", "--impl", "reference", "--steps", "2", "--warmup", "1"])
    b.main()
    assert seen.get("ref") and seen["warmup"] == 1


def test_measured_peaks_fallback(tmp_path, monkeypatch):
    b = _bench()
    monkeypatch.setattr(b, "ROOT", str(tmp_path))
    gbs, src, peaks = b.measured_peaks()
    assert gbs == b.FALLBACK_HBM_GBS and "fallback" in src and peaks == {}
