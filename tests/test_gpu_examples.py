"""The headless ports of the reference's twelve demo scripts (examples/*.py) run end to end on the
GPU with shortened loops; each test checks the quantity the reference's plot showed (log probability rising towards the
generating model's, recovered structure, bound increasing) rather than pixels."""
import importlib
import os
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
EX = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "examples")


def _load(name):
    if EX not in sys.path:
        sys.path.insert(0, EX)
    return importlib.import_module(name)


def test_demo_gaussian_er_gibbs():
    t = _load("demo_gaussian_er_gibbs").demo(seed=1234, N_samples=100, verbose=False)
    assert t["lps"].shape == (101,) and np.all(np.isfinite(t["lps"]))
    assert np.median(t["lps"][50:]) > t["lps"][0] - 5.0


def test_demo_gaussian_er_mf():
    t = _load("demo_gaussian_er_mf").demo(seed=1234, N_iters=20, verbose=False)
    assert np.all(np.isfinite(t["lps"])) and np.all(np.isfinite(t["vlbs"]))
    assert np.allclose(t["vlbs"], t["vlbs"][0])            # conjugate model, fixed network: one update is the optimum


def test_demo_gaussian_sbm_gibbs():
    t = _load("demo_gaussian_sbm_gibbs").demo(seed=2, N_samples=100, verbose=False)
    assert np.all(np.isfinite(t["lps"])) and t["lps"][-1] > t["lps"][0]
    assert t["pair_agreement"] > 0.7


def test_demo_distance_model_gibbs():
    t = _load("demo_distance_model_gibbs").demo(seed=1234, N_samples=300, verbose=False)
    assert np.all(np.isfinite(t["lps"])) and np.mean(t["lps"][-50:]) > t["lps"][0]
    assert 0.00005 <= t["step_sz"] <= 1.0


def test_demo_gaussian_distance_model():
    t = _load("demo_gaussian_distance_model").demo(seed=1234, N_samples=300, verbose=False)
    assert np.all(np.isfinite(t["lls"])) and t["lls"][-1] > t["lls"][0]
    assert t["distance_rmse"][-1] < t["distance_rmse"][0]


def test_demo_eigenmodel_gibbs():
    t = _load("demo_eigenmodel_gibbs").demo(seed=3, N_samples=200, verbose=False)
    assert np.all(np.isfinite(t["lps"])) and np.max(t["lps"][100:]) > t["lps"][0] - 1.0


def test_demo_eigenmodel_mf():
    t = _load("demo_eigenmodel_mf").demo(seed=11223344, N_iters=100, verbose=False)
    assert np.all(np.isfinite(t["vlbs"])) and t["vlbs"][-1] > t["vlbs"][0]


def test_demo_logistic_eigenmodel_gibbs():
    t = _load("demo_logistic_eigenmodel_gibbs").demo(seed=4, N_samples=200, verbose=False)
    assert np.all(np.isfinite(t["lps"])) and t["mu_w"].shape == (2,)


def test_demo_logistic_eigenmodel_mf():
    t = _load("demo_logistic_eigenmodel_mf").demo(seed=5, N_iters=40, verbose=False)
    assert np.all(np.isfinite(t["vlbs"])) and t["vlbs"][-1] > t["vlbs"][0]


def test_demo_model_comparison():
    t = _load("demo_model_comparison").demo(seed=1234, N_samples=60, verbose=False, M=20)
    names = sorted(k[:-5] for k in t if k.endswith("_plls"))
    assert len(names) == 6
    for nm in names:
        assert np.all(np.isfinite(t[nm + "_lps"])) and np.all(np.isfinite(t[nm + "_plls"])), nm
        assert np.mean(t[nm + "_lps"][-20:]) > t[nm + "_lps"][0], nm       # every sampler improves on its prior draw


def test_demo_ais():
    t = _load("demo_ais").demo(seed=6, N_samples=3, B=30, steps_per_B=2, verbose=False)
    for k in ("ais_er", "ais_sbm", "empirical_bayes_er"):
        assert np.isfinite(t[k])
    # the Bayesian ER evidence cannot exceed the likelihood at the empirical-Bayes rho
    assert t["ais_er"] <= t["empirical_bayes_er"] + 1.0


def test_demo_ais_batched():
    """the same demo with the particles co-resident on the device (gk_ais_*): the two drivers estimate the same evidence"""
    mod = _load("demo_ais")
    a = mod.demo(seed=6, N_samples=24, N=32, B=40, steps_per_B=2, verbose=False, batched=True)
    b = mod.demo(seed=6, N_samples=24, N=32, B=40, steps_per_B=2, verbose=False, batched=False)
    for k in ("ais_er", "ais_sbm"):
        assert np.isfinite(a[k]) and np.isfinite(b[k])
    assert a["ais_er"] <= a["empirical_bayes_er"] + 1.0
    se = np.sqrt(a["log_weights_sbm"].var(ddof=1) / 24 + b["log_weights_sbm"].var(ddof=1) / 24)
    assert abs(a["log_weights_sbm"].mean() - b["log_weights_sbm"].mean()) < 4 * se + 1.0
    assert abs(a["ais_er"] - b["ais_er"]) < 3.0


def test_demo_walkthrough():
    t = _load("demo").demo(seed=7, N_iter=30, verbose=False, geweke_iter=400)
    assert np.all(np.isfinite(t["lp_trace"])) and t["lp_trace"][-1] > t["lp_trace"][0]
    assert abs(t["geweke_b00_mean"] - t["geweke_b00_prior_mean"]) < 0.08
    assert np.all(np.abs(t["geweke_pZ0"] - 0.2) < 0.12)
