import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import tamcmc_b200_loader  # noqa: E402


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def tb():
    return tamcmc_b200_loader.load()


@pytest.fixture(scope="session")
def W(tb):
    return tb.workloads


@pytest.fixture(scope="session")
def oracle():
    from oracle.oracle import Oracle
    return Oracle()


@pytest.fixture(scope="session")
def reflib():
    from oracle.oracle import RefLib
    return RefLib()


@pytest.fixture(scope="session")
def golden():
    z = np.load(os.path.join(ROOT, "tests", "golden", "golden_ref.npz"), allow_pickle=False)
    out = {}
    for name in z["names"]:
        name = str(name)
        out[name] = {k.split("/", 1)[1]: z[k] for k in z.files if k.startswith(name + "/")}
    return out


class GoldenProblem:
    """Rebuilds a workloads.Problem from a golden entry."""

    def __init__(self, W, e):
        sig = e["sigma_y"]
        self.prob = W.Problem(name="golden", model_case=int(e["model_case"]), like_case=int(e["like_case"]),
                              prior_case=int(e["prior_case"]), plength=e["plength"].astype(np.int32),
                              params=e["params"].astype(np.float64), relax=e["relax"].astype(np.int32),
                              priors=e["priors"].astype(np.float64), priors_switch=e["priors_switch"].astype(np.int32),
                              extra_priors=e["extra_priors"].astype(np.float64), x=e["x"], y=e["y"],
                              sigma_y=(sig if sig.size else None), likelihood_params=float(e["likelihood_params"]))
        self.T = float(e["T"])
        self.ref_model, self.ref_logL, self.ref_logPrior = e["ref_model"], float(e["ref_logL"]), float(e["ref_logPrior"])


@pytest.fixture(scope="session")
def golden_problems(W, golden):
    return {k: GoldenProblem(W, e) for k, e in golden.items()}


def relerr(a, b):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    return float(np.max(np.abs(a - b) / np.maximum(np.abs(b), 1e-300)))
