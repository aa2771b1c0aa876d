import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def oracle():
    from oracle import oracle as O
    return O


@pytest.fixture(scope="session")
def coracle(oracle):
    return oracle.COracle()


@pytest.fixture(scope="session")
def pkg():
    import safe_bayesian_optimization_b200 as p
    return p


def make_ctx(pkg, p, precision=None, kernel="auto", max_obs=None, max_cells=None, options=None, **kw):
    """A context for problem p; `options` = {name: value} applied with sbo_set_option (the A/B switches)."""
    ctx = pkg.SboContext(sigma_f=p.sigma_f, length_scale=p.length, noise_var=p.noise_var, prior_mean=p.prior_mean,
                         precision=precision or p.precision, kernel=kernel, max_obs=max_obs or max(p.n, 1),
                         max_cells=max_cells or p.m, **kw)
    for name, value in (options or {}).items():
        ctx.set_option(name, value)
    return ctx
