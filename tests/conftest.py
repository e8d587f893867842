import json
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run with `pytest -m gpu` on the B200 box)")


@pytest.fixture(scope="session")
def golden():
    g = {}
    for name in ("points", "ref_tests", "ref_outputs"):
        with open(os.path.join(ROOT, "tests", "golden", name + ".json")) as f:
            g[name] = json.load(f)
    runs = {}
    for r in g["ref_outputs"]["runs"]:
        r = dict(r)
        for k in ("amax", "dmax"):
            if r[k] == "inf":
                r[k] = float("inf")
        runs.setdefault(r["tag"], []).append(r)
    g["runs"] = runs
    # the unmodified reference frozen on the instances bench.py times (tests/golden/make_heavy_golden.py)
    with open(os.path.join(ROOT, "tests", "golden", "heavy_outputs.json")) as f:
        heavy = json.load(f)["runs"]
    for r in heavy:
        if r["dmax"] == "inf":
            r["dmax"] = float("inf")
    g["heavy"] = {r["tag"]: r for r in heavy}
    with open(os.path.join(ROOT, "tests", "golden", "bench_n200.json")) as f:
        g["bench_n200"] = json.load(f)
    return g


@pytest.fixture(scope="session")
def built():
    """Everything native, compiled in-tree (nvcc cross-compiles without a GPU)."""
    import __graft_entry__ as ge
    ge.build()
    return ge


@pytest.fixture(scope="session")
def oracle(built):
    from oracle.oraclepy import Oracle
    return Oracle()


@pytest.fixture(scope="session")
def pkg(built):
    return built.load_package()


@pytest.fixture(scope="session")
def solver(pkg):
    # No skip and no fallback: on a GPU box a missing device or library is an error.
    s = pkg.Solver(0)
    yield s
    s.close()
