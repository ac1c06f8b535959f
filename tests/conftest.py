"""Shared fixtures.  `-m "not gpu"` = oracle vs the reference's golden vectors, host logic, ABI surface;
`-m gpu` = parity of the CUDA path (through the C ABI) against the oracle."""
import importlib.util
import json
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def load_pkg():
    """The package directory is named `zk-proof-of-assets_b200` (not an identifier): import it by path."""
    if "g16b200" in sys.modules:
        return sys.modules["g16b200"]
    path = os.path.join(ROOT, "zk-proof-of-assets_b200", "__init__.py")
    spec = importlib.util.spec_from_file_location("g16b200", path)
    mod = importlib.util.module_from_spec(spec)
    sys.modules["g16b200"] = mod
    spec.loader.exec_module(mod)
    return mod


@pytest.fixture(scope="session")
def pkg():
    return load_pkg()


@pytest.fixture(scope="session")
def golden():
    with open(os.path.join(ROOT, "tests", "golden", "reference_vectors.json")) as f:
        return json.load(f)


@pytest.fixture(scope="session")
def small_fixture():
    """Python-oracle SYNTH-1 circuit + known-trapdoor setup, small enough for big-int MSMs."""
    from oracle import binfile, factory
    circ = factory.SynthCircuit(n_vars=85, n_public=2, n_constraints=100, seed=0xB2000001)
    td = factory.Trapdoor(0xB2000002)
    zk, vkey = factory.setup(circ, td)
    w = circ.witness()
    assert circ.check_witness(w)
    return {"circ": circ, "td": td, "zk": zk, "vkey": vkey, "w": w,
            "zkey_bytes": binfile.write_zkey(zk), "wtns_bytes": binfile.write_wtns(w)}
