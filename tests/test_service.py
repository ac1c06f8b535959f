"""Resident-key service (`prover --serve`): the four-argument `prover` command line keeps its
contract (scripts/g16_prove.sh:246-252) when it is a client of the service.
CPU tests: wire protocol, error convention (non-zero exit + message, no output files), no silent CPU fallback.
GPU tests: bytes equal to the in-process prover and to the oracle; the key is opened once for many proofs."""
import os
import subprocess
import time

import pytest

from conftest import ROOT

EXE = os.path.join(ROOT, "zk-proof-of-assets_b200", "prover")


class Service:
    def __init__(self, tmp_path, *extra):
        self.sock = str(tmp_path / "g16.sock")
        self.proc = subprocess.Popen([EXE, "--serve", "--socket", self.sock, *extra], stderr=subprocess.PIPE, text=True)
        for _ in range(200):
            if self.ctl("--ping").returncode == 0:
                return
            assert self.proc.poll() is None, self.proc.stderr.read()
            time.sleep(0.05)
        raise AssertionError("service did not come up")

    def ctl(self, what):
        return subprocess.run([EXE, what, "--socket", self.sock], capture_output=True, text=True)

    def prove(self, *paths, **env):
        e = dict(os.environ, G16_SERVICE=self.sock, **env)
        return subprocess.run([EXE, *map(str, paths)], env=e, capture_output=True, text=True)

    def stop(self):
        if self.proc.poll() is None:
            self.ctl("--shutdown")
            try:
                self.proc.wait(timeout=30)
            except subprocess.TimeoutExpired:
                self.proc.kill()


@pytest.fixture
def service(tmp_path):
    s = Service(tmp_path)
    yield s
    s.stop()


def test_service_protocol_and_error_convention(service, small_fixture, tmp_path):
    r = service.ctl("--ping")
    assert r.returncode == 0 and r.stdout.startswith("OK g16b200")
    # missing key: non-zero exit, message on stderr, no outputs
    p = service.prove(tmp_path / "nope.zkey", tmp_path / "nope.wtns", tmp_path / "p.json", tmp_path / "u.json")
    assert p.returncode != 0 and "cannot open" in p.stderr
    assert not (tmp_path / "p.json").exists() and not (tmp_path / "u.json").exists()
    # malformed key: the library's own error text and status travel through the service
    (tmp_path / "bad.zkey").write_bytes(b"zkeyXXXXXXXXXXXXXXXXXXXX")
    (tmp_path / "w.wtns").write_bytes(bytes(small_fixture["wtns_bytes"]))
    p = service.prove(tmp_path / "bad.zkey", tmp_path / "w.wtns", tmp_path / "p.json", tmp_path / "u.json")
    assert p.returncode != 0 and p.stderr.startswith("prover: ")
    assert not (tmp_path / "p.json").exists()
    st = service.ctl("--stats")
    assert st.returncode == 0 and "proofs=0" in st.stdout and "resident=0" in st.stdout
    # a second service on the same socket refuses to start
    again = subprocess.run([EXE, "--serve", "--socket", service.sock], capture_output=True, text=True, timeout=20)
    assert again.returncode != 0 and "already listens" in again.stderr
    assert service.ctl("--shutdown").returncode == 0
    service.proc.wait(timeout=30)
    assert not os.path.exists(service.sock)
    assert service.ctl("--ping").returncode != 0


def test_client_without_service_still_proves_in_process_or_fails_loudly(small_fixture, tmp_path):
    """G16_SERVICE pointing nowhere: the client says so and runs the in-process CUDA path; without a GPU that path
    fails with a CUDA error -- it never produces a proof some other way."""
    (tmp_path / "c.zkey").write_bytes(bytes(small_fixture["zkey_bytes"]))
    (tmp_path / "w.wtns").write_bytes(bytes(small_fixture["wtns_bytes"]))
    env = dict(os.environ, G16_SERVICE=str(tmp_path / "absent.sock"))
    p = subprocess.run([EXE, str(tmp_path / "c.zkey"), str(tmp_path / "w.wtns"), str(tmp_path / "p.json"),
                        str(tmp_path / "u.json")], env=env, capture_output=True, text=True)
    assert "no service on" in p.stderr
    import torch
    if not torch.cuda.is_available():
        assert p.returncode != 0 and not (tmp_path / "p.json").exists()


@pytest.mark.gpu
def test_service_proofs_equal_in_process_and_oracle(service, small_fixture, tmp_path):
    from oracle import binfile, groth16 as g16
    f = small_fixture
    zp, wp = tmp_path / "c_final.zkey", tmp_path / "witness.wtns"
    zp.write_bytes(bytes(f["zkey_bytes"]))
    wp.write_bytes(bytes(f["wtns_bytes"]))
    want = g16.prove(f["zk"], f["w"], 1111, 2222)
    for i in range(3):
        p = service.prove(zp, wp, tmp_path / ("proof%d.json" % i), tmp_path / ("public%d.json" % i),
                          G16_R="1111", G16_S="2222", G16_VERBOSE="1")
        assert p.returncode == 0, p.stderr
        assert ((tmp_path / ("proof%d.json" % i)).read_text(), (tmp_path / ("public%d.json" % i)).read_text()) == want
        assert ("resident=%d" % (1 if i else 0)) in p.stderr
    st = service.ctl("--stats").stdout
    assert "proofs=3" in st and "opens=1" in st and "resident=1" in st
    # relative paths are resolved by the client, against ITS working directory
    e = dict(os.environ, G16_SERVICE=service.sock, G16_R="1111", G16_S="2222")
    p = subprocess.run([EXE, "c_final.zkey", "witness.wtns", "rel_proof.json", "rel_public.json"], env=e, cwd=tmp_path,
                       capture_output=True, text=True)
    assert p.returncode == 0, p.stderr
    assert (tmp_path / "rel_proof.json").read_text() == want[0]
    # a bad witness fails that request only; the key stays resident and the next proof is fine
    bad = tmp_path / "bad.wtns"
    bad.write_bytes(binfile.write_wtns(f["w"][:-2]))
    p = service.prove(zp, bad, tmp_path / "p2.json", tmp_path / "u2.json")
    assert p.returncode == 3 and "Invalid witness length" in p.stderr and not (tmp_path / "p2.json").exists()
    p = service.prove(zp, wp, tmp_path / "p3.json", tmp_path / "u3.json", G16_STYLE="snarkjs", G16_R="1111", G16_S="2222")
    assert p.returncode == 0 and '"curve": "bn128"' in (tmp_path / "p3.json").read_text()
    # a rewritten key file (new mtime/size) is a different key: it is opened again, not served from the stale copy
    time.sleep(0.01)
    zp.write_bytes(bytes(f["zkey_bytes"]))
    os.utime(zp, ns=(time.time_ns(), time.time_ns() + 10**9))
    p = service.prove(zp, wp, tmp_path / "p4.json", tmp_path / "u4.json", G16_R="1111", G16_S="2222", G16_VERBOSE="1")
    assert p.returncode == 0 and "resident=0" in p.stderr and (tmp_path / "p4.json").read_text() == want[0]


@pytest.mark.gpu
def test_service_concurrent_clients(service, small_fixture, tmp_path):
    """GNU-parallel style: several `prover` clients at once on one key (scripts/full_workflow.sh:552)."""
    from oracle import groth16 as g16
    f = small_fixture
    zp, wp = tmp_path / "c.zkey", tmp_path / "w.wtns"
    zp.write_bytes(bytes(f["zkey_bytes"]))
    wp.write_bytes(bytes(f["wtns_bytes"]))
    want = g16.prove(f["zk"], f["w"], 5, 7)
    e = dict(os.environ, G16_SERVICE=service.sock, G16_R="5", G16_S="7")
    procs = [subprocess.Popen([EXE, str(zp), str(wp), str(tmp_path / ("p%d.json" % i)), str(tmp_path / ("u%d.json" % i))],
                              env=e, stderr=subprocess.PIPE, text=True) for i in range(6)]
    for i, p in enumerate(procs):
        assert p.wait(timeout=120) == 0, p.stderr.read()
        assert ((tmp_path / ("p%d.json" % i)).read_text(), (tmp_path / ("u%d.json" % i)).read_text()) == want
    assert "opens=1" in service.ctl("--stats").stdout
