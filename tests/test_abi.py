"""The drop-in boundary on a machine without a GPU: the library loads, exports every symbol the headers declare,
and refuses (does not fall back) when asked to compute with no device."""
import ctypes
import os
import re
import subprocess

import pytest

from conftest import ROOT, load_pkg


def _declared(header):
    text = open(os.path.join(ROOT, "include", header)).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(g16_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    pkg = load_pkg()
    L = pkg.lib()
    declared = _declared("g16_b200.h")
    assert sorted(pkg.ABI_SYMBOLS) == declared
    for name in declared:
        assert hasattr(L, name), name
    S = pkg.synth_lib()
    declared = _declared("g16_synth.h")
    assert sorted(pkg.SYNTH_SYMBOLS) == declared
    for name in declared:
        assert hasattr(S, name), name


def test_no_torch_types_in_headers():
    for h in ("g16_b200.h", "g16_synth.h"):
        text = open(os.path.join(ROOT, "include", h)).read()
        assert "torch" not in text and "at::" not in text and "Tensor" not in text


def test_opts_struct_layout():
    pkg = load_pkg()
    o = pkg.make_opts(r=5, s=7, msm_c=13, shard_rank=1, shard_count=4)
    assert ctypes.sizeof(pkg.Opts) == 4 * 3 + 64 + 4 * 3 + 32
    assert (o.device, o.style, o.has_rs, o.msm_c, o.shard_rank, o.shard_count) == (-1, 0, 1, 13, 1, 4)
    assert bytes(o.r)[:2] == b"\x05\x00" and bytes(o.s)[:2] == b"\x07\x00"


def _has_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


@pytest.mark.skipif(_has_gpu(), reason="checks the no-device behaviour")
def test_no_cpu_fallback_without_device(small_fixture):
    pkg = load_pkg()
    with pytest.raises(pkg.G16Error) as e:
        pkg.ZKey(bytes(small_fixture["zkey_bytes"]))
    assert e.value.code == pkg.G16_ERR_CUDA
    with pytest.raises(pkg.G16Error) as e:
        pkg.msm_host(False, bytes(64), bytes(32), 1)
    assert e.value.code == pkg.G16_ERR_CUDA


def test_product_never_touches_the_oracle():
    """Nothing under the package (or the CLI) may import / link / execute oracle/."""
    pk = os.path.join(ROOT, "zk-proof-of-assets_b200")
    for dirpath, _, files in os.walk(pk):
        for fn in files:
            if fn.endswith((".py", ".cu", ".cuh", ".hpp", ".cpp", ".h")):
                text = open(os.path.join(dirpath, fn), errors="replace").read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", text, flags=re.M), fn
                assert not re.search(r'#include\s+"[^"]*oracle/', text), fn


def test_cli_argv_contract():
    exe = os.path.join(ROOT, "zk-proof-of-assets_b200", "prover")
    assert os.path.basename(exe) == "prover"  # scripts/g16_prove.sh:195-199
    p = subprocess.run([exe], capture_output=True, text=True)
    assert p.returncode != 0 and "Usage: prover <circuit.zkey> <witness.wtns> <proof.json> <public.json>" in p.stderr
    p = subprocess.run([exe, "/nonexistent.zkey", "/nonexistent.wtns", "/tmp/p.json", "/tmp/u.json"], capture_output=True, text=True)
    assert p.returncode != 0 and p.stderr.strip()


def test_msm_index_limits_are_enforced_before_any_sort():
    """Msm::reserve's limit arithmetic (host only): entry positions are 32-bit, base indices 31-bit, and cub's radix sort
    on the dense path takes an int item count -- a count above INT_MAX would come back from cub unsorted, i.e. a silently
    wrong proof.  Every case beyond a limit must be refused with 'shard it', every BASELINE shape must pass."""
    pkg = load_pkg()
    INT_MAX = (1 << 31) - 1
    # the reference's shapes: H MSM over a window table (dense scalars) at 2^21 .. 2^26
    for logn in (21, 24, 25, 26):
        sh = pkg.msm_shape(1 << logn, table=True)
        assert sh["entries"] == (1 << logn) * sh["windows"] <= INT_MAX
        assert sh["dense"] == (sh["entries"] > 80 << 20)
    assert not pkg.msm_shape(1 << 21, table=True)["dense"] and pkg.msm_shape(1 << 24, table=True)["dense"]
    # witness MSMs (sparse scalars, no table) never take the radix path
    assert not pkg.msm_shape(1 << 26, sparse=True)["dense"]
    # 2^28 points, dense over a table: c = 20 -> 13 windows -> 3.5e9 entries: above INT_MAX and above 2^31 - 1 base slots
    with pytest.raises(pkg.G16Error) as e:
        pkg.msm_shape(1 << 28, c=20, table=True)
    assert "shard it" in e.value.message
    # the exact boundary of the radix path's int count: n * windows just below / above INT_MAX (c = 16 -> 16 windows)
    w = pkg.msm_shape(1 << 20, c=16, table=True)["windows"]
    ok_n = INT_MAX // w
    assert pkg.msm_shape(ok_n, c=16, table=True)["entries"] == ok_n * w
    with pytest.raises(pkg.G16Error):
        pkg.msm_shape(ok_n + 1, c=16, table=True)
    # without a table (counting sort): 32-bit entry positions are the limit
    w = pkg.msm_shape(1 << 20, c=16)["windows"]
    ok_n = ((1 << 32) - 1) // w
    assert pkg.msm_shape(ok_n, c=16)["entries"] < 1 << 32
    with pytest.raises(pkg.G16Error) as e:
        pkg.msm_shape(ok_n + 1, c=16)
    assert "32-bit entry indices" in e.value.message
    # base indices carry the digit's sign in bit 31
    with pytest.raises(pkg.G16Error) as e:
        pkg.msm_shape(1 << 31, c=24)
    assert "31-bit base indices" in e.value.message


def test_headers_are_plain_c():
    """the boundary is a C ABI: both headers must compile as C99 on their own (what a cgo / N-API / ctypes binding includes)"""
    for h in ("g16_b200.h", "g16_synth.h"):
        r = subprocess.run(["gcc", "-x", "c", "-std=c99", "-Wall", "-Werror", "-fsyntax-only", os.path.join(ROOT, "include", h)],
                           capture_output=True, text=True)
        assert r.returncode == 0, r.stderr
