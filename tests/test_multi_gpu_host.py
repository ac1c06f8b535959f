"""Host-side logic of the multi-GPU path on CPU: shard ranges and the partial gather over a world_size-2 gloo group."""
import os
import socket
import subprocess
import sys

import pytest

from conftest import ROOT, load_pkg


def test_shard_ranges_partition_every_section():
    pkg = load_pkg()
    for n in (0, 1, 7, 2083343, 2097152):
        for count in (1, 2, 3, 4, 8):
            r = [pkg.shard_range(n, k, count) for k in range(count)]
            assert r[0][0] == 0 and r[-1][1] == n
            assert all(r[i][1] == r[i + 1][0] for i in range(count - 1))
            sizes = [hi - lo for lo, hi in r]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        pkg.shard_range(10, 2, 2)


def test_transform_owners_cover_the_three_polynomials():
    pkg = load_pkg()
    for world in (1, 2, 3, 4, 8):
        owners = [pkg.transform_owner(p, world) for p in range(3)]
        assert all(0 <= o < world for o in owners)
        assert len(set(owners)) == min(world, 3)          # spread over as many ranks as there are (at most three)
        masks = [sum(1 << p for p in range(3) if owners[p] == g) for g in range(world)]
        assert sum(masks) == 7 and all(m & k == 0 for i, m in enumerate(masks) for k in masks[i + 1:])


def test_split_transforms_start_at_six_ranks_and_can_be_switched_off(monkeypatch):
    """from six ranks on a sharded proof splits every coset transform in halves over two ranks (rank 2 * poly + parity);
    G16_SPLIT_TRANSFORMS=0 keeps whole transforms on ranks 0-2"""
    pkg = load_pkg()
    monkeypatch.delenv("G16_SPLIT_TRANSFORMS", raising=False)
    assert [pkg.split_transforms(w) for w in (1, 2, 3, 4, 5, 6, 7, 8, 16)] == [False] * 5 + [True] * 4
    owners = sorted((r // 2, r % 2) for r in range(6))
    assert owners == [(p, t) for p in range(3) for t in range(2)]        # every (polynomial, parity) has exactly one owner
    monkeypatch.setenv("G16_SPLIT_TRANSFORMS", "0")
    assert not pkg.split_transforms(8)


WORKER = r'''
import os, sys
sys.path.insert(0, sys.argv[1]); sys.path.insert(0, os.path.join(sys.argv[1], "tests"))
import torch.distributed as dist
from conftest import load_pkg
pkg = load_pkg()
dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
mine = bytes([(rank * 37 + i) % 251 for i in range(pkg.PARTIAL_BYTES)])
allp = pkg.all_gather_partials(mine)
assert len(allp) == world * pkg.PARTIAL_BYTES
for k in range(world):
    want = bytes([(k * 37 + i) % 251 for i in range(pkg.PARTIAL_BYTES)])
    assert allp[k * pkg.PARTIAL_BYTES:(k + 1) * pkg.PARTIAL_BYTES] == want, k
try:
    pkg.all_gather_partials(b"short")
    raise SystemExit("expected ValueError")
except ValueError:
    pass
dist.barrier()
dist.destroy_process_group()
print("rank", rank, "ok")
'''


def test_partial_gather_gloo_world2(tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(WORKER)
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    procs = []
    for rank in range(2):
        env = dict(os.environ, RANK=str(rank), WORLD_SIZE="2", LOCAL_RANK=str(rank), MASTER_ADDR="127.0.0.1",
                   MASTER_PORT=str(port))
        procs.append(subprocess.Popen([sys.executable, str(script), ROOT], env=env, stdout=subprocess.PIPE,
                                      stderr=subprocess.STDOUT, text=True))
    outs = [p.communicate(timeout=180)[0] for p in procs]
    for rank, (p, o) in enumerate(zip(procs, outs)):
        assert p.returncode == 0, o
        assert "rank %d ok" % rank in o


def test_prove_stream_orders_prefetch_commit_and_prove():
    """host logic of ZKey.prove_stream (no GPU): witness i+1 is handed to the copy stream BEFORE proof i is computed,
    every witness is committed exactly once, nothing is prefetched after the last one"""
    from conftest import load_pkg
    pkg = load_pkg()
    log = []

    class Fake(pkg.ZKey):
        def __init__(self):
            pass

        def __del__(self):
            pass

        def prefetch_witness(self, w):
            log.append(("prefetch", w))

        def commit_witness(self):
            log.append(("commit",))

        def prove_loaded(self, r=None, s=None, style=0):
            log.append(("prove",))
            return ("proof%d" % sum(1 for x in log if x[0] == "prove"), "public")

    out = list(Fake().prove_stream(["w0", "w1", "w2"]))
    assert [p for p, _ in out] == ["proof1", "proof2", "proof3"]
    assert log == [("prefetch", "w0"), ("commit",), ("prefetch", "w1"), ("prove",), ("commit",), ("prefetch", "w2"),
                   ("prove",), ("commit",), ("prove",)]
    log.clear()
    assert list(Fake().prove_stream([])) == [] and log == []
    log.clear()
    assert len(list(Fake().prove_stream(iter(["only"])))) == 1
    assert log == [("prefetch", "only"), ("commit",), ("prove",)]
