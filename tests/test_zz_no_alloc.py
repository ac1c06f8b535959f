"""g16_merkle_root_static: the no-allocation entry point (no counterpart in the reference; added
for callers that cannot call cudaMalloc).  Its output is compared, through the C ABI, with the existing
g16_merkle_build + g16_merkle_root on the same rows -- bit-exact, like every integer path here -- on calls in a row
that share the fixed device array.  Collected last on purpose (file name).

torch is used by the TEST for the caller's device buffers; the entry point itself allocates nothing."""
import random

import pytest

pytestmark = pytest.mark.gpu

R = 21888242871839275222246405745257275088548364400416034343698204186575808495617


def _rows(n, seed):
    rng = random.Random(seed)
    return b"".join(rng.randrange(R).to_bytes(32, "little") + rng.randrange(1 << 90).to_bytes(32, "little") for _ in range(n))


def _static_root(pkg, torch, rows, stream=None):
    n = len(rows) // 64
    d_rows = torch.frombuffer(bytearray(rows), dtype=torch.uint8).cuda()
    d_root = torch.full((32,), 0xEE, dtype=torch.uint8, device="cuda")
    torch.cuda.synchronize()
    pkg.merkle_root_static(d_rows.data_ptr(), n, d_root.data_ptr(), stream)
    torch.cuda.synchronize()
    return int.from_bytes(bytes(d_root.cpu().numpy().tobytes()), "little")


def _built_root(pkg, rows):
    with pkg.MerkleTree(rows) as t:
        return t.root()


def test_static_root_matches_built_tree_on_calls_in_a_row(pkg):
    import torch
    torch.cuda.set_device(0)
    torch.zeros(1, device="cuda")   # the caller's context on the current device
    # large, then small (stale upper levels and leaves of the first call must not show), then ragged, then the maximum
    for n, seed in [(5000, 1), (37, 2), (1, 3), (2, 4), (4097, 5), (pkg.MERKLE_STATIC_MAX_ROWS, 6), (3, 7)]:
        rows = _rows(n, seed)
        assert _static_root(pkg, torch, rows) == _built_root(pkg, rows), n
    # same rows twice in a row: identical
    rows = _rows(1000, 8)
    assert _static_root(pkg, torch, rows) == _static_root(pkg, torch, rows) == _built_root(pkg, rows)
    # on a caller's stream, back to back without a synchronise in between (one stream orders the two calls)
    s = torch.cuda.Stream()
    a, b = _rows(777, 9), _rows(64, 10)
    da = torch.frombuffer(bytearray(a), dtype=torch.uint8).cuda()
    db = torch.frombuffer(bytearray(b), dtype=torch.uint8).cuda()
    out = torch.zeros((2, 32), dtype=torch.uint8, device="cuda")
    torch.cuda.synchronize()
    pkg.merkle_root_static(da.data_ptr(), 777, out[0].data_ptr(), s.cuda_stream)
    pkg.merkle_root_static(db.data_ptr(), 64, out[1].data_ptr(), s.cuda_stream)
    s.synchronize()
    got = out.cpu().numpy()
    assert int.from_bytes(got[0].tobytes(), "little") == _built_root(pkg, a)
    assert int.from_bytes(got[1].tobytes(), "little") == _built_root(pkg, b)


def test_static_root_rejects_what_it_cannot_hold(pkg):
    import torch
    d = torch.zeros(64, dtype=torch.uint8, device="cuda")
    out = torch.zeros(32, dtype=torch.uint8, device="cuda")
    with pytest.raises(pkg.G16Error) as e:
        pkg.merkle_root_static(d.data_ptr(), pkg.MERKLE_STATIC_MAX_ROWS + 1, out.data_ptr())
    assert "exceed" in e.value.message and str(pkg.MERKLE_STATIC_MAX_ROWS) in e.value.message
    with pytest.raises(pkg.G16Error):
        pkg.merkle_root_static(d.data_ptr(), 0, out.data_ptr())
    with pytest.raises(pkg.G16Error):
        pkg.merkle_root_static(0, 1, out.data_ptr())
    with pytest.raises(pkg.G16Error):
        pkg.merkle_root_static(d.data_ptr(), 1, 0)
