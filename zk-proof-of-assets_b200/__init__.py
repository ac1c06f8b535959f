"""g16-b200: host-side mirror of the reference's prover interface over the C ABI of libg16b200.so.

The reference runs `snarkjs groth16 prove <zkey> <wtns> <proof.json> <public.json>` or rapidsnark's
`prover` with the same four arguments (scripts/g16_prove.sh:246-260) and, from TypeScript,
`snarkjs.groth16.prove(zkeyFileName, wtnsFileName) -> {proof, publicSignals}`
(git_modules/circom-ecdsa/scripts/groupsign_cli.ts:114).  node is not in this image, so the
host side above the C ABI is this Python module (ctypes, no torch types in any signature):

    groth16.prove(zkey_path, wtns_path)            -> (proof dict, publicSignals list)   [snarkjs JS API]
    prove_files(zkey, wtns, proof_json, public_json)                                     [the CLI contract]
    ZKey(path).prove(wtns_bytes)                   -> (proof.json text, public.json text) [resident key]

Everything computes on the GPU through libg16b200.so; there is NO CPU fallback: a missing library
or a missing device raises.  Nothing here imports oracle/.

The directory name contains '-', so import it by path:
    importlib.util.spec_from_file_location("g16b200", "<repo>/zk-proof-of-assets_b200/__init__.py")
(tests/conftest.py, bench.py and __graft_entry__.py do exactly that).
"""
import ctypes
import json
import os

# one hardware queue per stream of a proof (see csrc/prover.cu); must be in the environment before the process creates
# its CUDA context, e.g. before torch touches the device
os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libg16b200.so")
SYNTH_LIB_PATH = os.path.join(_HERE, "libg16synth.so")

G16_OK, G16_ERR_GENERIC, G16_ERR_SHORT_BUFFER, G16_ERR_WITNESS_LEN, G16_ERR_CURVE, G16_ERR_CUDA, G16_ERR_FORMAT = range(7)
STYLE_RAPIDSNARK, STYLE_SNARKJS = 0, 1
PARTIAL_BYTES = 768

# every symbol include/g16_b200.h declares (tests check the built library exports all of them)
ABI_SYMBOLS = [
    "g16_opts_default", "g16_last_error", "g16_version", "g16_zkey_open", "g16_zkey_open_mem", "g16_zkey_close",
    "g16_batch_apply_key", "g16_batch_apply_key_dev", "g16_zkey_info", "g16_prove", "g16_witness_load", "g16_witness_prefetch", "g16_witness_commit", "g16_prove_loaded", "g16_prove_files", "g16_prove_partial",
    "g16_prove_combine", "g16_msm_g1_host", "g16_msm_g2_host", "g16_msm_create", "g16_msm_destroy", "g16_msm_run_dev",
    "g16_msm_stats", "g16_h_scalars", "g16_coset_ntt_dev", "g16_coset_ntt_host", "g16_test_field_op",
    "g16_test_curve_op", "g16_last_timings", "g16_last_msm_stats", "g16_bench_int_pipe", "g16_h_pipeline_dev",
    "g16_last_accumulate_ms", "g16_launch_count", "g16_last_trace", "g16_h_transform", "g16_h_buffer",
    "g16_prove_partial_joined", "g16_prove_begin", "g16_h_transform_half", "g16_prove_partial_joined_halves", "g16_verify", "g16_sanitize",
    "g16_poseidon_hash2", "g16_merkle_build", "g16_merkle_close", "g16_merkle_info", "g16_merkle_root", "g16_merkle_node",
    "g16_merkle_proof", "g16_merkle_find", "g16_merkle_files", "g16_merkle_root_static",
    "g16_zkey_new", "g16_zkey_verify", "g16_wtns_check", "g16_zkey_contribute", "g16_zkey_beacon", "g16_zkey_export_vkey", "g16_test_hash", "g16_test_msm_shape",
]
SYNTH_SYMBOLS = [
    "g16_synth_create", "g16_synth_destroy", "g16_synth_sizes", "g16_synth_write", "g16_synth_expected",
    "g16_synth_vkey", "g16_synth_fixed_base", "g16_synth_last_error", "g16_synth_write_setup_inputs",
]


class G16Error(RuntimeError):
    def __init__(self, code, message):
        super().__init__("g16 error %d: %s" % (code, message))
        self.code = code
        self.message = message


class Opts(ctypes.Structure):
    _fields_ = [
        ("device", ctypes.c_int32), ("style", ctypes.c_int32), ("has_rs", ctypes.c_int32),
        ("r", ctypes.c_uint8 * 32), ("s", ctypes.c_uint8 * 32), ("msm_c", ctypes.c_int32),
        ("shard_rank", ctypes.c_int32), ("shard_count", ctypes.c_int32), ("h_table", ctypes.c_int32),
        ("reserved", ctypes.c_int32 * 7),
    ]


_lib = None
_synth = None


def lib():
    """The CUDA library.  Fails loudly if it has not been built (`make`, or __graft_entry__.build())."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError("libg16b200.so is not built (run `make` in the repo root); there is no CPU fallback")
        L = ctypes.CDLL(LIB_PATH)
        L.g16_last_error.restype = ctypes.c_char_p
        L.g16_version.restype = ctypes.c_char_p
        L.g16_zkey_close.restype = None
        L.g16_msm_destroy.restype = None
        L.g16_opts_default.restype = None
        for name in ("g16_msm_g1_host", "g16_msm_g2_host"):
            getattr(L, name).argtypes = [ctypes.c_char_p, ctypes.c_char_p, ctypes.c_size_t, ctypes.c_void_p, ctypes.c_void_p]
        L.g16_msm_create.argtypes = [ctypes.c_int, ctypes.c_size_t, ctypes.c_void_p, ctypes.c_void_p]
        L.g16_msm_destroy.argtypes = [ctypes.c_void_p]
        L.g16_msm_run_dev.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_size_t, ctypes.c_void_p]
        L.g16_msm_stats.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
        L.g16_zkey_open.argtypes = [ctypes.c_char_p, ctypes.c_void_p, ctypes.c_void_p]
        L.g16_zkey_open_mem.argtypes = [ctypes.c_void_p, ctypes.c_size_t, ctypes.c_void_p, ctypes.c_void_p]
        L.g16_zkey_close.argtypes = [ctypes.c_void_p]
        L.g16_zkey_info.argtypes = [ctypes.c_void_p] * 5
        L.g16_prove.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_size_t, ctypes.c_void_p, ctypes.c_void_p,
                                ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]
        L.g16_witness_load.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_size_t]
        L.g16_witness_prefetch.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_size_t]
        L.g16_witness_commit.argtypes = [ctypes.c_void_p]
        L.g16_prove_loaded.argtypes = [ctypes.c_void_p] * 6
        L.g16_prove_files.argtypes = [ctypes.c_char_p] * 4 + [ctypes.c_void_p]
        L.g16_prove_partial.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
        L.g16_prove_combine.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int] + [ctypes.c_void_p] * 5
        L.g16_h_scalars.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
        L.g16_h_pipeline_dev.argtypes = [ctypes.c_void_p]
        L.g16_coset_ntt_dev.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int]
        L.g16_batch_apply_key.argtypes = [ctypes.c_int, ctypes.c_void_p, ctypes.c_size_t, ctypes.c_char_p, ctypes.c_char_p,
                                          ctypes.c_void_p, ctypes.c_void_p]
        L.g16_batch_apply_key_dev.argtypes = [ctypes.c_int, ctypes.c_void_p, ctypes.c_size_t, ctypes.c_char_p,
                                              ctypes.c_char_p, ctypes.c_void_p]
        L.g16_coset_ntt_host.argtypes = [ctypes.c_void_p, ctypes.c_int]
        L.g16_test_field_op.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_char_p, ctypes.c_char_p, ctypes.c_void_p, ctypes.c_size_t]
        L.g16_test_curve_op.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_char_p, ctypes.c_size_t, ctypes.c_void_p]
        L.g16_last_timings.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
        L.g16_last_msm_stats.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
        L.g16_bench_int_pipe.argtypes = [ctypes.c_int, ctypes.c_void_p]
        L.g16_last_accumulate_ms.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
        L.g16_launch_count.restype = ctypes.c_ulonglong
        L.g16_last_trace.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
        L.g16_h_transform.argtypes = [ctypes.c_void_p, ctypes.c_int]
        L.g16_h_buffer.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]
        L.g16_prove_partial_joined.argtypes = [ctypes.c_void_p] * 5
        L.g16_prove_begin.argtypes = [ctypes.c_void_p]
        L.g16_test_msm_shape.argtypes = [ctypes.c_uint64, ctypes.c_int, ctypes.c_int, ctypes.c_int] + [ctypes.c_void_p] * 4
        L.g16_h_transform_half.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int]
        L.g16_prove_partial_joined_halves.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]
        L.g16_sanitize.argtypes = [ctypes.c_char_p, ctypes.c_size_t, ctypes.c_char_p, ctypes.c_size_t, ctypes.c_char_p,
                                   ctypes.c_size_t, ctypes.c_void_p, ctypes.c_void_p]
        L.g16_verify.argtypes = [ctypes.c_char_p, ctypes.c_size_t, ctypes.c_char_p, ctypes.c_size_t, ctypes.c_char_p,
                                 ctypes.c_size_t, ctypes.c_void_p]
        L.g16_poseidon_hash2.argtypes = [ctypes.c_void_p, ctypes.c_size_t, ctypes.c_void_p, ctypes.c_void_p]
        L.g16_merkle_build.argtypes = [ctypes.c_void_p, ctypes.c_size_t, ctypes.c_void_p, ctypes.c_void_p]
        L.g16_merkle_close.argtypes = [ctypes.c_void_p]
        L.g16_merkle_close.restype = None
        L.g16_merkle_info.argtypes = [ctypes.c_void_p] * 5
        L.g16_merkle_root.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
        L.g16_merkle_node.argtypes = [ctypes.c_void_p, ctypes.c_uint32, ctypes.c_uint64, ctypes.c_void_p]
        L.g16_merkle_proof.argtypes = [ctypes.c_void_p, ctypes.c_uint64, ctypes.c_void_p, ctypes.c_void_p]
        L.g16_merkle_find.argtypes = [ctypes.c_void_p, ctypes.c_char_p, ctypes.c_uint64, ctypes.c_void_p]
        L.g16_merkle_files.argtypes = [ctypes.c_char_p, ctypes.c_char_p, ctypes.c_char_p, ctypes.c_void_p]
        L.g16_merkle_root_static.argtypes = [ctypes.c_void_p, ctypes.c_size_t, ctypes.c_void_p, ctypes.c_void_p]
        L.g16_zkey_new.argtypes = [ctypes.c_char_p, ctypes.c_char_p, ctypes.c_char_p, ctypes.c_void_p, ctypes.c_void_p]
        L.g16_zkey_verify.argtypes = [ctypes.c_char_p, ctypes.c_char_p, ctypes.c_char_p, ctypes.c_void_p]
        L.g16_wtns_check.argtypes = [ctypes.c_char_p, ctypes.c_char_p, ctypes.c_void_p, ctypes.c_void_p]
        L.g16_zkey_contribute.argtypes = [ctypes.c_char_p, ctypes.c_char_p, ctypes.c_char_p, ctypes.c_char_p, ctypes.c_size_t,
                                          ctypes.c_void_p, ctypes.c_void_p]
        L.g16_zkey_beacon.argtypes = [ctypes.c_char_p, ctypes.c_char_p, ctypes.c_char_p, ctypes.c_char_p, ctypes.c_size_t,
                                      ctypes.c_uint32, ctypes.c_void_p, ctypes.c_void_p]
        L.g16_zkey_export_vkey.argtypes = [ctypes.c_char_p, ctypes.c_void_p, ctypes.c_void_p]
        L.g16_test_hash.argtypes = [ctypes.c_int, ctypes.c_char_p, ctypes.c_size_t, ctypes.c_void_p]
        _lib = L
    return _lib


def synth_lib():
    """The SYNTH-1 fixture generator (bench/test support; include/g16_synth.h)."""
    global _synth
    if _synth is None:
        if not os.path.exists(SYNTH_LIB_PATH):
            raise ImportError("libg16synth.so is not built (run `make` in the repo root)")
        S = ctypes.CDLL(SYNTH_LIB_PATH)
        S.g16_synth_last_error.restype = ctypes.c_char_p
        S.g16_synth_destroy.restype = None
        S.g16_synth_destroy.argtypes = [ctypes.c_void_p]
        S.g16_synth_create.argtypes = [ctypes.c_uint32, ctypes.c_uint32, ctypes.c_uint32, ctypes.c_uint64, ctypes.c_uint64, ctypes.c_void_p]
        S.g16_synth_sizes.argtypes = [ctypes.c_void_p] * 4
        S.g16_synth_write.argtypes = [ctypes.c_void_p] * 3
        S.g16_synth_expected.argtypes = [ctypes.c_void_p, ctypes.c_char_p, ctypes.c_char_p, ctypes.c_void_p]
        S.g16_synth_vkey.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
        S.g16_synth_write_setup_inputs.argtypes = [ctypes.c_void_p, ctypes.c_char_p, ctypes.c_char_p]
        S.g16_synth_fixed_base.argtypes = [ctypes.c_int, ctypes.c_char_p, ctypes.c_size_t, ctypes.c_void_p]
        _synth = S
    return _synth


def _check(rc):
    if rc != G16_OK:
        raise G16Error(rc, lib().g16_last_error().decode("utf-8", "replace"))


def make_opts(device=-1, style=STYLE_RAPIDSNARK, r=None, s=None, msm_c=0, shard_rank=0, shard_count=1, h_table=0):
    o = Opts()
    lib().g16_opts_default(ctypes.byref(o))
    o.device, o.style, o.msm_c, o.shard_rank, o.shard_count = device, style, msm_c, shard_rank, shard_count
    o.h_table = h_table
    if (r is None) != (s is None):
        raise ValueError("pin both r and s or neither")
    if r is not None:
        o.has_rs = 1
        o.r[:] = list(int(r).to_bytes(32, "little"))
        o.s[:] = list(int(s).to_bytes(32, "little"))
    return o


def _addr(buf):
    """address of a bytes / bytearray / numpy array / ctypes buffer, without copying"""
    if isinstance(buf, bytes):
        return ctypes.cast(ctypes.c_char_p(buf), ctypes.c_void_p).value
    if isinstance(buf, (bytearray, memoryview)):
        return ctypes.addressof((ctypes.c_char * len(buf)).from_buffer(buf))
    if hasattr(buf, "ctypes"):
        return buf.ctypes.data
    if hasattr(buf, "data_ptr"):  # a (pinned) host tensor used as a plain buffer
        return buf.data_ptr()
    return ctypes.addressof(buf)


class ZKey:
    """A proving key parsed once and kept resident on one GPU (g16_zkey_open).  Mirrors the role of the cached
    zkeys the workflow reuses across batches (scripts/full_workflow.sh:303-348)."""

    def __init__(self, source, device=-1, msm_c=0, shard_rank=0, shard_count=1, h_table=0):
        """h_table: fixed-base window table for the H section (1 = always, -1 = never, 0 = automatic)"""
        self._h = ctypes.c_void_p()
        self._opts = make_opts(device=device, msm_c=msm_c, shard_rank=shard_rank, shard_count=shard_count, h_table=h_table)
        if isinstance(source, (str, os.PathLike)):
            _check(lib().g16_zkey_open(os.fsencode(source), ctypes.byref(self._opts), ctypes.byref(self._h)))
        else:
            _check(lib().g16_zkey_open_mem(_addr(source), len(source), ctypes.byref(self._opts), ctypes.byref(self._h)))
        nv, npub, dom, nc = ctypes.c_uint32(), ctypes.c_uint32(), ctypes.c_uint32(), ctypes.c_uint64()
        _check(lib().g16_zkey_info(self._h, ctypes.byref(nv), ctypes.byref(npub), ctypes.byref(dom), ctypes.byref(nc)))
        self.n_vars, self.n_public, self.domain_size, self.n_coefs = nv.value, npub.value, dom.value, nc.value
        self._proof = ctypes.create_string_buffer(4096)
        self._public = ctypes.create_string_buffer(128 + 80 * max(1, self.n_public))

    def close(self):
        if self._h:
            lib().g16_zkey_close(self._h)
            self._h = ctypes.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _out(self, fn, *args):
        pl, ul = ctypes.c_size_t(len(self._proof)), ctypes.c_size_t(len(self._public))
        _check(fn(*args, self._proof, ctypes.byref(pl), self._public, ctypes.byref(ul)))
        return self._proof.raw[:pl.value].decode(), self._public.raw[:ul.value].decode()

    def prove(self, wtns, r=None, s=None, style=STYLE_RAPIDSNARK):
        """wtns = bytes of a .wtns file (host memory) -> (proof.json text, public.json text)."""
        o = make_opts(style=style, r=r, s=s)
        pl, ul = ctypes.c_size_t(len(self._proof)), ctypes.c_size_t(len(self._public))
        _check(lib().g16_prove(self._h, _addr(wtns), len(wtns), self._proof, ctypes.byref(pl), self._public,
                               ctypes.byref(ul), ctypes.byref(o)))
        return self._proof.raw[:pl.value].decode(), self._public.raw[:ul.value].decode()

    def load_witness(self, wtns):
        _check(lib().g16_witness_load(self._h, _addr(wtns), len(wtns)))

    def prefetch_witness(self, wtns):
        """start the host->device copy of the NEXT proof's witness (asynchronous when `wtns` is pinned host memory; keep
        the buffer alive until commit_witness)"""
        self._next_wtns = wtns
        _check(lib().g16_witness_prefetch(self._h, _addr(wtns), len(wtns)))

    def commit_witness(self):
        """wait for the prefetched copy and make it the witness prove_loaded uses"""
        _check(lib().g16_witness_commit(self._h))
        self._next_wtns = None

    def prove_stream(self, witnesses, r=None, s=None, style=STYLE_RAPIDSNARK):
        """Back-to-back proofs with this key, the batch loop of scripts/full_workflow.sh:552 for one zkey: yields
        (proof, public) per witness while the next witness is already being copied to the device."""
        it = iter(witnesses)
        try:
            self.prefetch_witness(next(it))
        except StopIteration:
            return
        more = True
        while more:
            self.commit_witness()
            try:
                self.prefetch_witness(next(it))
            except StopIteration:
                more = False
            yield self.prove_loaded(r=r, s=s, style=style)

    def prove_loaded(self, r=None, s=None, style=STYLE_RAPIDSNARK):
        o = make_opts(style=style, r=r, s=s)
        pl, ul = ctypes.c_size_t(len(self._proof)), ctypes.c_size_t(len(self._public))
        _check(lib().g16_prove_loaded(self._h, self._proof, ctypes.byref(pl), self._public, ctypes.byref(ul), ctypes.byref(o)))
        return self._proof.raw[:pl.value].decode(), self._public.raw[:ul.value].decode()

    def prove_partial(self):
        """This shard's five partial MSM sums (768 bytes) -- the unit the multi-GPU gather moves."""
        out = ctypes.create_string_buffer(PARTIAL_BYTES)
        _check(lib().g16_prove_partial(self._h, out))
        return out.raw

    def prove_begin(self):
        """start the witness-scalar MSMs of the loaded witness now (asynchronous); the next prove_partial* picks them up"""
        _check(lib().g16_prove_begin(self._h))

    def h_transform(self, which):
        """rows a4-a7 for the polynomials in bit mask `which` (1 A, 2 B, 4 C); results stay in the device buffers"""
        _check(lib().g16_h_transform(self._h, which))

    def h_buffer(self, poly):
        """-> (device pointer of polynomial buffer 0 A / 1 B / 2 C, this shard's H range lo, n)"""
        ptr, lo, n = ctypes.c_void_p(), ctypes.c_uint64(), ctypes.c_uint64()
        _check(lib().g16_h_buffer(self._h, poly, ctypes.byref(ptr), ctypes.byref(lo), ctypes.byref(n)))
        return ptr.value, lo.value, n.value

    def prove_partial_joined(self, d_a, d_b, d_c):
        out = ctypes.create_string_buffer(PARTIAL_BYTES)
        _check(lib().g16_prove_partial_joined(self._h, d_a, d_b, d_c, out))
        return out.raw

    def h_transform_half(self, poly, parity):
        """half of polynomial `poly`'s coset transform (the coefficients of one parity), left twice in that polynomial's buffer"""
        _check(lib().g16_h_transform_half(self._h, poly, parity))

    def prove_partial_joined_halves(self, ptrs):
        """ptrs: six device addresses -- this rank's slices of A even, A odd, B even, B odd, C even, C odd"""
        arr = (ctypes.c_void_p * 6)(*ptrs)
        out = ctypes.create_string_buffer(PARTIAL_BYTES)
        _check(lib().g16_prove_partial_joined_halves(self._h, arr, out))
        return out.raw

    def combine(self, partials, r=None, s=None, style=STYLE_RAPIDSNARK):
        assert len(partials) % PARTIAL_BYTES == 0
        o = make_opts(style=style, r=r, s=s)
        pl, ul = ctypes.c_size_t(len(self._proof)), ctypes.c_size_t(len(self._public))
        _check(lib().g16_prove_combine(self._h, _addr(partials), len(partials) // PARTIAL_BYTES, self._proof,
                                       ctypes.byref(pl), self._public, ctypes.byref(ul), ctypes.byref(o)))
        return self._proof.raw[:pl.value].decode(), self._public.raw[:ul.value].decode()

    def h_scalars(self):
        out = ctypes.create_string_buffer(self.domain_size * 32)
        _check(lib().g16_h_scalars(self._h, out))
        return out.raw

    def h_pipeline_dev(self):
        """Run rows a4-a8 on the device and wait; the scalars stay on the GPU (measurement helper)."""
        _check(lib().g16_h_pipeline_dev(self._h))

    def timings(self):
        t = (ctypes.c_double * 10)()
        _check(lib().g16_last_timings(self._h, t))
        keys = ["h2d_witness", "r1cs_eval", "ntt_pipeline", "msm_A", "msm_B1", "msm_B2", "msm_C", "msm_H", "device_region", "host_final"]
        return dict(zip(keys, list(t)))

    def accumulate_ms(self):
        t = (ctypes.c_double * 5)()
        _check(lib().g16_last_accumulate_ms(self._h, t))
        return dict(zip(["A", "B1", "B2", "C", "H"], list(t)))

    def trace(self):
        """device timeline of the last prove: per MSM, ms after the prove started"""
        t = (ctypes.c_double * 30)()
        _check(lib().g16_last_trace(self._h, t))
        keys = ["start", "digits_done", "sorted", "acc_start", "acc_done", "done"]
        return {n: dict(zip(keys, [round(x, 3) for x in t[6 * i:6 * i + 6]])) for i, n in enumerate(["A", "B1", "B2", "C", "H"])}

    def msm_stats(self):
        t = (ctypes.c_uint64 * 25)()
        _check(lib().g16_last_msm_stats(self._h, t))
        out = {}
        for i, name in enumerate(["A", "B1", "B2", "C", "H"]):
            out[name] = dict(zip(["additions", "tasks", "cut_buckets", "c", "W"], list(t[5 * i:5 * i + 5])))
        return out


class groth16:
    """`snarkjs.groth16` look-alike: prove(zkeyFileName, wtnsFileName) -> (proof, publicSignals)."""

    @staticmethod
    def prove(zkey_file_name, wtns_file_name, r=None, s=None, device=-1):
        with ZKey(zkey_file_name, device=device) as zk, open(wtns_file_name, "rb") as f:
            proof, public = zk.prove(f.read(), r=r, s=s, style=STYLE_SNARKJS)
        return json.loads(proof), json.loads(public)


def verify(vkey, public_signals, proof):
    """`snarkjs.groth16.verify(vkey, publicSignals, proof)` look-alike (scripts/g16_verify.sh:212-216): parsed JSON
    objects or JSON texts in, bool out.  Native host code (csrc/verify*.hpp), no GPU needed."""
    def text(x):
        if isinstance(x, (bytes, bytearray)):
            return bytes(x)
        return (x if isinstance(x, str) else json.dumps(x)).encode()
    v, u, p = text(vkey), text(public_signals), text(proof)
    ok = ctypes.c_int(0)
    _check(lib().g16_verify(v, len(v), u, len(u), p, len(p), ctypes.byref(ok)))
    return bool(ok.value)


def sanitize(vkey, public_signals, proof):
    """scripts/sanitize_groth16_proof.py as one call: -> the text of sanitized_proof.json (Python json.dump layout)"""
    def text(x):
        if isinstance(x, (bytes, bytearray)):
            return bytes(x)
        return (x if isinstance(x, str) else json.dumps(x)).encode()
    v, u, p = text(vkey), text(public_signals), text(proof)
    n = ctypes.c_size_t(1 << 16)
    buf = ctypes.create_string_buffer(n.value)
    rc = lib().g16_sanitize(v, len(v), u, len(u), p, len(p), buf, ctypes.byref(n))
    if rc == G16_ERR_SHORT_BUFFER:
        buf = ctypes.create_string_buffer(n.value + 1)
        n = ctypes.c_size_t(len(buf))
        rc = lib().g16_sanitize(v, len(v), u, len(u), p, len(p), buf, ctypes.byref(n))
    _check(rc)
    return buf.raw[:n.value].decode()


def prove_files(zkey_path, wtns_path, proof_path, public_path, r=None, s=None, style=STYLE_RAPIDSNARK, device=-1):
    """The four-argument command line as one call (scripts/g16_prove.sh:246-252)."""
    o = make_opts(device=device, style=style, r=r, s=s)
    _check(lib().g16_prove_files(os.fsencode(zkey_path), os.fsencode(wtns_path), os.fsencode(proof_path),
                                 os.fsencode(public_path), ctypes.byref(o)))


# ---- operators on the path --------------------------------------------------------------------------------
def msm_host(is_g2, bases, scalars, n, msm_c=0):
    """bases: n zkey records (64 B G1 / 128 B G2, Montgomery affine); scalars: n x 32 B LE standard form.
    -> affine Montgomery record of the sum (all-zero = infinity)."""
    out = ctypes.create_string_buffer(128 if is_g2 else 64)
    o = make_opts(msm_c=msm_c)
    fn = lib().g16_msm_g2_host if is_g2 else lib().g16_msm_g1_host
    _check(fn(bytes(bases), bytes(scalars), n, out, ctypes.byref(o)))
    return out.raw


class Msm:
    """Reusable workspace for device-resident MSMs (g16_msm_create / g16_msm_run_dev)."""

    def __init__(self, is_g2, max_n, msm_c=0, device=-1):
        self._h = ctypes.c_void_p()
        self.is_g2 = is_g2
        o = make_opts(device=device, msm_c=msm_c)
        _check(lib().g16_msm_create(int(is_g2), max_n, ctypes.byref(o), ctypes.byref(self._h)))

    def run(self, d_bases, d_scalars, n):
        out = ctypes.create_string_buffer(128 if self.is_g2 else 64)
        _check(lib().g16_msm_run_dev(self._h, d_bases, d_scalars, n, out))
        return out.raw

    def stats(self):
        t = (ctypes.c_uint64 * 6)()
        _check(lib().g16_msm_stats(self._h, t))
        return dict(zip(["additions", "tasks", "cut_buckets", "buckets", "c", "W"], list(t)))

    def close(self):
        if self._h:
            lib().g16_msm_destroy(self._h)
            self._h = ctypes.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def coset_ntt_host(data, logn):
    """In place on a copy: evaluations on the 2^logn domain -> evaluations on the odd coset (Montgomery Fr)."""
    buf = ctypes.create_string_buffer(bytes(data), len(data))
    _check(lib().g16_coset_ntt_host(buf, logn))
    return buf.raw


def batch_apply_key(is_g2, points, n, first, inc=1, device=-1):
    """out[i] = (first * inc^i) * P_i over n zkey-layout records in host memory -- ffjavascript `G.batchApplyKey`, the kernel
    of `snarkjs zkey contribute` / `beacon` (operator level).  first, inc: ints below r."""
    out = ctypes.create_string_buffer((128 if is_g2 else 64) * n)
    o = make_opts(device=device)
    _check(lib().g16_batch_apply_key(int(is_g2), _addr(points), n, int(first).to_bytes(32, "little"),
                                     int(inc).to_bytes(32, "little"), out, ctypes.byref(o)))
    return out.raw


def batch_apply_key_dev(is_g2, d_points, n, first, inc=1, d_out=None):
    """the same over device-resident records (in place unless d_out is given)"""
    _check(lib().g16_batch_apply_key_dev(int(is_g2), d_points, n, int(first).to_bytes(32, "little"),
                                         int(inc).to_bytes(32, "little"), d_points if d_out is None else d_out))


def coset_ntt_dev(d_ptr, logn, count=1):
    _check(lib().g16_coset_ntt_dev(d_ptr, logn, count))


def field_op(field, op, a, b=None):
    n = len(a) // 32
    out = ctypes.create_string_buffer(len(a))
    _check(lib().g16_test_field_op(field, op, bytes(a), None if b is None else bytes(b), out, n))
    return out.raw


def curve_op(is_g2, op, points, n):
    rec = 128 if is_g2 else 64
    out = ctypes.create_string_buffer(rec * (n if op == 2 else 1))
    _check(lib().g16_test_curve_op(int(is_g2), op, bytes(points), n, out))
    return out.raw


def launch_count():
    return int(lib().g16_launch_count())


def bench_int_pipe(op):
    v = ctypes.c_double()
    _check(lib().g16_bench_int_pipe(op, ctypes.byref(v)))
    return v.value


# ---- Poseidon Merkle tree of the anonymity set (scripts/merkle_tree.rs) -----------------------------------------
def poseidon_hash2(pairs, device=-1):
    """pairs = [(a, b)] ints below r -> [Poseidon(a, b)] (circomlib Poseidon(2)), hashed on the GPU"""
    n = len(pairs)
    buf = b"".join(int(a).to_bytes(32, "little") + int(b).to_bytes(32, "little") for a, b in pairs)
    out = ctypes.create_string_buffer(32 * max(n, 1))
    o = make_opts(device=device)
    _check(lib().g16_poseidon_hash2(_addr(buf) if n else None, n, out, ctypes.byref(o)))
    return [int.from_bytes(out.raw[32 * i:32 * i + 32], "little") for i in range(n)]


class MerkleTree:
    """The anonymity-set tree, built and kept on one GPU.  rows = [(address, balance)] ints, or the packed bytes
    (n x (address[32] | balance[32]), little-endian)."""

    def __init__(self, rows, device=-1):
        self._h = ctypes.c_void_p()
        if not isinstance(rows, (bytes, bytearray, memoryview)) and not hasattr(rows, "ctypes"):
            rows = b"".join(int(a).to_bytes(32, "little") + int(b).to_bytes(32, "little") for a, b in rows)
        n = (rows.nbytes if hasattr(rows, "nbytes") else len(rows)) // 64
        o = make_opts(device=device)
        _check(lib().g16_merkle_build(_addr(rows) if n else None, n, ctypes.byref(o), ctypes.byref(self._h)))
        nr, nl, lv, ms = ctypes.c_uint64(), ctypes.c_uint64(), ctypes.c_uint32(), ctypes.c_double()
        _check(lib().g16_merkle_info(self._h, ctypes.byref(nr), ctypes.byref(nl), ctypes.byref(lv), ctypes.byref(ms)))
        self.n_rows, self.n_leaves, self.levels, self.build_ms = nr.value, nl.value, lv.value, ms.value

    def root(self):
        out = ctypes.create_string_buffer(32)
        _check(lib().g16_merkle_root(self._h, out))
        return int.from_bytes(out.raw, "little")

    def node(self, level, index):
        out = ctypes.create_string_buffer(32)
        _check(lib().g16_merkle_node(self._h, level, index, out))
        return int.from_bytes(out.raw, "little")

    def proof(self, index):
        """-> (path_elements bottom-up, path_indices)"""
        k = self.levels - 1
        el, ix = ctypes.create_string_buffer(32 * max(k, 1)), ctypes.create_string_buffer(max(k, 1))
        _check(lib().g16_merkle_proof(self._h, index, el, ix))
        return [int.from_bytes(el.raw[32 * i:32 * i + 32], "little") for i in range(k)], list(ix.raw[:k])

    def find(self, leaf_hash, start=0):
        """first leaf index >= start holding leaf_hash, or None"""
        idx = ctypes.c_uint64()
        _check(lib().g16_merkle_find(self._h, int(leaf_hash).to_bytes(32, "little"), start, ctypes.byref(idx)))
        return None if idx.value == (1 << 64) - 1 else idx.value

    def close(self):
        if self._h:
            lib().g16_merkle_close(self._h)
            self._h = ctypes.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def merkle_files(anon_set_csv, poa_input_json, output_dir, device=-1):
    """`merkle-tree --anon-set .. --poa-input-data .. --output-dir ..` (scripts/full_workflow.sh:371-381) as one call"""
    o = make_opts(device=device)
    _check(lib().g16_merkle_files(os.fsencode(anon_set_csv), os.fsencode(poa_input_json) if poa_input_json else None,
                                  os.fsencode(output_dir), ctypes.byref(o)))


def msm_shape(n, c=0, table=False, sparse=False):
    """test hook (host arithmetic, no GPU needed): {entries, c, windows, dense} of an MSM of n points, or G16Error beyond
    the index limits Msm::reserve enforces"""
    e, co, w, d = ctypes.c_uint64(), ctypes.c_int(), ctypes.c_int(), ctypes.c_int()
    _check(lib().g16_test_msm_shape(n, c, int(table), int(sparse), ctypes.byref(e), ctypes.byref(co), ctypes.byref(w), ctypes.byref(d)))
    return {"entries": e.value, "c": co.value, "windows": w.value, "dense": bool(d.value)}


MERKLE_STATIC_MAX_ROWS = 65536


def merkle_root_static(d_rows, n_rows, d_root, stream=None):
    """g16_merkle_root_static: the tree's root with every buffer supplied -- d_rows / d_root are DEVICE addresses (ints) on
    the current device, stream a cudaStream_t handle (int) or None.  Allocates nothing, does not synchronise; all calls
    on a GPU share one fixed 4 MiB device array, so overlapping calls race.  At most MERKLE_STATIC_MAX_ROWS rows."""
    _check(lib().g16_merkle_root_static(ctypes.c_void_p(d_rows), n_rows, ctypes.c_void_p(d_root),
                                        ctypes.c_void_p(stream) if stream else None))


# ---- phase-2 ceremony steps + vkey export (scripts/g16_setup.sh:254-295) ------------------------------------------
def zkey_new(r1cs_path, ptau_path, zkey_out, device=-1):
    """`snarkjs zkey new <circuit.r1cs> <pot_final.ptau> <out.zkey>` (`groth16 setup`, scripts/g16_setup.sh:239-254); the
    sparse combinations of the ceremony's Lagrange points run on the GPU.  -> circuit hash"""
    o = make_opts(device=device)
    h = ctypes.create_string_buffer(64)
    _check(lib().g16_zkey_new(os.fsencode(r1cs_path), os.fsencode(ptau_path), os.fsencode(zkey_out), ctypes.byref(o), h))
    return h.raw


def zkey_verify(r1cs_path, ptau_path, zkey_path, device=-1):
    """`snarkjs zkey verify <circuit.r1cs> <pot.ptau> <zkey>` (scripts/g16_verify.sh:190-202): raises G16Error naming the first
    failed check; returns True for a key that is `zkey new` of the circuit plus a valid chain of this library's contributions"""
    o = make_opts(device=device)
    _check(lib().g16_zkey_verify(os.fsencode(r1cs_path), os.fsencode(ptau_path), os.fsencode(zkey_path), ctypes.byref(o)))
    return True


def wtns_check(r1cs_path, wtns_path, device=-1):
    """`snarkjs wtns check <circuit.r1cs> <witness.wtns>` (scripts/g16_verify.sh:205-209) on the GPU.
    -> None when every constraint holds, else the lowest index of a constraint that fails"""
    o = make_opts(device=device)
    bad = ctypes.c_int64(-1)
    _check(lib().g16_wtns_check(os.fsencode(r1cs_path), os.fsencode(wtns_path), ctypes.byref(o), ctypes.byref(bad)))
    return None if bad.value < 0 else bad.value


def zkey_contribute(zkey_in, zkey_out, name="", entropy=None, device=-1):
    """`snarkjs zkey contribute <in> <out> --name=.. -e=..`; sections 8/9 are updated on the GPU.  -> contribution hash"""
    o = make_opts(device=device)
    h = ctypes.create_string_buffer(64)
    e = None if entropy is None else (entropy.encode() if isinstance(entropy, str) else bytes(entropy))
    _check(lib().g16_zkey_contribute(os.fsencode(zkey_in), os.fsencode(zkey_out), name.encode(), e, len(e) if e else 0,
                                     ctypes.byref(o), h))
    return h.raw


def zkey_beacon(zkey_in, zkey_out, beacon_hash, num_iterations_exp, name="", device=-1):
    """`snarkjs zkey beacon <in> <out> <beaconHash hex> <numIterationsExp> -n=..`  -> contribution hash"""
    o = make_opts(device=device)
    h = ctypes.create_string_buffer(64)
    b = bytes.fromhex(beacon_hash) if isinstance(beacon_hash, str) else bytes(beacon_hash)
    _check(lib().g16_zkey_beacon(os.fsencode(zkey_in), os.fsencode(zkey_out), name.encode(), b, len(b), num_iterations_exp,
                                 ctypes.byref(o), h))
    return h.raw


def zkey_export_vkey(zkey_path):
    """`snarkjs zkey export verificationkey <zkey>` -> the text of verification_key.json (host only)"""
    n = ctypes.c_size_t(1 << 16)
    buf = ctypes.create_string_buffer(n.value)
    rc = lib().g16_zkey_export_vkey(os.fsencode(zkey_path), buf, ctypes.byref(n))
    if rc == G16_ERR_SHORT_BUFFER:
        buf = ctypes.create_string_buffer(n.value + 1)
        n = ctypes.c_size_t(len(buf))
        rc = lib().g16_zkey_export_vkey(os.fsencode(zkey_path), buf, ctypes.byref(n))
    _check(rc)
    return buf.raw[:n.value].decode()


def test_hash(which, data):
    out = ctypes.create_string_buffer({0: 64, 1: 32}.get(which, 128))
    _check(lib().g16_test_hash(which, bytes(data), len(data), out))
    return out.raw


# ---- SYNTH-1 fixtures (bench/test support) ------------------------------------------------------------------
class Synth:
    """Deterministic synthetic circuit + known-trapdoor setup (include/g16_synth.h)."""

    def __init__(self, n_vars, n_public, n_constraints, circuit_seed=1, trapdoor_seed=2):
        self._h = ctypes.c_void_p()
        S = synth_lib()
        if S.g16_synth_create(n_vars, n_public, n_constraints, circuit_seed, trapdoor_seed, ctypes.byref(self._h)):
            raise RuntimeError(S.g16_synth_last_error().decode())
        zb, wb, dom = ctypes.c_uint64(), ctypes.c_uint64(), ctypes.c_uint32()
        S.g16_synth_sizes(self._h, ctypes.byref(zb), ctypes.byref(wb), ctypes.byref(dom))
        self.zkey_bytes, self.wtns_bytes, self.domain_size = zb.value, wb.value, dom.value
        self.n_vars, self.n_public, self.n_constraints = n_vars, n_public, n_constraints

    def write(self, want_zkey=True, want_wtns=True):
        """-> (zkey bytearray | None, wtns bytearray | None); the zkey needs the GPU."""
        S = synth_lib()
        zk = bytearray(self.zkey_bytes) if want_zkey else None
        wt = bytearray(self.wtns_bytes) if want_wtns else None
        if S.g16_synth_write(self._h, _addr(zk) if zk is not None else None, _addr(wt) if wt is not None else None):
            raise RuntimeError(S.g16_synth_last_error().decode())
        return zk, wt

    def write_setup_inputs(self, r1cs_path, ptau_path):
        """the circuit as .r1cs + a (sparse) powers-of-tau file under this generator's tau, alpha, beta: `zkey new`'s inputs"""
        if synth_lib().g16_synth_write_setup_inputs(self._h, os.fsencode(r1cs_path), os.fsencode(ptau_path)):
            raise RuntimeError(synth_lib().g16_synth_last_error().decode())

    def expected(self, r, s):
        """closed-form proof for pinned r,s -> (pi_a (x,y), pi_b ((x0,x1),(y0,y1)), pi_c (x,y)) as ints"""
        out = ctypes.create_string_buffer(256)
        if synth_lib().g16_synth_expected(self._h, int(r).to_bytes(32, "little"), int(s).to_bytes(32, "little"), out):
            raise RuntimeError(synth_lib().g16_synth_last_error().decode())
        v = [int.from_bytes(out.raw[32 * i:32 * i + 32], "little") for i in range(8)]
        return (v[0], v[1]), ((v[2], v[3]), (v[4], v[5])), (v[6], v[7])

    def vkey(self):
        """snarkjs-style verification key dict (decimal strings)."""
        n = 64 + 3 * 128 + (self.n_public + 1) * 64
        out = ctypes.create_string_buffer(n)
        if synth_lib().g16_synth_vkey(self._h, out):
            raise RuntimeError(synth_lib().g16_synth_last_error().decode())
        v = [str(int.from_bytes(out.raw[32 * i:32 * i + 32], "little")) for i in range(n // 32)]

        def g2(o):
            return [[v[o], v[o + 1]], [v[o + 2], v[o + 3]], ["1", "0"]]

        ic = [[v[14 + 2 * i], v[15 + 2 * i], "1"] for i in range(self.n_public + 1)]
        return {"protocol": "groth16", "curve": "bn128", "nPublic": self.n_public, "vk_alpha_1": [v[0], v[1], "1"],
                "vk_beta_2": g2(2), "vk_gamma_2": g2(6), "vk_delta_2": g2(10), "IC": ic}

    def close(self):
        if self._h:
            synth_lib().g16_synth_destroy(self._h)
            self._h = ctypes.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def fixed_base(is_g2, scalars, n):
    out = ctypes.create_string_buffer((128 if is_g2 else 64) * n)
    if synth_lib().g16_synth_fixed_base(int(is_g2), bytes(scalars), n, out):
        raise RuntimeError(synth_lib().g16_synth_last_error().decode())
    return out.raw


# ---- multi-GPU host logic ----------------------------------------------------------
def shard_range(n, rank, count):
    """[lo, hi) of a section of n points held by shard `rank` of `count` -- the rule libg16b200 applies at
    g16_zkey_open (contiguous base ranges, sizes differing by at most one)."""
    if not 0 <= rank < count:
        raise ValueError("bad shard rank")
    return n * rank // count, n * (rank + 1) // count


def all_gather_partials(partial, group=None):
    """The one exchange step of a sharded proof: every rank contributes its 768-byte partial (five XYZZ partial
    sums), every rank receives all of them in rank order.  torch.distributed is plumbing here: NCCL on GPUs
    (768 B per rank over NVLink), gloo in the CPU tests."""
    import torch
    import torch.distributed as dist
    if len(partial) != PARTIAL_BYTES:
        raise ValueError("a partial is %d bytes" % PARTIAL_BYTES)
    world = dist.get_world_size(group)
    dev = torch.device("cuda", torch.cuda.current_device()) if dist.get_backend(group) == "nccl" else torch.device("cpu")
    mine = torch.frombuffer(bytearray(partial), dtype=torch.uint8).to(dev)
    out = torch.empty(world * PARTIAL_BYTES, dtype=torch.uint8, device=dev)
    dist.all_gather_into_tensor(out, mine, group=group)
    return out.cpu().numpy().tobytes()


def prove_sharded(zk, wtns, r=None, s=None, style=STYLE_RAPIDSNARK, group=None):
    """One proof over all ranks of `group`: zk was opened with shard_rank = rank, shard_count = world size.
    Every rank returns the same (proof.json, public.json); r, s must be equal on all ranks (pinned, or drawn by
    rank 0 and broadcast by the caller)."""
    zk.load_witness(wtns)
    partials = all_gather_partials(zk.prove_partial(), group)
    return zk.combine(partials, r=r, s=s, style=style)


def transform_owner(poly, world):
    """which rank transforms polynomial 0 A / 1 B / 2 C: spread over the first min(world, 3) ranks"""
    return poly % min(world, 3)


def split_transforms(world):
    """from six ranks on, each polynomial's transform is split in halves over two ranks: rank 2 * poly + parity"""
    return world >= 6 and os.environ.get("G16_SPLIT_TRANSFORMS", "1") != "0"


class _DevMem:
    """a device allocation owned by libg16b200, exposed to torch through __cuda_array_interface__ (no copy)"""

    def __init__(self, ptr, nbytes):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 2}


def exchange_h_slices(zk, group=None):
    """The second exchange step of a sharded proof: every rank receives, from each polynomial's owner, the slice of
    that polynomial's coset evaluations covering its own H range (one all-to-all per polynomial; the owner's input is
    its whole buffer split by the ranks' ranges, everyone else sends nothing).  -> three uint8 device tensors."""
    import torch
    import torch.distributed as dist
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    dev = torch.device("cuda", torch.cuda.current_device())
    n = zk.domain_size
    sizes = [(shard_range(n, g, world)[1] - shard_range(n, g, world)[0]) * 32 for g in range(world)]
    nccl = dist.get_backend(group) == "nccl"
    xdev = dev if nccl else torch.device("cpu")   # CPU tests (gloo): staged through host memory
    empty = torch.empty(0, dtype=torch.uint8, device=xdev)
    out = []
    for poly in range(3):
        owner = transform_owner(poly, world)
        ptr, lo, cnt = zk.h_buffer(poly)
        assert cnt * 32 == sizes[rank] and lo == shard_range(n, rank, world)[0]
        recv = torch.empty(sizes[rank], dtype=torch.uint8, device=xdev)
        if rank == owner:
            dist.all_to_all_single(recv, torch.as_tensor(_DevMem(ptr, n * 32), device=dev).to(xdev),
                                   [sizes[rank] if g == owner else 0 for g in range(world)], sizes, group=group)
        else:
            dist.all_to_all_single(recv, empty, [sizes[rank] if g == owner else 0 for g in range(world)],
                                   [0] * world, group=group)
        out.append(recv.to(dev))
    return out


def exchange_h_halves(zk, group=None):
    """exchange_h_slices for transforms split in halves: ranks 0..5 own (polynomial, parity) = (rank // 2, rank % 2); ONE
    all-to-all moves every owner's buffer, cut by the ranks' H ranges, to everyone.  -> a uint8 device tensor of six
    slices in the order A even, A odd, B even, B odd, C even, C odd."""
    import torch
    import torch.distributed as dist
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    dev = torch.device("cuda", torch.cuda.current_device())
    n = zk.domain_size
    sizes = [(shard_range(n, g, world)[1] - shard_range(n, g, world)[0]) * 32 for g in range(world)]
    xdev = dev if dist.get_backend(group) == "nccl" else torch.device("cpu")
    recv = torch.empty(6 * sizes[rank], dtype=torch.uint8, device=xdev)
    out_split = [sizes[rank] if g < 6 else 0 for g in range(world)]
    if rank < 6:
        ptr, lo, cnt = zk.h_buffer(rank // 2)
        assert cnt * 32 == sizes[rank] and lo == shard_range(n, rank, world)[0]
        dist.all_to_all_single(recv, torch.as_tensor(_DevMem(ptr, n * 32), device=dev).to(xdev), out_split, sizes, group=group)
    else:
        dist.all_to_all_single(recv, torch.empty(0, dtype=torch.uint8, device=xdev), out_split, [0] * world, group=group)
    return recv.to(dev)


def prove_sharded_distributed(zk, wtns=None, r=None, s=None, style=STYLE_RAPIDSNARK, group=None):
    """One proof over all ranks with the three coset transforms spread over the ranks as well: transform the
    polynomials this rank owns (from six ranks on: one HALF of one polynomial's transform), all-to-all the slices, MSMs on
    this rank's base ranges + join on its H range, all-gather the 768-byte partials, combine."""
    import torch
    import torch.distributed as dist
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    if wtns is not None:
        zk.load_witness(wtns)
    zk.prove_begin()      # the witness MSMs depend on neither the transforms nor the exchange: they run underneath both
    if split_transforms(world):
        if rank < 6:
            zk.h_transform_half(rank // 2, rank % 2)
        got = exchange_h_halves(zk, group)
        torch.cuda.current_stream().synchronize()
        step = got.numel() // 6
        partial = zk.prove_partial_joined_halves([got.data_ptr() + i * step for i in range(6)])
        return zk.combine(all_gather_partials(partial, group), r=r, s=s, style=style)
    mask = sum(1 << p for p in range(3) if transform_owner(p, world) == rank)
    zk.h_transform(mask)
    a, b, c = exchange_h_slices(zk, group)
    torch.cuda.current_stream().synchronize()
    partials = all_gather_partials(zk.prove_partial_joined(a.data_ptr(), b.data_ptr(), c.data_ptr()), group)
    return zk.combine(partials, r=r, s=s, style=style)
