/*
 * g16_b200.h -- C ABI of the B200-native Groth16/BN254 prover (libg16b200.so).
 *
 * Drop-in boundary for the `groth16 prove` step of silversixpence-crypto/zk-proof-of-assets:
 *   scripts/g16_prove.sh:246-252   "$rapidsnark_path" <zkey> <wtns> <proof.json> <public.json>
 *   scripts/g16_prove.sh:253-259   npx snarkjs groth16 prove <zkey> <wtns> <proof.json> <public.json>
 *   git_modules/circom-ecdsa/scripts/groupsign_cli.ts:114   snarkjs.groth16.prove(zkeyFileName, wtnsFileName)
 * The reference holds no prover source; the upstream homes of each replaced function are
 * cited per entry point.  Plain pointers and sizes only; every function returns a g16_status.
 *
 * Threading: distinct g16_zkey handles may be used from distinct threads; one prove in flight per handle.
 * All byte strings are little-endian; field elements are 32 bytes; "standard form" = not Montgomery.
 */
#ifndef G16_B200_H
#define G16_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum {
    G16_OK = 0,
    G16_ERR_GENERIC = 1,       /* message in g16_last_error() */
    G16_ERR_SHORT_BUFFER = 2,  /* required sizes written back to *proof_len / *public_len */
    G16_ERR_WITNESS_LEN = 3,   /* "Invalid witness length. Circuit: N, witness: M" (snarkjs groth16_prove.js) */
    G16_ERR_CURVE = 4,         /* "Curve of the witness does not match the curve of the proving key" */
    G16_ERR_CUDA = 5,          /* no device / CUDA failure: there is no CPU fallback */
    G16_ERR_FORMAT = 6         /* malformed .zkey / .wtns container */
} g16_status;

enum { G16_STYLE_RAPIDSNARK = 0, G16_STYLE_SNARKJS = 1 };

typedef struct g16_zkey g16_zkey; /* proving key parsed once and kept resident on one GPU */

typedef struct {
    int32_t device;      /* CUDA ordinal; -1 = current device */
    int32_t style;       /* output byte layout: G16_STYLE_RAPIDSNARK (what full_workflow.sh produces) or _SNARKJS */
    int32_t has_rs;      /* 1 = use r,s below (TEST HOOK; neither upstream prover can pin them) */
    uint8_t r[32];       /* blinding factors, standard form */
    uint8_t s[32];
    int32_t msm_c;       /* Pippenger window bits, 0 = automatic */
    int32_t shard_rank;  /* base-range sharding of every MSM across GPUs: this handle holds slice rank/count */
    int32_t shard_count; /* 0 or 1 = whole key */
    int32_t h_table;     /* fixed-base window table for the H section (rows 2^(c w) * H_i, built at open):
                            0 = automatic, 1 = always, -1 = never */
    int32_t reserved[7];
} g16_opts;

void g16_opts_default(g16_opts* o);
const char* g16_last_error(void); /* thread-local, valid until the next call on this thread */
const char* g16_version(void);

/* ---- proving key: replaces snarkjs zkey_utils.js readHeader + section reads / rapidsnark ZKeyUtils::loadHeader
 *      (rows a1, a2).  Uploads sections 4-9 to the GPU (only the shard's slice of 5-9). */
int g16_zkey_open(const char* path, const g16_opts* opts, g16_zkey** out);
int g16_zkey_open_mem(const uint8_t* data, size_t len, const g16_opts* opts, g16_zkey** out);
void g16_zkey_close(g16_zkey* zk);
int g16_zkey_info(const g16_zkey* zk, uint32_t* n_vars, uint32_t* n_public, uint32_t* domain_size, uint64_t* n_coefs);

/* ---- prove: replaces snarkjs groth16_prove.js groth16Prove / rapidsnark Groth16::Prover::prove (rows a3-a16).
 * wtns = the bytes of a .wtns file.  On success the JSON texts (no trailing NUL counted) are in the caller's
 * buffers and their lengths in *proof_len / *public_len (in: capacity, out: length). */
int g16_prove(g16_zkey* zk, const uint8_t* wtns, size_t wtns_len, char* proof_json, size_t* proof_len,
              char* public_json, size_t* public_len, const g16_opts* opts);

/* same, split so that a caller can keep the witness resident: load = parse + host->device copy */
int g16_witness_load(g16_zkey* zk, const uint8_t* wtns, size_t wtns_len);
int g16_prove_loaded(g16_zkey* zk, char* proof_json, size_t* proof_len, char* public_json, size_t* public_len,
                     const g16_opts* opts);

/* Double-buffered witness for back-to-back proofs with one key (the batch loop of scripts/full_workflow.sh:503,534,552
 * calls the prover once per batch with the same zkey): prefetch starts the host->device copy of the NEXT witness on its
 * own stream and returns (asynchronous when `wtns` is page-locked; the caller keeps the buffer alive until commit);
 * commit waits for that copy and makes it the witness g16_prove_loaded / g16_prove_partial use. */
int g16_witness_prefetch(g16_zkey* zk, const uint8_t* wtns, size_t wtns_len);
int g16_witness_commit(g16_zkey* zk);

/* the rapidsnark / snarkjs command line as one call: reads both files, writes both outputs */
int g16_prove_files(const char* zkey_path, const char* wtns_path, const char* proof_path, const char* public_path,
                    const g16_opts* opts);

/* ---- multi-GPU: every rank holds a base-range shard (opts.shard_rank/shard_count at open), computes its partial
 * sums of the five MSMs, and one rank combines.  A partial is G16_PARTIAL_BYTES: XYZZ Montgomery points
 * A, B1, C, H (4 x 128 B) then B2 (256 B).  The exchange (NCCL all_gather of 768 B) belongs to the caller. */
#define G16_PARTIAL_BYTES 768
int g16_prove_partial(g16_zkey* zk, uint8_t* partial_out);
int g16_prove_combine(g16_zkey* zk, const uint8_t* partials, int count, char* proof_json, size_t* proof_len,
                      char* public_json, size_t* public_len, const g16_opts* opts);

/* Start the four witness-scalar MSMs (A, B1, B2, C) of the loaded witness now and return without waiting.  The next
 * g16_prove_partial / g16_prove_partial_joined / g16_prove_loaded on this handle picks them up instead of starting its
 * own.  A sharded proof calls this before its transforms and the exchange of polynomial slices, which those MSMs do
 * not depend on, so that they run underneath both. */
int g16_prove_begin(g16_zkey* zk);
/* Transforms spread over the ranks (SURVEY.md 8(e) task placement): with G >= 2 ranks the three coset transforms
 * (rows a5-a7 for A, B, C) need not be repeated on every rank.  Rank o runs g16_h_transform(zk, mask of the
 * polynomials it owns: 1 = A, 2 = B, 4 = C): R1CS evaluation (row a4, cheap, every rank) + the odd-coset transform of
 * those polynomials, left in the handle's device buffers (domain_size x 32 B each, Montgomery).  The caller then moves
 * each polynomial's slice [h_lo, h_lo + h_n) from its owner to every rank (one NCCL all-to-all per polynomial;
 * g16_h_buffer gives the device pointers and this shard's H range) and calls g16_prove_partial_joined with device
 * pointers to its three received slices (h_n x 32 B each): the join (row a8) runs on the range only. */
int g16_h_transform(g16_zkey* zk, int which);
int g16_h_buffer(g16_zkey* zk, int poly /*0 A, 1 B, 2 C*/, void** d_ptr, uint64_t* h_lo, uint64_t* h_n);
int g16_prove_partial_joined(g16_zkey* zk, const void* d_a, const void* d_b, const void* d_c, uint8_t* partial_out);
/* With G >= 6 ranks each transform is split over TWO ranks (one polynomial's transform on one rank bounds a sharded
 * proof otherwise): X[k] = Y0[k mod n/2] + inc w^k Y1[k mod n/2], where Yt is the odd-coset transform, on the domain of
 * half the size, of the polynomial's coefficients of parity t (ffjavascript's engine_fft.js does the same radix-2 split
 * across its worker threads).  g16_h_transform_half(zk, poly, parity) leaves Yt (inc folded in for t = 1) TWICE in
 * polynomial `poly`'s buffer, so that slices [h_lo, h_lo + h_n) cut exactly as from a whole transform;
 * g16_prove_partial_joined_halves takes this rank's six slices, in the order A even, A odd, B even, B odd, C even,
 * C odd, and recombines the halves inside the join. */
int g16_h_transform_half(g16_zkey* zk, int poly /*0 A, 1 B, 2 C*/, int parity);
int g16_prove_partial_joined_halves(g16_zkey* zk, const void* const d_halves[6], uint8_t* partial_out);

/* ---- verification: replaces `snarkjs groth16 verify <vkey.json> <public.json> <proof.json>`
 * (scripts/g16_verify.sh:212-216), the consumer of the prover's output.  Host code only -- works without a GPU.
 * Inputs are the JSON texts; *ok = 1 iff e(-A,B) e(alpha,beta) e(vk_x,gamma) e(C,delta) == 1 and every point is on
 * its curve.  Returns G16_OK, or G16_ERR_FORMAT (message in g16_last_error) for malformed JSON. */
int g16_verify(const char* vkey_json, size_t vkey_len, const char* public_json, size_t public_len,
               const char* proof_json, size_t proof_len, int* ok);

/* proof + vkey + public -> the text of sanitized_proof.json, the 43-bit x 6 limb input of the recursive verifier
 * circuit incl. e(-alpha, beta): replaces scripts/sanitize_groth16_proof.py:39-123.  Host code only.  *out_len: in =
 * capacity, out = length (G16_ERR_SHORT_BUFFER writes the needed size back). */
int g16_sanitize(const char* vkey_json, size_t vkey_len, const char* public_json, size_t public_len,
                 const char* proof_json, size_t proof_len, char* out, size_t* out_len);

/* ---- operators on the path, usable on their own (metric: BN254 MSM Mpts/s; NTT GB/s) ------------------------
 * MSM: bases = zkey records (affine, Montgomery, all-zero = infinity; 64 B G1 / 128 B G2), scalars = 32 B standard
 * form.  *_host take host pointers (copies inside), *_dev take device pointers.  Result: affine, Montgomery,
 * zkey record layout (all-zero = infinity).  Replaces ffjavascript engine_multiexp.js multiExpAffine (rows a9-a13). */
int g16_msm_g1_host(const uint8_t* bases, const uint8_t* scalars, size_t n, uint8_t out[64], const g16_opts* opts);
int g16_msm_g2_host(const uint8_t* bases, const uint8_t* scalars, size_t n, uint8_t out[128], const g16_opts* opts);

typedef struct g16_msm g16_msm; /* reusable workspace for device-resident MSMs */
int g16_msm_create(int is_g2, size_t max_n, const g16_opts* opts, g16_msm** out);
void g16_msm_destroy(g16_msm* m);
int g16_msm_run_dev(g16_msm* m, const void* d_bases, const void* d_scalars, size_t n, uint8_t* out_affine);
/* counters of the last run: [0] non-zero digits (= mixed additions), [1] tasks, [2] cut buckets, [3] buckets,
 * [4] window bits, [5] windows */
int g16_msm_stats(const g16_msm* m, uint64_t out[6]);

/* H-polynomial pipeline on its own (rows a4-a8): witness already loaded; out = domain_size x 32 B standard form
 * (the scalars of the H MSM), host buffer. */
int g16_h_scalars(g16_zkey* zk, uint8_t* out);
/* same pipeline, result left on the device (measurement helper: no device->host copy in the call) */
int g16_h_pipeline_dev(g16_zkey* zk);

/* out[i] = (first * inc^i) * P_i, i < n -- ffjavascript 0.2.62 engine_applykey.js `G.batchApplyKey(buff, first, inc)`:
 * the kernel of `snarkjs zkey contribute` / `zkey beacon` (sections 8 and 9 times 1/delta'; reference call site
 * scripts/g16_setup.sh:240-283) and of the powers-of-tau steps.  SURVEY.md 8(f) rank 2 at operator level: the
 * ceremony bookkeeping (transcript hashes, section 10) stays with snarkjs.  points/out: n records in the zkey layout
 * (64 B G1 / 128 B G2, affine, Montgomery, all-zero = infinity); first, inc: 32 B little-endian, standard form, < r.
 * The _dev form takes device pointers (out may equal points). */
int g16_batch_apply_key(int is_g2, const uint8_t* points, size_t n, const uint8_t first[32], const uint8_t inc[32],
                        uint8_t* out, const g16_opts* opts);
int g16_batch_apply_key_dev(int is_g2, const void* d_points, size_t n, const uint8_t first[32], const uint8_t inc[32],
                            void* d_out);

/* odd-coset transform of `count` device-resident vectors of 2^logn Montgomery Fr elements, in place:
 * iNTT, multiply by inc^i, NTT (rows a5-a7).  d_data: count x 2^logn x 32 B. */
int g16_coset_ntt_dev(void* d_data, int logn, int count);
int g16_coset_ntt_host(uint8_t* data, int logn);

/* ---- field / curve micro-operators (parity tests against the oracle; n elements, host buffers) ------------- */
int g16_test_field_op(int field /*0 Fq, 1 Fr*/, int op /*0 mul, 1 add, 2 sub, 3 to_mont, 4 from_mont, 5 inv*/,
                      const uint8_t* a, const uint8_t* b, uint8_t* out, size_t n);
/* op 0: XYZZ acc (identity) += each affine point in turn via mixed add, over n points -> affine sum
 * op 1: tree of full XYZZ additions -> affine sum;  op 2: 2*P per point (out n affine points) */
int g16_test_curve_op(int is_g2, int op, const uint8_t* points, size_t n, uint8_t* out);

/* ---- measurement helpers ---------------------------------------------------------------------------------- */
/* per-phase device times of the last prove on this handle, milliseconds (CUDA events):
 * [0] h2d witness [1] r1cs eval [2] ntt pipeline [3] msm A [4] msm B1 [5] msm B2 [6] msm C [7] msm H
 * [8] whole device region [9] host finalisation */
int g16_last_timings(const g16_zkey* zk, double out_ms[10]);
/* counters of the last prove: per MSM (A,B1,B2,C,H) x {mixed additions, tasks, cut buckets, c, W} */
int g16_last_msm_stats(const g16_zkey* zk, uint64_t out[25]);
/* duration of the bucket-accumulation kernel (the dominant kernel) of each MSM of the last prove, milliseconds,
 * CUDA events on the stream the kernel ran on: A, B1, B2, C, H */
int g16_last_accumulate_ms(const g16_zkey* zk, double out[5]);
/* device timeline of the last prove, milliseconds after its start, per MSM (A, B1, B2, C, H) x { MSM starts, digits
 * done, sort + task counts done, accumulate starts, accumulate done, window sums on the host }; -1 = not reached */
int g16_last_trace(const g16_zkey* zk, double out[30]);
/* cumulative number of this library's own kernel launches in this process (library kernels such as cub's are
 * not counted) */
unsigned long long g16_launch_count(void);
/* integer-pipe microbenchmark: op 0 = 32-bit IMAD, 1 = IMAD.WIDE.U32 on ONE carry chain per thread, 2 = dependent
 * Montgomery products, 3 = IMAD.WIDE.U32 on FOUR independent carry chains per thread (the roofline denominator).
 * Returns achieved ops per second over the whole device (op 2: field products per second). */
int g16_bench_int_pipe(int op, double* ops_per_s);


/* ---- Poseidon Merkle tree of the anonymity set ------------------------------------------
 * Replaces the reference's scripts/merkle_tree.rs (`merkle-tree` binary run by scripts/full_workflow.sh:371-381):
 * hasher = light_poseidon Poseidon::<Fr>::new_circom(2) (merkle_tree.rs:150-181, the circomlib Poseidon(2) of
 * circuits/merkle.circom:39), leaf = hash(address, balance) + zero leaves up to a power of two (:206-269), proofs =
 * siblings bottom-up with the index bits (:290-396).  Field elements are 32-byte little-endian STANDARD form and must
 * be below r.  All hashing runs on the GPU; there is no CPU fallback. */
typedef struct g16_merkle g16_merkle;
/* out[i] = Poseidon(pairs[2i], pairs[2i+1]) for n pairs of host memory */
int g16_poseidon_hash2(const uint8_t* pairs, size_t n, uint8_t* out, const g16_opts* opts);
/* rows: n_rows x (address[32] | balance[32]); builds the whole tree and keeps it on the device */
int g16_merkle_build(const uint8_t* rows, size_t n_rows, const g16_opts* opts, g16_merkle** out);
void g16_merkle_close(g16_merkle* t);
/* levels = node levels including leaves and root (the number merkle_tree.rs prints as "height") */
int g16_merkle_info(const g16_merkle* t, uint64_t* n_rows, uint64_t* n_leaves, uint32_t* levels, double* build_ms);
int g16_merkle_root(const g16_merkle* t, uint8_t out[32]);
int g16_merkle_node(const g16_merkle* t, uint32_t level, uint64_t index, uint8_t out[32]);
/* path_elements: (levels-1) x 32 bytes, path_indices: levels-1 bytes (0 = the node is a left child) */
int g16_merkle_proof(const g16_merkle* t, uint64_t index, uint8_t* path_elements, uint8_t* path_indices);
/* first leaf index >= from whose hash equals `hash`, UINT64_MAX if none (the reference's forward scan, :315-335) */
int g16_merkle_find(const g16_merkle* t, const uint8_t hash[32], uint64_t from, uint64_t* index);
/* the whole `merkle-tree --anon-set <csv> --poa-input-data <json> --output-dir <dir>` command: writes
 * <dir>/merkle_root.json and <dir>/merkle_proofs.json in the reference's byte layout (scripts/README.md:633-673).
 * poa_input_json may be NULL (root only, empty proofs file). */
int g16_merkle_files(const char* anon_set_csv, const char* poa_input_json, const char* output_dir, const g16_opts* opts);
/* NO COUNTERPART IN THE REFERENCE: added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  The root of the same tree as
 * g16_merkle_build + g16_merkle_root, with every buffer supplied: d_rows = n_rows x (address[32] | balance[32]) in DEVICE
 * memory (standard form, below r: NOT checked here, the rows never visit the host), d_root = 32 bytes of DEVICE memory,
 * stream = a cudaStream_t (NULL = the default stream) of the CURRENT device.  The node levels live in a file-scope
 * `static __device__` array of 2 * G16_MERKLE_STATIC_MAX_ROWS * 32 bytes = 4 MiB (csrc/poseidon.cu), which every GPU
 * that loads the library holds until exit.  n_rows > G16_MERKLE_STATIC_MAX_ROWS is rejected (G16_ERR_GENERIC, message
 * in g16_last_error).  Calls no cudaMalloc/cudaFree or allocating library routine and does not synchronise: the result
 * is in d_root once `stream` reaches this point.  Every call on a GPU shares the array: calls that overlap on the GPU
 * (different streams, threads or processes sharing a context) RACE -- order them on one stream or with events.  A
 * device's first call uploads the Poseidon constants (cudaMemcpyToSymbol): make it outside graph capture. */
#define G16_MERKLE_STATIC_MAX_ROWS 65536
int g16_merkle_root_static(const void* d_rows, size_t n_rows, void* d_root, void* stream);


/* ---- phase-2 ceremony steps and verification-key export (file level) ----------------------
 * What scripts/g16_setup.sh:254-295 runs through snarkjs: `zkey contribute`, `zkey beacon`, `zkey export
 * verificationkey`.  A contribution multiplies delta by a secret delta' and sections 8 (C) and 9 (H) by 1/delta' (on
 * the GPU: g16_batch_apply_key), copies the other sections, and appends a section-10 record in snarkjs' layout
 * (deltaAfter, g1_s, g1_sx, g2_spx, Blake2b-512 transcript, name / beacon parameters).  The secrets and the
 * proof-of-knowledge point are derived with this library's own Blake2b construction (oracle/setup.py is the
 * specification); acceptance of that record by `snarkjs zkey verify` is unverified on this box (no node).
 * entropy may be NULL (64 bytes are drawn with getrandom(2)).  contribution_hash (optional) receives the Blake2b-512
 * hash snarkjs prints as "Contribution Hash". */
/* `snarkjs zkey new <circuit.r1cs> <pot_final.ptau> <circuit_0000.zkey>` (`groth16 setup`, scripts/g16_setup.sh:239-254):
 * the whole proving key -- header, IC, coefficients (x R^2), A, B1, B2, C, H, empty contributions section with the
 * circuit hash -- from a circuit and a ceremony file PREPARED FOR PHASE 2 (sections 12-15, as `snarkjs powersoftau
 * prepare phase2` writes them); gamma = delta = 1.  The sparse linear combinations of the ceremony's Lagrange points run
 * on the GPU (csrc/zkey_new.cu); only the needed ranges of the ceremony file are read.  cs_hash_out (optional) receives
 * the circuit hash (UNPINNED: restated from the published source, no zkey in the reference to pin it on).  Errors:
 * G16_ERR_CURVE for a circuit or ceremony over another field, G16_ERR_GENERIC "circuit too big for this power of tau
 * ceremony", G16_ERR_FORMAT for malformed files or a ceremony file that is not prepared. */
int g16_zkey_new(const char* r1cs_path, const char* ptau_path, const char* zkey_out, const g16_opts* opts,
                 uint8_t cs_hash_out[64]);
/* `snarkjs zkey verify <circuit.r1cs> <pot_final.ptau> <circuit_final.zkey>` (scripts/g16_verify.sh:190-202, behind the
 * reference's -z flag): g16_zkey_new makes the initial key again (GPU, temporary file under $TMPDIR), then the key under test
 * must equal it in the header's alpha/beta/gamma, the circuit hash and sections 3-7; every contribution record must carry
 * the right transcript, a valid proof of knowledge and a delta that follows from the previous one (pairing checks on the
 * host), a beacon record must be the beacon it names; and sections 8 and 9 must be the initial sections divided by delta
 * (random linear combinations, four MSMs on the GPU, two pairing checks).  G16_OK = valid; G16_ERR_GENERIC with the first
 * failed check in g16_last_error otherwise.  Transcript and proof-of-knowledge conventions are this library's own
 * (g16_zkey_contribute): a key contributed to by snarkjs itself is NOT expected to pass, nor the reverse. */
int g16_zkey_verify(const char* r1cs_path, const char* ptau_path, const char* zkey_path, const g16_opts* opts);
/* `snarkjs wtns check <circuit.r1cs> <witness.wtns>` (scripts/g16_verify.sh:205-209, behind the reference's -w flag):
 * <A_c, w> * <B_c, w> == <C_c, w> for every constraint, one GPU thread per constraint.  Returns G16_OK when the check RAN;
 * *first_bad_constraint is -1 for a satisfying witness, else the lowest index of a constraint that fails.  Errors:
 * G16_ERR_CURVE (another field), G16_ERR_WITNESS_LEN ("Invalid witness length. Circuit: N, witness: M"), G16_ERR_FORMAT
 * (malformed files, values not below r, wire 0 not the constant 1). */
int g16_wtns_check(const char* r1cs_path, const char* wtns_path, const g16_opts* opts, int64_t* first_bad_constraint);
int g16_zkey_contribute(const char* zkey_in, const char* zkey_out, const char* name, const uint8_t* entropy, size_t entropy_len,
                        const g16_opts* opts, uint8_t contribution_hash[64]);
/* delta' from 2^num_iterations_exp iterations of SHA-256 over beacon_hash (scripts/g16_setup.sh:272-277) */
int g16_zkey_beacon(const char* zkey_in, const char* zkey_out, const char* name, const uint8_t* beacon_hash, size_t hash_len,
                    uint32_t num_iterations_exp, const g16_opts* opts, uint8_t contribution_hash[64]);
/* the text of verification_key.json (JSON.stringify(vKey, null, 1), vk_alphabeta_12 included); host only, no GPU.
 * Returns G16_ERR_SHORT_BUFFER with the required size in *out_len when the buffer is too small. */
int g16_zkey_export_vkey(const char* zkey_path, char* out, size_t* out_len);
/* test hook: which = 0 Blake2b-512 (64 bytes out), 1 SHA-256 (32 bytes out) */
int g16_test_hash(int which, const uint8_t* data, size_t n, uint8_t* out);
/* Test hook: the entry count, window bits, window count and sort path an MSM of n points would use, and the limits
 * enforced when its workspace is reserved (32-bit entry positions, 31-bit base indices, cub's int item count on the
 * radix-sort path): G16_ERR_GENERIC with "... shard it" in g16_last_error beyond them.  Host arithmetic only; works
 * without a GPU.  c = 0 picks the window width from n as a proof would. */
int g16_test_msm_shape(uint64_t n, int c, int table, int sparse, uint64_t* entries, int* c_out, int* windows, int* dense);

#ifdef __cplusplus
}
#endif
#endif
