/*
 * g16_synth.h -- synthetic-workload generator (libg16synth.so).  BENCH / TEST SUPPORT, not part of the drop-in
 * boundary: the reference commits no .zkey/.wtns and none can be produced offline, so the
 * benchmark circuits are the deterministic "SYNTH-1" R1CS of oracle/factory.py under a trusted setup whose trapdoor
 * is known.  Known trapdoor => the proof for pinned r,s has a closed form, which is what g16_synth_expected returns.
 * The reference step this stands in for is `snarkjs zkey new/contribute` (scripts/g16_setup.sh:240-283) plus the
 * circom witness generator (scripts/g16_prove.sh:228-239).  Field scalars are computed on the host, the ~10^7
 * fixed-base scalar multiplications on the GPU.
 */
#ifndef G16_SYNTH_H
#define G16_SYNTH_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct g16_synth g16_synth;

int g16_synth_create(uint32_t n_vars, uint32_t n_public, uint32_t n_constraints, uint64_t circuit_seed,
                     uint64_t trapdoor_seed, g16_synth** out);
void g16_synth_destroy(g16_synth* s);
int g16_synth_sizes(const g16_synth* s, uint64_t* zkey_bytes, uint64_t* wtns_bytes, uint32_t* domain_size);
/* fills caller buffers with the bytes of a .zkey and a .wtns file (GPU needed for the zkey; either may be NULL) */
int g16_synth_write(g16_synth* s, uint8_t* zkey_out, uint8_t* wtns_out);
/* closed-form proof for blinding factors r,s: pi_a x,y | pi_b x.c0,x.c1,y.c0,y.c1 | pi_c x,y ; 32-byte LE standard form */
int g16_synth_expected(g16_synth* s, const uint8_t r[32], const uint8_t sblind[32], uint8_t out[256]);
/* verification key points, standard form LE: alpha1(64) beta2(128) gamma2(128) delta2(128) IC[(nPublic+1)](64 each) */
int g16_synth_vkey(g16_synth* s, uint8_t* out);
/* the circuit as an .r1cs file plus a powers-of-tau file under the generator's tau, alpha, beta: the inputs of `zkey new`
 * (g16_zkey_new).  The ceremony file (power = log2(domain) + 1, iden3 layout) is SPARSE: only the ranges `zkey new` reads
 * for this circuit hold points.  Sections 4-7 of the resulting key equal g16_synth_write's (they do not depend on gamma,
 * delta); IC, C and H differ by the factors 1/gamma, 1/delta that `zkey new` fixes at 1. */
int g16_synth_write_setup_inputs(g16_synth* s, const char* r1cs_path, const char* ptau_path);
/* k*G for n standard-form scalars; out = zkey records (Montgomery affine, all-zero = infinity) */
int g16_synth_fixed_base(int is_g2, const uint8_t* scalars, size_t n, uint8_t* out);
const char* g16_synth_last_error(void);

#ifdef __cplusplus
}
#endif
#endif
