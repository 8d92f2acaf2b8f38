/*
 * mfhe.h -- C ABI of the B200-native Matrix-FHE hot path (libmfhe_b200.so).
 *
 * Plain C: pointers, sizes and ints only.  Every function returns 0 on success and
 * a negative MFHE_E* code otherwise; it never throws, exits or prints.
 * mfhe_last_error() returns the message of the last failure on the calling thread.
 * Unless a name ends in _host, data pointers are DEVICE pointers owned by the caller
 * and the work is enqueued asynchronously on `stream` (a cudaStream_t passed as
 * void*; NULL = the legacy default stream, as the reference uses).  Nothing is
 * allocated on the device after plan/context creation.
 *
 * Each entry point cites the reference interface it replaces ("ph/" =
 * extern/phantom-fhe/).  INTEGRATION.md shows the binding a maintainer of the
 * reference adds on their side.
 */
#ifndef MFHE_H
#define MFHE_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MFHE_OK            0
#define MFHE_EINVAL       -1   /* bad argument (size, modulus without the needed root, ...) */
#define MFHE_ECUDA        -2   /* CUDA runtime error; see mfhe_last_error()                 */
#define MFHE_ENOMEM       -3
#define MFHE_EUNSUPPORTED -4

const char *mfhe_last_error(void);
/* library version string, e.g. "mfhe-b200 0.1 (sm_100a)" */
const char *mfhe_version(void);

/* ------------------------------------------------------------------------ *
 * Flat RNS negacyclic NTT over [batch][L][N] u64 (family A of SURVEY.md).
 * ------------------------------------------------------------------------ */
typedef struct mfhe_ntt_plan mfhe_ntt_plan;

/* X-axis conventions of the reference */
#define MFHE_NTT_PHANTOM_NEGACYCLIC 0 /* X^N + 1, natural in -> bit-reversed out
                                         ph/src/host/ntt.cu:11-56 tables, minimal primitive 2N-th root */
#define MFHE_NTT_GL_XN_MINUS_I      1 /* X^n - i, natural in/out: src/core/ntt_core.cu:462-481          */
#define MFHE_NTT_CYCLIC_NATURAL     2 /* X^n - 1, natural in/out: src/core/ntt_core.cu:394-431          */

/* Builds per-limb twiddle tables on the host and uploads them to the current device.
 * Replaces PhantomContext + DNTTTable construction (ph/src/context.cu:121-183,
 * src/core/HE.cu:327-336) and init_ntt_tables_manual / init_gl_twist_tables
 * (src/core/ntt_core.cu:75-198).  logn in 4..8 and 12..17; moduli < 2^60.
 * GL / cyclic conventions are available for logn <= 8 only. */
int mfhe_ntt_plan_create(int logn, int num_limbs, const uint64_t *moduli, int convention,
                         mfhe_ntt_plan **out);
void mfhe_ntt_plan_destroy(mfhe_ntt_plan *plan);
int mfhe_ntt_plan_logn(const mfhe_ntt_plan *plan);
int mfhe_ntt_plan_limbs(const mfhe_ntt_plan *plan);
/* psi used for limb l (the root the convention prescribes), for cross-checking */
uint64_t mfhe_ntt_plan_root(const mfhe_ntt_plan *plan, int limb);
/* copy the forward / inverse table words {w, w_shoup} of one limb to host arrays of N u64
 * each (index order tw[m + i]); inverse != 0 selects the inverse table.  Debug / parity aid. */
int mfhe_ntt_plan_get_table(const mfhe_ntt_plan *plan, int limb, int inverse, uint64_t *w,
                            uint64_t *w_shoup);

/* In-place transforms of `batch` polys of plan->L limbs each, layout [batch][L][N].
 * Replaces nwt_2d_radix8_forward_inplace / backward_inplace (ph/src/ntt/fntt_2d.cu:620-653,
 * ph/src/ntt/intt_2d.cu:724-757; one call per poly there, one call for the batch here)
 * and xy_ntt_forward/backward_phantom, xy_ntt_forward/backward_gl, custom_ntt_forward/
 * backward (src/core/ntt_core.cu:394-481). */
int mfhe_ntt_forward(const mfhe_ntt_plan *plan, uint64_t *d_data, size_t batch, void *stream);
int mfhe_ntt_inverse(const mfhe_ntt_plan *plan, uint64_t *d_data, size_t batch, void *stream);
/* Same on the limb window [limb_start, limb_start+limb_count): data is
 * [batch][limb_count][N] (Phantom's `start_modulus_idx` / `coeff_modulus_size`). */
int mfhe_ntt_forward_limbs(const mfhe_ntt_plan *plan, uint64_t *d_data, size_t batch,
                           int limb_start, int limb_count, void *stream);
int mfhe_ntt_inverse_limbs(const mfhe_ntt_plan *plan, uint64_t *d_data, size_t batch,
                           int limb_start, int limb_count, void *stream);
/* Host-buffer convenience (what a cgo/ctypes caller with host memory uses): H2D copy,
 * transform, D2H copy, synchronous.  Uses an internal device staging buffer that grows
 * to the largest request. */
int mfhe_ntt_forward_host(mfhe_ntt_plan *plan, uint64_t *h_data, size_t batch);
int mfhe_ntt_inverse_host(mfhe_ntt_plan *plan, uint64_t *h_data, size_t batch);

/* Tuning knob: number of limb-polys processed per column+row kernel pair so that the
 * intermediate stays in L2 (0 = whole batch in one pair of launches). */
int mfhe_ntt_plan_set_chunk(mfhe_ntt_plan *plan, size_t chunk_limb_polys);

/* Pointwise ops on [batch][L][N] in the plan's moduli (canonical in, canonical out).
 * Replaces pointwise_mul kernels of the reference (src/core/HE.cu:509-531 semantics with
 * a full-size second operand) and Phantom's multiply_rns_poly. */
int mfhe_poly_mul(const mfhe_ntt_plan *plan, const uint64_t *d_a, const uint64_t *d_b,
                  uint64_t *d_out, size_t batch, void *stream);
int mfhe_poly_add(const mfhe_ntt_plan *plan, const uint64_t *d_a, const uint64_t *d_b,
                  uint64_t *d_out, size_t batch, void *stream);
int mfhe_poly_sub(const mfhe_ntt_plan *plan, const uint64_t *d_a, const uint64_t *d_b,
                  uint64_t *d_out, size_t batch, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* MFHE_H */
