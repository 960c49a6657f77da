/* zecale_gpu.h — C ABI of the B200 (sm_100a) prover hot path for clearmatics/zecale.
 *
 * Drop-in boundary for the Groth16 prove step over BW6-761
 * (libsnark r1cs_gg_ppzksnark_prover, r1cs_gg_ppzksnark.tcc:464-603). The reference's
 * call sites are served by this library through the template specialisations in
 * zecale_b200/hooks/zecale_gpu_hooks.hpp:
 *
 *   libff::multi_exp<G1|G2, Fr, BDLO12_signed, special>            multiexp.tcc:648-688
 *   libff::multi_exp_filter_one_zero<G1, Fr, ...>                  multiexp.tcc:695-757
 *   libsnark::kc_multi_exp_with_mixed_addition<G2, G1, Fr, ...>    kc_multiexp.tcc:31-113
 *        -> zecale_gpu_msm_g1 / zecale_gpu_msm_g2
 *   libfqfft::basic_radix2_domain<Fr>::{FFT,iFFT,cosetFFT,icosetFFT}
 *        (basic_radix2_domain.tcc:55-88)                           -> zecale_gpu_ntt
 *
 * Conventions follow libff's existing FFI (libff/ffi/ffi.h:55-95): plain pointers and
 * sizes, no C++ types, no exceptions across the boundary. Return value 0 = success,
 * non-zero = error code (ZECALE_GPU_E_*); zecale_gpu_last_error() gives the text.
 *
 * Data formats are libff's in-memory formats on a little-endian host, read as-is:
 *   Fr  : 6  x u64 limbs, Montgomery form, R = 2^384          (48 bytes)
 *   Fq  : 12 x u64 limbs, Montgomery form, R = 2^768          (96 bytes)
 *   G1/G2 point : {X, Y, Z} of Fq, homogeneous projective     (288 bytes)
 *   MSM bases must be in libff "special" form: Z == Fq::one() or the zero point
 *   (0, one, 0) — which is what every proving-key query holds
 *   (r1cs_gg_ppzksnark.tcc:352-387 batch_to_special).
 *   Results are written in special form too, so they can be memcpy'd into a
 *   libff::bw6_761_G1 / G2 object.
 *
 * There is no CPU fallback: every entry point fails with ZECALE_GPU_E_CUDA when no
 * sm_100 device is usable.
 */
#ifndef ZECALE_GPU_H
#define ZECALE_GPU_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct zecale_gpu_ctx zecale_gpu_ctx;
typedef struct zecale_gpu_bases zecale_gpu_bases;
typedef struct zecale_gpu_r1cs zecale_gpu_r1cs;

enum {
    ZECALE_GPU_OK = 0,
    ZECALE_GPU_E_ARG = 1,      /* null pointer / bad size / bad flags                  */
    ZECALE_GPU_E_CUDA = 2,     /* CUDA runtime error (text in last_error)              */
    ZECALE_GPU_E_FORM = 3,     /* an MSM base is not in special form (Z != one, != 0)  */
    ZECALE_GPU_E_DOMAIN = 4,   /* log_m outside [1, 46 = Fr two-adicity] or too large  */
    ZECALE_GPU_E_NOMEM = 5     /* device allocation failed                             */
};

/* NTT flags (zecale_gpu_ntt). */
enum {
    ZECALE_GPU_NTT_INVERSE = 1u, /* omega^-1 and multiply by m^-1 (libfqfft iFFT)       */
    ZECALE_GPU_NTT_COSET = 2u    /* cosetFFT: a[i] *= g^i before the forward transform;
                                    icosetFFT: a[i] *= g^-i after the inverse transform */
};

/* Groups (zecale_gpu_pin_bases). */
enum { ZECALE_GPU_G1 = 1, ZECALE_GPU_G2 = 2 };

/* ---- lifetime ------------------------------------------------------------------ */

/* Create a context. devices == NULL means {current device}.
 * n_devices == 1: everything runs on that device.
 * n_devices > 1 (distinct devices of one node, at most 16): a multi-device context. Pinned base
 * vectors are sharded by points over the listed devices - the chunks of libff::multi_exp
 * (multiexp.tcc:663-676), the first device taking a slightly smaller one because it also runs the
 * witness map - and every zecale_gpu_msm_pinned / _pinned_async call runs one Pippenger per shard,
 * concurrently, the 384-byte partial sums travelling to devices[0] by peer copies (NVLink) where they
 * are summed and normalised; coefficients_for_H go the other way, one slice per device. The transforms,
 * the sparse mat-vecs and the un-pinned entry points run on devices[0]. The host API is unchanged, so
 * the hooked reference prover gains the extra devices by listing them ($ZECALE_GPU_DEVICES).
 * Test switch: with ZECALE_GPU_TEST_REPEAT_DEVICES=1 in the environment one ordinal may be listed several times (the
 * shards then share that GPU), which lets a one-GPU box run the sharded code; it has no use in production. */
int zecale_gpu_init(const int *devices, int n_devices, zecale_gpu_ctx **ctx);
/* Number of devices the context spans. */
int zecale_gpu_device_count(zecale_gpu_ctx *ctx);
void zecale_gpu_shutdown(zecale_gpu_ctx *ctx);
const char *zecale_gpu_last_error(zecale_gpu_ctx *ctx);

/* ---- MSM: replaces libff::multi_exp & friends ------------------------------------
 * out = sum_{i<n} scalar(i) * base(i), where
 *   base(i)   = *(point*)((char*)bases + i*base_stride)         (288 B read at that address)
 *   scalar(i) = *(Fr*)((char*)scalars + j*scalar_stride), j = scalar_index ? scalar_index[i] : i
 * base_stride is 288 for a libff::G1_vector and 576 for the values of a
 * knowledge_commitment_vector<G2,G1> (G2 at byte 0, G1 at byte 288: pass bases+288 for
 * the G1 half). scalar_index is how the sparse B_query walk of kc_multiexp.tcc:63-89
 * is expressed: scalar_index[i] = B_query.indices[i].
 * Zero scalars, scalars equal to one and zero points need no pre-filtering.
 * All pointers are HOST pointers; out_point receives 288 bytes. */
int zecale_gpu_msm_g1(zecale_gpu_ctx *ctx, const void *bases, size_t base_stride, const void *scalars,
                      size_t scalar_stride, const size_t *scalar_index, size_t n, void *out_point);
int zecale_gpu_msm_g2(zecale_gpu_ctx *ctx, const void *bases, size_t base_stride, const void *scalars,
                      size_t scalar_stride, const size_t *scalar_index, size_t n, void *out_point);

/* ---- NTT: replaces libfqfft::basic_radix2_domain<bw6_761_Fr> ---------------------
 * In place on a[0 .. 2^log_m), natural order in and out, Montgomery form in and out.
 * flags = 0: FFT; INVERSE: iFFT; COSET: cosetFFT(g); INVERSE|COSET: icosetFFT(g).
 * coset_g: 48-byte Fr (Montgomery), required when COSET is set. `a` is a HOST pointer. */
int zecale_gpu_ntt(zecale_gpu_ctx *ctx, void *a, unsigned log_m, unsigned flags, const void *coset_g);

/* ---- residency (optional): keep the constant proving-key queries on the device ----
 * The handle owns a device-side affine image of the n bases; the host memory is not
 * retained. group is ZECALE_GPU_G1 or ZECALE_GPU_G2 (same arithmetic, kept for checks). */
int zecale_gpu_pin_bases(zecale_gpu_ctx *ctx, const void *bases, size_t base_stride, size_t n, int group,
                         zecale_gpu_bases **handle);
/* Same with the expansion made explicit: expand = 0 keeps one copy of the bases (the MSM then
 * runs one bucket set per window and a Horner pass of c doublings per window, like an
 * un-pinned call); expand = 1 (what zecale_gpu_pin_bases does) also stores 2^(c*V*g) * P_i for
 * every window group g, so that all windows share V bucket sets (V = 1 when memory allows:
 * n * 192 B * ceil(378/c) bytes) and the doubling chain disappears from every later MSM. */
int zecale_gpu_pin_bases_ex(zecale_gpu_ctx *ctx, const void *bases, size_t base_stride, size_t n, int group,
                            int expand, zecale_gpu_bases **handle);
/* Same device image, made straight from the proving key's ON-DISK bytes: n points,
 * `stride` bytes apart (multiple of 4), each X || Y with the coordinate as a plain big-endian integer
 * (libzeth group_element_write_bytes, core/group_element_utils.tcc:93-100; zero = (0, 1)). For a
 * query written by group_elements_write_bytes (groth16_snark.tcc:105-109) pass file + 8 and stride
 * 192; for the sparse B query (knowledge_commitment_vector_write_bytes, serialization/
 * stream_utils.tcc:111-122: entries of index(8) | g(192) | h(192) after a 16-byte header) pass
 * file + 24 (G2 part) or file + 216 (G1 part) with stride 392. Byte order and the conversion to
 * Montgomery form (the host's Fp_model(bigint), fp.tcc:246-252) are done on the device. */
int zecale_gpu_pin_bases_bytes(zecale_gpu_ctx *ctx, const void *bytes, size_t stride, size_t n, int group, int expand,
                               zecale_gpu_bases **handle);
int zecale_gpu_unpin_bases(zecale_gpu_ctx *ctx, zecale_gpu_bases *handle);
/* MSM over the first n pinned bases. */
int zecale_gpu_msm_pinned(zecale_gpu_ctx *ctx, const zecale_gpu_bases *handle, const void *scalars,
                          size_t scalar_stride, const size_t *scalar_index, size_t n, void *out_point);

/* ---- device-resident variants (inputs already in HBM; used by bench.py `value`) ---
 * d_scalars / d_a / d_scalar_index are DEVICE pointers on the context's device;
 * d_scalar_index holds u64. out_point is a HOST pointer. `elapsed_ms`, if not NULL,
 * receives the CUDA-event time of the launched kernels on the context's stream. */
int zecale_gpu_msm_pinned_dev(zecale_gpu_ctx *ctx, const zecale_gpu_bases *handle, const void *d_scalars,
                              size_t scalar_stride, const uint64_t *d_scalar_index, size_t n, void *out_point,
                              float *elapsed_ms);
int zecale_gpu_ntt_dev(zecale_gpu_ctx *ctx, void *d_a, unsigned log_m, unsigned flags, const void *coset_g,
                       float *elapsed_ms);

/* ---- NTT for callers that may not allocate device memory ---------------------------
 * This entry point has no counterpart in the reference (libfqfft works in host memory and knows no
 * such restriction); it exists for two kinds of host program:
 *   - one whose own memory pool took (nearly) all of the GPU's memory when it started, so that a
 *     cudaMalloc made on its behalf by a library fails with "out of memory";
 *   - one that serves requests under a latency bound and forbids cudaMalloc / cudaFree on that path,
 *     because both wait for the whole device and stall every other stream of the process.
 * Neither can create a zecale_gpu_ctx (zecale_gpu_init allocates), so this call takes no context.
 * It performs the same transform as zecale_gpu_ntt_dev - same flags, same coset_g (HOST pointer to
 * one 48-byte Fr, needed with COSET), in place on the DEVICE vector d_a[0 .. 2^log_m), identical
 * result bits - enqueued on the caller's `stream` (a cudaStream_t; NULL = the default stream) and
 * NOT synchronised: the result is there once the stream has reached this point. The device that is
 * current when it is called must be the one that owns `stream` and `d_a`.
 *
 * Where the working memory comes from: the twiddle table, the scale table and the second vector the
 * passes alternate with sit in a fixed array inside the library's device image. It is
 * ZECALE_GPU_NTT_STATIC_BYTES (about 7.5 MiB) on every GPU that loads the library, from load until
 * the process ends, whether or not this function is ever called. That bounds the size:
 *   log_m <= ZECALE_GPU_NTT_STATIC_MAX_LOG_M (2^16 elements, the batch-1 aggregator domain);
 * anything larger is refused with ZECALE_GPU_E_DOMAIN and a message naming the limit. The call
 * itself runs no cudaMalloc, cudaFree, cudaMallocHost or other allocating routine.
 *
 * Sharing rules. There is ONE such array per GPU and every call uses it. A call does not depend on
 * what the previous call left in it: it recomputes the tables on `stream` before it reads them
 * (which makes it slower than zecale_gpu_ntt_dev on a warm context), and nothing in it has to be
 * zero beforehand. But two calls that execute on the same GPU at the same time - issued on two
 * streams with no event or synchronisation between them, or from two threads - overwrite each
 * other's tables and both results are garbage. Keep the calls of one GPU in one stream order.
 *
 * First call: the first call made on each device looks up where the driver put the array on that
 * device (and, with lazy module loading, makes the driver load the kernels). Make that first call
 * during start-up, on a stream that is not being captured.
 *
 * Errors: returns ZECALE_GPU_OK or a ZECALE_GPU_E_* code; because there is no context to hold the
 * text, the message is written to `err` (at most err_cap bytes, NUL-terminated) when err != NULL. */
#define ZECALE_GPU_NTT_STATIC_MAX_LOG_M 16u
#define ZECALE_GPU_NTT_STATIC_BYTES ((((size_t)1 << ZECALE_GPU_NTT_STATIC_MAX_LOG_M) * 5 / 2 + 4) * 48)
int zecale_gpu_ntt_static(void *stream, void *d_a, unsigned log_m, unsigned flags, const void *coset_g, char *err,
                          size_t err_cap);

/* ---- H polynomial: the device part of libsnark::r1cs_to_qap_witness_map ----------
 * (r1cs_to_qap.tcc:252-330 with d1 = d2 = d3 = 0, as r1cs_gg_ppzksnark_prover calls it,
 * r1cs_gg_ppzksnark.tcc:477-484). Inputs are the evaluations aA[i] = <A_i,w>,
 * aB[i] = <B_i,w>, aC[i] = <C_i,w> on the domain of size m = 2^log_m (zero padded, with
 * the input-consistency rows already in aA, r1cs_to_qap.tcc:236-248). Runs
 * 3 x (iFFT, cosetFFT), aA∘aB - aC, divide_by_Z_on_coset and icosetFFT on the device and
 * writes coefficients_for_H: m + 1 Fr elements (the last one is zero). coset_g is
 * Fr::multiplicative_generator in Montgomery form. HOST pointers. */
int zecale_gpu_qap_h(zecale_gpu_ctx *ctx, const void *aA, const void *aB, const void *aC, unsigned log_m,
                     const void *coset_g, void *out_h);
/* Device-resident variant: d_aA receives coefficients_for_H[0..m); d_aB, d_aC are clobbered. */
int zecale_gpu_qap_h_dev(zecale_gpu_ctx *ctx, void *d_aA, void *d_aB, void *d_aC, unsigned log_m,
                         const void *coset_g, float *elapsed_ms);

/* ---- H polynomial straight from the witness ---------------------------------
 * Device image of the constraint system of a proving key (libsnark r1cs_constraint_system,
 * relations/constraint_satisfaction_problems/r1cs/r1cs.hpp; pk.constraint_system in
 * r1cs_gg_ppzksnark.hpp:74-167): matrix k (0 = A, 1 = B, 2 = C) in CSR form, row i =
 * constraints[i].{a,b,c}.terms: col = variable index (0 = the constant one), val = coefficient
 * (Fr, Montgomery). row_ptr[k] has n_constraints + 1 entries. HOST pointers; nothing is retained. */
int zecale_gpu_r1cs_load(zecale_gpu_ctx *ctx, size_t n_constraints, size_t n_inputs, size_t n_variables,
                         const uint64_t *const row_ptr[3], const uint32_t *const col[3], const void *const val[3],
                         zecale_gpu_r1cs **handle);
int zecale_gpu_r1cs_free(zecale_gpu_ctx *ctx, zecale_gpu_r1cs *handle);
/* log2 of the evaluation domain r1cs_to_qap_witness_map uses for this system (power-of-two domain
 * of at least n_constraints + n_inputs + 1 points, r1cs_to_qap.tcc:224-227). */
unsigned zecale_gpu_r1cs_log_domain(const zecale_gpu_r1cs *handle);
/* The whole of r1cs_to_qap_witness_map (r1cs_to_qap.tcc:210-344, d1 = d2 = d3 = 0) on the device,
 * enqueued on `lane`: the evaluations <A_i,w>, <B_i,w>, <C_i,w> (+ the input-consistency rows) are
 * computed from the assignment by a sparse mat-vec kernel, then the seven transforms. assignment =
 * [1, primary..., auxiliary...] (n_variables + 1 Fr; HOST pointer, or device with
 * ZECALE_GPU_DEVICE_POINTERS). MSMs enqueued afterwards with scalars inside the same host range
 * reuse its upload. out_h: m + 1 Fr at zecale_gpu_sync, or NULL when only
 * ZECALE_GPU_SCALARS_FROM_QAP consumes the coefficients. */
int zecale_gpu_qap_h_from_witness_async(zecale_gpu_ctx *ctx, int lane, const zecale_gpu_r1cs *r1cs, const void *assignment,
                                        const void *coset_g, void *out_h, unsigned flags);

/* ---- asynchronous lanes ------------------------------------------------------------
 * The MSMs of one proof are data-independent (r1cs_gg_ppzksnark.tcc:531-580) and only the H
 * MSM waits for the H polynomial. A context owns zecale_gpu_lane_count() in-order lanes
 * (stream + scratch); the *_async calls enqueue on a lane and return, zecale_gpu_sync waits
 * for all lanes and then writes every out_point / out_h. Host buffers passed to an async call
 * must stay valid and unchanged until zecale_gpu_sync returns. Blocking calls use lane 0. */
enum {
    ZECALE_GPU_DEVICE_POINTERS = 1u, /* scalars / scalar_index are device pointers (index: u64)      */
    ZECALE_GPU_SCALARS_FROM_QAP = 2u /* scalars = coefficients_for_H of the last qap_h on this lane */
};
int zecale_gpu_lane_count(void);
int zecale_gpu_msm_pinned_async(zecale_gpu_ctx *ctx, int lane, const zecale_gpu_bases *handle, const void *scalars,
                                size_t scalar_stride, const void *scalar_index, size_t n, void *out_point,
                                unsigned flags);
/* out_h may be NULL when the coefficients are only consumed by ZECALE_GPU_SCALARS_FROM_QAP. */
int zecale_gpu_qap_h_async(zecale_gpu_ctx *ctx, int lane, const void *aA, const void *aB, const void *aC,
                           unsigned log_m, const void *coset_g, void *out_h);
int zecale_gpu_sync(zecale_gpu_ctx *ctx);
/* Enqueue what the async calls since the last sync/flush deferred (the latency-bound second halves of
 * their MSMs, issued together as one batch) without waiting for anything: a caller that keeps several
 * proofs in flight (one context each) calls this after the last *_async call of a proof, so that the
 * device can finish the proof while the host is already enqueueing the next one. zecale_gpu_sync
 * implies it. */
int zecale_gpu_flush(zecale_gpu_ctx *ctx);
/* Page-locked host memory for buffers handed to the async calls (a pageable buffer makes every
 * host-to-device copy synchronous). Not tied to a context; free with zecale_gpu_host_free. */
void *zecale_gpu_host_alloc(size_t bytes);
void zecale_gpu_host_free(void *p);

/* ---- multi-GPU building blocks (one process per GPU; the exchange itself is the caller's:
 * NCCL through torch.distributed in zecale_b200/dist.py) --------------------------------
 *
 * Point-sharded MSM: rank g runs zecale_gpu_msm_* / _pinned on bases [g*n/G, (g+1)*n/G) — the
 * chunking of libff::multi_exp (multiexp.tcc:663-685) — the G partial points are all-gathered
 * and summed in chunk order with zecale_gpu_point_sum (multiexp.tcc:680-685). `points`: n
 * projective points of 288 B (any Z), HOST pointer; out_point: special form. */
int zecale_gpu_point_sum(zecale_gpu_ctx *ctx, const void *points, size_t n, void *out_point);
/* 4-step NTT of m = 2^log_m elements over `world` ranks (a power of two) with ONE all-to-all.
 * With m1 = 2^ceil(log_m/2), m2 = m/m1, c1 = m1/world, c2 = m2/world:
 *   layout A: rank g holds index j = j1*m2 + j2 for j2 in [g*c2, (g+1)*c2), stored [j1][j2 - g*c2]
 *   layout B: rank h holds index k = k1 + m1*k2 for k1 in [h*c1, (h+1)*c1), stored [k1 - h*c1][k2]
 * FFT / cosetFFT take layout A to layout B; iFFT / icosetFFT take layout B to layout A (so the
 * iFFT -> cosetFFT -> pointwise -> icosetFFT chain of r1cs_to_qap_witness_map needs no other
 * exchange). phase 0: local sub-transforms + twiddles, d_out = `world` contiguous chunks, chunk d
 * for rank d; the caller then runs an all-to-all (chunk s of the result comes from rank s);
 * phase 1: local sub-transforms on the exchanged buffer. d_in / d_out: distinct DEVICE buffers of
 * m/world Fr. flags / coset_g as zecale_gpu_ntt. sync != 0 waits for the kernels (needed before
 * handing the buffer to a collective on another stream). */
int zecale_gpu_ntt_dist_phase(zecale_gpu_ctx *ctx, const void *d_in, void *d_out, unsigned log_m, unsigned flags,
                              const void *coset_g, int world, int rank, int phase, int sync);

/* The same phase enqueued on the CALLER's stream (a cudaStream_t, e.g. the stream the caller's NCCL
 * collectives are ordered on) and not synchronised: phase 0, the all-to-all and phase 1 then chain on the
 * device with no host round trip in between. Tables are built on first use (that call synchronises). */
int zecale_gpu_ntt_dist_phase_on(zecale_gpu_ctx *ctx, void *stream, const void *d_in, void *d_out, unsigned log_m,
                                 unsigned flags, const void *coset_g, int world, int rank, int phase);

/* The pointwise step of r1cs_to_qap_witness_map on `count` locally held elements of a domain of
 * size 2^log_m (r1cs_to_qap.tcc:283-315): aA[i] = (aA[i]*aB[i] - aC[i]) / (g^m - 1). DEVICE pointers. */
int zecale_gpu_qap_pointwise_dev(zecale_gpu_ctx *ctx, void *d_aA, const void *d_aB, const void *d_aC, unsigned log_m,
                                 const void *coset_g, size_t count);
/* ... on the caller's stream, not synchronised. */
int zecale_gpu_qap_pointwise_dev_on(zecale_gpu_ctx *ctx, void *stream, void *d_aA, const void *d_aB, const void *d_aC,
                                    unsigned log_m, const void *coset_g, size_t count);

/* ---- synthetic inputs (bench / tests): out_points[i] = k_i * generator of the group, in
 * libff special form (288 B each), k_i a non-zero 32-bit hash of (seed, i). HOST pointer. */
int zecale_gpu_synth_bases(zecale_gpu_ctx *ctx, int group, uint64_t seed, size_t n, void *out_points);

/* ---- fixed-base batch exponentiation (the trusted-setup generator's work) ----
 * out_points[i] = (coeff * scalars[i]) * base, i < n, in libff special form (288 B each; zero =
 * (0, one, 0)). Replaces libff::batch_exp / batch_exp_with_coeff (libff/algebra/
 * scalar_multiplication/multiexp.tcc:876-947) followed by batch_to_special (:949-974) as used by
 * r1cs_gg_ppzksnark_generator (libsnark r1cs_gg_ppzksnark.tcc:330-390) and kc_batch_exp
 * (knowledge_commitment/kc_multiexp.tcc:115-239). base: one point in special form; scalars:
 * Montgomery Fr, scalar_stride bytes apart; coeff: one Montgomery Fr or NULL. HOST pointers. */
int zecale_gpu_batch_exp(zecale_gpu_ctx *ctx, int group, const void *base_point, const void *scalars, size_t scalar_stride,
                         size_t n, const void *coeff, void *out_points);

/* ---- Lagrange basis at a point (the QAP instance evaluation of the trusted setup) ----
 * out[i] = L_i(t), i < m = 2^log_m, for the domain of m-th roots of unity: replaces
 * libfqfft::basic_radix2_domain<Fr>::evaluate_all_lagrange_polynomials (basic_radix2_domain.tcc:91-94 ->
 * _basic_radix2_evaluate_all_lagrange_polynomials, basic_radix2_domain_aux.tcc:186-239: m serial field
 * inversions), called by r1cs_to_qap_instance_map_with_evaluation (r1cs_to_qap.tcc:102-108) once per
 * setup. t: one Fr (Montgomery); out: m Fr. HOST pointers. Returns ZECALE_GPU_E_DOMAIN when t is itself a
 * point of the domain (the unit-vector case, :203-217) - the caller keeps the reference's code for that. */
int zecale_gpu_lagrange(zecale_gpu_ctx *ctx, unsigned log_m, const void *t, void *out);

/* ---- introspection ----------------------------------------------------------------*/
/* Number of kernel launches issued by this context so far (bench.py gpu_launches). */
uint64_t zecale_gpu_launch_count(zecale_gpu_ctx *ctx);
/* CUDA-event time spent in one kernel class since the last reset, and the number of timed
 * regions: class 0 = MSM bucket accumulation kernel (k_accumulate, one region per MSM),
 * class 1 = NTT transforms (one region per transform, tables excluded). */
int zecale_gpu_kernel_ms(zecale_gpu_ctx *ctx, int kernel_class, int reset, double *total_ms, uint64_t *count);
/* Shape of the last MSM this context ran: bucket entries (non-zero digits), accumulate tasks,
 * window bits, bucket sets V and base copies G. Mixed additions executed = entries - tasks. */
int zecale_gpu_msm_last_stats(zecale_gpu_ctx *ctx, uint64_t *entries, uint64_t *tasks, unsigned *c,
                              unsigned *bucket_sets, unsigned *base_copies);
/* Window size c an MSM over n points uses: the un-pinned calls (zecale_gpu_msm_g1/g2: bases uploaded per call,
 * one shared set of doublings at the end) and the pinned ones (zecale_gpu_pin_bases with expand = 1: pre-shifted
 * copies, cost-model choice; what zecale_gpu_msm_last_stats reports after such a call). */
unsigned zecale_gpu_msm_window_bits(size_t n);
unsigned zecale_gpu_msm_window_bits_pinned(size_t n);

/* ---- unit-test / microbenchmark hooks (not used by the prover) --------------------
 * Element-wise field and group operations on host arrays, executed by the same
 * device functions the MSM/NTT kernels use. op codes:
 *   field: 0 add, 1 sub, 2 mul, 3 sqr, 4 inverse, 5 neg, 6 from-Montgomery, 7 inverse by the Fermat ladder (Fq),
 *          8 mul by the interleaved CIOS body (Fq; the product path uses the Karatsuba body of op 2)
 *   group: 0 add (xyzz_add), 1 mixed add, 2 double, 3 normalise (to special)
 * Points are 288-byte libff projective points in and (special form) out. */
int zecale_gpu_test_fq(zecale_gpu_ctx *ctx, int op, const void *a, const void *b, void *out, size_t n);
int zecale_gpu_test_fr(zecale_gpu_ctx *ctx, int op, const void *a, const void *b, void *out, size_t n);
int zecale_gpu_test_group(zecale_gpu_ctx *ctx, int op, const void *a, const void *b, void *out, size_t n);
/* Fq Montgomery-product throughput: `threads` threads each run a dependent chain of
 * `iters` products; returns the CUDA-event time in ms in *elapsed_ms. */
int zecale_gpu_bench_fq_mul(zecale_gpu_ctx *ctx, unsigned blocks, unsigned threads_per_block, unsigned iters,
                            float *elapsed_ms);
/* Integer-multiply pipe ceiling (the denominator of the MSM roofline): every thread runs 8 independent dependent chains
 * of 32x32+64 -> 64 multiply-adds (IMAD.WIDE.U32), 8 * iters of them per thread; CUDA-event time in ms in *elapsed_ms. */
int zecale_gpu_bench_imad(zecale_gpu_ctx *ctx, unsigned blocks, unsigned threads_per_block, unsigned iters,
                          float *elapsed_ms);

#ifdef __cplusplus
}
#endif
#endif /* ZECALE_GPU_H */
