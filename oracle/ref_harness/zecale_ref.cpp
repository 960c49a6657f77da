// oracle/_ref harness — TEST INFRASTRUCTURE ONLY (never linked into the product).
//
// Builds the UNMODIFIED reference (libff / libfqfft / libsnark / libzeth /
// libzecale, from /root/reference) behind a small extern "C" surface so that
// tests, smoke() and bench.py's cpu_baseline / --impl reference legs can run the
// reference's own code on the same buffers the CUDA path sees.
//
// Everything below *calls* reference code; the only logic restated here is the
// ~40 lines of orchestration in r1cs_gg_ppzksnark_prover that must take (r, s)
// as parameters instead of sampling them from std::random_device
// (libsnark r1cs_gg_ppzksnark.tcc:502-503, 521-599).
//
// Memory formats are libff's own in-memory formats (no serialisation):
//   Fr  = bigint<6>  mont_repr, 48 B      Fq = bigint<12> mont_repr, 96 B
//   G1/G2 = {X, Y, Z} of Fq, 288 B        knowledge_commitment<G2,G1> = 576 B

#include <algorithm>
#include <array>
#include <chrono>
#include <cstdint>
#include <cstring>
#include <fstream>
#include <iostream>
#include <memory>
#include <sstream>
#include <string>
#include <vector>

#include <libff/algebra/curves/bw6_761/bw6_761_pp.hpp>
#include <thread>

#include <libff/algebra/scalar_multiplication/multiexp.hpp>
#include <libff/common/profiling.hpp>
#include <libff/common/rng.hpp>
#include <libfqfft/evaluation_domain/domains/basic_radix2_domain.hpp>
#include <libfqfft/evaluation_domain/get_evaluation_domain.hpp>
#include <libsnark/knowledge_commitment/kc_multiexp.hpp>
#include <libsnark/reductions/r1cs_to_qap/r1cs_to_qap.hpp>
#include <libsnark/zk_proof_systems/ppzksnark/r1cs_gg_ppzksnark/r1cs_gg_ppzksnark.hpp>
#include <libzeth/snarks/groth16/groth16_snark.hpp>

#ifdef ZREF_WITH_AGGREGATOR
#include "libzecale/circuits/aggregator_circuit.hpp"
#include "libzecale/circuits/groth16_verifier/groth16_verifier_parameters.hpp"
#include "libzecale/tests/circuits/dummy_application.hpp"

#include <libsnark/gadgetlib1/gadgets/pairing/bw6_761_bls12_377/bw6_761_pairing_params.hpp>
#endif

#ifdef _OPENMP
#include <omp.h>
#endif

using ppT = libff::bw6_761_pp;
using Fr = libff::Fr<ppT>;
using Fq = libff::bw6_761_Fq;
using G1 = libff::G1<ppT>;
using G2 = libff::G2<ppT>;
using KC = libsnark::knowledge_commitment<G2, G1>;
using snark = libzeth::groth16_snark<ppT>;

static_assert(sizeof(Fr) == 48, "Fr layout");
static_assert(sizeof(Fq) == 96, "Fq layout");
static_assert(sizeof(G1) == 288, "G1 layout");
static_assert(sizeof(G2) == 288, "G2 layout");
static_assert(sizeof(KC) == 576, "KC layout");

static const libff::multi_exp_method METHOD =
    libff::multi_exp_method_BDLO12_signed;
static const libff::multi_exp_base_form FORM =
    libff::multi_exp_base_form_special;

namespace
{

struct session_t {
    bool have_cs = false;
    libsnark::r1cs_constraint_system<Fr> cs;
    std::vector<Fr> primary;
    std::vector<Fr> aux;
    bool have_keys = false;
    std::unique_ptr<libsnark::r1cs_gg_ppzksnark_keypair<ppT>> kp_ptr;
    // last prove
    std::vector<Fr> h_coeffs;
    G1 At, Bt_h, Ht, Lt;
    G2 Bt_g;
    double t_witness_map = 0, t_msm_a = 0, t_msm_b = 0, t_msm_h = 0,
           t_msm_l = 0, t_total = 0;
    // aggregator sessions: seconds spent in the generate_r1cs_witness statements of
    // aggregator_circuit::prove (aggregator_circuit.tcc:136-157), i.e. SURVEY 8(f2); and the same
    // statements with the two per-nested-proof loops spread over host threads (prototype; -1 = the
    // parallel run did not reproduce the serial assignment)
    double t_witness_gen = 0, t_witness_gen_parallel = 0;
    int witness_gen_threads = 0;
};

std::unique_ptr<session_t> S_ptr(new session_t());
#define S (*S_ptr)
bool g_inited = false;

double now_s()
{
    return std::chrono::duration<double>(
               std::chrono::steady_clock::now().time_since_epoch())
        .count();
}

int threads()
{
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

template<typename T> void copy_out(void *dst, const std::vector<T> &v)
{
    if (!v.empty())
        std::memcpy(dst, (const void *)v.data(), v.size() * sizeof(T));
}

// Restatement of r1cs_gg_ppzksnark.tcc:464-603 with (r, s) injected and the
// intermediate results kept. All heavy lifting is the reference's own code.
libsnark::r1cs_gg_ppzksnark_proof<ppT> prove_with_rs(
    const libsnark::r1cs_gg_ppzksnark_proving_key<ppT> &pk,
    const std::vector<Fr> &primary,
    const std::vector<Fr> &aux,
    const Fr &r,
    const Fr &s,
    size_t chunks)
{
    const double t0 = now_s();
    const libsnark::qap_witness<Fr> qap_wit = libsnark::r1cs_to_qap_witness_map(
        pk.constraint_system,
        primary,
        aux,
        Fr::zero(),
        Fr::zero(),
        Fr::zero(),
        true);
    const double t1 = now_s();
    S.h_coeffs = qap_wit.coefficients_for_H;

    std::vector<Fr> padded(1, Fr::one());
    padded.insert(
        padded.end(),
        qap_wit.coefficients_for_ABCs.begin(),
        qap_wit.coefficients_for_ABCs.end());
    const size_t nv = qap_wit.num_variables();

    S.At = libff::multi_exp_filter_one_zero<G1, Fr, METHOD, FORM>(
        pk.A_query.begin(),
        pk.A_query.begin() + nv + 1,
        padded.begin(),
        padded.begin() + nv + 1,
        chunks);
    const double t2 = now_s();
    KC Bt = libsnark::kc_multi_exp_with_mixed_addition<G2, G1, Fr, METHOD, FORM>(
        pk.B_query, 0, nv + 1, padded.begin(), padded.begin() + nv + 1, chunks);
    S.Bt_g = Bt.g;
    S.Bt_h = Bt.h;
    const double t3 = now_s();
    S.Ht = libff::multi_exp<G1, Fr, METHOD, FORM>(
        pk.H_query.begin(),
        pk.H_query.begin() + (qap_wit.degree() - 1),
        qap_wit.coefficients_for_H.begin(),
        qap_wit.coefficients_for_H.begin() + (qap_wit.degree() - 1),
        chunks);
    const double t4 = now_s();
    S.Lt = libff::multi_exp_filter_one_zero<G1, Fr, METHOD, FORM>(
        pk.L_query.begin(),
        pk.L_query.end(),
        padded.begin() + qap_wit.num_inputs() + 1,
        padded.begin() + nv + 1,
        chunks);
    const double t5 = now_s();

    G1 g1_A = pk.alpha_g1 + S.At + r * pk.delta_g1;
    G1 g1_B = pk.beta_g1 + S.Bt_h + s * pk.delta_g1;
    G2 g2_B = pk.beta_g2 + S.Bt_g + s * pk.delta_g2;
    G1 g1_C = S.Ht + S.Lt + s * g1_A + r * g1_B - (r * s) * pk.delta_g1;
    const double t6 = now_s();

    S.t_witness_map = t1 - t0;
    S.t_msm_a = t2 - t1;
    S.t_msm_b = t3 - t2;
    S.t_msm_h = t4 - t3;
    S.t_msm_l = t5 - t4;
    S.t_total = t6 - t0;
    return libsnark::r1cs_gg_ppzksnark_proof<ppT>(
        std::move(g1_A), std::move(g2_B), std::move(g1_C));
}

void proof_bytes(const libsnark::r1cs_gg_ppzksnark_proof<ppT> &proof, void *out)
{
    std::ostringstream ss;
    snark::proof_write_bytes(proof, ss);
    const std::string b = ss.str();
    std::memcpy(out, b.data(), b.size());
}

void setup_from_seed(uint64_t seed)
{
    const Fr t = libff::SHA512_rng<Fr>(seed * 16 + 1);
    const Fr alpha = libff::SHA512_rng<Fr>(seed * 16 + 2);
    const Fr beta = libff::SHA512_rng<Fr>(seed * 16 + 3);
    const Fr delta = libff::SHA512_rng<Fr>(seed * 16 + 4);
    S.kp_ptr.reset(new libsnark::r1cs_gg_ppzksnark_keypair<ppT>(
        libsnark::r1cs_gg_ppzksnark_generator_from_secrets<ppT, FORM>(
            S.cs, t, alpha, beta, delta, G1::one(), G2::one(), true)));
    S.have_keys = true;
}

} // namespace

extern "C" int zref_init(void)
{
    if (!g_inited) {
        ppT::init_public_params();
#ifdef ZREF_WITH_AGGREGATOR
        libff::bls12_377_pp::init_public_params();
#endif
        libff::inhibit_profiling_info = true;
        libff::inhibit_profiling_counters = true;
        g_inited = true;
    }
    return threads();
}

extern "C" int zref_threads(void) { return threads(); }

extern "C" void zref_set_threads(int n)
{
#ifdef _OPENMP
    omp_set_num_threads(n);
#else
    (void)n;
#endif
}

// ---------------------------------------------------------------- fields --
// op: 0 add, 1 sub, 2 mul, 3 sqr, 4 inverse, 5 neg, 6 as_bigint (de-Montgomery),
//     7 from plain bigint (to Montgomery)
template<typename F> static void field_op(int op, const F *a, const F *b, F *o, size_t n)
{
    for (size_t i = 0; i < n; ++i) {
        switch (op) {
        case 0: o[i] = a[i] + b[i]; break;
        case 1: o[i] = a[i] - b[i]; break;
        case 2: o[i] = a[i] * b[i]; break;
        case 3: o[i] = a[i].squared(); break;
        case 4: o[i] = a[i].is_zero() ? F::zero() : a[i].inverse(); break;
        case 5: o[i] = -a[i]; break;
        case 6: {
            auto bi = a[i].as_bigint();
            std::memcpy((void *)&o[i], bi.data, sizeof(F));
            break;
        }
        case 7: {
            typename std::decay<decltype(a[i].mont_repr)>::type bi;
            std::memcpy(bi.data, (const void *)&a[i], sizeof(F));
            o[i] = F(bi);
            break;
        }
        }
    }
}

extern "C" void zref_fq_op(int op, const void *a, const void *b, void *o, size_t n)
{
    field_op<Fq>(op, (const Fq *)a, (const Fq *)b, (Fq *)o, n);
}
extern "C" void zref_fr_op(int op, const void *a, const void *b, void *o, size_t n)
{
    field_op<Fr>(op, (const Fr *)a, (const Fr *)b, (Fr *)o, n);
}

// libff::SHA512_rng<Fr>(seed + i), Montgomery form (rng.tcc:26-71).
extern "C" void zref_fr_rng(uint64_t seed, size_t n, void *out)
{
    Fr *o = (Fr *)out;
#pragma omp parallel for
    for (size_t i = 0; i < n; ++i)
        o[i] = libff::SHA512_rng<Fr>(seed + i);
}

// Signed digit of a plain (non-Montgomery) Fr bigint (field_utils.tcc:81-115).
extern "C" long zref_fr_signed_digit(const void *plain, size_t c, size_t digit_idx)
{
    libff::bigint<6> v;
    std::memcpy(v.data, plain, 48);
    return (long)libff::field_get_signed_digit(v, c, digit_idx);
}

extern "C" size_t zref_optimal_c(size_t n) { return libff::bdlo12_signed_optimal_c(n); }

// ---------------------------------------------------------------- groups --
// P_i = (i+1)*G, special (affine) form — test_multiexp.cpp:216-225 style bases.
template<typename G> static void bases_seq(size_t n, G *o)
{
    std::vector<G> v;
    v.reserve(n);
    G g = G::one();
    for (size_t i = 0; i < n; ++i) {
        v.push_back(g);
        g = g + G::one();
    }
    libff::batch_to_special<G>(v);
    for (size_t i = 0; i < n; ++i)
        o[i] = v[i];
}
extern "C" void zref_g1_bases_seq(size_t n, void *out) { bases_seq<G1>(n, (G1 *)out); }
extern "C" void zref_g2_bases_seq(size_t n, void *out) { bases_seq<G2>(n, (G2 *)out); }

// P_i = scalars[i]*G, special form ("random" bases).
template<typename G> static void bases_mul(const Fr *s, size_t n, G *o)
{
    std::vector<G> v(n);
#pragma omp parallel for
    for (size_t i = 0; i < n; ++i)
        v[i] = s[i] * G::one();
    libff::batch_to_special<G>(v);
    for (size_t i = 0; i < n; ++i)
        o[i] = v[i];
}
extern "C" void zref_g1_bases_mul(const void *s, size_t n, void *out)
{
    bases_mul<G1>((const Fr *)s, n, (G1 *)out);
}
extern "C" void zref_g2_bases_mul(const void *s, size_t n, void *out)
{
    bases_mul<G2>((const Fr *)s, n, (G2 *)out);
}

// The generator's fixed-base path as the reference runs it: get_exp_window_size / get_window_table /
// batch_exp or batch_exp_with_coeff (multiexp.tcc:776-947), then batch_to_special (:949-974).
template<typename G> static void ref_batch_exp(const G *base, const Fr *s, size_t n, const Fr *coeff, G *o)
{
    const size_t scalar_bits = Fr::size_in_bits();
    const size_t window = libff::get_exp_window_size<G>(n ? n : 1);
    libff::window_table<G> table = libff::get_window_table<G>(scalar_bits, window, *base);
    std::vector<Fr> v(s, s + n);
    std::vector<G> r = coeff ? libff::batch_exp_with_coeff<G, Fr>(scalar_bits, window, table, *coeff, v)
                             : libff::batch_exp<G, Fr>(scalar_bits, window, table, v);
    libff::batch_to_special<G>(r);
    for (size_t i = 0; i < n; ++i)
        o[i] = r[i];
}
extern "C" void zref_batch_exp(int group, const void *base, const void *s, size_t n, const void *coeff, void *out)
{
    if (group == 2)
        ref_batch_exp<G2>((const G2 *)base, (const Fr *)s, n, (const Fr *)coeff, (G2 *)out);
    else
        ref_batch_exp<G1>((const G1 *)base, (const Fr *)s, n, (const Fr *)coeff, (G1 *)out);
}

// The proving key's on-disk codec as libzeth writes it (groth16_snark.tcc:93-110): a query of n points
// through group_elements_write_bytes, a sparse G2/G1 query through knowledge_commitment_vector_write_bytes.
// Returns the number of bytes written (0 if `cap` is too small).
template<typename G> static size_t points_write_bytes(const G *pts, size_t n, void *out, size_t cap)
{
    std::vector<G> v(pts, pts + n);
    std::ostringstream ss;
    libzeth::group_elements_write_bytes(v, ss);
    const std::string b = ss.str();
    if (b.size() > cap)
        return 0;
    std::memcpy(out, b.data(), b.size());
    return b.size();
}
extern "C" size_t zref_points_write_bytes(int group, const void *pts, size_t n, void *out, size_t cap)
{
    return group == 2 ? points_write_bytes<G2>((const G2 *)pts, n, out, cap) : points_write_bytes<G1>((const G1 *)pts, n, out, cap);
}
extern "C" size_t zref_kc_write_bytes(const void *values, const size_t *indices, size_t n, size_t domain, void *out, size_t cap)
{
    typedef libsnark::knowledge_commitment<G2, G1> kc_t;
    libsnark::knowledge_commitment_vector<G2, G1> v;
    v.domain_size_ = domain;
    const kc_t *src = (const kc_t *)values;
    for (size_t i = 0; i < n; ++i) {
        v.values.push_back(src[i]);
        v.indices.push_back(indices[i]);
    }
    std::ostringstream ss;
    libzeth::knowledge_commitment_vector_write_bytes(v, ss);
    const std::string b = ss.str();
    if (b.size() > cap)
        return 0;
    std::memcpy(out, b.data(), b.size());
    return b.size();
}

// op: 0 operator+, 1 mixed_add (b special), 2 dbl(a), 3 to_special(a), 4 neg(a),
//     5 add(), 6 scalar-mul by Fr in b's first 48 bytes (b stride = 288)
template<typename G> static void group_op(int op, const G *a, const G *b, G *o, size_t n)
{
    for (size_t i = 0; i < n; ++i) {
        switch (op) {
        case 0: o[i] = a[i] + b[i]; break;
        case 1: o[i] = a[i].mixed_add(b[i]); break;
        case 2: o[i] = a[i].dbl(); break;
        case 3: { G t = a[i]; t.to_special(); o[i] = t; break; }
        case 4: o[i] = -a[i]; break;
        case 5: o[i] = a[i].add(b[i]); break;
        case 6: { Fr s; std::memcpy((void *)&s, (const void *)&b[i], 48); o[i] = s * a[i]; break; }
        }
    }
}
extern "C" void zref_g1_op(int op, const void *a, const void *b, void *o, size_t n)
{
    group_op<G1>(op, (const G1 *)a, (const G1 *)b, (G1 *)o, n);
}
extern "C" void zref_g2_op(int op, const void *a, const void *b, void *o, size_t n)
{
    group_op<G2>(op, (const G2 *)a, (const G2 *)b, (G2 *)o, n);
}
extern "C" int zref_g1_is_well_formed(const void *a) { return ((const G1 *)a)->is_well_formed(); }
extern "C" int zref_g2_is_well_formed(const void *a) { return ((const G2 *)a)->is_well_formed(); }

// ------------------------------------------------------------------- MSM --
// method: 0 = multi_exp, 1 = multi_exp_filter_one_zero; result in special form.
template<typename G>
static double msm(const G *bases, const Fr *scalars, size_t n, size_t chunks, int method, G *out)
{
    std::vector<G> b(bases, bases + n);
    std::vector<Fr> s(scalars, scalars + n);
    const double t0 = now_s();
    G r = method == 1
              ? libff::multi_exp_filter_one_zero<G, Fr, METHOD, FORM>(
                    b.begin(), b.end(), s.begin(), s.end(), chunks)
              : libff::multi_exp<G, Fr, METHOD, FORM>(
                    b.begin(), b.end(), s.begin(), s.end(), chunks);
    const double t1 = now_s();
    r.to_special();
    *out = r;
    return t1 - t0;
}
extern "C" double zref_msm_g1(const void *b, const void *s, size_t n, size_t chunks, int method, void *out)
{
    return msm<G1>((const G1 *)b, (const Fr *)s, n, chunks ? chunks : threads(), method, (G1 *)out);
}
extern "C" double zref_msm_g2(const void *b, const void *s, size_t n, size_t chunks, int method, void *out)
{
    return msm<G2>((const G2 *)b, (const Fr *)s, n, chunks ? chunks : threads(), method, (G2 *)out);
}

// Naive Σ s_i·P_i (independent cross-check of the oracle itself).
extern "C" double zref_msm_g1_naive(const void *b, const void *s, size_t n, void *out)
{
    std::vector<G1> bb((const G1 *)b, (const G1 *)b + n);
    std::vector<Fr> ss((const Fr *)s, (const Fr *)s + n);
    const double t0 = now_s();
    G1 r = libff::multi_exp<G1, Fr, libff::multi_exp_method_naive_plain, libff::multi_exp_base_form_normal>(
        bb.begin(), bb.end(), ss.begin(), ss.end(), 1);
    const double t1 = now_s();
    r.to_special();
    *(G1 *)out = r;
    return t1 - t0;
}

// kc_multi_exp_with_mixed_addition over a sparse (G2,G1) vector (kc_multiexp.tcc:31-113).
extern "C" double zref_msm_kc(
    const void *values, const uint64_t *indices, size_t nnz, size_t domain_size,
    const void *scalars, size_t n_scalars, size_t chunks, void *out)
{
    libsnark::knowledge_commitment_vector<G2, G1> v;
    v.values.assign((const KC *)values, (const KC *)values + nnz);
    v.indices.assign(indices, indices + nnz);
    v.domain_size_ = domain_size;
    std::vector<Fr> s((const Fr *)scalars, (const Fr *)scalars + n_scalars);
    const double t0 = now_s();
    KC r = libsnark::kc_multi_exp_with_mixed_addition<G2, G1, Fr, METHOD, FORM>(
        v, 0, n_scalars, s.begin(), s.end(), chunks ? chunks : threads());
    const double t1 = now_s();
    r.g.to_special();
    r.h.to_special();
    *(KC *)out = r;
    return t1 - t0;
}

// ------------------------------------------------------------------- NTT --
// flags bit0: inverse (iFFT incl. 1/m), bit1: coset. In place, natural order.
// Returns seconds, or -1 on a libfqfft exception.
extern "C" double zref_ntt(void *a, unsigned log_m, unsigned flags, const void *coset_g)
{
    const size_t m = (size_t)1 << log_m;
    try {
        libfqfft::basic_radix2_domain<Fr> dom(m);
        std::vector<Fr> v((const Fr *)a, (const Fr *)a + m);
        Fr g = Fr::multiplicative_generator;
        if (coset_g)
            std::memcpy((void *)&g, coset_g, 48);
        const double t0 = now_s();
        switch (flags & 3) {
        case 0: dom.FFT(v); break;
        case 1: dom.iFFT(v); break;
        case 2: dom.cosetFFT(v, g); break;
        case 3: dom.icosetFFT(v, g); break;
        }
        const double t1 = now_s();
        std::memcpy(a, (const void *)v.data(), m * 48);
        return t1 - t0;
    } catch (const std::exception &e) {
        return -1.0;
    }
}

// The transform part of r1cs_to_qap_witness_map (r1cs_to_qap.tcc:252-330, d1 = d2 = d3 = 0)
// on caller-supplied evaluation vectors, through the same evaluation_domain calls in the
// same order. h receives m + 1 elements. Returns seconds.
extern "C" double zref_qap_h(const void *aA_, const void *aB_, const void *aC_, unsigned log_m, void *h)
{
    const size_t m = (size_t)1 << log_m;
    const std::shared_ptr<libfqfft::evaluation_domain<Fr>> domain = libfqfft::get_evaluation_domain<Fr>(m);
    std::vector<Fr> aA((const Fr *)aA_, (const Fr *)aA_ + m), aB((const Fr *)aB_, (const Fr *)aB_ + m),
        aC((const Fr *)aC_, (const Fr *)aC_ + m);
    const double t0 = now_s();
    domain->iFFT(aA);
    domain->iFFT(aB);
    std::vector<Fr> coefficients_for_H(domain->m + 1, Fr::zero());
    domain->add_poly_Z(Fr::zero(), coefficients_for_H);
    const Fr g = Fr::multiplicative_generator;
    domain->cosetFFT(aA, g);
    domain->cosetFFT(aB, g);
    std::vector<Fr> &H_tmp = aA;
#ifdef MULTICORE
#pragma omp parallel for
#endif
    for (size_t i = 0; i < domain->m; ++i)
        H_tmp[i] = aA[i] * aB[i];
    domain->iFFT(aC);
    domain->cosetFFT(aC, g);
#ifdef MULTICORE
#pragma omp parallel for
#endif
    for (size_t i = 0; i < domain->m; ++i)
        H_tmp[i] = (H_tmp[i] - aC[i]);
    domain->divide_by_Z_on_coset(H_tmp);
    domain->icosetFFT(H_tmp, g);
#ifdef MULTICORE
#pragma omp parallel for
#endif
    for (size_t i = 0; i < domain->m; ++i)
        coefficients_for_H[i] += H_tmp[i];
    const double t1 = now_s();
    std::memcpy(h, (const void *)coefficients_for_H.data(), (m + 1) * 48);
    return t1 - t0;
}

// evaluate_all_lagrange_polynomials(t) of the size-2^log_m domain through get_evaluation_domain (so the
// hooked build exercises the device version); out receives m elements. Returns seconds.
extern "C" double zref_lagrange(unsigned log_m, const void *t48, void *out)
{
    const size_t m = (size_t)1 << log_m;
    Fr t;
    std::memcpy((void *)&t, t48, 48);
    const std::shared_ptr<libfqfft::evaluation_domain<Fr>> domain = libfqfft::get_evaluation_domain<Fr>(m);
    const double t0 = now_s();
    const std::vector<Fr> u = domain->evaluate_all_lagrange_polynomials(t);
    const double t1 = now_s();
    copy_out(out, u);
    return t1 - t0;
}

// omega for the size-2^log_m domain (field_utils.tcc:190-206); (g^m - 1)^-1 helper.
extern "C" void zref_domain_params(unsigned log_m, void *omega, void *z_inv_on_coset)
{
    const size_t m = (size_t)1 << log_m;
    libfqfft::basic_radix2_domain<Fr> dom(m);
    Fr w = dom.get_domain_element(1);
    std::memcpy(omega, (const void *)&w, 48);
    Fr z = dom.compute_vanishing_polynomial(Fr::multiplicative_generator).inverse();
    std::memcpy(z_inv_on_coset, (const void *)&z, 48);
}

// ----------------------------------------------------- Groth16 sessions --
// Deterministic synthetic R1CS in the spirit of r1cs_examples.tcc:25-100:
// a chain x_{i+2} = x_i (+|*) x_{i+1} with full-width values, num_inputs
// primary inputs, plus `n_bits` boolean variables (b*b=b) to mimic the
// aggregator witness' 0/1 entries.
extern "C" int zref_session_synthetic(size_t num_constraints, size_t num_inputs, size_t n_bits, uint64_t seed)
{
    using namespace libsnark;
    S_ptr.reset(new session_t());
    r1cs_constraint_system<Fr> cs;
    cs.primary_input_size = num_inputs;
    const size_t chain = num_constraints - n_bits;
    const size_t nvars = 2 + chain + n_bits;
    cs.auxiliary_input_size = nvars - num_inputs;
    std::vector<Fr> full(nvars);
    full[0] = libff::SHA512_rng<Fr>(seed * 1000003 + 0);
    full[1] = libff::SHA512_rng<Fr>(seed * 1000003 + 1);
    for (size_t i = 0; i < chain; ++i) {
        linear_combination<Fr> A, B, C;
        const Fr k = libff::SHA512_rng<Fr>(seed * 1000003 + 2 + i);
        if (i % 2) {
            A.add_term(i + 1, 1);
            A.add_term(i + 2, k);
            B.add_term(0, 1);
            C.add_term(i + 3, 1);
            full[i + 2] = full[i] + k * full[i + 1];
        } else {
            A.add_term(i + 1, 1);
            B.add_term(i + 2, 1);
            B.add_term(0, k);
            C.add_term(i + 3, 1);
            full[i + 2] = full[i] * (full[i + 1] + k);
        }
        cs.add_constraint(r1cs_constraint<Fr>(A, B, C));
    }
    for (size_t j = 0; j < n_bits; ++j) {
        const size_t v = 2 + chain + j;
        linear_combination<Fr> A, B, C;
        A.add_term(v + 1, 1);
        B.add_term(v + 1, 1);
        C.add_term(v + 1, 1);
        cs.add_constraint(r1cs_constraint<Fr>(A, B, C));
        full[v] = ((seed + j * 7) % 3 == 0) ? Fr::one() : Fr::zero();
    }
    S.cs = cs;
    S.primary.assign(full.begin(), full.begin() + num_inputs);
    S.aux.assign(full.begin() + num_inputs, full.end());
    S.have_cs = true;
    return S.cs.is_satisfied(S.primary, S.aux) ? 0 : 1;
}

#ifdef ZREF_WITH_AGGREGATOR
template<size_t N> static int aggregator_session(uint64_t seed)
{
    using namespace libzecale;
    using npp = libsnark::other_curve<ppT>;
    using nverifier = groth16_verifier_parameters<ppT>;
    using nsnark = typename nverifier::snark;
    S_ptr.reset(new session_t());
    test::dummy_app_wrapper<npp, nsnark> dummy_app;
    const typename nsnark::keypair nkp = dummy_app.generate_keypair();
    std::vector<libzeth::extended_proof<npp, nsnark>> nproofs;
    for (size_t i = 0; i < N; ++i)
        nproofs.push_back(dummy_app.prove(5 + i + seed, nkp.pk));
    std::array<const libzeth::extended_proof<npp, nsnark> *, N> batch;
    for (size_t i = 0; i < N; ++i)
        batch[i] = &nproofs[i];

    aggregator_circuit<ppT, snark, nverifier, N> agg(1);
    // Witness generation: the statements of aggregator_circuit::prove
    // (aggregator_circuit.tcc:136-157) without the final generate_proof call
    // (built with -fno-access-control so the protoboard can be read back).
    std::array<const libsnark::r1cs_primary_input<libff::Fr<npp>> *, N> nested_inputs{};
    const double tw0 = now_s();
    for (size_t i = 0; i < N; ++i) {
        nested_inputs[i] = &batch[i]->get_primary_inputs();
        agg._nested_proofs[i]->generate_r1cs_witness(batch[i]->get_proof());
    }
    agg._nested_vk->generate_r1cs_witness(nkp.vk);
    agg._nested_vk_hash_gadget->generate_r1cs_witness();
    agg._aggregator_gadget->generate_r1cs_witness(nested_inputs);
    agg._nested_proof_results_packer->generate_r1cs_witness_from_bits();
    const double tw1 = now_s();

    S.t_witness_gen = tw1 - tw0;
    // SURVEY 8(f2) prototype: the N nested proofs are witnessed independently (aggregator_circuit.tcc:136-146,
    // aggregator_gadget.tcc:94-111: disjoint protoboard variables, the shared processed verification key is
    // only read), so both loops run one nested proof per host thread. Must reproduce the serial assignment.
    {
        const std::vector<Fr> serial = agg._pb.full_variable_assignment();
        auto &gad = *agg._aggregator_gadget;
        const int nt = std::min<int>((int)N, threads());
        const double tp0 = now_s();
#pragma omp parallel for num_threads(nt) schedule(static, 1)
        for (size_t i = 0; i < N; ++i)
            agg._nested_proofs[i]->generate_r1cs_witness(batch[i]->get_proof());
        agg._nested_vk->generate_r1cs_witness(nkp.vk);
        agg._nested_vk_hash_gadget->generate_r1cs_witness();
        gad.vk_processor.generate_r1cs_witness();
#pragma omp parallel for num_threads(nt) schedule(static, 1)
        for (size_t i = 0; i < N; ++i) {
            const libff::bit_vector input_bits =
                libff::convert_field_element_vector_to_bit_vector<libff::Fr<npp>>(*(nested_inputs[i]));
            gad.nested_primary_inputs_bits[i].fill_with_bits(agg._pb, input_bits);
            gad.nested_primary_input_packers[i]->generate_r1cs_witness_from_bits();
            gad.verifiers[i]->generate_r1cs_witness();
        }
        agg._nested_proof_results_packer->generate_r1cs_witness_from_bits();
        const double tp1 = now_s();
        S.witness_gen_threads = nt;
        S.t_witness_gen_parallel = agg._pb.full_variable_assignment() == serial ? tp1 - tp0 : -1.0;
    }
    S.cs = agg.get_constraint_system();
    S.primary = agg._pb.primary_input();
    S.aux = agg._pb.auxiliary_input();
    S.have_cs = true;
    return S.cs.is_satisfied(S.primary, S.aux) ? 0 : 1;
}
#endif

extern "C" int zref_session_aggregator(size_t batch, uint64_t seed)
{
#ifdef ZREF_WITH_AGGREGATOR
    switch (batch) {
    case 1: return aggregator_session<1>(seed);
    case 2: return aggregator_session<2>(seed);
    case 4: return aggregator_session<4>(seed);
    case 8: return aggregator_session<8>(seed);
    case 16: return aggregator_session<16>(seed);
    case 32: return aggregator_session<32>(seed);
    default: return -2;
    }
#else
    (void)batch;
    (void)seed;
    return -3;
#endif
}

extern "C" double zref_session_witness_gen_seconds(void) { return S.t_witness_gen; }
extern "C" double zref_session_witness_gen_parallel_seconds(int *threads_used)
{
    if (threads_used)
        *threads_used = S.witness_gen_threads;
    return S.t_witness_gen_parallel;
}

// A session whose constraint system is given as three CSR matrices (row i of matrix k =
// constraints[i].{a,b,c}.terms; col 0 = the constant one) and whose assignment is `full` =
// [1, primary..., aux...]: lets bench.py hand the SAME synthetic batch-8-shaped system to the
// reference's own r1cs_to_qap_witness_map and to the device path.
extern "C" int zref_session_from_csr(
    size_t n_constraints, size_t n_inputs, size_t n_variables,
    const uint64_t *const row_ptr[3], const uint32_t *const col[3], const void *const val[3],
    const void *full)
{
    S_ptr.reset(new session_t());
    S.cs.primary_input_size = n_inputs;
    S.cs.auxiliary_input_size = n_variables - n_inputs;
    S.cs.constraints.reserve(n_constraints);
    for (size_t i = 0; i < n_constraints; ++i) {
        libsnark::linear_combination<Fr> lc[3];
        for (int k = 0; k < 3; ++k) {
            const Fr *v = (const Fr *)val[k];
            for (uint64_t t = row_ptr[k][i]; t < row_ptr[k][i + 1]; ++t)
                lc[k].terms.emplace_back(libsnark::linear_term<Fr>(libsnark::variable<Fr>(col[k][t]), v[t]));
        }
        S.cs.constraints.emplace_back(libsnark::r1cs_constraint<Fr>(lc[0], lc[1], lc[2]));
    }
    const Fr *f = (const Fr *)full;
    S.primary.assign(f + 1, f + 1 + n_inputs);
    S.aux.assign(f + 1 + n_inputs, f + 1 + n_variables);
    S.have_cs = true;
    return 0;
}

// The reference's own r1cs_to_qap_witness_map (r1cs_to_qap.tcc:210-344: serial mat-vecs + 7 transforms)
// on the session's system and assignment, d1 = d2 = d3 = 0 as the Groth16 prover calls it. h (may be
// NULL) receives the m + 1 coefficients. Returns seconds.
extern "C" double zref_session_witness_map(void *h)
{
    if (!S.have_cs)
        return -1;
    const double t0 = now_s();
    const libsnark::qap_witness<Fr> w = libsnark::r1cs_to_qap_witness_map(
        S.cs, S.primary, S.aux, Fr::zero(), Fr::zero(), Fr::zero(), true);
    const double t1 = now_s();
    if (h)
        copy_out(h, w.coefficients_for_H);
    return t1 - t0;
}

// The session's keypair to / from a file in libzeth's on-disk format (groth16_snark::keypair_write_bytes /
// keypair_read_bytes, groth16_snark.tcc:178-191): lets the plain build prove with a key the hooked
// build generated (large circuits: the CPU generator takes minutes). The reader
// needs a session with the same constraint system. Returns seconds, or a negative code.
extern "C" double zref_session_keypair_write(const char *path)
{
    if (!S.have_keys)
        return -1;
    const double t0 = now_s();
    std::ofstream f(path, std::ios::binary);
    if (!f)
        return -2;
    snark::keypair_write_bytes(*S.kp_ptr, f);
    f.close();
    return now_s() - t0;
}

extern "C" double zref_session_keypair_read(const char *path)
{
    if (!S.have_cs)
        return -1;
    const double t0 = now_s();
    std::ifstream f(path, std::ios::binary);
    if (!f)
        return -2;
    try {
        S.kp_ptr.reset(new libsnark::r1cs_gg_ppzksnark_keypair<ppT>());
        snark::keypair_read_bytes(*S.kp_ptr, f);
    } catch (const std::exception &e) {
        std::cerr << "zref_session_keypair_read: " << e.what() << std::endl;
        return -3;
    }
    S.have_keys = true;
    return now_s() - t0;
}

extern "C" double zref_session_setup(uint64_t seed)
{
    if (!S.have_cs)
        return -1;
    const double t0 = now_s();
    setup_from_seed(seed);
    return now_s() - t0;
}

// sizes[0..15]: num_constraints, num_variables, num_inputs, domain m,
// |A_query|, nnz(B_query), B domain size, |H_query|, |L_query|,
// nnz(A), nnz(B), nnz(C) of pk.constraint_system (post swap_AB_if_beneficial)
extern "C" int zref_session_sizes(uint64_t *sz)
{
    if (!S.have_keys)
        return -1;
    const auto &pk = S.kp_ptr->pk;
    const auto &cs = pk.constraint_system;
    size_t m = 1;
    while (m < cs.num_constraints() + cs.num_inputs() + 1)
        m <<= 1;
    sz[0] = cs.num_constraints();
    sz[1] = cs.num_variables();
    sz[2] = cs.num_inputs();
    sz[3] = m;
    sz[4] = pk.A_query.size();
    sz[5] = pk.B_query.values.size();
    sz[6] = pk.B_query.domain_size();
    sz[7] = pk.H_query.size();
    sz[8] = pk.L_query.size();
    size_t nz[3] = {0, 0, 0};
    for (const auto &c : cs.constraints) {
        nz[0] += c.a.terms.size();
        nz[1] += c.b.terms.size();
        nz[2] += c.c.terms.size();
    }
    sz[9] = nz[0];
    sz[10] = nz[1];
    sz[11] = nz[2];
    return 0;
}

// which: 0 A_query (288 B each), 1 B_query.values (576 B), 2 B_query.indices (u64),
// 3 H_query, 4 L_query, 5 {alpha_g1, beta_g1, delta_g1} (3x288),
// 6 {beta_g2, delta_g2} (2x288), 7 full assignment [1, primary, aux] (48 B each)
extern "C" int zref_session_get(int which, void *out)
{
    if (!S.have_keys)
        return -1;
    const auto &pk = S.kp_ptr->pk;
    switch (which) {
    case 0: copy_out(out, pk.A_query); break;
    case 1: copy_out(out, pk.B_query.values); break;
    case 2: {
        uint64_t *o = (uint64_t *)out;
        for (size_t i = 0; i < pk.B_query.indices.size(); ++i)
            o[i] = pk.B_query.indices[i];
        break;
    }
    case 3: copy_out(out, pk.H_query); break;
    case 4: copy_out(out, pk.L_query); break;
    case 5: {
        G1 *o = (G1 *)out;
        o[0] = pk.alpha_g1;
        o[1] = pk.beta_g1;
        o[2] = pk.delta_g1;
        break;
    }
    case 6: {
        G2 *o = (G2 *)out;
        o[0] = pk.beta_g2;
        o[1] = pk.delta_g2;
        break;
    }
    case 7: {
        Fr *o = (Fr *)out;
        o[0] = Fr::one();
        std::memcpy((void *)(o + 1), (const void *)S.primary.data(), S.primary.size() * 48);
        std::memcpy((void *)(o + 1 + S.primary.size()), (const void *)S.aux.data(), S.aux.size() * 48);
        break;
    }
    default: return -2;
    }
    return 0;
}

// CSR export of pk.constraint_system matrix `which` (0 A, 1 B, 2 C):
// row_ptr[num_constraints+1] (u64), col[nnz] (u32, variable index, 0 = constant one),
// val[nnz] (Fr Montgomery).
extern "C" int zref_session_csr(int which, uint64_t *row_ptr, uint32_t *col, void *val)
{
    if (!S.have_keys)
        return -1;
    const auto &cs = S.kp_ptr->pk.constraint_system;
    Fr *v = (Fr *)val;
    size_t k = 0;
    for (size_t i = 0; i < cs.constraints.size(); ++i) {
        row_ptr[i] = k;
        const auto &lc = which == 0 ? cs.constraints[i].a
                         : which == 1 ? cs.constraints[i].b
                                      : cs.constraints[i].c;
        for (const auto &t : lc.terms) {
            col[k] = (uint32_t)t.index;
            v[k] = t.coeff;
            ++k;
        }
    }
    row_ptr[cs.constraints.size()] = k;
    return 0;
}

// The evaluation vectors r1cs_to_qap_witness_map builds before its first transform
// (r1cs_to_qap.tcc:224-248, 293-296): aA[i] = <A_i, w>, aB[i] = <B_i, w>, aC[i] = <C_i, w> for
// i < num_constraints, aA[num_constraints + i] = (1, primary...)[i], zero padded to the
// power-of-two domain. Each buffer holds m Fr; returns m.
extern "C" long zref_session_abc(void *aA_, void *aB_, void *aC_)
{
    if (!S.have_keys)
        return -1;
    const auto &cs = S.kp_ptr->pk.constraint_system;
    const std::shared_ptr<libfqfft::evaluation_domain<Fr>> domain =
        libfqfft::get_evaluation_domain<Fr>(1ull << libff::log2(cs.num_constraints() + cs.num_inputs() + 1));
    std::vector<Fr> full = S.primary;
    full.insert(full.end(), S.aux.begin(), S.aux.end());
    std::vector<Fr> aA(domain->m, Fr::zero()), aB(domain->m, Fr::zero()), aC(domain->m, Fr::zero());
    for (size_t i = 0; i <= cs.num_inputs(); ++i)
        aA[i + cs.num_constraints()] = (i > 0 ? full[i - 1] : Fr::one());
    for (size_t i = 0; i < cs.num_constraints(); ++i) {
        aA[i] += cs.constraints[i].a.evaluate(full);
        aB[i] += cs.constraints[i].b.evaluate(full);
        aC[i] += cs.constraints[i].c.evaluate(full);
    }
    copy_out(aA_, aA);
    copy_out(aB_, aB);
    copy_out(aC_, aC);
    return (long)domain->m;
}

// Reference prove with fixed (r, s). Outputs: 576 proof bytes, the five MSM
// results in special form (At, Bt.g, Bt.h, Ht, Lt: 5 x 288 B) and
// coefficients_for_H (m+1 Fr). timings[6] = witness map, A, B, H, L, total.
extern "C" int zref_session_prove(const void *r48, const void *s48, size_t chunks, void *proof576, void *msm5, void *hcoeffs, double *timings)
{
    if (!S.have_keys)
        return -1;
    Fr r, s;
    std::memcpy((void *)&r, r48, 48);
    std::memcpy((void *)&s, s48, 48);
    auto proof = prove_with_rs(S.kp_ptr->pk, S.primary, S.aux, r, s, chunks ? chunks : threads());
    if (proof576)
        proof_bytes(proof, proof576);
    if (msm5) {
        G1 a = S.At, bh = S.Bt_h, h = S.Ht, l = S.Lt;
        G2 bg = S.Bt_g;
        a.to_special(); bh.to_special(); h.to_special(); l.to_special(); bg.to_special();
        char *o = (char *)msm5;
        std::memcpy(o, (const void *)&a, 288);
        std::memcpy(o + 288, (const void *)&bg, 288);
        std::memcpy(o + 576, (const void *)&bh, 288);
        std::memcpy(o + 864, (const void *)&h, 288);
        std::memcpy(o + 1152, (const void *)&l, 288);
    }
    if (hcoeffs)
        copy_out(hcoeffs, S.h_coeffs);
    if (timings) {
        timings[0] = S.t_witness_map; timings[1] = S.t_msm_a; timings[2] = S.t_msm_b;
        timings[3] = S.t_msm_h; timings[4] = S.t_msm_l; timings[5] = S.t_total;
    }
    const bool ok = snark::verify(S.primary, proof, S.kp_ptr->vk);
    return ok ? 0 : 1;
}

// Assemble + serialise a proof from externally computed MSM results (5 x 288 B,
// same order as above) with the reference's own group ops and codec
// (r1cs_gg_ppzksnark.tcc:582-599, groth16_snark.tcc:160-166). Returns 0 if it verifies.
extern "C" int zref_session_assemble(const void *r48, const void *s48, const void *msm5, void *proof576)
{
    if (!S.have_keys)
        return -1;
    const auto &pk = S.kp_ptr->pk;
    Fr r, s;
    std::memcpy((void *)&r, r48, 48);
    std::memcpy((void *)&s, s48, 48);
    G1 At, Bt_h, Ht, Lt;
    G2 Bt_g;
    const char *p = (const char *)msm5;
    std::memcpy((void *)&At, p, 288);
    std::memcpy((void *)&Bt_g, p + 288, 288);
    std::memcpy((void *)&Bt_h, p + 576, 288);
    std::memcpy((void *)&Ht, p + 864, 288);
    std::memcpy((void *)&Lt, p + 1152, 288);
    G1 g1_A = pk.alpha_g1 + At + r * pk.delta_g1;
    G1 g1_B = pk.beta_g1 + Bt_h + s * pk.delta_g1;
    G2 g2_B = pk.beta_g2 + Bt_g + s * pk.delta_g2;
    G1 g1_C = Ht + Lt + s * g1_A + r * g1_B - (r * s) * pk.delta_g1;
    libsnark::r1cs_gg_ppzksnark_proof<ppT> proof(std::move(g1_A), std::move(g2_B), std::move(g1_C));
    proof_bytes(proof, proof576);
    return snark::verify(S.primary, proof, S.kp_ptr->vk) ? 0 : 1;
}

// Unmodified reference prover (random r, s): timing only. Returns seconds.
extern "C" double zref_session_prove_stock(void)
{
    if (!S.have_keys)
        return -1;
    const double t0 = now_s();
    auto proof = snark::generate_proof(S.kp_ptr->pk, S.primary, S.aux);
    const double t1 = now_s();
    return snark::verify(S.primary, proof, S.kp_ptr->vk) ? (t1 - t0) : -2.0;
}


// ---------------------------------------------------------------- hooked build
// When this file is compiled with `-include zecale_gpu_hooks.hpp` (make ref-hooked) every
// multi_exp / kc_multi_exp / evaluation-domain call above — including the ones inside the
// unmodified r1cs_gg_ppzksnark_prover behind zref_session_prove_stock — runs on the GPU through
// libzecale_gpu.so. That build is the drop-in demonstration, not an oracle.
extern "C" int zref_is_hooked(void)
{
#ifdef ZECALE_GPU_HOOKS_HPP
    return 1;
#else
    return 0;
#endif
}

extern "C" int zref_hooks_pin_pk(void)
{
#ifdef ZECALE_GPU_HOOKS_HPP
    if (!S.have_keys)
        return -1;
    try {
        zecale_gpu_hooks::pin_proving_key(S.kp_ptr->pk);
    } catch (const std::exception &e) {
        std::cerr << "zref_hooks_pin_pk: " << e.what() << std::endl;
        return 1;
    }
    return 0;
#else
    return -3;
#endif
}

extern "C" int zref_hooks_unpin(void)
{
#ifdef ZECALE_GPU_HOOKS_HPP
    zecale_gpu_hooks::unpin_all();
    return 0;
#else
    return -3;
#endif
}

// Replace the session's witness with `full` = [1, primary..., aux...] (num_variables + 1 Fr), as
// exported by zref_session_get(7): lets two builds of this harness prove the same statement even
// though nested-proof generation draws from std::random_device. Returns 0 if it satisfies the cs.
extern "C" int zref_session_set_assignment(const void *full, size_t count)
{
    if (!S.have_cs || count != S.cs.num_variables() + 1)
        return -1;
    const Fr *f = (const Fr *)full;
    S.primary.assign(f + 1, f + 1 + S.cs.num_inputs());
    S.aux.assign(f + 1 + S.cs.num_inputs(), f + count);
    return S.cs.is_satisfied(S.primary, S.aux) ? 0 : 1;
}

// Hooked build only: zecale_gpu_hooks::prove_with_rs — what the specialised
// libsnark::r1cs_gg_ppzksnark_prover<bw6_761_pp> runs — with (r, s) given. Uses the fused device
// path when the key is resident (zref_hooks_pin_pk). Outputs as zref_session_prove; timings[0] =
// seconds for the whole call. Returns 0 if the proof verifies.
extern "C" int zref_session_prove_fused(const void *r48, const void *s48, void *proof576, void *msm5, double *timings)
{
#ifdef ZECALE_GPU_HOOKS_HPP
    if (!S.have_keys)
        return -1;
    Fr r, s;
    std::memcpy((void *)&r, r48, 48);
    std::memcpy((void *)&s, s48, 48);
    zecale_gpu_hooks::evaluations_t e;
    const double t0 = now_s();
    auto proof = zecale_gpu_hooks::prove_with_rs(S.kp_ptr->pk, S.primary, S.aux, true, r, s, &e);
    const double t1 = now_s();
    if (timings)
        timings[0] = t1 - t0;
    if (proof576)
        proof_bytes(proof, proof576);
    if (msm5) {
        e.At.to_special(); e.Bt_g.to_special(); e.Bt_h.to_special(); e.Ht.to_special(); e.Lt.to_special();
        char *o = (char *)msm5;
        std::memcpy(o, (const void *)&e.At, 288);
        std::memcpy(o + 288, (const void *)&e.Bt_g, 288);
        std::memcpy(o + 576, (const void *)&e.Bt_h, 288);
        std::memcpy(o + 864, (const void *)&e.Ht, 288);
        std::memcpy(o + 1152, (const void *)&e.Lt, 288);
    }
    return snark::verify(S.primary, proof, S.kp_ptr->vk) ? 0 : 1;
#else
    (void)r48; (void)s48; (void)proof576; (void)msm5; (void)timings;
    return -3;
#endif
}

// Hooked build only: `count` proofs of the session's statement back to back through
// zecale_gpu_hooks::prove_async / collect (up to pipeline_depth() in flight), proof i using the
// randomness pair rs[i % n_rs] (each 96 bytes: r then s). proofs receives count x 576 bytes.
// Returns the wall-clock seconds for all of them (host assembly and serialisation included),
// or a negative code.
extern "C" double zref_session_prove_pipeline(size_t count, const void *rs, size_t n_rs, void *proofs)
{
#ifdef ZECALE_GPU_HOOKS_HPP
    if (!S.have_keys || !n_rs)
        return -1;
    namespace H = zecale_gpu_hooks;
    const Fr *rsv = (const Fr *)rs;
    const size_t depth = H::pipeline_depth();
    std::vector<H::ticket_t *> tickets(count, nullptr);
    try {
        const double t0 = now_s();
        for (size_t i = 0; i < count + depth; ++i) {
            if (i >= depth && i - depth < count) {
                const size_t j = i - depth;
                auto proof = H::collect(tickets[j]);
                tickets[j] = nullptr;
                proof_bytes(proof, (char *)proofs + 576 * j);
            }
            if (i < count) {
                tickets[i] = H::prove_async(S.kp_ptr->pk, S.primary, S.aux, true, rsv[2 * (i % n_rs)], rsv[2 * (i % n_rs) + 1]);
                if (!tickets[i])
                    return -4; // key not resident
            }
        }
        return now_s() - t0;
    } catch (const std::exception &e) {
        std::cerr << "zref_session_prove_pipeline: " << e.what() << std::endl;
        return -2;
    }
#else
    (void)count; (void)rs; (void)n_rs; (void)proofs;
    return -3;
#endif
}

// Hooked build only: `n_threads` host threads prove the session's statement at the same time, each
// `per_thread` times, through the specialised prover body with (r, s) given — what concurrent gRPC
// workers would do (aggregator_server.cpp:399-415). proofs: n_threads * per_thread x 576 bytes.
extern "C" double zref_session_prove_threads(size_t n_threads, size_t per_thread, const void *r48, const void *s48, void *proofs)
{
#ifdef ZECALE_GPU_HOOKS_HPP
    if (!S.have_keys)
        return -1;
    Fr r, s;
    std::memcpy((void *)&r, r48, 48);
    std::memcpy((void *)&s, s48, 48);
    std::vector<std::thread> th;
    std::vector<int> rc(n_threads, 0);
    const double t0 = now_s();
    for (size_t t = 0; t < n_threads; ++t)
        th.emplace_back([&, t] {
            try {
                for (size_t k = 0; k < per_thread; ++k) {
                    auto proof = zecale_gpu_hooks::prove_with_rs(S.kp_ptr->pk, S.primary, S.aux, true, r, s, nullptr);
                    proof_bytes(proof, (char *)proofs + 576 * (t * per_thread + k));
                }
            } catch (const std::exception &e) {
                std::cerr << "zref_session_prove_threads: " << e.what() << std::endl;
                rc[t] = 1;
            }
        });
    for (auto &x : th)
        x.join();
    for (int v : rc)
        if (v)
            return -2;
    return now_s() - t0;
#else
    (void)n_threads; (void)per_thread; (void)r48; (void)s48; (void)proofs;
    return -3;
#endif
}
