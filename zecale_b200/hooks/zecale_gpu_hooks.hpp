// zecale_gpu_hooks.hpp — routes the reference prover's UNCHANGED call sites to libzecale_gpu.so.
//
// Force-include this header ahead of libsnark in every translation unit that proves
// (g++ -include zecale_gpu_hooks.hpp ...). It declares explicit full
// specialisations of the function templates that libsnark::r1cs_gg_ppzksnark_prover
// (libsnark r1cs_gg_ppzksnark.tcc:460-603) and r1cs_to_qap_witness_map (r1cs_to_qap.tcc:210-344)
// instantiate for ppT = bw6_761_pp, so that, with no edit to libsnark / libzeth / libzecale:
//
//   libff::multi_exp<G1|G2, Fr, BDLO12_signed, special>                 multiexp.tcc:648-688
//   libff::multi_exp_filter_one_zero<G1|G2, Fr, BDLO12_signed, special> multiexp.tcc:695-757
//        -> zecale_gpu_msm_g1 / zecale_gpu_msm_g2 (or the pinned image of the same vector)
//   libsnark::kc_multi_exp_with_mixed_addition<G2, G1, Fr, ...>         kc_multiexp.tcc:31-113
//        -> one G2 and one G1 MSM over the sparse B query (stride 576, scalar_index = indices)
//   libfqfft::get_evaluation_domain<bw6_761_Fr>                         get_evaluation_domain.tcc:33-69
//        -> gpu_radix2_domain, whose FFT / iFFT / cosetFFT / icosetFFT call zecale_gpu_ntt
//
// Semantics kept: same return types and values (group elements are canonical once normalised,
// so proofs serialise to the same bytes), same exceptions on size mismatch
// (libfqfft::DomainSizeException), std::runtime_error carrying zecale_gpu_last_error() on any
// device failure (the aggregator server maps std::exception to INVALID_ARGUMENT,
// aggregator_server.cpp:338-341). There is no CPU fallback: if the library cannot reach a B200
// the prover throws.
#ifndef ZECALE_GPU_HOOKS_HPP
#define ZECALE_GPU_HOOKS_HPP

#include <cstdlib>
#include <cstring>
#include <map>
#include <memory>
#include <mutex>
#include <stdexcept>
#include <string>
#include <vector>

#include <libff/algebra/curves/bw6_761/bw6_761_pp.hpp>
#include <libff/algebra/scalar_multiplication/multiexp.hpp>
#include <libfqfft/evaluation_domain/domains/basic_radix2_domain.hpp>
#include <libfqfft/evaluation_domain/get_evaluation_domain.hpp>
#include <libsnark/knowledge_commitment/kc_multiexp.hpp>

#include "zecale_gpu.h"

namespace zecale_gpu_hooks
{

using Fr = libff::bw6_761_Fr;
using G1 = libff::bw6_761_G1;
using G2 = libff::bw6_761_G2;
using KC = libsnark::knowledge_commitment<G2, G1>;

static_assert(sizeof(Fr) == 48, "bw6_761_Fr must be 6 x u64 (bigint<6> mont_repr)");
static_assert(sizeof(G1) == 288 && sizeof(G2) == 288, "bw6_761 G1/G2 must be {X,Y,Z} of 12 x u64");
static_assert(sizeof(KC) == 576, "knowledge_commitment<G2,G1> must be {G2, G1}");
static_assert(sizeof(size_t) == 8, "sparse_vector indices are passed as u64");

struct state_t {
    std::mutex mu;
    zecale_gpu_ctx *ctx = nullptr;
    // explicit residency registry: first-base address -> pinned image (see pin_bases below)
    struct pin_t {
        zecale_gpu_bases *handle;
        size_t n;
    };
    std::map<const void *, pin_t> pins;
};

inline state_t &state()
{
    static state_t s;
    return s;
}

/// The process-wide context, created on first use on device $ZECALE_GPU_DEVICE (default: current).
inline zecale_gpu_ctx *context()
{
    state_t &s = state();
    std::lock_guard<std::mutex> lk(s.mu);
    if (!s.ctx) {
        const char *dev = std::getenv("ZECALE_GPU_DEVICE");
        int d = dev ? std::atoi(dev) : 0;
        const int rc = zecale_gpu_init(dev ? &d : nullptr, dev ? 1 : 0, &s.ctx);
        if (rc != ZECALE_GPU_OK) {
            s.ctx = nullptr;
            throw std::runtime_error("zecale_gpu_init failed (code " + std::to_string(rc) + "): no usable sm_100 device");
        }
    }
    return s.ctx;
}

inline void check(int rc, const char *what)
{
    if (rc != ZECALE_GPU_OK) {
        throw std::runtime_error(std::string(what) + ": " + zecale_gpu_last_error(state().ctx));
    }
}

/// Optional residency: upload a constant base vector (a proving-key query) once. Later MSM calls
/// whose first base has the same address (and n <= the pinned count) use the device image. The
/// caller promises the vector is neither modified nor freed until unpin_all(). Off by default
/// because a content-agnostic pointer cache would be unsafe (SURVEY.md §8b "Ownership").
inline void pin_bases(const void *bases, size_t stride, size_t n, int group)
{
    if (n == 0) return;
    zecale_gpu_ctx *ctx = context();
    zecale_gpu_bases *h = nullptr;
    check(zecale_gpu_pin_bases(ctx, bases, stride, n, group, &h), "zecale_gpu_pin_bases");
    state_t &s = state();
    std::lock_guard<std::mutex> lk(s.mu);
    s.pins[bases] = {h, n};
}

/// Pin the four queries of a Groth16 proving key (A, B as G2 and G1 halves, H, L).
template<typename PK> inline void pin_proving_key(const PK &pk)
{
    pin_bases(pk.A_query.data(), sizeof(G1), pk.A_query.size(), ZECALE_GPU_G1);
    pin_bases(pk.H_query.data(), sizeof(G1), pk.H_query.size(), ZECALE_GPU_G1);
    pin_bases(pk.L_query.data(), sizeof(G1), pk.L_query.size(), ZECALE_GPU_G1);
    if (!pk.B_query.values.empty()) {
        const char *b = reinterpret_cast<const char *>(pk.B_query.values.data());
        pin_bases(b, sizeof(KC), pk.B_query.values.size(), ZECALE_GPU_G2);
        pin_bases(b + sizeof(G2), sizeof(KC), pk.B_query.values.size(), ZECALE_GPU_G1);
    }
}

inline void unpin_all()
{
    state_t &s = state();
    std::lock_guard<std::mutex> lk(s.mu);
    for (auto &kv : s.pins) zecale_gpu_unpin_bases(s.ctx, kv.second.handle);
    s.pins.clear();
}

inline void shutdown()
{
    unpin_all();
    state_t &s = state();
    std::lock_guard<std::mutex> lk(s.mu);
    if (s.ctx) zecale_gpu_shutdown(s.ctx);
    s.ctx = nullptr;
}

/// out = sum_i scalars[index ? index[i] : i] * base(i); group 1 = G1, 2 = G2.
template<typename GroupT>
inline GroupT msm(int group, const void *bases, size_t stride, const Fr *scalars, const size_t *index, size_t n)
{
    zecale_gpu_ctx *ctx = context();
    GroupT out; // any valid object; every byte is overwritten below
    unsigned char buf[288];
    zecale_gpu_bases *pinned = nullptr;
    {
        state_t &s = state();
        std::lock_guard<std::mutex> lk(s.mu);
        auto it = s.pins.find(bases);
        if (it != s.pins.end() && n <= it->second.n) pinned = it->second.handle;
    }
    if (pinned) {
        check(zecale_gpu_msm_pinned(ctx, pinned, scalars, sizeof(Fr), index, n, buf), "zecale_gpu_msm_pinned");
    } else if (group == ZECALE_GPU_G1) {
        check(zecale_gpu_msm_g1(ctx, bases, stride, scalars, sizeof(Fr), index, n, buf), "zecale_gpu_msm_g1");
    } else {
        check(zecale_gpu_msm_g2(ctx, bases, stride, scalars, sizeof(Fr), index, n, buf), "zecale_gpu_msm_g2");
    }
    std::memcpy(static_cast<void *>(&out), buf, sizeof buf); // libff special form: (x, y, one) or (0, one, 0)
    return out;
}

template<typename GroupT>
inline GroupT msm_iter(
    int group,
    typename std::vector<GroupT>::const_iterator vec_start,
    typename std::vector<GroupT>::const_iterator vec_end,
    typename std::vector<Fr>::const_iterator scalar_start,
    typename std::vector<Fr>::const_iterator scalar_end)
{
    const size_t n = vec_end - vec_start;
    if (static_cast<size_t>(scalar_end - scalar_start) != n) {
        throw std::invalid_argument("multi_exp: base and scalar ranges differ in length");
    }
    if (n == 0) return GroupT::zero();
    return msm<GroupT>(group, &*vec_start, sizeof(GroupT), &*scalar_start, nullptr, n);
}

/// evaluation_domain<Fr> of size m = 2^k whose transforms run on the device. The scalar helpers the
/// witness map also calls (add_poly_Z, divide_by_Z_on_coset, ...) are the reference's own code on a
/// contained basic_radix2_domain (basic_radix2_domain.tcc:91-126).
class gpu_radix2_domain : public libfqfft::evaluation_domain<Fr>
{
public:
    explicit gpu_radix2_domain(const size_t m)
        : libfqfft::evaluation_domain<Fr>(m), cpu_(m), log_m_(libff::log2(m))
    {
    }

    void FFT(std::vector<Fr> &a) override { run(a, 0, nullptr); }
    void iFFT(std::vector<Fr> &a) override { run(a, ZECALE_GPU_NTT_INVERSE, nullptr); }
    void cosetFFT(std::vector<Fr> &a, const Fr &g) override { run(a, ZECALE_GPU_NTT_COSET, &g); }
    void icosetFFT(std::vector<Fr> &a, const Fr &g) override
    {
        run(a, ZECALE_GPU_NTT_INVERSE | ZECALE_GPU_NTT_COSET, &g);
    }
    std::vector<Fr> evaluate_all_lagrange_polynomials(const Fr &t) override
    {
        return cpu_.evaluate_all_lagrange_polynomials(t);
    }
    Fr get_domain_element(const size_t idx) override { return cpu_.get_domain_element(idx); }
    Fr compute_vanishing_polynomial(const Fr &t) override { return cpu_.compute_vanishing_polynomial(t); }
    void add_poly_Z(const Fr &coeff, std::vector<Fr> &H) override { cpu_.add_poly_Z(coeff, H); }
    void divide_by_Z_on_coset(std::vector<Fr> &P) override { cpu_.divide_by_Z_on_coset(P); }

private:
    void run(std::vector<Fr> &a, unsigned flags, const Fr *g)
    {
        if (a.size() != this->m) {
            throw libfqfft::DomainSizeException("gpu_radix2: expected a.size() == this->m");
        }
        check(zecale_gpu_ntt(context(), a.data(), (unsigned)log_m_, flags, g), "zecale_gpu_ntt");
    }
    libfqfft::basic_radix2_domain<Fr> cpu_;
    size_t log_m_;
};

} // namespace zecale_gpu_hooks

// ---------------------------------------------------------------- the specialisations
namespace libfqfft
{

/// get_evaluation_domain<bw6_761_Fr>: the reference's selection order (get_evaluation_domain.tcc:33-69)
/// with the first choice — the only one Groth16 reaches, since r1cs_to_qap_witness_map is called
/// with force_pow_2_domain (r1cs_to_qap.tcc:224-227) — served by the device.
template<>
inline std::shared_ptr<evaluation_domain<libff::bw6_761_Fr>> get_evaluation_domain<libff::bw6_761_Fr>(
    const size_t min_size)
{
    using FieldT = libff::bw6_761_Fr;
    std::shared_ptr<evaluation_domain<FieldT>> result;
    const size_t big = 1ul << (libff::log2(min_size) - 1);
    const size_t small = min_size - big;
    const size_t rounded_small = (1ul << libff::log2(small));
    if (basic_radix2_domain<FieldT>::valid_for_size(min_size)) {
        result.reset(new zecale_gpu_hooks::gpu_radix2_domain(min_size));
    } else if (extended_radix2_domain<FieldT>::valid_for_size(min_size)) {
        result.reset(new extended_radix2_domain<FieldT>(min_size));
    } else if (step_radix2_domain<FieldT>::valid_for_size(min_size)) {
        result.reset(new step_radix2_domain<FieldT>(min_size));
    } else if (basic_radix2_domain<FieldT>::valid_for_size(big + rounded_small)) {
        result.reset(new zecale_gpu_hooks::gpu_radix2_domain(big + rounded_small));
    } else if (extended_radix2_domain<FieldT>::valid_for_size(big + rounded_small)) {
        result.reset(new extended_radix2_domain<FieldT>(big + rounded_small));
    } else if (step_radix2_domain<FieldT>::valid_for_size(big + rounded_small)) {
        result.reset(new step_radix2_domain<FieldT>(big + rounded_small));
    } else if (geometric_sequence_domain<FieldT>::valid_for_size(min_size)) {
        result.reset(new geometric_sequence_domain<FieldT>(min_size));
    } else if (arithmetic_sequence_domain<FieldT>::valid_for_size(min_size)) {
        result.reset(new arithmetic_sequence_domain<FieldT>(min_size));
    } else {
        throw DomainSizeException("get_evaluation_domain: no matching domain");
    }
    return result;
}

} // namespace libfqfft

namespace libff
{

#define ZECALE_GPU_MSM_SPECIALISE(FN, GROUP, GID)                                                    \
    template<>                                                                                       \
    inline GROUP FN<GROUP, bw6_761_Fr, multi_exp_method_BDLO12_signed, multi_exp_base_form_special>( \
        std::vector<GROUP>::const_iterator vec_start,                                                \
        std::vector<GROUP>::const_iterator vec_end,                                                  \
        std::vector<bw6_761_Fr>::const_iterator scalar_start,                                        \
        std::vector<bw6_761_Fr>::const_iterator scalar_end,                                          \
        const size_t /*chunks: the device shards the sum itself*/)                                   \
    {                                                                                                \
        return zecale_gpu_hooks::msm_iter<GROUP>(GID, vec_start, vec_end, scalar_start, scalar_end); \
    }

// multi_exp (multiexp.tcc:648-688) and multi_exp_filter_one_zero (:695-757). The 0/1 pre-filter is
// not needed on the device: a zero scalar has only zero digits and a one is an ordinary bucket entry.
ZECALE_GPU_MSM_SPECIALISE(multi_exp, bw6_761_G1, ZECALE_GPU_G1)
ZECALE_GPU_MSM_SPECIALISE(multi_exp, bw6_761_G2, ZECALE_GPU_G2)
ZECALE_GPU_MSM_SPECIALISE(multi_exp_filter_one_zero, bw6_761_G1, ZECALE_GPU_G1)
ZECALE_GPU_MSM_SPECIALISE(multi_exp_filter_one_zero, bw6_761_G2, ZECALE_GPU_G2)
#undef ZECALE_GPU_MSM_SPECIALISE

} // namespace libff

namespace libsnark
{

/// kc_multi_exp_with_mixed_addition<G2, G1, Fr, BDLO12_signed, special> (kc_multiexp.tcc:31-113):
/// the sparse walk over vec.indices in [min_idx, max_idx) becomes the scalar_index argument; the G2
/// and G1 halves of the 576-byte pairs are two strided MSMs over the same index list.
template<>
inline knowledge_commitment<libff::bw6_761_G2, libff::bw6_761_G1> kc_multi_exp_with_mixed_addition<
    libff::bw6_761_G2,
    libff::bw6_761_G1,
    libff::bw6_761_Fr,
    libff::multi_exp_method_BDLO12_signed,
    libff::multi_exp_base_form_special>(
    const knowledge_commitment_vector<libff::bw6_761_G2, libff::bw6_761_G1> &vec,
    const size_t min_idx,
    const size_t max_idx,
    std::vector<libff::bw6_761_Fr>::const_iterator scalar_start,
    std::vector<libff::bw6_761_Fr>::const_iterator scalar_end,
    const size_t /*chunks*/)
{
    using namespace zecale_gpu_hooks;
    auto first = std::lower_bound(vec.indices.begin(), vec.indices.end(), min_idx);
    auto last = std::lower_bound(first, vec.indices.end(), max_idx);
    const size_t offset = first - vec.indices.begin();
    const size_t n = last - first;
    if (n == 0) return KC::zero();
    const size_t scalar_length = scalar_end - scalar_start;
    if (*(last - 1) - min_idx >= scalar_length) {
        throw std::invalid_argument("kc_multi_exp_with_mixed_addition: index beyond the scalar range");
    }
    std::vector<size_t> shifted;
    const size_t *index = vec.indices.data() + offset;
    if (min_idx != 0) {
        shifted.assign(first, last);
        for (size_t &i : shifted) i -= min_idx;
        index = shifted.data();
    }
    const char *base = reinterpret_cast<const char *>(vec.values.data() + offset);
    const Fr *scalars = &*scalar_start;
    KC out;
    out.g = msm<G2>(ZECALE_GPU_G2, base, sizeof(KC), scalars, index, n);
    out.h = msm<G1>(ZECALE_GPU_G1, base + sizeof(G2), sizeof(KC), scalars, index, n);
    return out;
}

} // namespace libsnark

#endif // ZECALE_GPU_HOOKS_HPP
