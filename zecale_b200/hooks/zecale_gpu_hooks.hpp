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

#include <condition_variable>
#include <cstdlib>
#include <cstring>
#include <future>
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
#include <libsnark/zk_proof_systems/ppzksnark/r1cs_gg_ppzksnark/r1cs_gg_ppzksnark.hpp>

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
    // Proof slots of the fused prover: each in-flight proof owns one context (its own lanes, streams
    // and scratch; the pinned key images are shared) and a pinned staging buffer for its assignment.
    // A thread that proves takes a free slot for the whole enqueue + synchronise sequence, so
    // concurrent provers (the aggregator server's gRPC workers, aggregator_server.cpp:399-415) never
    // interleave on a context, and prove_async() keeps several proofs in flight from one thread.
    struct slot_t {
        zecale_gpu_ctx *ctx = nullptr;
        void *staging = nullptr;
        size_t staging_bytes = 0;
        bool busy = false;
    };
    std::vector<slot_t> slots;
    std::condition_variable slot_free;
    // explicit residency registry: first-base address -> pinned image (see pin_bases below)
    struct pin_t {
        zecale_gpu_bases *handle;
        size_t n;
    };
    std::map<const void *, pin_t> pins;
    // whole-key images for the fused prover (pin_proving_key): key = address of pk.A_query[0]
    struct key_t {
        zecale_gpu_bases *A, *B2, *B1, *H, *L;
        zecale_gpu_r1cs *r1cs;
    };
    std::map<const void *, key_t> keys;
};

inline state_t &state()
{
    static state_t s;
    return s;
}

/// Devices of every context the hooks create: $ZECALE_GPU_DEVICES = "0,1,2,3" (a multi-device context:
/// the MSMs of a resident key are sharded over them), else $ZECALE_GPU_DEVICE, else the current device.
inline std::vector<int> device_list()
{
    std::vector<int> out;
    if (const char *l = std::getenv("ZECALE_GPU_DEVICES")) {
        const char *p = l;
        while (*p) {
            char *end = nullptr;
            const long v = std::strtol(p, &end, 10);
            if (end == p) break;
            out.push_back((int)v);
            p = *end == ',' ? end + 1 : end;
        }
    } else if (const char *d = std::getenv("ZECALE_GPU_DEVICE")) {
        out.push_back(std::atoi(d));
    }
    return out;
}

/// The process-wide context, created on first use (see device_list).
inline zecale_gpu_ctx *context()
{
    state_t &s = state();
    std::lock_guard<std::mutex> lk(s.mu);
    if (!s.ctx) {
        const std::vector<int> devs = device_list();
        const int rc = zecale_gpu_init(devs.empty() ? nullptr : devs.data(), (int)devs.size(), &s.ctx);
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

inline zecale_gpu_bases *pinned_handle(const void *bases)
{
    state_t &s = state();
    std::lock_guard<std::mutex> lk(s.mu);
    auto it = s.pins.find(bases);
    return it == s.pins.end() ? nullptr : it->second.handle;
}

/// Make a Groth16 proving key device-resident: the four queries (A; B as its G2 and G1 halves; H; L)
/// and the CSR image of pk.constraint_system. Afterwards r1cs_gg_ppzksnark_prover on this key runs
/// the fused path (see prove_fused). The key must outlive unpin_all().
template<typename PK> inline void pin_proving_key(const PK &pk)
{
    pin_bases(pk.A_query.data(), sizeof(G1), pk.A_query.size(), ZECALE_GPU_G1);
    pin_bases(pk.H_query.data(), sizeof(G1), pk.H_query.size(), ZECALE_GPU_G1);
    pin_bases(pk.L_query.data(), sizeof(G1), pk.L_query.size(), ZECALE_GPU_G1);
    const char *b = reinterpret_cast<const char *>(pk.B_query.values.data());
    if (!pk.B_query.values.empty()) {
        pin_bases(b, sizeof(KC), pk.B_query.values.size(), ZECALE_GPU_G2);
        pin_bases(b + sizeof(G2), sizeof(KC), pk.B_query.values.size(), ZECALE_GPU_G1);
    }
    // CSR image of the constraint system (linear_combination terms: index 0 = the constant one)
    const auto &cs = pk.constraint_system;
    std::vector<uint64_t> rp[3];
    std::vector<uint32_t> col[3];
    std::vector<Fr> val[3];
    for (int k = 0; k < 3; k++) rp[k].reserve(cs.constraints.size() + 1);
    for (const auto &c : cs.constraints) {
        const libsnark::linear_combination<Fr> *lc[3] = {&c.a, &c.b, &c.c};
        for (int k = 0; k < 3; k++) {
            rp[k].push_back(col[k].size());
            for (const auto &t : lc[k]->terms) {
                col[k].push_back((uint32_t)t.index);
                val[k].push_back(t.coeff);
            }
        }
    }
    for (int k = 0; k < 3; k++) rp[k].push_back(col[k].size());
    const uint64_t *rpp[3] = {rp[0].data(), rp[1].data(), rp[2].data()};
    const uint32_t *cp[3] = {col[0].data(), col[1].data(), col[2].data()};
    const void *vp[3] = {val[0].data(), val[1].data(), val[2].data()};
    zecale_gpu_r1cs *r1cs = nullptr;
    check(zecale_gpu_r1cs_load(context(), cs.num_constraints(), cs.num_inputs(), cs.num_variables(), rpp, cp, vp, &r1cs),
          "zecale_gpu_r1cs_load");
    state_t::key_t key{pinned_handle(pk.A_query.data()), pinned_handle(b), pinned_handle(b + sizeof(G2)),
                       pinned_handle(pk.H_query.data()), pinned_handle(pk.L_query.data()), r1cs};
    state_t &s = state();
    std::lock_guard<std::mutex> lk(s.mu);
    s.keys[pk.A_query.data()] = key;
}

inline void unpin_all()
{
    state_t &s = state();
    std::lock_guard<std::mutex> lk(s.mu);
    // the proof slots prove on their own contexts but read the images pinned on s.ctx: unpin only waits for the
    // pinning context, so drain every slot first
    for (auto &sl : s.slots)
        if (sl.ctx) zecale_gpu_sync(sl.ctx);
    for (auto &kv : s.pins) zecale_gpu_unpin_bases(s.ctx, kv.second.handle);
    s.pins.clear();
    for (auto &kv : s.keys) zecale_gpu_r1cs_free(s.ctx, kv.second.r1cs);
    s.keys.clear();
}

inline void shutdown()
{
    unpin_all();
    state_t &s = state();
    std::lock_guard<std::mutex> lk(s.mu);
    for (auto &sl : s.slots) {
        if (sl.staging) zecale_gpu_host_free(sl.staging);
        if (sl.ctx) zecale_gpu_shutdown(sl.ctx);
    }
    s.slots.clear();
    if (s.ctx) zecale_gpu_shutdown(s.ctx);
    s.ctx = nullptr;
}

/// Number of proofs the fused prover keeps in flight ($ZECALE_GPU_PIPELINE, default 3).
inline size_t pipeline_depth()
{
    const char *v = std::getenv("ZECALE_GPU_PIPELINE");
    const int d = v ? std::atoi(v) : 3;
    return d < 1 ? 1 : (d > 8 ? 8 : (size_t)d);
}

/// Take a free proof slot (waits while all are busy); the slot's context is created on first use, on
/// the device of the process-wide context.
inline size_t acquire_slot(size_t staging_bytes)
{
    context();
    state_t &s = state();
    std::unique_lock<std::mutex> lk(s.mu);
    if (s.slots.empty()) s.slots.resize(pipeline_depth());
    size_t k = 0;
    s.slot_free.wait(lk, [&] {
        for (k = 0; k < s.slots.size(); k++)
            if (!s.slots[k].busy) return true;
        return false;
    });
    state_t::slot_t &sl = s.slots[k];
    if (!sl.ctx) {
        const std::vector<int> devs = device_list();
        if (zecale_gpu_init(devs.empty() ? nullptr : devs.data(), (int)devs.size(), &sl.ctx) != ZECALE_GPU_OK) {
            sl.ctx = nullptr;
            throw std::runtime_error("zecale_gpu_init failed for a proof slot");
        }
    }
    if (sl.staging_bytes < staging_bytes) {
        if (sl.staging) zecale_gpu_host_free(sl.staging);
        sl.staging = zecale_gpu_host_alloc(staging_bytes);
        sl.staging_bytes = sl.staging ? staging_bytes : 0;
        if (!sl.staging) throw std::runtime_error("zecale_gpu_host_alloc failed");
    }
    sl.busy = true;
    return k;
}

inline void release_slot(size_t k)
{
    state_t &s = state();
    {
        std::lock_guard<std::mutex> lk(s.mu);
        s.slots[k].busy = false;
    }
    s.slot_free.notify_one();
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
    /// The m Lagrange coefficients at t (the QAP instance evaluation of the setup): batched inversions on
    /// the device instead of m serial ones; t on the domain itself keeps the reference's code.
    std::vector<Fr> evaluate_all_lagrange_polynomials(const Fr &t) override
    {
        std::vector<Fr> u(this->m, Fr::zero());
        const int rc = zecale_gpu_lagrange(context(), (unsigned)log_m_, &t, u.data());
        if (rc == ZECALE_GPU_E_DOMAIN) return cpu_.evaluate_all_lagrange_polynomials(t);
        check(rc, "zecale_gpu_lagrange");
        return u;
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

// The base point of a libff window table: powers_of_g[0][1] = 1 * 2^0 * g (multiexp.tcc:826-842), present in the
// reference's full table and in the two-entry one of ZECALE_GPU_SLIM_WINDOW_TABLE alike. Any other shape is refused.
template<typename GroupT>
inline const GroupT &table_base(const std::vector<std::vector<GroupT>> &table)
{
    if (table.empty() || table[0].size() < 2) {
        throw std::invalid_argument("zecale_gpu_hooks: window table without a base point (table[0][1])");
    }
    return table[0][1];
}

// out[i] = (coeff * v[i]) * g on the device (zecale_gpu_batch_exp), special-form points.
template<typename GroupT>
inline std::vector<GroupT> batch_exp(int group, const GroupT &g, const std::vector<Fr> &v, size_t n, const Fr *coeff)
{
    static_assert(sizeof(GroupT) == 288 && sizeof(Fr) == 48, "libff in-memory layout");
    if (n > v.size()) {
        throw std::invalid_argument("batch_exp: num_entries exceeds the scalar vector");
    }
    std::vector<GroupT> out(n, GroupT::zero());
    if (n == 0) {
        return out;
    }
    GroupT base = g;
    base.to_special();
    check(
        zecale_gpu_batch_exp(context(), group, &base, v.data(), sizeof(Fr), n, coeff, out.data()),
        "zecale_gpu_batch_exp");
    return out;
}


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

// Fixed-base exponentiations of the trusted setup (r1cs_gg_ppzksnark_generator, r1cs_gg_ppzksnark.tcc:
// 349-398: A, H, L queries and gamma_ABC): batch_exp (multiexp.tcc:887-916) and batch_exp_with_coeff
// (:918-947). The device builds its own window table; the specialised readers below (batch_exp,
// batch_exp_with_coeff, kc_batch_exp<G2,G1,Fr>) only take the base point table[0][1] = 1 * 2^0 * g from the
// CPU table (get_window_table, multiexp.tcc:809-845), which is there whatever the table's shape.
//
// ZECALE_GPU_SLIM_WINDOW_TABLE (opt-in, -D on the command line): also specialise get_window_table to the
// two-entry table {0, g} and windowed_exp to pow * g, which removes the millions of serial point additions
// the CPU spends on tables the device never reads (batch-8 setup 7.9 s -> 1.2 s). It is NOT the default
// because these are weak inline symbols: if any translation unit of the final binary instantiates the
// GENERIC batch_exp / windowed_exp / kc_batch_exp for bw6_761 without this header (e.g. the -include was
// added to one target only), the linker may pair that generic reader with the two-entry table and it then
// reads past its end - wrong keys, silently. And a reader this header does not specialise (kc_batch_exp<G1,
// G1,Fr> of the PGHR13 generator) would fall back to a double-and-add per element. Define it only when
// every TU that touches libff's multiexp code for bw6_761 is compiled with the -include.
#ifdef ZECALE_GPU_SLIM_WINDOW_TABLE
#define ZECALE_GPU_WINDOW_TABLE_SPECIALISE(GROUP)                                                       \
    template<>                                                                                         \
    inline window_table<GROUP> get_window_table<GROUP>(const size_t, const size_t, const GROUP &g)     \
    {                                                                                                  \
        return window_table<GROUP>(1, std::vector<GROUP>{GROUP::zero(), g});                           \
    }                                                                                                  \
    template<>                                                                                         \
    inline GROUP windowed_exp<GROUP, bw6_761_Fr>(                                                      \
        const size_t, const size_t, const window_table<GROUP> &table, const bw6_761_Fr &pow)           \
    {                                                                                                  \
        return pow * zecale_gpu_hooks::table_base(table);                                              \
    }
ZECALE_GPU_WINDOW_TABLE_SPECIALISE(bw6_761_G1)
ZECALE_GPU_WINDOW_TABLE_SPECIALISE(bw6_761_G2)
#undef ZECALE_GPU_WINDOW_TABLE_SPECIALISE
#endif

#define ZECALE_GPU_BATCH_EXP_SPECIALISE(GROUP, GID)                                                    \
    template<>                                                                                         \
    inline std::vector<GROUP> batch_exp<GROUP, bw6_761_Fr>(                                            \
        const size_t, const size_t, const window_table<GROUP> &table, const std::vector<bw6_761_Fr> &v, \
        size_t num_entries)                                                                            \
    {                                                                                                  \
        return zecale_gpu_hooks::batch_exp<GROUP>(GID, zecale_gpu_hooks::table_base(table), v, num_entries, nullptr);          \
    }                                                                                                  \
    template<>                                                                                         \
    inline std::vector<GROUP> batch_exp_with_coeff<GROUP, bw6_761_Fr>(                                 \
        const size_t, const size_t, const window_table<GROUP> &table, const bw6_761_Fr &coeff,         \
        const std::vector<bw6_761_Fr> &v)                                                              \
    {                                                                                                  \
        return zecale_gpu_hooks::batch_exp<GROUP>(GID, zecale_gpu_hooks::table_base(table), v, v.size(), &coeff);              \
    }
ZECALE_GPU_BATCH_EXP_SPECIALISE(bw6_761_G1, ZECALE_GPU_G1)
ZECALE_GPU_BATCH_EXP_SPECIALISE(bw6_761_G2, ZECALE_GPU_G2)
#undef ZECALE_GPU_BATCH_EXP_SPECIALISE

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

// The B query of the generator (r1cs_gg_ppzksnark.tcc:360-373): kc_batch_exp (kc_multiexp.tcc:148-239)
// keeps the non-zero scalars only; their G2 and G1 powers are two device batch exponentiations.
template<>
inline knowledge_commitment_vector<libff::bw6_761_G2, libff::bw6_761_G1> kc_batch_exp<
    libff::bw6_761_G2,
    libff::bw6_761_G1,
    libff::bw6_761_Fr>(
    const size_t,
    const size_t,
    const size_t,
    const libff::window_table<libff::bw6_761_G2> &T1_table,
    const libff::window_table<libff::bw6_761_G1> &T2_table,
    const libff::bw6_761_Fr &T1_coeff,
    const libff::bw6_761_Fr &T2_coeff,
    const std::vector<libff::bw6_761_Fr> &v,
    const size_t,
    bool)
{
    knowledge_commitment_vector<libff::bw6_761_G2, libff::bw6_761_G1> res;
    res.domain_size_ = v.size();
    std::vector<libff::bw6_761_Fr> nz;
    for (size_t i = 0; i < v.size(); ++i) {
        if (!v[i].is_zero()) {
            nz.push_back(v[i]);
            res.indices.push_back(i);
        }
    }
    const std::vector<libff::bw6_761_G2> g2 = zecale_gpu_hooks::batch_exp<libff::bw6_761_G2>(
        ZECALE_GPU_G2, zecale_gpu_hooks::table_base(T1_table), nz, nz.size(), &T1_coeff);
    const std::vector<libff::bw6_761_G1> g1 = zecale_gpu_hooks::batch_exp<libff::bw6_761_G1>(
        ZECALE_GPU_G1, zecale_gpu_hooks::table_base(T2_table), nz, nz.size(), &T2_coeff);
    res.values.reserve(nz.size());
    for (size_t i = 0; i < nz.size(); ++i) {
        res.values.emplace_back(knowledge_commitment<libff::bw6_761_G2, libff::bw6_761_G1>(g2[i], g1[i]));
    }
    return res;
}

} // namespace libsnark

namespace zecale_gpu_hooks
{

// ---------------------------------------------------------------- the prover itself
using ppT = libff::bw6_761_pp;
using proof_t = libsnark::r1cs_gg_ppzksnark_proof<ppT>;
using proving_key_t = libsnark::r1cs_gg_ppzksnark_proving_key<ppT>;

/// The five multi-exponentiation results of r1cs_gg_ppzksnark_prover (r1cs_gg_ppzksnark.tcc:531-580).
struct evaluations_t {
    G1 At;
    G2 Bt_g;
    G1 Bt_h;
    G1 Ht;
    G1 Lt;
};

/// One proof in flight on the fused path (prove_async .. collect).
/// The part of the proof assembly that does not depend on the device results: four of the six scalar
/// multiplications of r1cs_gg_ppzksnark.tcc:582-599 (about 1.5 ms each on the host). Computed between enqueue
/// and collect, i.e. while the device works.
struct blinding_t {
    G1 r_delta, s_delta, rs_delta;
    G2 s_delta2;
    bool ready = false;
};

struct ticket_t {
    size_t slot;
    const proving_key_t *pk;
    size_t nnz_b;
    libff::bw6_761_Fr r, s;
    blinding_t blind;
    unsigned char pt[5][288];
};

/// Fused path for a key made resident with pin_proving_key: the assignment is staged in pinned memory
/// and uploaded once, the sparse mat-vecs and the seven transforms of r1cs_to_qap_witness_map run on
/// the device, the H MSM reads coefficients_for_H from device memory, the five MSMs overlap on the
/// lanes of the proof's own context. Enqueues everything and returns a ticket (nullptr if the key is
/// not resident or the reference would not use the power-of-two domain for it); collect_evaluations
/// waits for the 5 x 288 result bytes and frees the slot.
inline ticket_t *evaluate_fused_async(
    const proving_key_t &pk,
    const std::vector<Fr> &primary_input,
    const std::vector<Fr> &auxiliary_input,
    bool force_pow_2_domain)
{
    state_t::key_t key;
    {
        state_t &s = state();
        std::lock_guard<std::mutex> lk(s.mu);
        auto it = s.keys.find(pk.A_query.data());
        if (it == s.keys.end()) return nullptr;
        key = it->second;
    }
    const auto &cs = pk.constraint_system;
    const size_t nv = cs.num_variables(), ni = cs.num_inputs();
    const size_t min_n = cs.num_constraints() + ni + 1;
    const size_t m = (size_t)1 << zecale_gpu_r1cs_log_domain(key.r1cs);
    if (!force_pow_2_domain && min_n != m) return nullptr;
    if (primary_input.size() != ni || primary_input.size() + auxiliary_input.size() != nv ||
        pk.A_query.size() != nv + 1 || pk.H_query.size() != m - 1 || pk.L_query.size() != nv - ni ||
        !key.A || !key.H || !key.L) {
        return nullptr;
    }
    const size_t slot = acquire_slot((nv + 1) * sizeof(Fr));
    std::unique_ptr<ticket_t> t(new ticket_t());
    t->slot = slot;
    t->pk = &pk;
    t->nnz_b = pk.B_query.indices.size();
    try {
        zecale_gpu_ctx *ctx;
        Fr *full;
        {
            state_t &s = state();
            std::lock_guard<std::mutex> lk(s.mu);
            ctx = s.slots[slot].ctx;
            full = static_cast<Fr *>(s.slots[slot].staging);
        }
        // const_padded_assignment of the reference (r1cs_gg_ppzksnark.tcc:523-528), in pinned memory
        const Fr one = Fr::one();
        std::memcpy(static_cast<void *>(full), static_cast<const void *>(&one), sizeof(Fr));
        if (ni) std::memcpy(static_cast<void *>(full + 1), static_cast<const void *>(primary_input.data()), ni * sizeof(Fr));
        if (nv > ni)
            std::memcpy(static_cast<void *>(full + 1 + ni), static_cast<const void *>(auxiliary_input.data()), (nv - ni) * sizeof(Fr));
        const Fr g = Fr::multiplicative_generator; // the coset of r1cs_to_qap_witness_map (r1cs_to_qap.tcc:270)
        auto ck = [&](int rc, const char *what) {
            if (rc != ZECALE_GPU_OK) throw std::runtime_error(std::string(what) + ": " + zecale_gpu_last_error(ctx));
        };
        ck(zecale_gpu_qap_h_from_witness_async(ctx, 0, key.r1cs, full, &g, nullptr, 0), "zecale_gpu_qap_h_from_witness_async");
        ck(zecale_gpu_msm_pinned_async(ctx, 1, key.A, full, sizeof(Fr), nullptr, nv + 1, t->pt[0], 0), "msm A");
        ck(zecale_gpu_msm_pinned_async(ctx, 4, key.L, full + ni + 1, sizeof(Fr), nullptr, nv - ni, t->pt[4], 0), "msm L");
        if (t->nnz_b) {
            ck(zecale_gpu_msm_pinned_async(ctx, 2, key.B2, full, sizeof(Fr), pk.B_query.indices.data(), t->nnz_b, t->pt[1], 0), "msm B (G2)");
            ck(zecale_gpu_msm_pinned_async(ctx, 3, key.B1, full, sizeof(Fr), pk.B_query.indices.data(), t->nnz_b, t->pt[2], 0), "msm B (G1)");
        }
        ck(zecale_gpu_msm_pinned_async(ctx, 0, key.H, nullptr, sizeof(Fr), nullptr, m - 1, t->pt[3], ZECALE_GPU_SCALARS_FROM_QAP), "msm H");
        ck(zecale_gpu_flush(ctx), "zecale_gpu_flush");
    } catch (...) {
        zecale_gpu_ctx *ctx = nullptr;
        {
            state_t &s = state();
            std::lock_guard<std::mutex> lk(s.mu);
            ctx = s.slots[slot].ctx;
        }
        if (ctx) zecale_gpu_sync(ctx);
        release_slot(slot);
        throw;
    }
    return t.release();
}

inline void collect_evaluations(ticket_t *ticket, evaluations_t &out)
{
    std::unique_ptr<ticket_t> t(ticket);
    zecale_gpu_ctx *ctx;
    {
        state_t &s = state();
        std::lock_guard<std::mutex> lk(s.mu);
        ctx = s.slots[t->slot].ctx;
    }
    const int rc = zecale_gpu_sync(ctx);
    const std::string err = rc == ZECALE_GPU_OK ? std::string() : std::string(zecale_gpu_last_error(ctx));
    release_slot(t->slot);
    if (rc != ZECALE_GPU_OK) throw std::runtime_error("zecale_gpu_sync: " + err);
    std::memcpy(static_cast<void *>(&out.At), t->pt[0], 288);
    if (t->nnz_b) {
        std::memcpy(static_cast<void *>(&out.Bt_g), t->pt[1], 288);
        std::memcpy(static_cast<void *>(&out.Bt_h), t->pt[2], 288);
    } else {
        out.Bt_g = G2::zero();
        out.Bt_h = G1::zero();
    }
    std::memcpy(static_cast<void *>(&out.Ht), t->pt[3], 288);
    std::memcpy(static_cast<void *>(&out.Lt), t->pt[4], 288);
}

inline bool evaluate_fused(
    const proving_key_t &pk,
    const std::vector<Fr> &primary_input,
    const std::vector<Fr> &auxiliary_input,
    bool force_pow_2_domain,
    evaluations_t &out)
{
    ticket_t *t = evaluate_fused_async(pk, primary_input, auxiliary_input, force_pow_2_domain);
    if (!t) return false;
    collect_evaluations(t, out);
    return true;
}

/// The reference's own sequence of calls (r1cs_gg_ppzksnark.tcc:477-580), each of which resolves
/// to a specialisation below: used when the key is not resident.
inline void evaluate_sequential(
    const proving_key_t &pk,
    const std::vector<Fr> &primary_input,
    const std::vector<Fr> &auxiliary_input,
    bool force_pow_2_domain,
    evaluations_t &out)
{
    const libsnark::qap_witness<Fr> qap_wit = libsnark::r1cs_to_qap_witness_map(
        pk.constraint_system, primary_input, auxiliary_input, Fr::zero(), Fr::zero(), Fr::zero(), force_pow_2_domain);
    std::vector<Fr> padded(1, Fr::one());
    padded.insert(padded.end(), qap_wit.coefficients_for_ABCs.begin(), qap_wit.coefficients_for_ABCs.end());
    const size_t nv = qap_wit.num_variables();
    constexpr libff::multi_exp_method M = libff::multi_exp_method_BDLO12_signed;
    constexpr libff::multi_exp_base_form F = libff::multi_exp_base_form_special;
    out.At = libff::multi_exp_filter_one_zero<G1, Fr, M, F>(
        pk.A_query.begin(), pk.A_query.begin() + nv + 1, padded.begin(), padded.begin() + nv + 1, 1);
    const KC bt = libsnark::kc_multi_exp_with_mixed_addition<G2, G1, Fr, M, F>(
        pk.B_query, 0, nv + 1, padded.begin(), padded.begin() + nv + 1, 1);
    out.Bt_g = bt.g;
    out.Bt_h = bt.h;
    out.Ht = libff::multi_exp<G1, Fr, M, F>(
        pk.H_query.begin(), pk.H_query.begin() + (qap_wit.degree() - 1), qap_wit.coefficients_for_H.begin(),
        qap_wit.coefficients_for_H.begin() + (qap_wit.degree() - 1), 1);
    out.Lt = libff::multi_exp_filter_one_zero<G1, Fr, M, F>(
        pk.L_query.begin(), pk.L_query.end(), padded.begin() + qap_wit.num_inputs() + 1, padded.begin() + nv + 1, 1);
}

inline void make_blinding(const proving_key_t &pk, const Fr &r, const Fr &s, blinding_t &b)
{
    b.r_delta = r * pk.delta_g1;
    b.s_delta = s * pk.delta_g1;
    b.s_delta2 = s * pk.delta_g2;
    b.rs_delta = (r * s) * pk.delta_g1;
    b.ready = true;
}

/// Proof assembly: the reference's formulae (r1cs_gg_ppzksnark.tcc:582-599) on its own CPU group arithmetic. The same
/// group elements whatever the order of evaluation, so the serialised proof is the reference's. The two scalar
/// multiplications that need the device results (s * A, r * B) run on two host threads.
inline proof_t assemble(const proving_key_t &pk, const evaluations_t &e, const Fr &r, const Fr &s, const blinding_t *pre = nullptr)
{
    blinding_t local;
    if (!pre || !pre->ready) {
        make_blinding(pk, r, s, local);
        pre = &local;
    }
    G1 g1_A = pk.alpha_g1 + e.At + pre->r_delta;
    G1 g1_B = pk.beta_g1 + e.Bt_h + pre->s_delta;
    G2 g2_B = pk.beta_g2 + e.Bt_g + pre->s_delta2;
    std::future<G1> r_b = std::async(std::launch::async, [&]() { return r * g1_B; });
    const G1 s_a = s * g1_A;
    G1 g1_C = e.Ht + e.Lt + s_a + r_b.get() - pre->rs_delta;
    return proof_t(std::move(g1_A), std::move(g2_B), std::move(g1_C));
}

/// r1cs_gg_ppzksnark_prover with the prover randomness as parameters (the reference samples r, s
/// from std::random_device, r1cs_gg_ppzksnark.tcc:502-503). Safe to call from several threads: each
/// call owns a proof slot until its results are back.
inline proof_t prove_with_rs(
    const proving_key_t &pk,
    const std::vector<Fr> &primary_input,
    const std::vector<Fr> &auxiliary_input,
    bool force_pow_2_domain,
    const Fr &r,
    const Fr &s,
    evaluations_t *evaluations_out = nullptr)
{
    evaluations_t e;
    blinding_t blind;
    ticket_t *t = evaluate_fused_async(pk, primary_input, auxiliary_input, force_pow_2_domain);
    if (t) {
        make_blinding(pk, r, s, blind); // overlaps the device work
        collect_evaluations(t, e);
    } else {
        evaluate_sequential(pk, primary_input, auxiliary_input, force_pow_2_domain, e);
    }
    if (evaluations_out) *evaluations_out = e;
    return assemble(pk, e, r, s, &blind);
}

/// Pipelined use of the fused prover from one thread: prove_async enqueues a proof on a free slot and
/// returns at once (nullptr if the key is not resident: use prove_with_rs), collect waits for it and
/// assembles. Up to pipeline_depth() proofs are in flight; the host-side assembly of proof i overlaps
/// the device work of proof i+1. pk must outlive the ticket.
inline ticket_t *prove_async(
    const proving_key_t &pk,
    const std::vector<Fr> &primary_input,
    const std::vector<Fr> &auxiliary_input,
    bool force_pow_2_domain,
    const Fr &r,
    const Fr &s)
{
    ticket_t *t = evaluate_fused_async(pk, primary_input, auxiliary_input, force_pow_2_domain);
    if (t) {
        t->r = r;
        t->s = s;
        make_blinding(pk, r, s, t->blind);
    }
    return t;
}

inline proof_t collect(ticket_t *t, evaluations_t *evaluations_out = nullptr)
{
    const proving_key_t &pk = *t->pk;
    const Fr r = t->r, s = t->s;
    const blinding_t blind = t->blind; // the ticket is freed by collect_evaluations
    evaluations_t e;
    collect_evaluations(t, e);
    if (evaluations_out) *evaluations_out = e;
    return assemble(pk, e, r, s, &blind);
}

} // namespace zecale_gpu_hooks

namespace libsnark
{

/// r1cs_gg_ppzksnark_prover<bw6_761_pp> (r1cs_gg_ppzksnark.tcc:460-603): same signature, same
/// distribution of (r, s); the evaluations come from the device.
template<>
inline r1cs_gg_ppzksnark_proof<libff::bw6_761_pp> r1cs_gg_ppzksnark_prover<
    libff::bw6_761_pp,
    libff::multi_exp_method_BDLO12_signed,
    libff::multi_exp_base_form_special>(
    const r1cs_gg_ppzksnark_proving_key<libff::bw6_761_pp> &pk,
    const r1cs_gg_ppzksnark_primary_input<libff::bw6_761_pp> &primary_input,
    const r1cs_gg_ppzksnark_auxiliary_input<libff::bw6_761_pp> &auxiliary_input,
    bool force_pow_2_domain)
{
    const libff::bw6_761_Fr r = libff::bw6_761_Fr::random_element();
    const libff::bw6_761_Fr s = libff::bw6_761_Fr::random_element();
    return zecale_gpu_hooks::prove_with_rs(pk, primary_input, auxiliary_input, force_pow_2_domain, r, s);
}

} // namespace libsnark

#endif // ZECALE_GPU_HOOKS_HPP
