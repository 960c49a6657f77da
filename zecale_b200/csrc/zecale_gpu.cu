// C ABI of libzecale_gpu.so (see include/zecale_gpu.h). One translation unit: the
// device functions are header-defined so that each kernel gets a single out-of-line
// copy of fq_mul & co.
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/zecale_gpu.h"
#include "msm.cuh"
#include "ntt.cuh"

using namespace zecale;

struct DevBuf {
    void *p = nullptr;
    size_t cap = 0;
};

struct zecale_gpu_bases {
    Affine *d = nullptr;      // G copies of n points: copy g holds 2^(c*V*g) * P_i
    uint8_t *is_inf = nullptr; // n flags
    size_t n = 0;
    int group = 0;
    uint32_t c = 0, V = 0;    // window bits / bucket sets the image was expanded for (V = 0: not expanded)
    // multi-device context: the points are sharded by libff::multi_exp's chunk arithmetic
    // (multiexp.tcc:663-676); shard g = bases [lo[g], lo[g+1]) pinned on worker g (nullptr if empty)
    std::vector<zecale_gpu_bases *> shards;
    std::vector<size_t> lo;
};

// Device image of an R1CS (three CSR matrices over Fr), loaded once per circuit.
struct zecale_gpu_r1cs {
    size_t n_constraints = 0, n_inputs = 0, n_variables = 0;
    unsigned log_m = 0;
    uint64_t *row_ptr[3] = {nullptr, nullptr, nullptr};
    uint32_t *col[3] = {nullptr, nullptr, nullptr};
    Fr *val[3] = {nullptr, nullptr, nullptr};
};

// One in-order work queue of the context: a stream, its scratch buffers and the pinned
// staging area results are read back into. Lane 0 serves the blocking entry points; the
// *_async entry points let the caller put independent MSMs / the H polynomial on different
// lanes so that one MSM's latency-bound tail overlaps another's bucket accumulation.
static constexpr int ZECALE_GPU_LANES = 6;
struct Lane {
    cudaStream_t stream = nullptr;
    // High-priority twin of `stream`: the latency-bound parts (NTT passes, the fold / running-sum /
    // Horner tail of an MSM: long chains of small dependent kernels) run here so that their blocks
    // are placed ahead of the queued blocks of other lanes' throughput-bound bucket accumulation.
    cudaStream_t hi = nullptr;
    cudaEvent_t ev_hand = nullptr, ev_back = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    DevBuf ws_msm, ws_sc, ws_idx, ws_ntt, ws_qap;
    struct Staging {
        Proj result;
        uint32_t stats[2];
    } *h = nullptr;              // pinned
    void *user_out = nullptr;    // where h->result goes at the next sync (async MSM)
    void *user_h = nullptr;      // async qap_h: host destination ...
    size_t user_h_bytes = 0;     // ... and its size (a zero Fr is appended)
    uint32_t last_stats[2] = {0, 0};
    MsmPlan last_plan{};
    // Asynchronous MSM whose second half (fold / running sums / Horner) has not been enqueued yet:
    // the tails of all MSMs of a proof are issued together, after every bucket accumulation of the
    // proof, so that the latency-bound tails neither starve behind nor displace the accumulation
    // kernels of sibling lanes (see flush_tails).
    bool tail_pending = false;
    MsmPlan tail_plan{};
    MsmBuffers tail_buf{};
    cudaEvent_t ev_acc = nullptr; // bucket accumulation of the pending MSM finished
    // the last witness map enqueued on THIS lane has written its coefficients_for_H into ws_qap (what a
    // ZECALE_GPU_SCALARS_FROM_QAP MSM of this lane reads, possibly from another stream or device); the context-wide
    // ev_qap only says that the most recent witness map of any lane is done and is a scheduling hint
    cudaEvent_t ev_qap = nullptr;
    // shard of a multi-device MSM: the un-normalised partial sum (XYZZ) is copied to this address on
    // device partial_dev instead of being normalised and sent to the host
    XYZZ *partial_dst = nullptr;
    int partial_dev = 0;
    bool busy = false;            // an asynchronous MSM of this lane has not been synchronised yet
};

struct zecale_gpu_ctx {
    int device = 0;
    std::vector<int> devices;
    Lane lanes[ZECALE_GPU_LANES];
    Lane *cur = nullptr;
    std::mutex mu;
    std::string err;
    uint64_t launches = 0;
    DevBuf ws_in, ws_tmp;        // import / test scratch (lane 0 only)
    int *d_status = nullptr;
    NttTables ntt;
    // per-kernel-class timers (class 0: MSM bucket accumulation, class 1: NTT transforms)
    std::vector<cudaEvent_t> ev_pool;
    std::vector<std::pair<cudaEvent_t, cudaEvent_t>> ev_pending[2];
    double kernel_ms[2] = {0, 0};
    uint64_t kernel_n[2] = {0, 0};
    std::map<std::pair<uint32_t, std::vector<uint32_t>>, Fr *> zinv; // (log_m, g) -> (g^m - 1)^-1
    // host scalar vectors uploaded since the last zecale_gpu_sync: the prover passes the same
    // assignment to A, B and L (r1cs_gg_ppzksnark.tcc:531-580), so it crosses PCIe once
    struct EpochUpload {
        const char *host;
        size_t bytes;
        int buf;
        cudaEvent_t ev;
    };
    std::vector<EpochUpload> uploads;
    DevBuf up_pool[8];
    // Set by qap_h_async until the next sync: asynchronous MSMs enqueued after it on other lanes
    // start once the seven transforms are done. The transforms are short (~1 ms) but sit on the
    // proof's critical path (the H MSM consumes them); left to compete with four bucket
    // accumulations for SM space they finish ten times later.
    cudaEvent_t ev_qap = nullptr;
    bool qap_in_epoch = false;
    // Multi-device context (zecale_gpu_init with n_devices > 1): this object coordinates and owns the
    // witness map on devices[0]; workers[g] is a complete single-device context on devices[g] (g = 0
    // shares the first device) that runs shard g of every pinned MSM. The partial sums (384 B each) are
    // peer-copied into d_gather on devices[0], summed and normalised there (msm_launch_gather_batch).
    std::vector<zecale_gpu_ctx *> workers;
    static constexpr int MAX_DEVICES = 16;
    XYZZ *d_gather = nullptr;        // [lane][MAX_DEVICES]
    Proj *d_gather_res = nullptr;    // [lane]
    cudaStream_t gstream = nullptr;  // high priority, devices[0]
    cudaEvent_t ev_gdone = nullptr;
    cudaEvent_t ev_parts = nullptr;  // (worker) partial sums of the last flush have been sent
    struct Gather {
        bool pending = false, flushed = false;
        uint32_t n_parts = 0;
    } gather[ZECALE_GPU_LANES];
    // ZECALE_GPU_TRACE=1: phase timestamps of every lane, printed to stderr at each sync
    bool trace = false;
    struct TraceMark {
        const char *label;
        int lane;
        cudaEvent_t ev;
    };
    std::vector<TraceMark> marks;
};

namespace {

int fail(zecale_gpu_ctx *ctx, int code, const char *fmt, ...)
{
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    if (ctx) ctx->err = buf;
    return code;
}

#define CK(call)                                                                                   \
    do {                                                                                           \
        cudaError_t e_ = (call);                                                                   \
        if (e_ != cudaSuccess)                                                                     \
            return fail(ctx, e_ == cudaErrorMemoryAllocation ? ZECALE_GPU_E_NOMEM : ZECALE_GPU_E_CUDA, \
                        "%s:%d %s: %s", __FILE__, __LINE__, #call, cudaGetErrorString(e_));       \
    } while (0)

int reserve(zecale_gpu_ctx *ctx, DevBuf &b, size_t bytes)
{
    if (bytes <= b.cap) return 0;
    if (b.p) CK(cudaFree(b.p));
    b.p = nullptr;
    b.cap = 0;
    size_t want = bytes + bytes / 8 + 4096;
    CK(cudaMalloc(&b.p, want));
    b.cap = want;
    return 0;
}

cudaEvent_t take_event(zecale_gpu_ctx *ctx)
{
    if (!ctx->ev_pool.empty()) {
        cudaEvent_t e = ctx->ev_pool.back();
        ctx->ev_pool.pop_back();
        return e;
    }
    cudaEvent_t e = nullptr;
    cudaEventCreate(&e);
    return e;
}

void trace_mark(zecale_gpu_ctx *ctx, const char *label, cudaStream_t st, int lane = -1)
{
    if (!ctx->trace) return;
    cudaEvent_t e = take_event(ctx);
    cudaEventRecord(e, st);
    ctx->marks.push_back({label, lane >= 0 ? lane : (int)(ctx->cur - ctx->lanes), e});
}

// Enqueue the second half of every pending asynchronous MSM as ONE batch (each launch of the tail
// serves all of them, msm_launch_tail_batch) on the first pending lane's high-priority stream, gated
// on the accumulation kernels of all pending MSMs, then joined back into every lane.
int flush_tails(zecale_gpu_ctx *ctx)
{
    TailJob jobs[ZECALE_GPU_LANES];
    Lane *owner[ZECALE_GPU_LANES];
    uint32_t n = 0;
    for (Lane &L : ctx->lanes)
        if (L.tail_pending) {
            jobs[n] = msm_tail_job(L.tail_plan, L.tail_buf);
            if (L.partial_dst) jobs[n].result = nullptr; // a shard: leave the XYZZ sum in fin[0]
            owner[n++] = &L;
        }
    if (n == 0) return 0;
    for (Lane &L : ctx->lanes) L.tail_pending = false;
    Lane &F = *owner[0];
    const int lane = (int)(&F - ctx->lanes);
    for (uint32_t i = 0; i < n; i++) CK(cudaStreamWaitEvent(F.hi, owner[i]->ev_acc, 0));
    trace_mark(ctx, "tail0", F.hi, lane);
    uint64_t nl = 0;
    CK(msm_launch_tail_batch(jobs, n, F.hi, &nl));
    ctx->launches += nl;
    bool sent = false;
    for (uint32_t i = 0; i < n; i++) {
        Lane &L = *owner[i];
        if (L.partial_dst) {
            CK(cudaMemcpyPeerAsync(L.partial_dst, L.partial_dev, L.tail_buf.fin, ctx->device, sizeof(XYZZ), F.hi));
            L.partial_dst = nullptr;
            sent = true;
        } else {
            CK(cudaMemcpyAsync(&L.h->result, L.tail_buf.result, sizeof(Proj), cudaMemcpyDeviceToHost, F.hi));
        }
    }
    if (sent) CK(cudaEventRecord(ctx->ev_parts, F.hi));
    CK(cudaEventRecord(F.ev_back, F.hi));
    for (uint32_t i = 0; i < n; i++) CK(cudaStreamWaitEvent(owner[i]->stream, F.ev_back, 0));
    trace_mark(ctx, "tail1", F.stream, lane);
    return 0;
}

void trace_dump(zecale_gpu_ctx *ctx)
{
    if (ctx->marks.empty()) return;
    fprintf(stderr, "[zecale_gpu trace ctx=%p] ms since first mark (lane: label=t ...)\n", (void *)ctx);
    for (int lane = 0; lane < ZECALE_GPU_LANES; lane++) {
        bool any = false;
        for (auto &m : ctx->marks) {
            if (m.lane != lane) continue;
            float ms = 0.f;
            cudaEventElapsedTime(&ms, ctx->marks[0].ev, m.ev);
            if (!any) fprintf(stderr, "  lane %d:", lane);
            any = true;
            fprintf(stderr, " %s=%.2f", m.label, ms);
        }
        if (any) fprintf(stderr, "\n");
    }
    for (auto &m : ctx->marks) ctx->ev_pool.push_back(m.ev);
    ctx->marks.clear();
}

// Fold finished event pairs into the per-class totals (synchronises the stream).
// Wait for every lane and deliver the results of asynchronous calls to the caller's buffers.
int sync_lanes(zecale_gpu_ctx *ctx)
{
    int rc = flush_tails(ctx);
    for (Lane &L : ctx->lanes) {
        if (!L.stream) continue;
        cudaError_t e = cudaStreamSynchronize(L.stream);
        if (e != cudaSuccess && rc == 0)
            rc = fail(ctx, ZECALE_GPU_E_CUDA, "stream synchronize: %s", cudaGetErrorString(e));
        L.busy = false;
        if (L.user_out) {
            if (e == cudaSuccess) {
                memcpy(L.user_out, &L.h->result, sizeof(Proj));
                L.last_stats[0] = L.h->stats[0];
                L.last_stats[1] = L.h->stats[1];
            }
            L.user_out = nullptr;
        }
        if (&L == &ctx->lanes[ZECALE_GPU_LANES - 1]) {
            for (auto &u : ctx->uploads) ctx->ev_pool.push_back(u.ev);
            ctx->uploads.clear();
        }
        if (L.user_h) { // the copy itself was enqueued by qap_h_async
            if (e == cudaSuccess) memset((char *)L.user_h + L.user_h_bytes, 0, sizeof(Fr));
            L.user_h = nullptr;
        }
    }
    ctx->qap_in_epoch = false;
    if (ctx->trace) trace_dump(ctx);
    return rc;
}

// Fold event pairs into the per-class totals. block = true (zecale_gpu_kernel_ms, shutdown) waits
// for all lanes first; block = false (called from the enqueue paths when many pairs are pending)
// only harvests the pairs that have already completed and never synchronises.
static constexpr size_t MAX_PENDING_TIMERS = 4096;
void drain_timers(zecale_gpu_ctx *ctx, bool block = true)
{
    if (block) sync_lanes(ctx);
    for (int k = 0; k < 2; k++) {
        auto &v = ctx->ev_pending[k];
        // Pairs finish roughly in the order they were recorded: the non-blocking harvest stops at the first pair that
        // is still running instead of querying the whole backlog (a pipelined prover that never blocks calls this on
        // every enqueue, so the cost per call has to stay O(harvested), not O(pending)).
        size_t done = 0;
        for (; done < v.size(); done++) {
            if (!block && cudaEventQuery(v[done].second) != cudaSuccess) break;
            float ms = 0.f;
            if (cudaEventElapsedTime(&ms, v[done].first, v[done].second) == cudaSuccess) {
                ctx->kernel_ms[k] += ms;
                ctx->kernel_n[k]++;
            }
            ctx->ev_pool.push_back(v[done].first);
            ctx->ev_pool.push_back(v[done].second);
        }
        v.erase(v.begin(), v.begin() + done);
        // A caller that never synchronises and whose oldest pair never completes would grow the backlog without bound:
        // beyond MAX_PENDING_TIMERS the oldest pairs are given up (their events go back to the pool untimed).
        if (v.size() > MAX_PENDING_TIMERS) {
            const size_t drop = v.size() - MAX_PENDING_TIMERS;
            for (size_t i = 0; i < drop; i++) {
                ctx->ev_pool.push_back(v[i].first);
                ctx->ev_pool.push_back(v[i].second);
            }
            v.erase(v.begin(), v.begin() + drop);
        }
    }
}

int import_bases(zecale_gpu_ctx *ctx, const void *bases, size_t stride, size_t n, Affine *d_out, uint8_t *d_inf)
{
    // stage the raw libff array on the device, then convert (coalesced 16 B accesses)
    size_t raw = (n - 1) * stride + sizeof(Proj);
    int rc = reserve(ctx, ctx->ws_in, raw);
    if (rc) return rc;
    CK(cudaMemcpyAsync(ctx->ws_in.p, bases, raw, cudaMemcpyHostToDevice, ctx->cur->stream));
    CK(cudaMemsetAsync(ctx->d_status, 0, sizeof(int), ctx->cur->stream));
    k_import_special<<<(unsigned)((n + 127) / 128), 128, 0, ctx->cur->stream>>>((const uint8_t *)ctx->ws_in.p, stride,
                                                                            (uint32_t)n, d_out, d_inf, ctx->d_status);
    ctx->launches++;
    int st = 0;
    CK(cudaMemcpyAsync(&st, ctx->d_status, sizeof(int), cudaMemcpyDeviceToHost, ctx->cur->stream));
    CK(cudaStreamSynchronize(ctx->cur->stream));
    if (st & 1) return fail(ctx, ZECALE_GPU_E_FORM, "MSM base not in special form (Z must be one or zero)");
    return 0;
}

int import_bases_bytes(zecale_gpu_ctx *ctx, const void *bytes, size_t stride, size_t n, Affine *d_out, uint8_t *d_inf)
{
    // the host buffer may sit at any address of a key file: stage it from a 4-byte aligned copy point
    size_t raw = (n - 1) * stride + sizeof(Affine);
    int rc = reserve(ctx, ctx->ws_in, raw);
    if (rc) return rc;
    CK(cudaMemcpyAsync(ctx->ws_in.p, bytes, raw, cudaMemcpyHostToDevice, ctx->cur->stream));
    k_import_bytes<<<(unsigned)((n + 127) / 128), 128, 0, ctx->cur->stream>>>((const uint8_t *)ctx->ws_in.p, stride, (uint32_t)n,
                                                                          d_out, d_inf);
    ctx->launches++;
    CK(cudaStreamSynchronize(ctx->cur->stream));
    return 0;
}

// d_bases: affine image on device; scalars: device pointer. Result -> host out (288 B).
int msm_device(zecale_gpu_ctx *ctx, const Affine *d_bases, const uint8_t *d_inf, uint32_t c, uint32_t V,
               uint32_t copy_stride, const uint8_t *d_scalars, size_t sstride, const uint64_t *d_index, size_t n,
               void *out_point, float *elapsed_ms, bool nosync = false, XYZZ *partial_dst = nullptr, int partial_dev = 0)
{
    if (n == 0) {
        // empty sum = zero point (0, one, 0), as libff::multi_exp returns GroupT::zero()
        Proj z;
        memset(&z, 0, sizeof z);
        memcpy(z.Y.l, FQ_ONE_HOST, sizeof z.Y.l);
        memcpy(out_point, &z, sizeof z);
        if (elapsed_ms) *elapsed_ms = 0.f;
        return 0;
    }
    if (ctx->cur->user_out || ctx->cur->busy) { // an asynchronous MSM of this lane is still in flight
        int rc0 = sync_lanes(ctx);
        if (rc0) return rc0;
    }
    const uint32_t c_eff = c ? c : msm_default_c((uint32_t)n);
    if (c_eff < MSM_MIN_C || c_eff > MSM_MAX_C)
        return fail(ctx, ZECALE_GPU_E_ARG, "MSM window size %u outside [%u, %u]", c_eff, MSM_MIN_C, MSM_MAX_C);
    MsmPlan plan = msm_make_plan((uint32_t)n, c_eff, V, copy_stride);
    MsmBuffers lay = msm_layout(plan, nullptr);
    int rc = reserve(ctx, ctx->cur->ws_msm, lay.total);
    if (rc) return rc;
    MsmBuffers b = msm_layout(plan, (uint8_t *)ctx->cur->ws_msm.p);
    if (elapsed_ms) CK(cudaEventRecord(ctx->cur->ev0, ctx->cur->stream));
    cudaEvent_t a0 = take_event(ctx), a1 = take_event(ctx);
    trace_mark(ctx, "msm", ctx->cur->stream);
    Lane &L = *ctx->cur;
    CK(msm_launch_front(plan, b, d_bases, d_inf, d_scalars, sstride, d_index, L.stream, a0, a1));
    if (ctx->trace) { // the accumulate events become trace marks (the per-class timer is off while tracing)
        ctx->marks.push_back({"acc0", (int)(ctx->cur - ctx->lanes), a0});
        ctx->marks.push_back({"acc1", (int)(ctx->cur - ctx->lanes), a1});
    } else {
        ctx->ev_pending[0].emplace_back(a0, a1);
    }
    if (ctx->ev_pending[0].size() > 256) drain_timers(ctx, false);
    ctx->launches += msm_front_launch_count(plan);
    CK(cudaMemcpyAsync(&L.h->stats[0], b.offsets + plan.n_buckets, 4, cudaMemcpyDeviceToHost, L.stream));
    CK(cudaMemcpyAsync(&L.h->stats[1], b.task_off + plan.n_buckets, 4, cudaMemcpyDeviceToHost, L.stream));
    if (nosync) { // the tail is issued by flush_tails (at the next sync), with its siblings
        CK(cudaEventRecord(L.ev_acc, L.stream));
        L.tail_pending = true;
        L.tail_plan = plan;
        L.tail_buf = b;
        L.partial_dst = partial_dst;
        L.partial_dev = partial_dev;
    } else {
        TailJob job = msm_tail_job(plan, b);
        uint64_t nl = 0;
        CK(msm_launch_tail_batch(&job, 1, L.stream, &nl));
        ctx->launches += nl;
        if (elapsed_ms) CK(cudaEventRecord(L.ev1, L.stream));
        CK(cudaMemcpyAsync(&L.h->result, b.result, sizeof(Proj), cudaMemcpyDeviceToHost, L.stream));
    }
    L.last_plan = plan;
    L.user_out = out_point;
    L.busy = nosync;
    if (nosync) return 0;
    CK(cudaStreamSynchronize(L.stream));
    memcpy(out_point, &L.h->result, sizeof(Proj));
    L.last_stats[0] = L.h->stats[0];
    L.last_stats[1] = L.h->stats[1];
    L.user_out = nullptr;
    if (elapsed_ms) CK(cudaEventElapsedTime(elapsed_ms, L.ev0, L.ev1));
    return 0;
}

int upload_scalars(zecale_gpu_ctx *ctx, const void *scalars, size_t sstride, const size_t *index, size_t n,
                   const uint8_t **d_sc, const uint64_t **d_idx)
{
    // number of scalars touched: with an index the caller's scalar array extends to max(index)+1
    size_t n_sc = n;
    if (index) {
        size_t mx = 0;
        for (size_t i = 0; i < n; i++)
            if (index[i] > mx) mx = index[i];
        n_sc = mx + 1;
    }
    size_t bytes = (n_sc - 1) * sstride + sizeof(Fr);
    int rc = reserve(ctx, ctx->cur->ws_sc, bytes);
    if (rc) return rc;
    CK(cudaMemcpyAsync(ctx->cur->ws_sc.p, scalars, bytes, cudaMemcpyHostToDevice, ctx->cur->stream));
    *d_sc = (const uint8_t *)ctx->cur->ws_sc.p;
    *d_idx = nullptr;
    if (index) {
        static_assert(sizeof(size_t) == 8, "size_t");
        rc = reserve(ctx, ctx->cur->ws_idx, n * 8);
        if (rc) return rc;
        CK(cudaMemcpyAsync(ctx->cur->ws_idx.p, index, n * 8, cudaMemcpyHostToDevice, ctx->cur->stream));
        *d_idx = (const uint64_t *)ctx->cur->ws_idx.p;
    }
    return 0;
}

int msm_host(zecale_gpu_ctx *ctx, const void *bases, size_t bstride, const void *scalars, size_t sstride,
             const size_t *index, size_t n, void *out)
{
    if (!ctx) return ZECALE_GPU_E_ARG;
    std::lock_guard<std::mutex> lk(ctx->mu);
    ctx->cur = &ctx->lanes[0];
    if (!out || (n && (!bases || !scalars)) || bstride < sizeof(Proj) || sstride < sizeof(Fr) || n >= (1ull << 26))
        return fail(ctx, ZECALE_GPU_E_ARG, "zecale_gpu_msm: bad argument");
    CK(cudaSetDevice(ctx->device));
    if (n == 0) return msm_device(ctx, nullptr, nullptr, 0, 0, 0, nullptr, 0, nullptr, 0, out, nullptr);
    int rc = reserve(ctx, ctx->ws_tmp, n * sizeof(Affine) + n);
    if (rc) return rc;
    Affine *d_aff = (Affine *)ctx->ws_tmp.p;
    uint8_t *d_inf = (uint8_t *)(d_aff + n);
    rc = import_bases(ctx, bases, bstride, n, d_aff, d_inf);
    if (rc) return rc;
    const uint8_t *d_sc;
    const uint64_t *d_idx;
    rc = upload_scalars(ctx, scalars, sstride, index, n, &d_sc, &d_idx);
    if (rc) return rc;
    return msm_device(ctx, d_aff, d_inf, 0, 0, (uint32_t)n, d_sc, sstride, d_idx, n, out, nullptr);
}

// NTT on a device vector (or `batch` vectors back to back), no locking, no sync; timed as kernel class 1.
int ntt_device(zecale_gpu_ctx *ctx, Fr *d_a, unsigned log_m, unsigned flags, const Fr *coset_g, uint32_t batch = 1)
{
    size_t m = (size_t)1 << log_m;
    int rc = reserve(ctx, ctx->cur->ws_ntt, (size_t)batch * m * sizeof(Fr));
    if (rc) return rc;
    unsigned nl = 0;
    // tables are built (and synchronised) before the timed region
    const Fr *tw = nullptr;
    CK(ntt_get_twiddles(ctx->ntt, log_m, ctx->cur->stream, &tw, &nl));
    cudaEvent_t e0 = take_event(ctx), e1 = take_event(ctx);
    CK(cudaEventRecord(e0, ctx->cur->stream));
    CK(ntt_launch(ctx->ntt, d_a, (Fr *)ctx->cur->ws_ntt.p, log_m, flags, coset_g, ctx->cur->stream, &nl, batch));
    CK(cudaEventRecord(e1, ctx->cur->stream));
    ctx->ev_pending[1].emplace_back(e0, e1);
    if (ctx->ev_pending[1].size() > 256) drain_timers(ctx, false);
    ctx->launches += nl;
    return 0;
}

// (g^m - 1)^-1 on the device, cached per (log_m, g).
int get_zinv(zecale_gpu_ctx *ctx, unsigned log_m, const Fr *g, const Fr **out)
{
    auto key = std::make_pair((uint32_t)log_m, std::vector<uint32_t>(g->l, g->l + 12));
    auto it = ctx->zinv.find(key);
    if (it != ctx->zinv.end()) {
        *out = it->second;
        return 0;
    }
    Fr *d = nullptr;
    CK(cudaMalloc((void **)&d, sizeof(Fr)));
    k_fr_zinv<<<1, 1, 0, ctx->cur->stream>>>(d, *g, log_m);
    ctx->launches++;
    CK(cudaGetLastError());
    CK(cudaStreamSynchronize(ctx->cur->stream)); // other lanes may read it right away
    ctx->zinv[key] = d;
    *out = d;
    return 0;
}

// The device part of r1cs_to_qap_witness_map (libsnark r1cs_to_qap.tcc:252-330) for d1 = d2 = d3 = 0:
// aA, aB, aC hold the evaluations <A_i,w>, <B_i,w>, <C_i,w> on the domain; on return aA holds
// coefficients_for_H[0..m) (entry m is zero). aB and aC are clobbered.
int qap_h_device(zecale_gpu_ctx *ctx, Fr *aA, Fr *aB, Fr *aC, unsigned log_m, const Fr *g)
{
    const size_t m = (size_t)1 << log_m;
    const Fr *zinv = nullptr;
    int rc = get_zinv(ctx, log_m, g, &zinv);
    if (rc) return rc;
    // the seven transforms are on the proof's critical path (the H MSM waits for them): run them on
    // the lane's high-priority stream, in order with the rest of the lane
    Lane &L = *ctx->cur;
    trace_mark(ctx, "qap0", L.stream);
    CK(cudaEventRecord(L.ev_hand, L.stream));
    CK(cudaStreamWaitEvent(L.hi, L.ev_hand, 0));
    std::swap(L.stream, L.hi);
    if (aB == aA + m && aC == aB + m) {
        // the three chains are independent and sit back to back: one launch per pass for all of them
        rc = ntt_device(ctx, aA, log_m, ZECALE_GPU_NTT_INVERSE, nullptr, 3);
        if (rc == 0) rc = ntt_device(ctx, aA, log_m, ZECALE_GPU_NTT_COSET, g, 3);
    } else {
        Fr *v[3] = {aA, aB, aC};
        for (int k = 0; k < 3 && rc == 0; k++) {
            rc = ntt_device(ctx, v[k], log_m, ZECALE_GPU_NTT_INVERSE, nullptr);
            if (rc == 0) rc = ntt_device(ctx, v[k], log_m, ZECALE_GPU_NTT_COSET, g);
        }
    }
    if (rc == 0) {
        k_qap_pointwise<<<(unsigned)((m + 255) / 256), 256, 0, L.stream>>>(aA, aB, aC, zinv, m);
        ctx->launches++;
        rc = ntt_device(ctx, aA, log_m, ZECALE_GPU_NTT_INVERSE | ZECALE_GPU_NTT_COSET, g);
    }
    std::swap(L.stream, L.hi);
    if (rc) return rc;
    CK(cudaGetLastError());
    CK(cudaEventRecord(L.ev_back, L.hi));
    CK(cudaStreamWaitEvent(L.stream, L.ev_back, 0));
    trace_mark(ctx, "qap1", L.stream);
    return 0;
}

// ------------------------------------------------------------------ test kernels
__global__ void __launch_bounds__(64, 1) k_test_fq(int op, const Fq *a, const Fq *b, Fq *out, size_t n)
{
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    FqSlots S;
    vs_set(S, VS_QX, a + i);
    vs_set(S, VS_QY, b ? b + i : a + i);
    if (op == 4 || op == 7) { // inverse: 4 = binary Euclid (product path), 7 = Fermat ladder (cross-check); 0 -> 0
        vs_set(S, VS_K1, a + i);
        vs_set(S, VS_T1, a + i);
        fqvm_run(S, op == 4 ? VPC_INVPROD : VPC_INV_FERMAT, [&](uint32_t, uint32_t &, FqSlots &) { return false; });
        vs_get(out + i, S, VS_T0);
        return;
    }
    uint32_t x[24], y[24], r[24];
    vs_load(x, S, VS_QX);
    vs_load(y, S, VS_QY);
    switch (op) {
    case 0: fq_add_regs(r, x, y); break;
    case 1: fq_sub_regs(r, x, y); break;
    case 2: fq_mul_regs(r, x, y); break;
    case 3: fq_sqr_regs(r, x); break;
    case 8: fq_mul_cios_regs(r, x, y); break; // the interleaved r01 body: cross-check of the Karatsuba product on the device
    case 5: {
#pragma unroll
        for (int k = 0; k < 24; k++) y[k] = 0;
        fq_sub_regs(r, y, x);
        break;
    }
    case 6: {
#pragma unroll
        for (int k = 0; k < 24; k++) y[k] = k == 0 ? 1u : 0u;
        fq_mul_regs(r, x, y);
        break;
    }
    default:
#pragma unroll
        for (int k = 0; k < 24; k++) r[k] = 0;
    }
    vs_store(S, VS_T0, r);
    vs_get(out + i, S, VS_T0);
}

__global__ void k_test_fr(int op, const Fr *a, const Fr *b, Fr *out, size_t n)
{
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint32_t x[12], y[12], r[12];
    fr_load(x, a + i);
    if (b)
        fr_load(y, b + i);
    else
        fr_load(y, a + i);
    switch (op) {
    case 0: fr_add_regs(r, x, y); break;
    case 1: fr_sub_regs(r, x, y); break;
    case 2: fr_mul_regs(r, x, y); break;
    case 3: fr_sqr_regs(r, x); break;
    case 4: fr_inv(r, x); break;
    case 5: {
        uint32_t z[12];
#pragma unroll
        for (int k = 0; k < 12; k++) z[k] = 0;
        fr_sub_regs(r, z, x);
        break;
    }
    case 6: fr_from_mont(r, x); break;
    default:
#pragma unroll
        for (int k = 0; k < 12; k++) r[k] = 0;
    }
    fr_store(out + i, r);
}

// acc = a[i] (any projective), then op with b[i], result normalised to special form.
__global__ void __launch_bounds__(64, 1) k_test_group(int op, const Proj *a, const Proj *b, Proj *out, size_t n)
{
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    FqSlots S;
    auto load_proj = [&](const Proj *p) {
        vs_set(S, VS_QX, &p->X);
        vs_set(S, VS_QY, &p->Y);
        vs_set(S, VS_QZZ, &p->Z);
    };
    load_proj(a + i);
    uint32_t stage = 0;
    fqvm_run(S, VPC_FROMPROJ, [&](uint32_t id, uint32_t &pc, FqSlots &S_) {
        switch (stage++) {
        case 0: pc = VPC_MOVQA; return true;       // acc = a
        case 1:
            if (op == 2) {
                pc = VPC_DBL;
                stage = 3;
                return true;
            }
            if (op == 3) {
                pc = VPC_TOAFF;
                stage = 4;
                return true;
            }
            load_proj(b + i);
            if (op == 1) { // mixed: b is special (Z = one or 0); infinity is filtered by the callers of madd
                if (vs_is_zero(S_, VS_QZZ)) {
                    pc = VPC_TOAFF;
                    stage = 4;
                } else {
                    pc = VPC_MADD;
                    stage = 3;
                }
                return true;
            }
            pc = VPC_FROMPROJ;
            return true;
        case 2: pc = VPC_ADD; return true;
        case 3: pc = VPC_TOAFF; return true;
        default: break;
        }
        Proj *o = out + i;
        if (id == 2) {
            vs_zero(S_, VS_T0);
            vs_get(&o->X, S_, VS_T0);
            vs_get(&o->Z, S_, VS_T0);
            vs_one(S_, VS_T0);
            vs_get(&o->Y, S_, VS_T0);
        } else {
            vs_get(&o->X, S_, VS_QX);
            vs_get(&o->Y, S_, VS_QY);
            vs_one(S_, VS_T0);
            vs_get(&o->Z, S_, VS_T0);
        }
        return false;
    });
}

// `iters` dependent Montgomery products per thread through the VM (local-memory operands,
// same code path as the MSM kernels).
__global__ void __launch_bounds__(128, 2) k_bench_fq_mul(Fq *sink, unsigned iters)
{
    FqSlots S;
    unsigned t = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t x[24], y[24];
#pragma unroll
    for (int i = 0; i < 24; i++) {
        x[i] = FQ_R2[i] ^ (t * 2654435761u >> (i & 7));
        y[i] = FQ_ONE[i] + t;
    }
    x[23] &= 0xffffff;
    y[23] &= 0xffffff;
    vs_store(S, VS_ZZZ, x);
    vs_store(S, VS_K1, y);
    unsigned k = 0;
    fqvm_run(S, VPC_PRODSTEP, [&](uint32_t, uint32_t &pc, FqSlots &) {
        pc = VPC_PRODSTEP;
        return ++k < iters;
    });
    vs_load(x, S, VS_K1);
    if (x[0] == 0x12345678u && x[5] == 77u) vs_get(sink, S, VS_K1); // keep the chain alive
}

// Pipe-rate probe: 8 independent chains of IMAD.WIDE.U32 per thread (the operand is perturbed by a LOP3 so the
// chain cannot be strength-reduced); 8 * iters wide multiply-adds per thread.
__global__ void __launch_bounds__(256) k_bench_imad(uint64_t *sink, uint32_t a, uint32_t b, unsigned iters)
{
    uint64_t x[8];
#pragma unroll
    for (int k = 0; k < 8; k++) x[k] = threadIdx.x + k;
    for (unsigned i = 0; i < iters; i++)
#pragma unroll
        for (int k = 0; k < 8; k++) x[k] = (uint64_t)((uint32_t)x[k] ^ a) * b + x[k];
    uint64_t s = 0;
#pragma unroll
    for (int k = 0; k < 8; k++) s += x[k];
    if (s == 0x123456789abcdefull) sink[0] = s; // keep the chains alive
}

// out = sum of n projective points (any Z), normalised to libff special form. One thread: n is
// the number of ranks / chunks (multi_exp's serial "final = final + partial[i]", multiexp.tcc:680-685).
__global__ void __launch_bounds__(32, 1) k_point_sum(const Proj *pts, uint32_t n, Proj *out)
{
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    FqSlots S;
    vs_acc_inf(S);
    auto load = [&](FqSlots &S_, uint32_t i) {
        vs_set(S_, VS_QX, &pts[i].X);
        vs_set(S_, VS_QY, &pts[i].Y);
        vs_set(S_, VS_QZZ, &pts[i].Z);
    };
    uint32_t i = 0, stage = 0; // 0: conversion of point i finished, 1: its addition finished, 2: normalised
    load(S, 0);
    fqvm_run(S, VPC_FROMPROJ, [&](uint32_t id, uint32_t &pc, FqSlots &S_) {
        if (stage == 2) {
            if (id == 2) { // infinity
                vs_zero(S_, VS_T0);
                vs_get(&out->X, S_, VS_T0);
                vs_get(&out->Z, S_, VS_T0);
                vs_one(S_, VS_T0);
                vs_get(&out->Y, S_, VS_T0);
            } else {
                vs_get(&out->X, S_, VS_QX);
                vs_get(&out->Y, S_, VS_QY);
                vs_one(S_, VS_T0);
                vs_get(&out->Z, S_, VS_T0);
            }
            return false;
        }
        if (stage == 0) {
            stage = 1;
            pc = VPC_ADD;
            return true;
        }
        if (++i < n) {
            load(S_, i);
            stage = 0;
            pc = VPC_FROMPROJ;
            return true;
        }
        stage = 2;
        pc = VPC_TOAFF;
        return true;
    });
}

template <typename T>
int test_elementwise(zecale_gpu_ctx *ctx, void (*launch)(int, const T *, const T *, T *, size_t, cudaStream_t), int op,
                     const void *a, const void *b, void *out, size_t n)
{
    if (!ctx) return ZECALE_GPU_E_ARG;
    std::lock_guard<std::mutex> lk(ctx->mu);
    ctx->cur = &ctx->lanes[0];
    if (!a || !out || !n) return fail(ctx, ZECALE_GPU_E_ARG, "test hook: bad argument");
    CK(cudaSetDevice(ctx->device));
    size_t bytes = n * sizeof(T);
    int rc = reserve(ctx, ctx->ws_tmp, bytes * 3);
    if (rc) return rc;
    T *da = (T *)ctx->ws_tmp.p, *db = da + n, *dout = db + n;
    CK(cudaMemcpyAsync(da, a, bytes, cudaMemcpyHostToDevice, ctx->cur->stream));
    if (b) CK(cudaMemcpyAsync(db, b, bytes, cudaMemcpyHostToDevice, ctx->cur->stream));
    launch(op, da, b ? db : nullptr, dout, n, ctx->cur->stream);
    ctx->launches++;
    CK(cudaGetLastError());
    CK(cudaMemcpyAsync(out, dout, bytes, cudaMemcpyDeviceToHost, ctx->cur->stream));
    CK(cudaStreamSynchronize(ctx->cur->stream));
    return 0;
}

void launch_fq(int op, const Fq *a, const Fq *b, Fq *o, size_t n, cudaStream_t s)
{
    k_test_fq<<<(unsigned)((n + 63) / 64), 64, 0, s>>>(op, a, b, o, n);
}
void launch_fr(int op, const Fr *a, const Fr *b, Fr *o, size_t n, cudaStream_t s)
{
    k_test_fr<<<(unsigned)((n + 63) / 64), 64, 0, s>>>(op, a, b, o, n);
}
void launch_group(int op, const Proj *a, const Proj *b, Proj *o, size_t n, cudaStream_t s)
{
    k_test_group<<<(unsigned)((n + 63) / 64), 64, 0, s>>>(op, a, b, o, n);
}

// Device copy of the host byte range [host, host+bytes) for the current lane: reuses an upload
// made earlier in this epoch (any lane) if it covers the range, else uploads on this lane.
int epoch_upload(zecale_gpu_ctx *ctx, const void *host, size_t bytes, const uint8_t **d_out)
{
    const char *h = (const char *)host;
    for (auto &u : ctx->uploads)
        if (h >= u.host && h + bytes <= u.host + u.bytes) {
            CK(cudaStreamWaitEvent(ctx->cur->stream, u.ev, 0));
            *d_out = (const uint8_t *)ctx->up_pool[u.buf].p + (h - u.host);
            return 0;
        }
    if (ctx->uploads.size() >= 8) return 1; // pool exhausted: caller falls back to a lane-private upload
    int slot = (int)ctx->uploads.size();
    int rc = reserve(ctx, ctx->up_pool[slot], bytes);
    if (rc) return rc;
    CK(cudaMemcpyAsync(ctx->up_pool[slot].p, host, bytes, cudaMemcpyHostToDevice, ctx->cur->stream));
    cudaEvent_t ev = take_event(ctx);
    CK(cudaEventRecord(ev, ctx->cur->stream));
    ctx->uploads.push_back({h, bytes, slot, ev});
    *d_out = (const uint8_t *)ctx->up_pool[slot].p;
    return 0;
}


// ------------------------------------------------------------------ multi-device contexts
struct DevGuard { // cudaSetDevice for the scope of a call into a worker
    int prev = 0;
    explicit DevGuard(int dev)
    {
        cudaGetDevice(&prev);
        cudaSetDevice(dev);
    }
    ~DevGuard() { cudaSetDevice(prev); }
};

bool is_multi(const zecale_gpu_ctx *ctx) { return !ctx->workers.empty(); }

// Shard g of n points over G workers: libff::multi_exp's equal chunks (multiexp.tcc:663-676), except
// that the first worker - whose device also runs the witness map of every proof - takes `w0` of an
// equal share ($ZECALE_GPU_SHARD0_WEIGHT, default 1 - 0.05 (G - 1)).
void shard_bounds(size_t n, size_t G, std::vector<size_t> &lo)
{
    lo.assign(G + 1, n);
    lo[0] = 0;
    if (n < G || G == 1) return; // too few points: the first worker takes them all
    const char *e = getenv("ZECALE_GPU_SHARD0_WEIGHT");
    double w0 = e ? atof(e) : 1.0 - 0.05 * (double)(G - 1);
    if (w0 < 0.05) w0 = 0.05;
    if (w0 > 1.0) w0 = 1.0;
    const double one = (double)n / (w0 + (double)(G - 1));
    for (size_t g = 1; g < G; g++) lo[g] = (size_t)(one * (w0 + (double)(g - 1)));
}

// Second halves of the workers' pending shard MSMs, then (once per epoch) the gather on devices[0]:
// sum of the partial sums of every sharded MSM, normalisation, results to the lanes' staging areas.
int multi_flush(zecale_gpu_ctx *ctx)
{
    for (zecale_gpu_ctx *w : ctx->workers) {
        DevGuard dg(w->device);
        int rc = flush_tails(w);
        if (rc) return fail(ctx, rc, "worker on device %d: %s", w->device, w->err.c_str());
    }
    TailJob jobs[ZECALE_GPU_LANES];
    int lanes[ZECALE_GPU_LANES];
    uint32_t n = 0;
    for (int l = 0; l < ZECALE_GPU_LANES; l++) {
        auto &g = ctx->gather[l];
        if (!g.pending || g.flushed) continue;
        g.flushed = true;
        TailJob j{};
        j.fin = ctx->d_gather + (size_t)l * zecale_gpu_ctx::MAX_DEVICES;
        j.n_fin = g.n_parts;
        j.result = ctx->d_gather_res + l;
        jobs[n] = j;
        lanes[n++] = l;
    }
    if (n == 0) return 0;
    for (zecale_gpu_ctx *w : ctx->workers) CK(cudaStreamWaitEvent(ctx->gstream, w->ev_parts, 0));
    CK(msm_launch_gather_batch(jobs, n, ctx->gstream));
    ctx->launches += 2;
    for (uint32_t i = 0; i < n; i++)
        CK(cudaMemcpyAsync(&ctx->lanes[lanes[i]].h->result, ctx->d_gather_res + lanes[i], sizeof(Proj), cudaMemcpyDeviceToHost,
                           ctx->gstream));
    CK(cudaEventRecord(ctx->ev_gdone, ctx->gstream));
    for (uint32_t i = 0; i < n; i++) CK(cudaStreamWaitEvent(ctx->lanes[lanes[i]].stream, ctx->ev_gdone, 0));
    return 0;
}

int multi_sync(zecale_gpu_ctx *ctx)
{
    int rc = multi_flush(ctx);
    int rc1 = sync_lanes(ctx); // delivers the gathered results (and the coordinator's own lanes)
    if (rc == 0) rc = rc1;
    for (zecale_gpu_ctx *w : ctx->workers) {
        DevGuard dg(w->device);
        int rcw = sync_lanes(w);
        if (rcw && rc == 0) rc = fail(ctx, rcw, "worker on device %d: %s", w->device, w->err.c_str());
    }
    for (auto &g : ctx->gather) g = zecale_gpu_ctx::Gather();
    return rc;
}

// The whole assignment on every worker (epoch upload table), so that the shard MSMs of A, B and L find
// their scalar ranges without further copies.
int multi_broadcast_host(zecale_gpu_ctx *ctx, int lane, const void *host, size_t bytes)
{
    for (zecale_gpu_ctx *w : ctx->workers) {
        DevGuard dg(w->device);
        w->cur = &w->lanes[lane];
        const uint8_t *d = nullptr;
        int rc = epoch_upload(w, host, bytes, &d);
        if (rc > 1) return fail(ctx, rc, "worker on device %d: %s", w->device, w->err.c_str());
    }
    return 0;
}

// Shard g of an MSM over the first n bases of a sharded image on lane `lane` of worker g, partial sums
// routed to the coordinator's gather buffer. Scalars: host pointers, or (from_qap) the coefficients of
// the H polynomial in the coordinator's lane on devices[0], sent to the other devices over NVLink.
int multi_msm_async(zecale_gpu_ctx *ctx, int lane, const zecale_gpu_bases *h, const void *scalars, size_t sstride,
                    const size_t *index, size_t n, void *out_point, bool from_qap)
{
    Lane &PL = ctx->lanes[lane];
    if (PL.user_out || ctx->gather[lane].pending) {
        int rc0 = multi_sync(ctx);
        if (rc0) return rc0;
    }
    if (h->shards.size() != ctx->workers.size()) return fail(ctx, ZECALE_GPU_E_ARG, "pinned image was not made by this multi-device context");
    uint32_t parts = 0;
    for (size_t g = 0; g < ctx->workers.size(); g++) {
        const size_t lo = h->lo[g], hi = h->lo[g + 1] < n ? h->lo[g + 1] : n;
        const zecale_gpu_bases *hs = h->shards[g];
        if (lo >= hi || !hs) continue;
        const size_t ng = hi - lo;
        zecale_gpu_ctx *w = ctx->workers[g];
        DevGuard dg(w->device);
        w->cur = &w->lanes[lane];
        Lane &WL = *w->cur;
        if (WL.busy) return fail(ctx, ZECALE_GPU_E_ARG, "lane %d already holds an MSM of this proof", lane);
        const uint8_t *d_sc = nullptr;
        const uint64_t *d_idx = nullptr;
        size_t stride = sstride;
        int rc = 0;
        if (from_qap) {
            const uint8_t *src = (const uint8_t *)PL.ws_qap.p + lo * sizeof(Fr);
            stride = sizeof(Fr);
            CK(cudaStreamWaitEvent(WL.stream, PL.ev_qap, 0)); // this lane's witness map, not whichever ran last
            if (w->device == ctx->device) {
                d_sc = src;
            } else {
                rc = reserve(w, WL.ws_sc, ng * sizeof(Fr));
                if (rc) return fail(ctx, rc, "worker on device %d: %s", w->device, w->err.c_str());
                CK(cudaMemcpyPeerAsync(WL.ws_sc.p, w->device, src, ctx->device, ng * sizeof(Fr), WL.stream));
                d_sc = (const uint8_t *)WL.ws_sc.p;
            }
        } else {
            const size_t *idx = index ? index + lo : nullptr;
            const char *sc = (const char *)scalars;
            size_t first = lo, count = ng; // scalar range touched
            if (idx) {
                size_t mx = 0;
                for (size_t i = 0; i < ng; i++)
                    if (idx[i] > mx) mx = idx[i];
                first = 0;
                count = mx + 1;
            }
            rc = epoch_upload(w, sc + first * sstride, (count - 1) * sstride + sizeof(Fr), &d_sc);
            if (rc == 1) { // upload table full: lane-private copy
                rc = reserve(w, WL.ws_sc, (count - 1) * sstride + sizeof(Fr));
                if (rc == 0) {
                    CK(cudaMemcpyAsync(WL.ws_sc.p, sc + first * sstride, (count - 1) * sstride + sizeof(Fr), cudaMemcpyHostToDevice, WL.stream));
                    d_sc = (const uint8_t *)WL.ws_sc.p;
                }
            }
            if (rc) return fail(ctx, rc, "worker on device %d: %s", w->device, w->err.c_str());
            if (idx) {
                const uint8_t *d_i = nullptr;
                rc = epoch_upload(w, idx, ng * 8, &d_i);
                if (rc == 1) {
                    rc = reserve(w, WL.ws_idx, ng * 8);
                    if (rc == 0) {
                        CK(cudaMemcpyAsync(WL.ws_idx.p, idx, ng * 8, cudaMemcpyHostToDevice, WL.stream));
                        d_i = (const uint8_t *)WL.ws_idx.p;
                    }
                }
                if (rc) return fail(ctx, rc, "worker on device %d: %s", w->device, w->err.c_str());
                d_idx = (const uint64_t *)d_i;
            }
            // on the witness-map device the accumulations start once the seven transforms are done
            if (ctx->qap_in_epoch && w->device == ctx->device) CK(cudaStreamWaitEvent(WL.stream, ctx->ev_qap, 0));
        }
        rc = msm_device(w, hs->d, hs->is_inf, hs->c, hs->V, (uint32_t)hs->n, d_sc, stride, d_idx, ng, nullptr, nullptr, true,
                        ctx->d_gather + (size_t)lane * zecale_gpu_ctx::MAX_DEVICES + parts, ctx->device);
        if (rc) return fail(ctx, rc, "worker on device %d: %s", w->device, w->err.c_str());
        parts++;
    }
    if (parts == 0) { // no point, or none below n: the zero point
        Proj z;
        memset(&z, 0, sizeof z);
        memcpy(z.Y.l, FQ_ONE_HOST, sizeof z.Y.l);
        memcpy(out_point, &z, sizeof z);
        return 0;
    }
    PL.user_out = out_point;
    ctx->gather[lane].pending = true;
    ctx->gather[lane].flushed = false;
    ctx->gather[lane].n_parts = parts;
    return 0;
}

} // namespace

// ====================================================================== C ABI
extern "C" {

int zecale_gpu_init(const int *devices, int n_devices, zecale_gpu_ctx **out)
{
    if (!out) return ZECALE_GPU_E_ARG;
    *out = nullptr;
    // one hardware work queue per stream (the default of 8 makes lanes share queues, which
    // serialises independent streams); only effective if CUDA is not initialised yet
    setenv("CUDA_DEVICE_MAX_CONNECTIONS", "32", 0);
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0) return ZECALE_GPU_E_CUDA;
    zecale_gpu_ctx *ctx = new zecale_gpu_ctx();
    if (devices && n_devices > 0) {
        ctx->devices.assign(devices, devices + n_devices);
    } else {
        int cur = 0;
        cudaGetDevice(&cur);
        ctx->devices.push_back(cur);
    }
    ctx->device = ctx->devices[0];
    cudaDeviceProp prop;
    if (cudaSetDevice(ctx->device) != cudaSuccess || cudaGetDeviceProperties(&prop, ctx->device) != cudaSuccess ||
        prop.major < 10) {
        delete ctx;
        return ZECALE_GPU_E_CUDA; // sm_100a only: no fallback path
    }
    bool ok = cudaMalloc((void **)&ctx->d_status, sizeof(int)) == cudaSuccess &&
              cudaEventCreateWithFlags(&ctx->ev_qap, cudaEventDisableTiming) == cudaSuccess &&
              cudaEventCreateWithFlags(&ctx->ev_parts, cudaEventDisableTiming) == cudaSuccess;
    int prio_lo = 0, prio_hi = 0;
    cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi); // numerically lower = higher priority
    for (Lane &L : ctx->lanes)
        ok = ok && cudaStreamCreateWithPriority(&L.stream, cudaStreamNonBlocking, prio_lo) == cudaSuccess &&
             cudaStreamCreateWithPriority(&L.hi, cudaStreamNonBlocking, prio_hi) == cudaSuccess &&
             cudaEventCreateWithFlags(&L.ev_hand, cudaEventDisableTiming) == cudaSuccess &&
             cudaEventCreateWithFlags(&L.ev_back, cudaEventDisableTiming) == cudaSuccess &&
             cudaEventCreateWithFlags(&L.ev_acc, cudaEventDisableTiming) == cudaSuccess &&
             cudaEventCreateWithFlags(&L.ev_qap, cudaEventDisableTiming) == cudaSuccess &&
             cudaEventCreate(&L.ev0) == cudaSuccess && cudaEventCreate(&L.ev1) == cudaSuccess &&
             cudaMallocHost((void **)&L.h, sizeof(Lane::Staging)) == cudaSuccess;
    ctx->cur = &ctx->lanes[0];
    const char *tr = getenv("ZECALE_GPU_TRACE");
    ctx->trace = tr && tr[0] == '1';
    if (ok && ctx->devices.size() > 1) {
        // multi-device context: one worker per listed device (the first shares devices[0] with this
        // coordinator), peer access between all pairs so that the partial sums and the H coefficients
        // travel over NVLink, a gather buffer and a high-priority gather stream on devices[0]
        ok = ctx->devices.size() <= (size_t)zecale_gpu_ctx::MAX_DEVICES;
        for (size_t a = 0; ok && a < ctx->devices.size(); a++) {
            // test switch: with ZECALE_GPU_TEST_REPEAT_DEVICES=1 the same ordinal may be listed more than once, so that the
            // sharding, the per-shard Pippenger and the gather can be exercised on a one-GPU box (tests/test_multi_gpu.py)
            const char *rep = getenv("ZECALE_GPU_TEST_REPEAT_DEVICES");
            for (size_t b = 0; b < a && !(rep && rep[0] == '1'); b++) ok = ok && ctx->devices[a] != ctx->devices[b];
            ok = ok && ctx->devices[a] >= 0 && ctx->devices[a] < count;
        }
        for (size_t a = 0; ok && a < ctx->devices.size(); a++) {
            cudaSetDevice(ctx->devices[a]);
            for (size_t b = 0; b < ctx->devices.size(); b++) {
                if (a == b || ctx->devices[a] == ctx->devices[b]) continue;
                cudaError_t pe = cudaDeviceEnablePeerAccess(ctx->devices[b], 0);
                if (pe != cudaSuccess && pe != cudaErrorPeerAccessAlreadyEnabled) {
                    // no direct path: cudaMemcpyPeerAsync still works (staged), keep going
                }
                cudaGetLastError();
            }
        }
        for (size_t a = 0; ok && a < ctx->devices.size(); a++) {
            zecale_gpu_ctx *w = nullptr;
            ok = zecale_gpu_init(&ctx->devices[a], 1, &w) == 0;
            if (ok) ctx->workers.push_back(w);
        }
        cudaSetDevice(ctx->device);
        ok = ok && cudaMalloc((void **)&ctx->d_gather, sizeof(XYZZ) * ZECALE_GPU_LANES * zecale_gpu_ctx::MAX_DEVICES) == cudaSuccess &&
             cudaMalloc((void **)&ctx->d_gather_res, sizeof(Proj) * ZECALE_GPU_LANES) == cudaSuccess &&
             cudaStreamCreateWithPriority(&ctx->gstream, cudaStreamNonBlocking, prio_hi) == cudaSuccess &&
             cudaEventCreateWithFlags(&ctx->ev_gdone, cudaEventDisableTiming) == cudaSuccess;
    }
    if (!ok) {
        zecale_gpu_shutdown(ctx);
        return ZECALE_GPU_E_CUDA;
    }
    *out = ctx;
    return 0;
}

void zecale_gpu_shutdown(zecale_gpu_ctx *ctx)
{
    if (!ctx) return;
    for (zecale_gpu_ctx *w : ctx->workers) zecale_gpu_shutdown(w);
    ctx->workers.clear();
    cudaSetDevice(ctx->device);
    ctx->cur = &ctx->lanes[0];
    drain_timers(ctx);
    if (ctx->d_gather) cudaFree(ctx->d_gather);
    if (ctx->d_gather_res) cudaFree(ctx->d_gather_res);
    if (ctx->gstream) cudaStreamDestroy(ctx->gstream);
    if (ctx->ev_gdone) cudaEventDestroy(ctx->ev_gdone);
    if (ctx->ev_parts) cudaEventDestroy(ctx->ev_parts);
    for (Lane &L : ctx->lanes) {
        for (DevBuf *b : {&L.ws_msm, &L.ws_sc, &L.ws_idx, &L.ws_ntt, &L.ws_qap})
            if (b->p) cudaFree(b->p);
        if (L.h) cudaFreeHost(L.h);
        if (L.ev0) cudaEventDestroy(L.ev0);
        if (L.ev1) cudaEventDestroy(L.ev1);
        if (L.ev_hand) cudaEventDestroy(L.ev_hand);
        if (L.ev_back) cudaEventDestroy(L.ev_back);
        if (L.ev_acc) cudaEventDestroy(L.ev_acc);
        if (L.ev_qap) cudaEventDestroy(L.ev_qap);
        if (L.hi) cudaStreamDestroy(L.hi);
        if (L.stream) cudaStreamDestroy(L.stream);
    }
    for (DevBuf *b : {&ctx->ws_in, &ctx->ws_tmp})
        if (b->p) cudaFree(b->p);
    for (DevBuf &b : ctx->up_pool)
        if (b.p) cudaFree(b.p);
    ntt_tables_free(ctx->ntt);
    for (cudaEvent_t e : ctx->ev_pool) cudaEventDestroy(e);
    for (auto &kv : ctx->zinv) cudaFree(kv.second);
    if (ctx->d_status) cudaFree(ctx->d_status);
    if (ctx->ev_qap) cudaEventDestroy(ctx->ev_qap);
    delete ctx;
}

const char *zecale_gpu_last_error(zecale_gpu_ctx *ctx) { return ctx ? ctx->err.c_str() : "null context"; }

int zecale_gpu_msm_g1(zecale_gpu_ctx *ctx, const void *bases, size_t base_stride, const void *scalars,
                      size_t scalar_stride, const size_t *scalar_index, size_t n, void *out_point)
{
    return msm_host(ctx, bases, base_stride, scalars, scalar_stride, scalar_index, n, out_point);
}

// G2 lives over the same Fq with a = 0: identical arithmetic (tools/gen_fqvm_prog.py docstring).
int zecale_gpu_msm_g2(zecale_gpu_ctx *ctx, const void *bases, size_t base_stride, const void *scalars,
                      size_t scalar_stride, const size_t *scalar_index, size_t n, void *out_point)
{
    return msm_host(ctx, bases, base_stride, scalars, scalar_stride, scalar_index, n, out_point);
}

// Window size for a fully expanded image (one bucket set shared by all W = ceil(378/c) windows):
// the c that minimises the Fq products of one MSM over n points,
//     10 * (n W - non-empty buckets)   mixed additions (the first entry of a bucket is a copy), divided by the
//                                      fraction of 1.3 waves of accumulate threads (148 SM x 384) the non-empty
//                                      buckets (= tasks) fill: measured, 33 k tasks run at 0.46 of the pipe's peak
//   + 14 * 2 * 2^(c-1) * 1.5           additions of the tail (main tree + bit trees; weighted 1.5 because the
//                                      tail runs latency-bound, below the accumulation's pipe utilisation),
// with a surcharge when the top window holds so few bits that its entries pile up in a handful of
// buckets (those are cut into tasks and folded, which lengthens the latency-bound tail). For the
// batch-8 queries (n = 88 k ... 262 k) this gives c = 18: 378 = 21 * 18, no ragged top window, 131 072 tasks.
static uint32_t pinned_c(size_t n)
{
    uint32_t best_c = 8;
    double best = 1e300;
    for (uint32_t c = 8; c <= 20; c++) {
        const double W = (double)((FR_SCALAR_BITS + 1 + c - 1) / c), NB = (double)(1u << (c - 1));
        const double entries = (double)n * W;
        const double nonempty = NB * (1.0 - exp(-entries / NB));
        const double fill = nonempty / (1.3 * 148.0 * 384.0);
        double cost = 10.0 * (entries - nonempty) / (fill < 1.0 ? fill : 1.0) + 42.0 * NB;
        const uint32_t top_bits = FR_SCALAR_BITS + 1 - ((uint32_t)W - 1) * c; // data bits of the top window (incl. the carry bit)
        if (top_bits + 4 < c && (double)n / (double)(1u << top_bits) > 8.0 * (entries / NB + 8.0)) cost *= 1.03;
        if (cost < best) {
            best = cost;
            best_c = c;
        }
    }
    return best_c;
}

// format 0: libff in-memory special-form points (288 B); 1: libzeth key-file bytes (X || Y, 192 B big-endian plain)
static int pin_bases_impl(zecale_gpu_ctx *ctx, const void *bases, size_t base_stride, size_t n, int group, int expand,
                          zecale_gpu_bases **handle, int format);
static int pin_bases_impl(zecale_gpu_ctx *ctx, const void *bases, size_t base_stride, size_t n, int group, int expand,
                          zecale_gpu_bases **handle, int format)
{
    if (!ctx) return ZECALE_GPU_E_ARG;
    std::lock_guard<std::mutex> lk(ctx->mu);
    ctx->cur = &ctx->lanes[0];
    if (!bases || !handle || !n || base_stride < (format ? sizeof(Affine) : sizeof(Proj)) || (format && (base_stride & 3)) ||
        n >= (1ull << 26) || (group != ZECALE_GPU_G1 && group != ZECALE_GPU_G2))
        return fail(ctx, ZECALE_GPU_E_ARG, "zecale_gpu_pin_bases: bad argument");
    if (is_multi(ctx)) { // shard the points over the workers, each pins (and expands) its own range
        zecale_gpu_bases *h = new zecale_gpu_bases();
        h->n = n;
        h->group = group;
        shard_bounds(n, ctx->workers.size(), h->lo);
        h->shards.assign(ctx->workers.size(), nullptr);
        for (size_t g = 0; g < ctx->workers.size(); g++) {
            const size_t ng = h->lo[g + 1] - h->lo[g];
            if (!ng) continue;
            int rc = pin_bases_impl(ctx->workers[g], (const char *)bases + h->lo[g] * base_stride, base_stride, ng, group, expand,
                                    &h->shards[g], format);
            if (rc) {
                fail(ctx, rc, "worker on device %d: %s", ctx->workers[g]->device, ctx->workers[g]->err.c_str());
                for (size_t k = 0; k < g; k++)
                    if (h->shards[k]) zecale_gpu_unpin_bases(ctx->workers[k], h->shards[k]);
                delete h;
                cudaSetDevice(ctx->device);
                return rc;
            }
        }
        cudaSetDevice(ctx->device);
        *handle = h;
        return 0;
    }
    CK(cudaSetDevice(ctx->device));
    zecale_gpu_bases *h = new zecale_gpu_bases();
    h->n = n;
    h->group = group;
    // choose the expansion: G copies 2^(c*V*g) P, limited by a memory budget (<= 1/3 of free HBM)
    uint32_t c = 0, V = 0, G = 1;
    if (expand) {
        size_t free_b = 0, total_b = 0;
        CK(cudaMemGetInfo(&free_b, &total_b));
        c = pinned_c(n);
        const uint32_t W = (FR_SCALAR_BITS + 1 + c - 1) / c;
        size_t max_copies = (free_b / 3) / (n * sizeof(Affine));
        if (max_copies < 1) max_copies = 1;
        V = 1;
        while ((W + V - 1) / V > max_copies || (uint64_t)((W + V - 1) / V) * n >= (1ull << 31)) V++;
        G = (W + V - 1) / V;
        if (G == 1) { // nothing to expand: fall back to the default per-call plan
            c = 0;
            V = 0;
        }
    }
    h->c = c;
    h->V = V;
    cudaError_t e = cudaMalloc((void **)&h->d, (size_t)G * n * sizeof(Affine));
    if (e == cudaSuccess) e = cudaMalloc((void **)&h->is_inf, n);
    if (e != cudaSuccess) {
        cudaFree(h->d);
        delete h;
        return fail(ctx, ZECALE_GPU_E_NOMEM, "pin_bases: %s", cudaGetErrorString(e));
    }
    int rc = format ? import_bases_bytes(ctx, bases, base_stride, n, h->d, h->is_inf)
                    : import_bases(ctx, bases, base_stride, n, h->d, h->is_inf);
    if (rc == 0 && G > 1) {
        rc = reserve(ctx, ctx->ws_tmp, n * (sizeof(XYZZ) + sizeof(Fq)));
        if (rc == 0) {
            XYZZ *sp = (XYZZ *)ctx->ws_tmp.p;
            Fq *pf = (Fq *)(sp + n);
            unsigned threads = (unsigned)((n + EXPAND_K - 1) / EXPAND_K);
            for (uint32_t g = 1; g < G; g++) {
                k_expand_copy<<<(threads + 127) / 128, 128, 0, ctx->cur->stream>>>(h->d + (size_t)(g - 1) * n, h->d + (size_t)g * n,
                                                                               (uint32_t)n, c * V, sp, pf);
                ctx->launches++;
            }
            cudaError_t e2 = cudaStreamSynchronize(ctx->cur->stream);
            if (e2 != cudaSuccess) rc = fail(ctx, ZECALE_GPU_E_CUDA, "pin_bases expand: %s", cudaGetErrorString(e2));
        }
    }
    if (rc) {
        cudaFree(h->d);
        cudaFree(h->is_inf);
        delete h;
        return rc;
    }
    *handle = h;
    return 0;
}

int zecale_gpu_pin_bases_ex(zecale_gpu_ctx *ctx, const void *bases, size_t base_stride, size_t n, int group, int expand,
                            zecale_gpu_bases **handle)
{
    return pin_bases_impl(ctx, bases, base_stride, n, group, expand, handle, 0);
}

int zecale_gpu_pin_bases_bytes(zecale_gpu_ctx *ctx, const void *bytes, size_t stride, size_t n, int group, int expand,
                               zecale_gpu_bases **handle)
{
    return pin_bases_impl(ctx, bytes, stride, n, group, expand, handle, 1);
}

int zecale_gpu_pin_bases(zecale_gpu_ctx *ctx, const void *bases, size_t base_stride, size_t n, int group,
                         zecale_gpu_bases **handle)
{
    return zecale_gpu_pin_bases_ex(ctx, bases, base_stride, n, group, 1, handle);
}

int zecale_gpu_unpin_bases(zecale_gpu_ctx *ctx, zecale_gpu_bases *h)
{
    if (!ctx || !h) return ZECALE_GPU_E_ARG;
    if (!h->shards.empty()) {
        if (h->shards.size() != ctx->workers.size()) return ZECALE_GPU_E_ARG;
        {
            std::lock_guard<std::mutex> lk(ctx->mu);
            multi_sync(ctx);
        }
        for (size_t g = 0; g < h->shards.size(); g++)
            if (h->shards[g]) zecale_gpu_unpin_bases(ctx->workers[g], h->shards[g]);
        cudaSetDevice(ctx->device);
        delete h;
        return 0;
    }
    std::lock_guard<std::mutex> lk(ctx->mu);
    ctx->cur = &ctx->lanes[0];
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->cur->stream);
    cudaFree(h->d);
    cudaFree(h->is_inf);
    delete h;
    return 0;
}

int zecale_gpu_msm_pinned(zecale_gpu_ctx *ctx, const zecale_gpu_bases *h, const void *scalars, size_t scalar_stride,
                          const size_t *scalar_index, size_t n, void *out_point)
{
    if (!ctx) return ZECALE_GPU_E_ARG;
    std::lock_guard<std::mutex> lk(ctx->mu);
    ctx->cur = &ctx->lanes[0];
    if (!h || !out_point || n > h->n || (n && !scalars) || scalar_stride < sizeof(Fr))
        return fail(ctx, ZECALE_GPU_E_ARG, "zecale_gpu_msm_pinned: bad argument");
    CK(cudaSetDevice(ctx->device));
    if (is_multi(ctx)) {
        int rc = multi_msm_async(ctx, 0, h, scalars, scalar_stride, scalar_index, n, out_point, false);
        int rc1 = multi_sync(ctx);
        return rc ? rc : rc1;
    }
    if (!h->shards.empty()) // a sharded image has no device array of its own (d == nullptr): only its maker can use it
        return fail(ctx, ZECALE_GPU_E_ARG, "zecale_gpu_msm_pinned: the pinned image is sharded over a multi-device context");
    if (n == 0) return msm_device(ctx, nullptr, nullptr, 0, 0, 0, nullptr, 0, nullptr, 0, out_point, nullptr);
    const uint8_t *d_sc;
    const uint64_t *d_idx;
    int rc = upload_scalars(ctx, scalars, scalar_stride, scalar_index, n, &d_sc, &d_idx);
    if (rc) return rc;
    return msm_device(ctx, h->d, h->is_inf, h->c, h->V, (uint32_t)h->n, d_sc, scalar_stride, d_idx, n, out_point,
                      nullptr);
}

int zecale_gpu_msm_pinned_dev(zecale_gpu_ctx *ctx, const zecale_gpu_bases *h, const void *d_scalars,
                              size_t scalar_stride, const uint64_t *d_scalar_index, size_t n, void *out_point,
                              float *elapsed_ms)
{
    if (!ctx) return ZECALE_GPU_E_ARG;
    std::lock_guard<std::mutex> lk(ctx->mu);
    ctx->cur = &ctx->lanes[0];
    if (!h || !out_point || n > h->n || (n && !d_scalars) || scalar_stride < sizeof(Fr))
        return fail(ctx, ZECALE_GPU_E_ARG, "zecale_gpu_msm_pinned_dev: bad argument");
    if (is_multi(ctx)) return fail(ctx, ZECALE_GPU_E_ARG, "zecale_gpu_msm_pinned_dev: device-resident scalars belong to one device; use a single-device context");
    if (!h->shards.empty())
        return fail(ctx, ZECALE_GPU_E_ARG, "zecale_gpu_msm_pinned_dev: the pinned image is sharded over a multi-device context");
    CK(cudaSetDevice(ctx->device));
    return msm_device(ctx, h->d, h->is_inf, h->c, h->V, (uint32_t)h->n, (const uint8_t *)d_scalars, scalar_stride,
                      d_scalar_index, n, out_point, elapsed_ms);
}

int zecale_gpu_ntt_dev(zecale_gpu_ctx *ctx, void *d_a, unsigned log_m, unsigned flags, const void *coset_g,
                       float *elapsed_ms)
{
    if (!ctx) return ZECALE_GPU_E_ARG;
    std::lock_guard<std::mutex> lk(ctx->mu);
    ctx->cur = &ctx->lanes[0];
    if (!d_a || (flags & ~3u) || ((flags & ZECALE_GPU_NTT_COSET) && !coset_g))
        return fail(ctx, ZECALE_GPU_E_ARG, "zecale_gpu_ntt: bad argument");
    if (log_m < 1 || log_m > 28) return fail(ctx, ZECALE_GPU_E_DOMAIN, "zecale_gpu_ntt: log_m %u out of range", log_m);
    CK(cudaSetDevice(ctx->device));
    if (elapsed_ms) CK(cudaEventRecord(ctx->cur->ev0, ctx->cur->stream));
    int rc = ntt_device(ctx, (Fr *)d_a, log_m, flags, (const Fr *)coset_g);
    if (rc) return rc;
    if (elapsed_ms) CK(cudaEventRecord(ctx->cur->ev1, ctx->cur->stream));
    CK(cudaStreamSynchronize(ctx->cur->stream));
    if (elapsed_ms) CK(cudaEventElapsedTime(elapsed_ms, ctx->cur->ev0, ctx->cur->ev1));
    return 0;
}

int zecale_gpu_ntt(zecale_gpu_ctx *ctx, void *a, unsigned log_m, unsigned flags, const void *coset_g)
{
    if (!ctx) return ZECALE_GPU_E_ARG;
    std::lock_guard<std::mutex> lk(ctx->mu);
    ctx->cur = &ctx->lanes[0];
    if (!a || (flags & ~3u) || ((flags & ZECALE_GPU_NTT_COSET) && !coset_g))
        return fail(ctx, ZECALE_GPU_E_ARG, "zecale_gpu_ntt: bad argument");
    if (log_m < 1 || log_m > 28) return fail(ctx, ZECALE_GPU_E_DOMAIN, "zecale_gpu_ntt: log_m %u out of range", log_m);
    size_t bytes = ((size_t)1 << log_m) * sizeof(Fr);
    CK(cudaSetDevice(ctx->device));
    int rc = reserve(ctx, ctx->ws_in, bytes);
    if (rc) return rc;
    void *d = ctx->ws_in.p;
    CK(cudaMemcpyAsync(d, a, bytes, cudaMemcpyHostToDevice, ctx->cur->stream));
    rc = ntt_device(ctx, (Fr *)d, log_m, flags, (const Fr *)coset_g);
    if (rc) return rc;
    CK(cudaMemcpyAsync(a, d, bytes, cudaMemcpyDeviceToHost, ctx->cur->stream));
    CK(cudaStreamSynchronize(ctx->cur->stream));
    return 0;
}

int zecale_gpu_qap_h_dev(zecale_gpu_ctx *ctx, void *d_aA, void *d_aB, void *d_aC, unsigned log_m, const void *coset_g,
                         float *elapsed_ms)
{
    if (!ctx) return ZECALE_GPU_E_ARG;
    std::lock_guard<std::mutex> lk(ctx->mu);
    ctx->cur = &ctx->lanes[0];
    if (!d_aA || !d_aB || !d_aC || !coset_g) return fail(ctx, ZECALE_GPU_E_ARG, "zecale_gpu_qap_h: bad argument");
    if (log_m < 1 || log_m > 28) return fail(ctx, ZECALE_GPU_E_DOMAIN, "zecale_gpu_qap_h: log_m %u out of range", log_m);
    CK(cudaSetDevice(ctx->device));
    if (elapsed_ms) CK(cudaEventRecord(ctx->cur->ev0, ctx->cur->stream));
    int rc = qap_h_device(ctx, (Fr *)d_aA, (Fr *)d_aB, (Fr *)d_aC, log_m, (const Fr *)coset_g);
    if (rc) return rc;
    if (elapsed_ms) CK(cudaEventRecord(ctx->cur->ev1, ctx->cur->stream));
    CK(cudaStreamSynchronize(ctx->cur->stream));
    if (elapsed_ms) CK(cudaEventElapsedTime(elapsed_ms, ctx->cur->ev0, ctx->cur->ev1));
    return 0;
}

int zecale_gpu_qap_h(zecale_gpu_ctx *ctx, const void *aA, const void *aB, const void *aC, unsigned log_m,
                     const void *coset_g, void *out_h)
{
    if (!ctx) return ZECALE_GPU_E_ARG;
    std::lock_guard<std::mutex> lk(ctx->mu);
    ctx->cur = &ctx->lanes[0];
    if (!aA || !aB || !aC || !coset_g || !out_h) return fail(ctx, ZECALE_GPU_E_ARG, "zecale_gpu_qap_h: bad argument");
    if (log_m < 1 || log_m > 28) return fail(ctx, ZECALE_GPU_E_DOMAIN, "zecale_gpu_qap_h: log_m %u out of range", log_m);
    CK(cudaSetDevice(ctx->device));
    const size_t m = (size_t)1 << log_m, bytes = m * sizeof(Fr);
    int rc = reserve(ctx, ctx->cur->ws_qap, 3 * bytes);
    if (rc) return rc;
    Fr *d = (Fr *)ctx->cur->ws_qap.p;
    CK(cudaMemcpyAsync(d, aA, bytes, cudaMemcpyHostToDevice, ctx->cur->stream));
    CK(cudaMemcpyAsync(d + m, aB, bytes, cudaMemcpyHostToDevice, ctx->cur->stream));
    CK(cudaMemcpyAsync(d + 2 * m, aC, bytes, cudaMemcpyHostToDevice, ctx->cur->stream));
    rc = qap_h_device(ctx, d, d + m, d + 2 * m, log_m, (const Fr *)coset_g);
    if (rc) return rc;
    CK(cudaMemcpyAsync(out_h, d, bytes, cudaMemcpyDeviceToHost, ctx->cur->stream));
    memset((char *)out_h + bytes, 0, sizeof(Fr)); // coefficients_for_H[m] = 0 when d1 = d2 = d3 = 0
    CK(cudaStreamSynchronize(ctx->cur->stream));
    return 0;
}

int zecale_gpu_point_sum(zecale_gpu_ctx *ctx, const void *points, size_t n, void *out_point)
{
    if (!ctx) return ZECALE_GPU_E_ARG;
    std::lock_guard<std::mutex> lk(ctx->mu);
    ctx->cur = &ctx->lanes[0];
    if (!out_point || (n && !points) || n >= (1u << 20)) return fail(ctx, ZECALE_GPU_E_ARG, "zecale_gpu_point_sum: bad argument");
    CK(cudaSetDevice(ctx->device));
    if (n == 0) return msm_device(ctx, nullptr, nullptr, 0, 0, 0, nullptr, 0, nullptr, 0, out_point, nullptr);
    int rc = reserve(ctx, ctx->ws_tmp, (n + 1) * sizeof(Proj));
    if (rc) return rc;
    Proj *d = (Proj *)ctx->ws_tmp.p;
    CK(cudaMemcpyAsync(d, points, n * sizeof(Proj), cudaMemcpyHostToDevice, ctx->cur->stream));
    k_point_sum<<<1, 32, 0, ctx->cur->stream>>>(d, (uint32_t)n, d + n);
    ctx->launches++;
    CK(cudaGetLastError());
    CK(cudaMemcpyAsync(out_point, d + n, sizeof(Proj), cudaMemcpyDeviceToHost, ctx->cur->stream));
    CK(cudaStreamSynchronize(ctx->cur->stream));
    return 0;
}

static int ntt_dist_phase_impl(zecale_gpu_ctx *ctx, cudaStream_t user_stream, bool on_user_stream, const void *d_in, void *d_out,
                               unsigned log_m, unsigned flags, const void *coset_g, int world, int rank, int phase, int sync)
{
    if (!ctx) return ZECALE_GPU_E_ARG;
    std::lock_guard<std::mutex> lk(ctx->mu);
    ctx->cur = &ctx->lanes[0];
    if (!d_in || !d_out || d_in == d_out || (flags & ~3u) || ((flags & ZECALE_GPU_NTT_COSET) && !coset_g) || world < 1 ||
        rank < 0 || rank >= world || (phase != 0 && phase != 1))
        return fail(ctx, ZECALE_GPU_E_ARG, "zecale_gpu_ntt_dist_phase: bad argument");
    if (log_m < 4 || log_m > 24) return fail(ctx, ZECALE_GPU_E_DOMAIN, "zecale_gpu_ntt_dist_phase: log_m %u out of range", log_m);
    CK(cudaSetDevice(ctx->device));
    unsigned nl = 0;
    const Fr *tw = nullptr;
    CK(ntt_get_twiddles(ctx->ntt, log_m, ctx->cur->stream, &tw, &nl)); // tables outside the timed region (built once, synchronised)
    if ((flags & ZECALE_GPU_NTT_COSET) || (flags & ZECALE_GPU_NTT_INVERSE)) {
        Fr one_g;
        memcpy(one_g.l, FR_ONE_HOST, sizeof(Fr));
        const Fr *ctab = nullptr;
        CK(ntt_get_coset(ctx->ntt, log_m, (flags & ZECALE_GPU_NTT_COSET) ? *(const Fr *)coset_g : one_g, flags & 1u, ctx->cur->stream,
                         &ctab, &nl));
    }
    cudaStream_t st = on_user_stream ? user_stream : ctx->cur->stream;
    cudaEvent_t e0 = take_event(ctx), e1 = take_event(ctx);
    CK(cudaEventRecord(e0, st));
    cudaError_t e = ntt_dist_phase(ctx->ntt, (const Fr *)d_in, (Fr *)d_out, log_m, flags, (const Fr *)coset_g, (uint32_t)world,
                                   (uint32_t)rank, (uint32_t)phase, st, &nl);
    if (e == cudaErrorInvalidValue)
        return fail(ctx, ZECALE_GPU_E_DOMAIN, "zecale_gpu_ntt_dist_phase: world %d does not divide the 4-step factors of 2^%u", world, log_m);
    CK(e);
    CK(cudaEventRecord(e1, st));
    ctx->ev_pending[1].emplace_back(e0, e1);
    if (ctx->ev_pending[1].size() > 256) drain_timers(ctx, false);
    ctx->launches += nl;
    if (sync) CK(cudaStreamSynchronize(st));
    return 0;
}

int zecale_gpu_ntt_dist_phase(zecale_gpu_ctx *ctx, const void *d_in, void *d_out, unsigned log_m, unsigned flags,
                              const void *coset_g, int world, int rank, int phase, int sync)
{
    return ntt_dist_phase_impl(ctx, nullptr, false, d_in, d_out, log_m, flags, coset_g, world, rank, phase, sync);
}

int zecale_gpu_ntt_dist_phase_on(zecale_gpu_ctx *ctx, void *stream, const void *d_in, void *d_out, unsigned log_m, unsigned flags,
                                 const void *coset_g, int world, int rank, int phase)
{
    return ntt_dist_phase_impl(ctx, (cudaStream_t)stream, true, d_in, d_out, log_m, flags, coset_g, world, rank, phase, 0);
}

int zecale_gpu_qap_pointwise_dev_on(zecale_gpu_ctx *ctx, void *stream, void *d_aA, const void *d_aB, const void *d_aC, unsigned log_m,
                                    const void *coset_g, size_t count)
{
    if (!ctx) return ZECALE_GPU_E_ARG;
    std::lock_guard<std::mutex> lk(ctx->mu);
    ctx->cur = &ctx->lanes[0];
    if (!d_aA || !d_aB || !d_aC || !coset_g || !count) return fail(ctx, ZECALE_GPU_E_ARG, "zecale_gpu_qap_pointwise_dev_on: bad argument");
    if (log_m < 1 || log_m > 28) return fail(ctx, ZECALE_GPU_E_DOMAIN, "zecale_gpu_qap_pointwise_dev_on: log_m %u out of range", log_m);
    CK(cudaSetDevice(ctx->device));
    const Fr *zinv = nullptr;
    int rc = get_zinv(ctx, log_m, (const Fr *)coset_g, &zinv); // cached per (log_m, g); built (and synchronised) on first use
    if (rc) return rc;
    k_qap_pointwise<<<(unsigned)((count + 255) / 256), 256, 0, (cudaStream_t)stream>>>((Fr *)d_aA, (const Fr *)d_aB, (const Fr *)d_aC,
                                                                                    zinv, count);
    ctx->launches++;
    CK(cudaGetLastError());
    return 0;
}

int zecale_gpu_qap_pointwise_dev(zecale_gpu_ctx *ctx, void *d_aA, const void *d_aB, const void *d_aC, unsigned log_m,
                                 const void *coset_g, size_t count)
{
    if (!ctx) return ZECALE_GPU_E_ARG;
    std::lock_guard<std::mutex> lk(ctx->mu);
    ctx->cur = &ctx->lanes[0];
    if (!d_aA || !d_aB || !d_aC || !coset_g || !count) return fail(ctx, ZECALE_GPU_E_ARG, "zecale_gpu_qap_pointwise_dev: bad argument");
    if (log_m < 1 || log_m > 28) return fail(ctx, ZECALE_GPU_E_DOMAIN, "zecale_gpu_qap_pointwise_dev: log_m %u out of range", log_m);
    CK(cudaSetDevice(ctx->device));
    const Fr *zinv = nullptr;
    int rc = get_zinv(ctx, log_m, (const Fr *)coset_g, &zinv);
    if (rc) return rc;
    k_qap_pointwise<<<(unsigned)((count + 255) / 256), 256, 0, ctx->cur->stream>>>((Fr *)d_aA, (const Fr *)d_aB, (const Fr *)d_aC,
                                                                                zinv, count);
    ctx->launches++;
    CK(cudaGetLastError());
    CK(cudaStreamSynchronize(ctx->cur->stream));
    return 0;
}

int zecale_gpu_lagrange(zecale_gpu_ctx *ctx, unsigned log_m, const void *t, void *out)
{
    if (!ctx) return ZECALE_GPU_E_ARG;
    std::lock_guard<std::mutex> lk(ctx->mu);
    ctx->cur = &ctx->lanes[0];
    if (!t || !out) return fail(ctx, ZECALE_GPU_E_ARG, "zecale_gpu_lagrange: bad argument");
    if (log_m < 1 || log_m > 28) return fail(ctx, ZECALE_GPU_E_DOMAIN, "zecale_gpu_lagrange: log_m %u out of range", log_m);
    CK(cudaSetDevice(ctx->device));
    const size_t m = (size_t)1 << log_m;
    cudaStream_t st = ctx->cur->stream;
    unsigned nl = 0;
    const Fr *tw = nullptr;
    CK(ntt_get_twiddles(ctx->ntt, log_m, st, &tw, &nl));
    int rc = reserve(ctx, ctx->ws_in, m * sizeof(Fr) + 2 * sizeof(Fr));
    if (rc) return rc;
    Fr *d_out = (Fr *)ctx->ws_in.p, *d_seed = d_out + m;
    Fr tv;
    memcpy(tv.l, t, sizeof(Fr));
    k_lagrange_seed<<<1, 1, 0, st>>>(tv, log_m, d_seed, ctx->d_status);
    int flag = 0;
    CK(cudaMemcpyAsync(&flag, ctx->d_status, sizeof(int), cudaMemcpyDeviceToHost, st));
    const size_t threads = (m + LAGRANGE_CHUNK - 1) / LAGRANGE_CHUNK;
    k_lagrange<<<(unsigned)((threads + 127) / 128), 128, 0, st>>>(tw, log_m, d_seed, d_out);
    ctx->launches += nl + 2;
    CK(cudaGetLastError());
    CK(cudaStreamSynchronize(st));
    if (flag) return fail(ctx, ZECALE_GPU_E_DOMAIN, "zecale_gpu_lagrange: t is a point of the domain");
    CK(cudaMemcpy(out, d_out, m * sizeof(Fr), cudaMemcpyDeviceToHost));
    return 0;
}

int zecale_gpu_lane_count(void) { return ZECALE_GPU_LANES; }

int zecale_gpu_msm_pinned_async(zecale_gpu_ctx *ctx, int lane, const zecale_gpu_bases *h, const void *scalars,
                                size_t scalar_stride, const void *scalar_index, size_t n, void *out_point, unsigned flags)
{
    if (!ctx) return ZECALE_GPU_E_ARG;
    std::lock_guard<std::mutex> lk(ctx->mu);
    if (lane < 0 || lane >= ZECALE_GPU_LANES || !h || !out_point || !n || n > h->n || (flags & ~3u) ||
        (!(flags & ZECALE_GPU_SCALARS_FROM_QAP) && (!scalars || scalar_stride < sizeof(Fr))))
        return fail(ctx, ZECALE_GPU_E_ARG, "zecale_gpu_msm_pinned_async: bad argument");
    CK(cudaSetDevice(ctx->device));
    ctx->cur = &ctx->lanes[lane];
    if (is_multi(ctx)) {
        if (flags & ZECALE_GPU_DEVICE_POINTERS)
            return fail(ctx, ZECALE_GPU_E_ARG, "msm_pinned_async: device-resident scalars belong to one device; use a single-device context");
        if ((flags & ZECALE_GPU_SCALARS_FROM_QAP) && !ctx->cur->ws_qap.p)
            return fail(ctx, ZECALE_GPU_E_ARG, "msm_pinned_async: no H polynomial on this lane");
        return multi_msm_async(ctx, lane, h, scalars, scalar_stride, (const size_t *)scalar_index, n, out_point,
                               (flags & ZECALE_GPU_SCALARS_FROM_QAP) != 0);
    }
    if (!h->shards.empty())
        return fail(ctx, ZECALE_GPU_E_ARG, "msm_pinned_async: the pinned image is sharded over a multi-device context");
    const uint8_t *d_sc = (const uint8_t *)scalars;
    const uint64_t *d_idx = (const uint64_t *)scalar_index;
    if (flags & ZECALE_GPU_SCALARS_FROM_QAP) {
        if (!ctx->cur->ws_qap.p) return fail(ctx, ZECALE_GPU_E_ARG, "msm_pinned_async: no H polynomial on this lane");
        d_sc = (const uint8_t *)ctx->cur->ws_qap.p;
        scalar_stride = sizeof(Fr);
        d_idx = nullptr;
    } else if (!(flags & ZECALE_GPU_DEVICE_POINTERS)) {
        if (ctx->cur->user_out) { // the staging buffers of this lane are still in flight
            int rc0 = sync_lanes(ctx);
            if (rc0) return rc0;
        }
        const size_t *idx = (const size_t *)scalar_index;
        size_t n_sc = n;
        if (idx) {
            size_t mx = 0;
            for (size_t i = 0; i < n; i++)
                if (idx[i] > mx) mx = idx[i];
            n_sc = mx + 1;
        }
        int rc = epoch_upload(ctx, scalars, (n_sc - 1) * scalar_stride + sizeof(Fr), &d_sc);
        if (rc == 1) rc = upload_scalars(ctx, scalars, scalar_stride, nullptr, n_sc, &d_sc, &d_idx);
        if (rc) return rc;
        d_idx = nullptr;
        if (idx) {
            const uint8_t *d_i = nullptr;
            rc = epoch_upload(ctx, idx, n * 8, &d_i);
            if (rc == 1) {
                rc = reserve(ctx, ctx->cur->ws_idx, n * 8);
                if (rc == 0) {
                    CK(cudaMemcpyAsync(ctx->cur->ws_idx.p, idx, n * 8, cudaMemcpyHostToDevice, ctx->cur->stream));
                    d_i = (const uint8_t *)ctx->cur->ws_idx.p;
                }
            }
            if (rc) return rc;
            d_idx = (const uint64_t *)d_i;
        }
    }
    if (ctx->qap_in_epoch && !(flags & ZECALE_GPU_SCALARS_FROM_QAP))
        CK(cudaStreamWaitEvent(ctx->cur->stream, ctx->ev_qap, 0));
    return msm_device(ctx, h->d, h->is_inf, h->c, h->V, (uint32_t)h->n, d_sc, scalar_stride, d_idx, n, out_point, nullptr,
                      true);
}

int zecale_gpu_qap_h_async(zecale_gpu_ctx *ctx, int lane, const void *aA, const void *aB, const void *aC, unsigned log_m,
                           const void *coset_g, void *out_h)
{
    if (!ctx) return ZECALE_GPU_E_ARG;
    std::lock_guard<std::mutex> lk(ctx->mu);
    if (lane < 0 || lane >= ZECALE_GPU_LANES || !aA || !aB || !aC || !coset_g)
        return fail(ctx, ZECALE_GPU_E_ARG, "zecale_gpu_qap_h_async: bad argument");
    if (log_m < 1 || log_m > 28) return fail(ctx, ZECALE_GPU_E_DOMAIN, "zecale_gpu_qap_h: log_m %u out of range", log_m);
    CK(cudaSetDevice(ctx->device));
    ctx->cur = &ctx->lanes[lane];
    if (ctx->cur->user_out || ctx->cur->user_h) {
        int rc0 = sync_lanes(ctx);
        if (rc0) return rc0;
    }
    const size_t m = (size_t)1 << log_m, bytes = m * sizeof(Fr);
    int rc = reserve(ctx, ctx->cur->ws_qap, 3 * bytes);
    if (rc) return rc;
    Fr *d = (Fr *)ctx->cur->ws_qap.p;
    // cudaMemcpyDefault: the three vectors may be host or device memory
    CK(cudaMemcpyAsync(d, aA, bytes, cudaMemcpyDefault, ctx->cur->stream));
    CK(cudaMemcpyAsync(d + m, aB, bytes, cudaMemcpyDefault, ctx->cur->stream));
    CK(cudaMemcpyAsync(d + 2 * m, aC, bytes, cudaMemcpyDefault, ctx->cur->stream));
    trace_mark(ctx, "abc_h2d", ctx->cur->stream);
    rc = qap_h_device(ctx, d, d + m, d + 2 * m, log_m, (const Fr *)coset_g);
    if (rc) return rc;
    CK(cudaEventRecord(ctx->cur->ev_qap, ctx->cur->stream));
    CK(cudaEventRecord(ctx->ev_qap, ctx->cur->stream));
    ctx->qap_in_epoch = true;
    // coefficients_for_H go back on the lane's stream (copy engine; overlaps the MSMs)
    if (out_h) CK(cudaMemcpyAsync(out_h, d, bytes, cudaMemcpyDeviceToHost, ctx->cur->stream));
    trace_mark(ctx, "h_d2h", ctx->cur->stream);
    ctx->cur->user_h = out_h; // may be NULL: the coefficients then only feed ZECALE_GPU_SCALARS_FROM_QAP
    ctx->cur->user_h_bytes = bytes;
    return 0;
}

int zecale_gpu_r1cs_load(zecale_gpu_ctx *ctx, size_t n_constraints, size_t n_inputs, size_t n_variables,
                         const uint64_t *const row_ptr[3], const uint32_t *const col[3], const void *const val[3],
                         zecale_gpu_r1cs **handle)
{
    if (!ctx) return ZECALE_GPU_E_ARG;
    std::lock_guard<std::mutex> lk(ctx->mu);
    ctx->cur = &ctx->lanes[0];
    if (!handle || !row_ptr || !col || !val || !n_constraints || n_inputs > n_variables || n_variables >= (1ull << 31))
        return fail(ctx, ZECALE_GPU_E_ARG, "zecale_gpu_r1cs_load: bad argument");
    CK(cudaSetDevice(ctx->device));
    zecale_gpu_r1cs *h = new zecale_gpu_r1cs();
    h->n_constraints = n_constraints;
    h->n_inputs = n_inputs;
    h->n_variables = n_variables;
    // domain of r1cs_to_qap_witness_map with force_pow_2_domain (r1cs_to_qap.tcc:224-227)
    while (((size_t)1 << h->log_m) < n_constraints + n_inputs + 1) h->log_m++;
    cudaError_t e = cudaSuccess;
    for (int k = 0; k < 3 && e == cudaSuccess; k++) {
        if (!row_ptr[k] || row_ptr[k][0] != 0) {
            e = cudaErrorInvalidValue;
            break;
        }
        const size_t nnz = row_ptr[k][n_constraints];
        for (size_t t = 0; t < nnz; t++)
            if (col[k][t] > n_variables) e = cudaErrorInvalidValue; // index 0 = one, 1..n_variables
        if (e != cudaSuccess) break;
        e = cudaMalloc((void **)&h->row_ptr[k], (n_constraints + 1) * 8);
        if (e == cudaSuccess) e = cudaMalloc((void **)&h->col[k], (nnz + 1) * 4);
        if (e == cudaSuccess) e = cudaMalloc((void **)&h->val[k], (nnz + 1) * sizeof(Fr));
        if (e == cudaSuccess) e = cudaMemcpy(h->row_ptr[k], row_ptr[k], (n_constraints + 1) * 8, cudaMemcpyHostToDevice);
        if (e == cudaSuccess && nnz) e = cudaMemcpy(h->col[k], col[k], nnz * 4, cudaMemcpyHostToDevice);
        if (e == cudaSuccess && nnz) e = cudaMemcpy(h->val[k], val[k], nnz * sizeof(Fr), cudaMemcpyHostToDevice);
    }
    if (e != cudaSuccess) {
        for (int k = 0; k < 3; k++) {
            cudaFree(h->row_ptr[k]);
            cudaFree(h->col[k]);
            cudaFree(h->val[k]);
        }
        delete h;
        return fail(ctx, e == cudaErrorInvalidValue ? ZECALE_GPU_E_ARG : ZECALE_GPU_E_NOMEM, "zecale_gpu_r1cs_load: %s",
                    cudaGetErrorString(e));
    }
    *handle = h;
    return 0;
}

int zecale_gpu_r1cs_free(zecale_gpu_ctx *ctx, zecale_gpu_r1cs *h)
{
    if (!ctx || !h) return ZECALE_GPU_E_ARG;
    std::lock_guard<std::mutex> lk(ctx->mu);
    cudaSetDevice(ctx->device);
    sync_lanes(ctx);
    for (int k = 0; k < 3; k++) {
        cudaFree(h->row_ptr[k]);
        cudaFree(h->col[k]);
        cudaFree(h->val[k]);
    }
    delete h;
    return 0;
}

unsigned zecale_gpu_r1cs_log_domain(const zecale_gpu_r1cs *h) { return h ? h->log_m : 0; }

int zecale_gpu_qap_h_from_witness_async(zecale_gpu_ctx *ctx, int lane, const zecale_gpu_r1cs *r1cs, const void *assignment,
                                        const void *coset_g, void *out_h, unsigned flags)
{
    if (!ctx) return ZECALE_GPU_E_ARG;
    std::lock_guard<std::mutex> lk(ctx->mu);
    if (lane < 0 || lane >= ZECALE_GPU_LANES || !r1cs || !assignment || !coset_g || (flags & ~ZECALE_GPU_DEVICE_POINTERS))
        return fail(ctx, ZECALE_GPU_E_ARG, "zecale_gpu_qap_h_from_witness_async: bad argument");
    CK(cudaSetDevice(ctx->device));
    ctx->cur = &ctx->lanes[lane];
    if (ctx->cur->user_out || ctx->cur->user_h) {
        int rc0 = sync_lanes(ctx);
        if (rc0) return rc0;
    }
    const unsigned log_m = r1cs->log_m;
    const size_t m = (size_t)1 << log_m, bytes = m * sizeof(Fr);
    int rc = reserve(ctx, ctx->cur->ws_qap, 3 * bytes);
    if (rc) return rc;
    Fr *d = (Fr *)ctx->cur->ws_qap.p;
    // the assignment [1, primary, aux] crosses PCIe once per proof: the MSM calls that pass the
    // same host range (A, B, L) find it in the epoch's upload table
    const uint8_t *d_w = (const uint8_t *)assignment;
    if (is_multi(ctx) && (ctx->gather[lane].pending || ctx->cur->user_out)) {
        int rc0 = multi_sync(ctx);
        if (rc0) return rc0;
    }
    if (!(flags & ZECALE_GPU_DEVICE_POINTERS)) {
        const size_t wbytes = (r1cs->n_variables + 1) * sizeof(Fr);
        if (is_multi(ctx)) { // every worker gets the assignment once; the shard MSMs of A, B and L find it there
            rc = multi_broadcast_host(ctx, lane, assignment, wbytes);
            if (rc) return rc;
            CK(cudaSetDevice(ctx->device));
        }
        rc = epoch_upload(ctx, assignment, wbytes, &d_w);
        if (rc == 1) { // upload table full: lane-private copy
            rc = reserve(ctx, ctx->cur->ws_sc, wbytes);
            if (rc) return rc;
            CK(cudaMemcpyAsync(ctx->cur->ws_sc.p, assignment, wbytes, cudaMemcpyHostToDevice, ctx->cur->stream));
            d_w = (const uint8_t *)ctx->cur->ws_sc.p;
        } else if (rc) {
            return rc;
        }
    }
    trace_mark(ctx, "w_h2d", ctx->cur->stream);
    const unsigned grid = (unsigned)((m + 127) / 128);
    for (int k = 0; k < 3; k++) {
        k_r1cs_eval<<<grid, 128, 0, ctx->cur->stream>>>(r1cs->row_ptr[k], r1cs->col[k], r1cs->val[k], (const Fr *)d_w,
                                                         (uint32_t)r1cs->n_constraints, k == 0 ? (uint32_t)r1cs->n_inputs + 1 : 0u,
                                                         d + (size_t)k * m, m);
        ctx->launches++;
    }
    CK(cudaGetLastError());
    rc = qap_h_device(ctx, d, d + m, d + 2 * m, log_m, (const Fr *)coset_g);
    if (rc) return rc;
    CK(cudaEventRecord(ctx->cur->ev_qap, ctx->cur->stream));
    CK(cudaEventRecord(ctx->ev_qap, ctx->cur->stream));
    ctx->qap_in_epoch = true;
    if (out_h) CK(cudaMemcpyAsync(out_h, d, bytes, cudaMemcpyDeviceToHost, ctx->cur->stream));
    ctx->cur->user_h = out_h;
    ctx->cur->user_h_bytes = bytes;
    return 0;
}

int zecale_gpu_sync(zecale_gpu_ctx *ctx)
{
    if (!ctx) return ZECALE_GPU_E_ARG;
    std::lock_guard<std::mutex> lk(ctx->mu);
    CK(cudaSetDevice(ctx->device));
    return is_multi(ctx) ? multi_sync(ctx) : sync_lanes(ctx);
}

int zecale_gpu_flush(zecale_gpu_ctx *ctx)
{
    if (!ctx) return ZECALE_GPU_E_ARG;
    std::lock_guard<std::mutex> lk(ctx->mu);
    CK(cudaSetDevice(ctx->device));
    return is_multi(ctx) ? multi_flush(ctx) : flush_tails(ctx);
}

void *zecale_gpu_host_alloc(size_t bytes)
{
    void *p = nullptr;
    if (!bytes || cudaMallocHost(&p, bytes) != cudaSuccess) return nullptr;
    return p;
}

void zecale_gpu_host_free(void *p)
{
    if (p) cudaFreeHost(p);
}

int zecale_gpu_synth_bases(zecale_gpu_ctx *ctx, int group, uint64_t seed, size_t n, void *out_points)
{
    if (!ctx) return ZECALE_GPU_E_ARG;
    std::lock_guard<std::mutex> lk(ctx->mu);
    ctx->cur = &ctx->lanes[0];
    if (!out_points || !n || n >= (1ull << 31) || (group != ZECALE_GPU_G1 && group != ZECALE_GPU_G2))
        return fail(ctx, ZECALE_GPU_E_ARG, "zecale_gpu_synth_bases: bad argument");
    CK(cudaSetDevice(ctx->device));
    int rc = reserve(ctx, ctx->ws_tmp, n * sizeof(Proj));
    if (rc) return rc;
    k_synth_bases<<<(unsigned)((n + 127) / 128), 128, 0, ctx->cur->stream>>>(group, seed, (uint32_t)n, (Proj *)ctx->ws_tmp.p);
    ctx->launches++;
    CK(cudaGetLastError());
    CK(cudaMemcpyAsync(out_points, ctx->ws_tmp.p, n * sizeof(Proj), cudaMemcpyDeviceToHost, ctx->cur->stream));
    CK(cudaStreamSynchronize(ctx->cur->stream));
    return 0;
}

int zecale_gpu_batch_exp(zecale_gpu_ctx *ctx, int group, const void *base_point, const void *scalars, size_t scalar_stride,
                         size_t n, const void *coeff, void *out_points)
{
    if (!ctx) return ZECALE_GPU_E_ARG;
    std::lock_guard<std::mutex> lk(ctx->mu);
    ctx->cur = &ctx->lanes[0];
    if (!base_point || !out_points || (n && !scalars) || n >= (1ull << 31) || scalar_stride < sizeof(Fr) || (scalar_stride & 7) ||
        (group != ZECALE_GPU_G1 && group != ZECALE_GPU_G2))
        return fail(ctx, ZECALE_GPU_E_ARG, "zecale_gpu_batch_exp: bad argument");
    if (n == 0) return 0;
    CK(cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->cur->stream;
    // workspace: [base affine + flag | powers XYZZ/affine | table XYZZ/affine | scalars | coeff | out XYZZ | prefix | out Proj]
    const size_t n_tab = (size_t)BEXP_W * BEXP_ROW, n_max = n > n_tab ? n : n_tab;
    size_t off = 0;
    auto take = [&](size_t bytes) {
        size_t o = off;
        off += align_up(bytes, 256);
        return o;
    };
    const size_t o_base = take(sizeof(Affine)), o_inf = take(256), o_pw_x = take(BEXP_W * sizeof(XYZZ)),
                 o_pw_a = take(BEXP_W * sizeof(Affine)), o_tab_a = take(n_tab * sizeof(Affine)),
                 o_sc = take((n - 1) * scalar_stride + sizeof(Fr)), o_cf = take(sizeof(Fr)), o_x = take(n_max * sizeof(XYZZ)),
                 o_pre = take(n_max * sizeof(Fq)), o_out = take(n * sizeof(Proj));
    int rc = reserve(ctx, ctx->ws_tmp, off);
    if (rc) return rc;
    uint8_t *ws = (uint8_t *)ctx->ws_tmp.p;
    Affine *d_base = (Affine *)(ws + o_base), *d_pw_a = (Affine *)(ws + o_pw_a), *d_tab = (Affine *)(ws + o_tab_a);
    XYZZ *d_pw_x = (XYZZ *)(ws + o_pw_x), *d_x = (XYZZ *)(ws + o_x);
    Fq *d_pre = (Fq *)(ws + o_pre);
    rc = import_bases(ctx, base_point, sizeof(Proj), 1, d_base, ws + o_inf);
    if (rc) return rc;
    uint8_t base_inf = 0;
    CK(cudaMemcpyAsync(&base_inf, ws + o_inf, 1, cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(ws + o_sc, scalars, (n - 1) * scalar_stride + sizeof(Fr), cudaMemcpyHostToDevice, st));
    if (coeff) CK(cudaMemcpyAsync(ws + o_cf, coeff, sizeof(Fr), cudaMemcpyHostToDevice, st));
    CK(cudaStreamSynchronize(st));
    if (base_inf) { // zero base: every power is zero
        CK(cudaMemsetAsync(d_x, 0, n * sizeof(XYZZ), st)); // zz == 0 marks infinity
    } else {
        k_bexp_powers<<<1, 32, 0, st>>>(d_base, d_pw_x);
        k_xyzz_to_affine<<<(BEXP_W + TOAFF_K * 128 - 1) / (TOAFF_K * 128), 128, 0, st>>>(d_pw_x, BEXP_W, d_pw_a, nullptr, d_pre);
        k_bexp_table<<<(unsigned)((n_tab + 127) / 128), 128, 0, st>>>(d_pw_a, d_x);
        k_xyzz_to_affine<<<(unsigned)((n_tab + TOAFF_K * 128 - 1) / (TOAFF_K * 128)), 128, 0, st>>>(d_x, (uint32_t)n_tab, d_tab,
                                                                                                       nullptr, d_pre);
        k_bexp_eval<<<(unsigned)((n + 127) / 128), 128, 0, st>>>(d_tab, ws + o_sc, scalar_stride, (uint32_t)n,
                                                                   coeff ? (const Fr *)(ws + o_cf) : nullptr, d_x);
        ctx->launches += 5;
    }
    k_xyzz_to_affine<<<(unsigned)((n + TOAFF_K * 128 - 1) / (TOAFF_K * 128)), 128, 0, st>>>(d_x, (uint32_t)n, nullptr,
                                                                                          (Proj *)(ws + o_out), d_pre);
    ctx->launches++;
    CK(cudaGetLastError());
    CK(cudaMemcpyAsync(out_points, ws + o_out, n * sizeof(Proj), cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    return 0;
}

int zecale_gpu_kernel_ms(zecale_gpu_ctx *ctx, int kernel_class, int reset, double *total_ms, uint64_t *count)
{
    if (!ctx) return ZECALE_GPU_E_ARG;
    std::lock_guard<std::mutex> lk(ctx->mu);
    ctx->cur = &ctx->lanes[0];
    if (kernel_class < 0 || kernel_class > 1) return fail(ctx, ZECALE_GPU_E_ARG, "zecale_gpu_kernel_ms: bad class");
    CK(cudaSetDevice(ctx->device));
    if (is_multi(ctx)) { // deliver what is in flight the regular way before the workers' lanes are drained below
        int rc = multi_sync(ctx);
        if (rc) return rc;
    }
    drain_timers(ctx);
    double ms = ctx->kernel_ms[kernel_class];
    uint64_t cnt = ctx->kernel_n[kernel_class];
    if (reset) {
        ctx->kernel_ms[kernel_class] = 0;
        ctx->kernel_n[kernel_class] = 0;
    }
    // multi-device context: the shard MSMs (and their accumulate kernels) are timed by the workers; report them
    // too, as zecale_gpu_launch_count does. The sum is device time added over devices, not elapsed time.
    for (zecale_gpu_ctx *w : ctx->workers) {
        DevGuard dg(w->device);
        drain_timers(w);
        ms += w->kernel_ms[kernel_class];
        cnt += w->kernel_n[kernel_class];
        if (reset) {
            w->kernel_ms[kernel_class] = 0;
            w->kernel_n[kernel_class] = 0;
        }
    }
    if (total_ms) *total_ms = ms;
    if (count) *count = cnt;
    return 0;
}

int zecale_gpu_msm_last_stats(zecale_gpu_ctx *ctx, uint64_t *entries, uint64_t *tasks, unsigned *c, unsigned *bucket_sets,
                              unsigned *base_copies)
{
    if (!ctx) return ZECALE_GPU_E_ARG;
    std::lock_guard<std::mutex> lk(ctx->mu);
    ctx->cur = &ctx->lanes[0];
    if (entries) *entries = ctx->cur->last_stats[0];
    if (tasks) *tasks = ctx->cur->last_stats[1];
    if (c) *c = ctx->cur->last_plan.c;
    if (bucket_sets) *bucket_sets = ctx->cur->last_plan.V;
    if (base_copies) *base_copies = ctx->cur->last_plan.G;
    return 0;
}

uint64_t zecale_gpu_launch_count(zecale_gpu_ctx *ctx)
{
    if (!ctx) return 0;
    uint64_t n = ctx->launches;
    for (zecale_gpu_ctx *w : ctx->workers) n += w->launches;
    return n;
}

int zecale_gpu_device_count(zecale_gpu_ctx *ctx) { return ctx ? (int)ctx->devices.size() : 0; }

unsigned zecale_gpu_msm_window_bits(size_t n) { return n ? msm_default_c((uint32_t)n) : 0; }
unsigned zecale_gpu_msm_window_bits_pinned(size_t n) { return n ? pinned_c(n) : 0; }

int zecale_gpu_test_fq(zecale_gpu_ctx *ctx, int op, const void *a, const void *b, void *out, size_t n)
{
    return test_elementwise<Fq>(ctx, launch_fq, op, a, b, out, n);
}
int zecale_gpu_test_fr(zecale_gpu_ctx *ctx, int op, const void *a, const void *b, void *out, size_t n)
{
    return test_elementwise<Fr>(ctx, launch_fr, op, a, b, out, n);
}
int zecale_gpu_test_group(zecale_gpu_ctx *ctx, int op, const void *a, const void *b, void *out, size_t n)
{
    return test_elementwise<Proj>(ctx, launch_group, op, a, b, out, n);
}

int zecale_gpu_bench_fq_mul(zecale_gpu_ctx *ctx, unsigned blocks, unsigned tpb, unsigned iters, float *elapsed_ms)
{
    if (!ctx) return ZECALE_GPU_E_ARG;
    std::lock_guard<std::mutex> lk(ctx->mu);
    ctx->cur = &ctx->lanes[0];
    if (!blocks || !tpb || tpb > 128 || !elapsed_ms || !iters) return fail(ctx, ZECALE_GPU_E_ARG, "bench_fq_mul: bad argument");
    CK(cudaSetDevice(ctx->device));
    int rc = reserve(ctx, ctx->ws_tmp, sizeof(Fq));
    if (rc) return rc;
    k_bench_fq_mul<<<blocks, tpb, 0, ctx->cur->stream>>>((Fq *)ctx->ws_tmp.p, 8); // warm-up
    CK(cudaEventRecord(ctx->cur->ev0, ctx->cur->stream));
    k_bench_fq_mul<<<blocks, tpb, 0, ctx->cur->stream>>>((Fq *)ctx->ws_tmp.p, iters);
    CK(cudaEventRecord(ctx->cur->ev1, ctx->cur->stream));
    ctx->launches += 2;
    CK(cudaStreamSynchronize(ctx->cur->stream));
    CK(cudaEventElapsedTime(elapsed_ms, ctx->cur->ev0, ctx->cur->ev1));
    return 0;
}

int zecale_gpu_bench_imad(zecale_gpu_ctx *ctx, unsigned blocks, unsigned tpb, unsigned iters, float *elapsed_ms)
{
    if (!ctx) return ZECALE_GPU_E_ARG;
    std::lock_guard<std::mutex> lk(ctx->mu);
    ctx->cur = &ctx->lanes[0];
    if (!blocks || !tpb || tpb > 256 || !elapsed_ms || !iters) return fail(ctx, ZECALE_GPU_E_ARG, "bench_imad: bad argument");
    CK(cudaSetDevice(ctx->device));
    int rc = reserve(ctx, ctx->ws_tmp, sizeof(Fq));
    if (rc) return rc;
    k_bench_imad<<<blocks, tpb, 0, ctx->cur->stream>>>((uint64_t *)ctx->ws_tmp.p, 0x9e3779b9u, 0x85ebca6bu, 8); // warm-up
    CK(cudaEventRecord(ctx->cur->ev0, ctx->cur->stream));
    k_bench_imad<<<blocks, tpb, 0, ctx->cur->stream>>>((uint64_t *)ctx->ws_tmp.p, 0x9e3779b9u, 0x85ebca6bu, iters);
    CK(cudaEventRecord(ctx->cur->ev1, ctx->cur->stream));
    ctx->launches += 2;
    CK(cudaStreamSynchronize(ctx->cur->stream));
    CK(cudaEventElapsedTime(elapsed_ms, ctx->cur->ev0, ctx->cur->ev1));
    return 0;
}

} // extern "C"

// =====================================================================================================
// Allocation-free NTT: zecale_gpu_ntt_static. Nothing above this line is used differently because of it.
//
// Some programs that want this transform are not allowed to obtain device memory while they run:
//   * a prover embedded in a service whose own pool reserved (almost) all of the HBM at start-up, so a
//     cudaMalloc issued by anybody else comes back "out of memory";
//   * a request path with a latency budget, where cudaMalloc / cudaFree are banned because they
//     synchronise the whole device and stall every other stream of the process.
// Such a caller cannot even create a zecale_gpu_ctx (zecale_gpu_init allocates), and the context's NTT
// needs three pieces of global memory that are neither input nor output: the twiddle table, the coset /
// 1/m scale table and a second vector the passes ping-pong through. Here those three live in one
// file-scope device array that is part of the library image: the driver places it when the module is
// loaded on a device and it stays there until the process exits (ZG_STATIC_BYTES on EVERY device that
// loads the library, used or not), which is why the largest size is kept modest.
//
// The array is shared by every call on a device and is never cleared: a call first rewrites, on the
// caller's stream, every entry it is going to read (root, seeds, both tables) and only then runs the
// passes, so what an earlier call left behind is never consumed and nothing has to start at zero.
// The price is that the tables are recomputed per call. Two calls that are in flight on one device at
// the same time (different streams without an ordering between them, or two host threads) overwrite
// each other's tables: results are then undefined. Serialise them.
// =====================================================================================================
#define ZG_STATIC_MAX_LOG_M ZECALE_GPU_NTT_STATIC_MAX_LOG_M
#define ZG_STATIC_M ((size_t)1 << ZG_STATIC_MAX_LOG_M)
// [ second vector: M Fr | twiddles w^k: M/2 Fr | scale table: M Fr | root, g^(+-1), 1 or 1/m, spare: 4 Fr ]
#define ZG_STATIC_BYTES ((ZG_STATIC_M * 5 / 2 + 4) * sizeof(Fr))
static_assert(ZG_STATIC_BYTES == ZECALE_GPU_NTT_STATIC_BYTES, "header and library disagree on the scratch size");

static __device__ __align__(256) unsigned char zg_static_scratch[ZG_STATIC_BYTES];

// table[i] = scale * base^i for i < n, with base and scale read from device memory (scale == nullptr: one),
// so the seeds computed by a preceding kernel on the same stream never travel through the host.
__global__ void k_fr_powers_from(const Fr *base, const Fr *scale, Fr *table, size_t n, uint32_t chunk)
{
    size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    size_t i0 = t * chunk;
    if (i0 >= n) return;
    uint32_t b[12], cur[12], s[12];
    fr_load(b, base);
    if (scale)
        fr_load(s, scale);
    else
        fr_set_const(s, FR_ONE);
    fr_pow_u64(cur, b, (uint64_t)i0);
    fr_mul_regs(cur, cur, s);
    for (uint32_t k = 0; k < chunk && i0 + k < n; k++) {
        fr_store(table + i0 + k, cur);
        fr_mul_regs(cur, cur, b);
    }
}

namespace {

std::mutex zg_static_mu;
constexpr int ZG_STATIC_MAX_DEVICES = 64;
// Device address of zg_static_scratch on each device, filled on the first call made with that device current.
// In host code the name `zg_static_scratch` is only a handle for the runtime, not a pointer a kernel could use:
// the real address is a per-device fact and comes from cudaGetSymbolAddress.
unsigned char *zg_static_addr[ZG_STATIC_MAX_DEVICES] = {};

int static_fail(char *err, size_t cap, int code, const char *fmt, ...)
{
    if (err && cap) {
        va_list ap;
        va_start(ap, fmt);
        vsnprintf(err, cap, fmt, ap);
        va_end(ap);
    }
    return code;
}

#define CKS(call)                                                                                              \
    do {                                                                                                       \
        cudaError_t e_ = (call);                                                                               \
        if (e_ != cudaSuccess)                                                                                 \
            return static_fail(err, err_cap, ZECALE_GPU_E_CUDA, "zecale_gpu_ntt_static: %s:%d %s: %s", __FILE__, \
                               __LINE__, #call, cudaGetErrorString(e_));                                       \
    } while (0)

} // namespace

extern "C" int zecale_gpu_ntt_static(void *stream, void *d_a, unsigned log_m, unsigned flags, const void *coset_g, char *err,
                                     size_t err_cap)
{
    if (err && err_cap) err[0] = 0;
    if (!d_a || (flags & ~3u) || ((flags & ZECALE_GPU_NTT_COSET) && !coset_g))
        return static_fail(err, err_cap, ZECALE_GPU_E_ARG, "zecale_gpu_ntt_static: bad argument (null vector, unknown flag bits, "
                                                           "or COSET without coset_g)");
    if (log_m < 1 || log_m > ZG_STATIC_MAX_LOG_M)
        return static_fail(err, err_cap, ZECALE_GPU_E_DOMAIN,
                           "zecale_gpu_ntt_static: log_m = %u is outside [1, %u]; the built-in scratch array holds tables for at most "
                           "2^%u elements (use zecale_gpu_ntt_dev on a context for larger transforms)",
                           log_m, ZG_STATIC_MAX_LOG_M, ZG_STATIC_MAX_LOG_M);
    int dev = -1;
    CKS(cudaGetDevice(&dev));
    if (dev < 0 || dev >= ZG_STATIC_MAX_DEVICES)
        return static_fail(err, err_cap, ZECALE_GPU_E_ARG, "zecale_gpu_ntt_static: device ordinal %d not supported (max %d)", dev,
                           ZG_STATIC_MAX_DEVICES - 1);
    unsigned char *base = nullptr;
    {
        // set-up, once per device, with that device current
        std::lock_guard<std::mutex> lk(zg_static_mu);
        if (!zg_static_addr[dev]) {
            void *p = nullptr;
            CKS(cudaGetSymbolAddress(&p, zg_static_scratch));
            if (!p) return static_fail(err, err_cap, ZECALE_GPU_E_CUDA, "zecale_gpu_ntt_static: cudaGetSymbolAddress returned null");
            zg_static_addr[dev] = (unsigned char *)p;
        }
        base = zg_static_addr[dev];
    }
    cudaStream_t st = (cudaStream_t)stream;
    const uint32_t inverse = flags & 1u, coset = (flags >> 1) & 1u;
    const size_t m = (size_t)1 << log_m;
    Fr *a = (Fr *)d_a;
    Fr *tmp = (Fr *)base;
    Fr *tw = tmp + ZG_STATIC_M;
    Fr *ctab = tw + ZG_STATIC_M / 2;
    Fr *seed = ctab + ZG_STATIC_M;

    // 1. seeds and tables, rebuilt by every call (see the header of this section)
    const uint32_t chunk = 64;
    k_fr_root<<<1, 1, 0, st>>>(seed, log_m, 0);
    {
        size_t half = m >> 1, nthreads = (half + chunk - 1) / chunk;
        k_fr_powers_from<<<(unsigned)((nthreads + 127) / 128), 128, 0, st>>>(seed, nullptr, tw, half, chunk);
    }
    if (coset || inverse) {
        Fr g; // a plain iFFT still needs the 1/m scaling: the table for g = 1
        memcpy(g.l, coset ? coset_g : (const void *)FR_ONE_HOST, sizeof(Fr));
        k_fr_coset_seed<<<1, 1, 0, st>>>(seed + 1, g, log_m, inverse); // seed[1] = g^(+-1), seed[2] = 1 or 1/m
        size_t nthreads = (m + chunk - 1) / chunk;
        k_fr_powers_from<<<(unsigned)((nthreads + 127) / 128), 128, 0, st>>>(seed + 1, seed + 2, ctab, m, chunk);
    }
    CKS(cudaGetLastError());

    // 2. the passes: the factorisations ntt_launch (ntt.cuh) uses for these sizes, on the static tables
    const Fr *in_tab = (coset && !inverse) ? ctab : nullptr;
    const Fr *out_tab = inverse ? ctab : nullptr;
    NttPass p{};
    p.log_m = log_m;
    p.inverse = inverse;
    p.batch_stride = m;
    if (log_m <= 11) {
        p.logR = log_m;
        p.T = 1;
        p.n_batch = 1;
        p.n_inner = 1;
        p.isr = 1;
        p.osr = 1;
        p.sub_tw_stride = 1;
        p.in_scale_tab = in_tab; p.isc_sr = 1;
        p.scale_tab = out_tab; p.sc_sr = 1;
        CKS(ntt_run_pass(a, tmp, tw, p, st));
        CKS(cudaMemcpyAsync(a, tmp, m * sizeof(Fr), cudaMemcpyDeviceToDevice, st));
    } else {
        const uint32_t l1 = (log_m + 1) / 2, l2 = log_m - l1;
        const uint64_t m1 = 1ull << l1, m2 = 1ull << l2;
        // columns: vector u = j2 < m2, element r = j1, index r*m2 + u; output times w_m^(u*k)
        p.logR = l1;
        p.n_batch = 1;
        p.n_inner = (uint32_t)m2;
        p.T = ntt_pick_T(l1, p.n_inner);
        p.isv = 1; p.isr = m2;
        p.osv = 1; p.osr = m2;
        p.in_fast_v = 1;
        p.tw_post = 1;
        p.tw_mul = 1;
        p.sub_tw_stride = m2;
        p.in_scale_tab = in_tab; p.isc_sv = 1; p.isc_sr = m2;
        CKS(ntt_run_pass(a, tmp, tw, p, st));
        // rows: vector beta = k1 < m1, element r = j2: in beta*m2 + r, out beta + k*m1
        p.in_scale_tab = nullptr;
        p.logR = l2;
        p.n_batch = (uint32_t)m1;
        p.n_inner = 1;
        p.T = ntt_pick_T(l2, p.n_batch);
        p.tile_over_batch = 1;
        p.ibs = m2; p.isv = 0; p.isr = 1;
        p.obs = 1; p.osv = 0; p.osr = m1;
        p.in_fast_v = 0;
        p.tw_post = 0;
        p.sub_tw_stride = m1;
        p.scale_tab = out_tab; p.sc_bs = 1; p.sc_sr = m1;
        CKS(ntt_run_pass(tmp, a, tw, p, st));
    }
    return 0;
}
