// sequential.cu — scheduling cycles with state: Filter + Score + selectHost for ONE pod against the current node
// table, then the reference's `assume` (NodeInfo.AddPod on the chosen node) and FlexGPU.Reserve (device index),
// then the next pod. One persistent, cooperative kernel runs every cycle of a queue; SURVEY.md §8(f1).
//
//   scheduleOne            vendor/k8s.io/kubernetes/pkg/scheduler/scheduler.go:441-540 (Schedule → assume → Reserve;
//                          a failed Reserve forgets the pod :512-521)
//   assume / AddPod        scheduler.go:353-370; framework/types.go:583-609,715-745
//   FlexGPU.Reserve        pkg/flexgpu/flex_gpu.go:178-223; GPUAssumeFitIndexes / MemAssumeFitIndexes
//                          pkg/flexgpu/gpu_node.go:122-177 (first fitting device in index order)
//   per-(pod,node) math    the same as kernels.cu (flex_gpu.go:41-166, fit.go:256-329, least/most_allocated.go,
//                          allocatable.go:146-171, normalize_score.go:26-54, generic_scheduler.go:146-167)
//
// Mapping: the grid is G <= #SM CTAs of 256 threads; every thread OWNS the same nodes for the whole kernel
// (node = cta·256·K + j·256 + tid), so a node row is only ever read and rewritten by one thread: its derived
// columns live in that CTA's shared memory (global memory when the cluster is too large for that) and no
// cross-thread visibility of node state is needed. What the CTAs exchange per cycle is two small reductions:
// (A) feasible count, the maxima NormalizeScore needs and the filter-panic flag, (B) the selectHost key and the
// score-time error bits. Each CTA posts its partial result as self-tagged 64-bit words (payload 32 bits, cycle
// tag 32 bits) into a global mailbox and one warp per CTA polls all mailboxes — no fences, no atomics, no grid
// barrier: a word is valid exactly when its tag equals the current cycle. Pods stream through shared memory in
// tiles of 64 by TMA bulk copy (cp.async.bulk + mbarrier), two stages.
//
// Latency-bound by design (one pod × N nodes per cycle is a few hundred KB of L2/shared-memory traffic); the
// figure of merit is cycles per second, not HBM bandwidth.
#include <cuda_runtime.h>

#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <cstdio>
#include <string>
#include <vector>

#include "derive.cuh"
#include "engine.h"
#include "kernels.cuh"
#include "pack.h"

namespace fgs {

typedef long long i64;
typedef unsigned long long u64;

#define FGS_CUDA(e, call)                                                    \
    do {                                                                     \
        cudaError_t _err = (call);                                           \
        if (_err != cudaSuccess) {                                           \
            (e)->err = std::string(#call) + ": " + cudaGetErrorString(_err); \
            return FGS_E_CUDA;                                               \
        }                                                                    \
    } while (0)

namespace {

constexpr i64 kMin64 = (i64)0x8000000000000000ull;
constexpr int kSeqThreads = 256;              // node-owning threads per CTA
constexpr int kSeqWarps = kSeqThreads / 32;  // node-owning warps
constexpr int kSeqBlock = kSeqThreads + 32;   // + the communication warp (warp 0)
constexpr int kSeqTileP = 64;
constexpr int kMaxK = 32;        // nodes per thread (bit mask of feasible nodes)
constexpr int kWordsA = 8;       // mailbox words per CTA, phase A (7 used)
constexpr int kWordsB = 4;       // phase B (3 used)
constexpr int kMaxCluster = 16;  // CTAs of the single-cluster variant

enum SeqExtra : int { SX_ACCT_CPU = 0, SX_ACCT_MEM, SX_ANNOT, SX_NS, SX_COLS };
enum SeqOut : int { SO_COUNT = 0, SO_NODE, SO_SCORE, SO_ERR, SO_DEVICE, SO_COMMITTED, SO_QUOTA, SO_COLS };

struct SeqParams {
    DeriveParams dv;       // raw + derived node tables, plugin args of the derivation
    const i64* pods;       // [pd_cols(S)][p_pad]
    const i64* extra;      // [SX_COLS][p_pad]
    i64* out;              // [SO_COLS][p_pad]
    u64* mail_a;           // [2][G][kWordsA] self-tagged words
    u64* mail_b;           // [2][G][kWordsB]
    u64* prof;             // [G][16] clock sums per section (FGS_SEQ_PROFILE=1), else NULL
    // ElasticQuota tables (CapacityScheduling in the profile), device memory; q_has_eq == NULL: not in the profile
    const uint8_t* q_has_eq; // [NS]
    const i64* q_min;        // [NS][2+S]
    const i64* q_max;
    i64* q_used;             // rewritten by CapacityScheduling.Reserve
    uint32_t* q_used_present;
    i64 n_ns;
    int* status;           // [0]: index of the first pod whose commit left the supported range, INT_MAX none
    i64 n_pods, p_pad;
    int K, state_smem;
    uint32_t filter_mask, score_mask;
    int flex_filter_first;
    i64 w_flex, w_fit, w_alloc;
    int fit_strategy, fit_n_res;
    int fit_res[FGS_MAX_SCORE_RES];
    i64 fit_w[FGS_MAX_SCORE_RES];
    uint32_t fit_ignored;
    u64 seed;
    int key_shift, reserve_flex;
};

// ---- small PTX wrappers -------------------------------------------------------------------------
__device__ __forceinline__ uint32_t s_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(s_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(s_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    uint32_t ok = 0;
    while (!ok) {
        asm volatile(
            "{\n"
            ".reg .pred p;\n"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
            "selp.u32 %0, 1, 0, p;\n"
            "}\n"
            : "=r"(ok)
            : "r"(s_u32(bar)), "r"(parity)
            : "memory");
    }
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(s_u32(dst)),
                 "l"(__cvta_generic_to_global(src)), "r"(bytes), "r"(s_u32(bar))
                 : "memory");
}
// mailbox words: relaxed, gpu scope, never cached in L1
__device__ __forceinline__ void st_word(u64* p, uint32_t payload, uint32_t tag) {
    const u64 w = ((u64)tag << 32) | payload;
    asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(w) : "memory");
}
__device__ __forceinline__ u64 ld_word(const u64* p) {
    u64 w;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(w) : "l"(p) : "memory");
    return w;
}

// the same words through distributed shared memory (one thread-block cluster): remote store into a peer CTA's
// mailbox, local polling
__device__ __forceinline__ uint32_t mapa_u32(uint32_t saddr, uint32_t rank) {
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(saddr), "r"(rank));
    return r;
}
__device__ __forceinline__ void st_word_peer(uint32_t raddr, uint32_t payload, uint32_t tag) {
    const u64 w = ((u64)tag << 32) | payload;
    asm volatile("st.relaxed.cluster.shared::cluster.u64 [%0], %1;" ::"r"(raddr), "l"(w) : "memory");
}
__device__ __forceinline__ u64 ld_word_local(const u64* p) {
    u64 w;
    asm volatile("ld.relaxed.cluster.shared::cta.u64 %0, [%1];" : "=l"(w) : "r"(s_u32(p)) : "memory");
    return w;
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// a real consumer of `v` that cannot be removed: the clock read after it cannot issue before v has arrived
__device__ __forceinline__ void consume(i64 v) {
    asm volatile("{\n.reg .pred q;\nsetp.eq.s64 q, %0, 0x7ff0123456789abc;\n@q trap;\n}\n" ::"l"(v));
}

__device__ __forceinline__ i64 wmul(i64 a, i64 b) { return (i64)((u64)a * (u64)b); }
__device__ __forceinline__ i64 wadd(i64 a, i64 b) { return (i64)((u64)a + (u64)b); }
__device__ __forceinline__ i64 wsub(i64 a, i64 b) { return (i64)((u64)a - (u64)b); }

__device__ __forceinline__ i64 warp_max_i64(i64 v) {
    const int hi = (int)(v >> 32);
    const uint32_t lo = (uint32_t)v;
    const int mhi = __reduce_max_sync(0xffffffffu, hi);
    const uint32_t mlo = __reduce_max_sync(0xffffffffu, hi == mhi ? lo : 0u);
    return (i64)(((u64)(uint32_t)mhi << 32) | mlo);
}

__device__ __forceinline__ float u64_to_float(u64 a) {
    return fmaf(__uint2float_rn((uint32_t)(a >> 32)), 4294967296.0f, __uint2float_rn((uint32_t)a));
}

// Go's int64 x / d (truncation toward zero, wrapping), d != 0. The common case — both positive, quotient below
// 2^18 — avoids the 64-bit division routine: a float estimate is within 1/8 of the true quotient there, so its
// rounding k is the quotient or one above, and the sign of the exact remainder x − k·d decides.
__device__ __forceinline__ i64 go_div(i64 x, i64 d) {
    if (x >= 0 && d > 0 && x < ((i64)1 << 62) && d < ((i64)1 << 62)) {
        const float est = __fdividef(u64_to_float((u64)x), u64_to_float((u64)d));
        if (est < 262144.0f) {
            const int k = __float2int_rn(est);
            const i64 r = x - (i64)((u64)(uint32_t)k * (u64)d);
            return (i64)k + (r >> 63);
        }
    }
    if (d == -1) return (i64)(0ull - (u64)x);
    return x / d;
}

} // namespace

// =================================================================================================
// seq_kernel
// =================================================================================================
namespace {

struct SeqSmem {
    int bars, ring, part_a, fin_a, part_b, fin_b, mail_a, mail_b, fit, state, total;
    int npc; // pod columns per tile
};

__host__ __device__ inline SeqSmem seq_smem(int S, int D, int K, int state_smem) {
    SeqSmem L;
    L.npc = pd_cols(S) + SX_COLS;
    int off = 0;
    L.bars = off;
    off += 64;
    L.ring = off;
    off += 2 * L.npc * kSeqTileP * 8;
    L.part_a = off;
    off += kSeqWarps * 4 * 8;
    L.fin_a = off;
    off += 4 * 8;
    L.part_b = off;
    off += kSeqWarps * 2 * 8;
    L.fin_b = off;
    off += 2 * 8;
    L.mail_a = off; // cluster variant: u64 [2][kMaxCluster][kWordsA]
    off += 2 * kMaxCluster * kWordsA * 8;
    L.mail_b = off;
    off += 2 * kMaxCluster * kWordsB * 8;
    L.fit = off;
    off += K * kSeqThreads * 8;
    off = (off + 127) & ~127;
    L.state = off;
    if (state_smem) off += (nd_cols(S) + nr_cols(S, D)) * kSeqThreads * K * 8;
    L.total = off;
    return L;
}

// FlexGPU.Reserve on one raw node row (flex_gpu.go:178-223): the device index, or FGS_RESERVE_*.
// mem_each = memCnt / gpuCnt as NewGPUNode computes it (gpu_node.go:55), taken from the derived row.
__device__ __forceinline__ int reserve_on_row(const i64* raw, i64 rs, i64 mem_each, int S, int D, uint32_t kind, i64 mem_limit) {
    if (kind == 0) return FGS_RESERVE_NONE;     // :187-189
    if (kind == 3) return FGS_RESERVE_CONFLICT; // :191-193
    auto R = [&](int c) { return raw[(i64)c * rs]; };
    const u64 meta = (u64)R(NR_META);
    const uint32_t flags = (uint32_t)meta, mono = (uint32_t)(meta >> 32);
    if (flags & FGS_NF_REF_PANIC) return FGS_RESERVE_PANIC; // NewGPUNode, :195
    const i64 gpuCnt = R(nr_alloc_scalar(FGS_SCALAR_FLEX_GPU)); // > 0: REF_PANIC covers the rest
    int dev = FGS_RESERVE_FAIL;
    for (int d = D - 1; d >= 0; --d) { // lowest fitting index wins; no early exit so that the loads overlap
        const i64 used = R(nr_used(S, d));
        const bool m = (mono >> d) & 1u;
        // GPUAssumeFitIndexes, gpu_node.go:163-177 / MemAssumeFitIndexes, :122-161 (indexes[0] is the first fit)
        const bool fit = kind == 1 ? (!m && used == 0) : (!m && mem_each >= used + mem_limit);
        if (fit && (i64)d < gpuCnt) dev = d;
    }
    return dev;
}

template <bool SMEM, bool CLUSTER>
__global__ void __launch_bounds__(kSeqBlock, 1) seq_kernel(const __grid_constant__ SeqParams prm) {
    extern __shared__ __align__(128) unsigned char smem[];
    const int S = prm.dv.S, D = prm.dv.D;
    const int K = prm.K;
    const SeqSmem L = seq_smem(S, D, K, SMEM ? 1 : 0);
    uint64_t* bars = (uint64_t*)(smem + L.bars);
    i64* s_ring = (i64*)(smem + L.ring);
    i64* s_part_a = (i64*)(smem + L.part_a);
    i64* s_fin_a = (i64*)(smem + L.fin_a);
    u64* s_part_b = (u64*)(smem + L.part_b);
    u64* s_fin_b = (u64*)(smem + L.fin_b);
    i64* s_fit = (i64*)(smem + L.fit);
    u64* s_mail_a = (u64*)(smem + L.mail_a);
    u64* s_mail_b = (u64*)(smem + L.mail_b);
    int* s_dirty = (int*)(smem + L.bars + 32); // local index of the node whose row the last commit rewrote, -1 none

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const bool comm = warp == 0;        // warp 0 talks to the other CTAs; warps 1..8 own the nodes
    const int ct = tid - 32, cw = warp - 1;
    const int G = (int)gridDim.x, cta = (int)blockIdx.x;
    const i64 np = prm.dv.n_pad, n_nodes = prm.dv.n_nodes;
    const i64 cta_nodes = (i64)kSeqThreads * K;
    const i64 node0 = (i64)cta * cta_nodes;
    const int ncols = nd_cols(S);
    const int nrc = nr_cols(S, D);
    const int npc = L.npc;
    const i64 n_tiles = (prm.n_pods + kSeqTileP - 1) / kSeqTileP;

    auto issue_pod_tile = [&](i64 it) { // one elected thread
        const int stage = (int)(it & 1);
        const i64 p0 = it * kSeqTileP;
        i64* dst = s_ring + (size_t)stage * npc * kSeqTileP;
        mbar_expect_tx(&bars[stage], (uint32_t)(npc * kSeqTileP * 8));
        for (int c = 0; c < pd_cols(S); ++c) bulk_g2s(dst + c * kSeqTileP, prm.pods + (i64)c * prm.p_pad + p0, kSeqTileP * 8, &bars[stage]);
        for (int c = 0; c < SX_COLS; ++c)
            bulk_g2s(dst + (pd_cols(S) + c) * kSeqTileP, prm.extra + (i64)c * prm.p_pad + p0, kSeqTileP * 8, &bars[stage]);
    };

    if (tid == 0) {
        mbar_init(&bars[0], 1);
        mbar_init(&bars[1], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        *s_dirty = -1;
        s_dirty[1] = 0;
    }
    __syncthreads();
    if (tid == 0 && n_tiles > 0) issue_pod_tile(0);

    // ---- node state of the nodes this CTA owns: st = derived columns, rw = raw columns; element (c, i) at
    // [c * stride + i], i = node − node0 ----------------------------------------------------------------------
    i64 *st, *rw;
    i64 st_stride;
    if (SMEM) {
        st = (i64*)(smem + L.state);
        rw = st + (i64)ncols * cta_nodes;
        st_stride = cta_nodes;
        for (int c = 0; c < ncols + nrc; ++c)
            for (i64 i = tid; i < cta_nodes; i += kSeqBlock) {
                const i64 n = node0 + i;
                const i64* src = c < ncols ? (const i64*)prm.dv.nd + (i64)c * np : (const i64*)prm.dv.raw + (i64)(c - ncols) * np;
                st[(i64)c * st_stride + i] = n < n_nodes ? src[n] : 0;
            }
    } else {
        st = (i64*)prm.dv.nd + node0;
        rw = (i64*)prm.dv.raw + node0;
        st_stride = np;
    }
    if (CLUSTER) {
        for (int k = tid; k < 2 * kMaxCluster * (kWordsA + kWordsB); k += kSeqBlock) s_mail_a[k] = 0; // mail_b follows mail_a
        cluster_sync_all(); // no peer writes into a mailbox that is not cleared yet
    }
    __syncthreads();

    const bool use_flex_filter = prm.filter_mask & FGS_PL_FLEXGPU;
    const bool use_fit_filter = prm.filter_mask & FGS_PL_FIT;
    const bool sc_flex = prm.score_mask & FGS_PL_FLEXGPU;
    const bool sc_fit = prm.score_mask & FGS_PL_FIT;
    const bool sc_alloc = prm.score_mask & FGS_PL_ALLOCATABLE;

    long long pc[6] = {0, 0, 0, 0, 0, 0}, pt = 0;
    const bool prof = prm.prof != nullptr && tid == 32;
    auto tick = [&](int k) {
        if (prof) {
            const long long now = clock64();
            pc[k] += now - pt;
            pt = now;
        }
    };
    const bool cprof = prm.prof != nullptr && tid == 0;
    long long cc[4] = {0, 0, 0, 0}, cpt = 0;
    auto ctick = [&](int k) {
        if (cprof) {
            const long long now = clock64();
            cc[k] += now - cpt;
            cpt = now;
        }
    };
    if (prof) pt = clock64();
    if (cprof) cpt = clock64();
    // communication warp, shared-memory state only: copy the row the last commit rewrote to the global tables
    // (off the critical path: the owner thread only flags it)
    auto flush_dirty = [&]() {
        const int di = *(volatile int*)s_dirty;
        if (di >= 0) {
            const i64 n = node0 + di;
            for (int c = lane; c < ncols + nrc; c += 32) {
                if (c < ncols) ((i64*)prm.dv.nd)[(i64)c * np + n] = st[(i64)c * st_stride + di];
                else ((i64*)prm.dv.raw)[(i64)(c - ncols) * np + n] = rw[(i64)(c - ncols) * st_stride + di];
            }
            __syncwarp();
            if (lane == 0) *s_dirty = -1;
        }
    };

    // ---- ElasticQuota state: owned by the communication warp of CTA 0 (lane c holds resource column c) ---------
    const bool quota_on = prm.q_has_eq != nullptr;
    const bool quota_warp = quota_on && comm && cta == 0;
    const int QW = 2 + S;
    i64 agg_used = 0, agg_min = 0; // Σ over namespaces with a quota (aggregatedUsedOverMinWith, elasticquota.go:40-51)
    uint32_t agg_present = 0;
    bool pend_has = false, pend_key = false; // the previous pod's reservation, applied once its commit is known
    int pend_ns = -1;
    i64 pend_v = 0;
    uint32_t pend_keys = 0;
    if (quota_warp) {
        for (i64 k = 0; k < prm.n_ns; ++k) {
            if (!prm.q_has_eq[k]) continue;
            const uint32_t up = prm.q_used_present[k];
            agg_present |= up;
            if (lane < QW) {
                if (lane < 2 || ((up >> (lane - 2)) & 1u)) agg_used += prm.q_used[k * QW + lane];
                agg_min += prm.q_min[k * QW + lane];
            }
        }
    }
    int* s_commit = s_dirty + 1; // 1: the previous cycle's pod was committed by a thread of this CTA

    // with quotas one more (drain) iteration carries the last pod's commit flag to the quota warp
    const i64 n_iter = prm.n_pods + (quota_on ? 1 : 0);
    for (i64 p = 0; p < n_iter; ++p) {
        const bool drain = p >= prm.n_pods;
        const i64 it = p / kSeqTileP;
        const int jp = (int)(p % kSeqTileP), stage = (int)(it & 1);
        const uint32_t tag = (uint32_t)(p + 1);
        const int par = (int)(p & 1);
        const i64* praw = s_ring + (size_t)stage * npc * kSeqTileP;
        auto PC = [&](int c) { return praw[c * kSeqTileP + jp]; };

        if (comm) {
            // ================= the communication warp =====================================================
            // CapacityScheduling.PreFilter of this pod (capacity_scheduling.go:201-275, no nominated pods), while
            // the compute warps filter: once against the quota table as it is and once with the previous pod's
            // reservation added — which of the two holds is known when that pod's commit flag arrives with (A)
            uint32_t qbits = 0;
            bool q_has = false, q_key = false;
            int q_ns = -1;
            i64 q_v = 0;
            uint32_t q_keys = 0;
            if (quota_warp && !drain) {
                if (jp == 0) mbar_wait(&bars[stage], (uint32_t)((it >> 1) & 1));
                q_ns = (int)PC(pd_cols(S) + SX_NS);
                q_keys = (uint32_t)((u64)PC(PD_META) >> 32);
                q_v = lane == 0 ? PC(PD_REQ_CPU) : (lane == 1 ? PC(PD_REQ_MEM) : (lane < QW ? PC(PD_FIXED_COLS + lane - 2) : 0));
                q_has = q_ns >= 0 && q_ns < prm.n_ns && prm.q_has_eq[q_ns];
                q_key = lane < 2 || (lane < QW && ((q_keys >> (lane - 2)) & 1u));
                if (q_has) { // a namespace without an ElasticQuota passes (:210-217)
                    const uint32_t up = *(volatile uint32_t*)(prm.q_used_present + q_ns);
                    const bool u_key = lane < 2 || (lane < QW && ((up >> (lane - 2)) & 1u));
                    const i64 u = u_key ? *(volatile i64*)(prm.q_used + (i64)q_ns * QW + lane) : 0;
                    const i64 mx = lane < QW ? prm.q_max[(i64)q_ns * QW + lane] : 0;
                    const i64 d_same = (pend_has && pend_ns == q_ns && pend_key) ? pend_v : 0;
                    const i64 d_all = (pend_has && pend_key) ? pend_v : 0;
                    // usedOverMaxWith = cmp2(request, Used, Max): the request's scalar keys (elasticquota.go:94-96,165-181)
                    const uint32_t m0 = __ballot_sync(0xffffffffu, q_key && q_v + u > mx);
                    const uint32_t m1 = __ballot_sync(0xffffffffu, q_key && q_v + u + d_same > mx);
                    // aggregatedUsedOverMinWith: keys of Σused ∪ the request's (:40-51)
                    const uint32_t ak0 = agg_present | q_keys, ak1 = ak0 | (pend_has ? pend_keys : 0u);
                    const bool a0 = lane < 2 || (lane < QW && ((ak0 >> (lane - 2)) & 1u));
                    const bool a1 = lane < 2 || (lane < QW && ((ak1 >> (lane - 2)) & 1u));
                    const uint32_t n0 = __ballot_sync(0xffffffffu, a0 && agg_used + q_v > agg_min);
                    const uint32_t n1 = __ballot_sync(0xffffffffu, a1 && agg_used + d_all + q_v > agg_min);
                    const uint32_t code0 = m0 ? 1u : (n0 ? 2u : 0u), code1 = m1 ? 1u : (n1 ? 2u : 0u);
                    qbits = (code0 << 1) | (code1 << 3);
                }
            }
            __syncthreads(); // (1) the compute warps' phase-A partials are in shared memory
            if (lane == 0 && !drain && jp == 0 && it + 1 < n_tiles) { // every thread is past its last read of the previous tile
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                issue_pod_tile(it + 1);
            }
            const bool w = lane < kSeqWarps;
            const i64 cf = w ? s_part_a[lane * 4 + 0] : 0;
            if (cprof) consume(cf);
            ctick(0); // commit + Filter of the slowest thread of this CTA
            uint32_t c2 = __reduce_add_sync(0xffffffffu, (uint32_t)cf & 0xffffffu);
            uint32_t f2 = __reduce_or_sync(0xffffffffu, (uint32_t)(cf >> 24));
            if (quota_on) { // bits 1-4: the two PreFilter verdicts (CTA 0); bit 5: the previous pod was committed here
                const int cm = *(volatile int*)s_commit;
                f2 |= qbits | (cm ? 32u : 0u);
                __syncwarp();
                if (lane == 0 && cm) *s_commit = 0;
            }
            i64 a1 = warp_max_i64(w ? s_part_a[lane * 4 + 1] : kMin64);
            i64 a2 = warp_max_i64(w ? s_part_a[lane * 4 + 2] : kMin64);
            i64 a3 = warp_max_i64(w ? s_part_a[lane * 4 + 3] : kMin64);
            if (G > 1) {
                uint32_t pay = 0;
                if (lane == 0) pay = c2 | (f2 << 24);
                else if (lane < 7) {
                    const i64 v = lane <= 2 ? a1 : (lane <= 4 ? a2 : a3);
                    pay = (lane & 1) ? (uint32_t)v : (uint32_t)((u64)v >> 32);
                }
                if (CLUSTER) {
                    const uint32_t slot = s_u32(s_mail_a + ((size_t)par * kMaxCluster + cta) * kWordsA + lane);
                    if (lane < 7)
                        for (int dst = 0; dst < G; ++dst) st_word_peer(mapa_u32(slot, (uint32_t)dst), pay, tag);
                } else if (lane < 7) {
                    st_word(prm.mail_a + ((size_t)par * G + cta) * kWordsA + lane, pay, tag);
                }
                uint32_t tc = 0, tf = 0;
                i64 t1 = kMin64, t2 = kMin64, t3 = kMin64;
                for (int c = lane; c < G; c += 32) {
                    const u64* m = CLUSTER ? s_mail_a + ((size_t)par * kMaxCluster + c) * kWordsA : prm.mail_a + ((size_t)par * G + c) * kWordsA;
                    u64 wv[7];
                    bool ok;
                    do {
                        ok = true;
#pragma unroll
                        for (int k = 0; k < 7; ++k) {
                            wv[k] = CLUSTER ? ld_word_local(m + k) : ld_word(m + k);
                            ok = ok && (uint32_t)(wv[k] >> 32) == tag;
                        }
                    } while (!ok);
                    tc += (uint32_t)wv[0] & 0xffffffu;
                    tf |= ((uint32_t)wv[0] >> 24) & 0xffu;
                    const i64 x1 = (i64)((wv[2] << 32) | (wv[1] & 0xffffffffull));
                    const i64 x2 = (i64)((wv[4] << 32) | (wv[3] & 0xffffffffull));
                    const i64 x3 = (i64)((wv[6] << 32) | (wv[5] & 0xffffffffull));
                    t1 = x1 > t1 ? x1 : t1;
                    t2 = x2 > t2 ? x2 : t2;
                    t3 = x3 > t3 ? x3 : t3;
                }
                c2 = __reduce_add_sync(0xffffffffu, tc);
                f2 = __reduce_or_sync(0xffffffffu, tf);
                a1 = warp_max_i64(t1);
                a2 = warp_max_i64(t2);
                a3 = warp_max_i64(t3);
            }
            if (cprof) consume((i64)c2);
            ctick(1); // exchange A, including the wait for the slowest CTA
            uint32_t qcode = 0;
            if (quota_on) {
                const bool committed_prev = (f2 >> 5) & 1u;
                qcode = committed_prev ? (f2 >> 3) & 3u : (f2 >> 1) & 3u;
                if (quota_warp) {
                    if (pend_has && committed_prev) { // CapacityScheduling.Reserve → reserveResource (elasticquota.go:74-80)
                        if (pend_key) {
                            const uint32_t up = *(volatile uint32_t*)(prm.q_used_present + pend_ns);
                            const bool u_key = lane < 2 || ((up >> (lane - 2)) & 1u);
                            i64* up_ = prm.q_used + (i64)pend_ns * QW + lane;
                            *(volatile i64*)up_ = (u_key ? *(volatile i64*)up_ : 0) + pend_v;
                            agg_used += pend_v;
                        }
                        __syncwarp();
                        if (lane == 0) *(volatile uint32_t*)(prm.q_used_present + pend_ns) |= pend_keys;
                        __syncwarp();
                        agg_present |= pend_keys;
                    }
                    pend_has = q_has && qcode == 0 && !drain;
                    pend_ns = q_ns;
                    pend_v = q_v;
                    pend_keys = q_keys;
                    pend_key = q_key;
                }
                if (qcode) c2 = 0; // rejected in PreFilter: no Filter/Score results, nothing to select
                f2 = (f2 & 1u) | (qcode << 8);
            }
            if (lane == 0) {
                s_fin_a[0] = (i64)(((u64)f2 << 32) | c2);
                s_fin_a[1] = a1;
                s_fin_a[2] = a2;
                s_fin_a[3] = a3;
            }
            __syncthreads(); // (2)
            if (SMEM) flush_dirty();
            if (c2 > 0) {    // uniform over the grid: phase B runs
                __syncthreads(); // (3) the compute warps' keys are in shared memory
                u64 k2 = (u64)warp_max_i64(w ? (i64)s_part_b[lane * 2 + 0] : 0);
                if (cprof) consume((i64)k2);
                ctick(2); // Fit score + NormalizeScore of the slowest thread of this CTA
                uint32_t e2 = __reduce_or_sync(0xffffffffu, w ? (uint32_t)s_part_b[lane * 2 + 1] : 0u);
                if (G > 1) {
                    const uint32_t pay = lane == 0 ? (uint32_t)k2 : (lane == 1 ? (uint32_t)(k2 >> 32) : e2);
                    if (CLUSTER) {
                        const uint32_t slot = s_u32(s_mail_b + ((size_t)par * kMaxCluster + cta) * kWordsB + lane);
                        if (lane < 3)
                            for (int dst = 0; dst < G; ++dst) st_word_peer(mapa_u32(slot, (uint32_t)dst), pay, tag);
                    } else if (lane < 3) {
                        st_word(prm.mail_b + ((size_t)par * G + cta) * kWordsB + lane, pay, tag);
                    }
                    u64 tk = 0;
                    uint32_t te = 0;
                    for (int c = lane; c < G; c += 32) {
                        const u64* m = CLUSTER ? s_mail_b + ((size_t)par * kMaxCluster + c) * kWordsB : prm.mail_b + ((size_t)par * G + c) * kWordsB;
                        u64 w0, w1, w2;
                        do {
                            w0 = CLUSTER ? ld_word_local(m + 0) : ld_word(m + 0);
                            w1 = CLUSTER ? ld_word_local(m + 1) : ld_word(m + 1);
                            w2 = CLUSTER ? ld_word_local(m + 2) : ld_word(m + 2);
                        } while ((uint32_t)(w0 >> 32) != tag || (uint32_t)(w1 >> 32) != tag || (uint32_t)(w2 >> 32) != tag);
                        const u64 x = (w1 << 32) | (w0 & 0xffffffffull);
                        tk = x > tk ? x : tk;
                        te |= (uint32_t)w2;
                    }
                    k2 = (u64)warp_max_i64((i64)tk);
                    e2 = __reduce_or_sync(0xffffffffu, te);
                }
                if (cprof) consume((i64)k2);
                ctick(3); // exchange B
                if (lane == 0) {
                    s_fin_b[0] = k2;
                    s_fin_b[1] = e2;
                }
                __syncthreads(); // (4)
            }
            continue;
        }

        // ================= the compute warps ===============================================================
        if (!drain && jp == 0) mbar_wait(&bars[stage], (uint32_t)((it >> 1) & 1));
        // per-pod values (broadcast reads)
        const u64 pmeta = (u64)PC(PD_META);
        const uint32_t pf = (uint32_t)pmeta;
        const uint32_t kind = pf & 3u;
        const bool req_none = pf & FGS_PF_REQ_NONE;
        const uint32_t smask = (uint32_t)(pmeta >> 32) & ~prm.fit_ignored;
        const i64 gpu_lim = PC(PD_GPU_LIM), mem_lim = PC(PD_MEM_LIM);
        const i64 lim = kind == 1 ? gpu_lim : mem_lim;
        const i64 rc = PC(PD_REQ_CPU), rm = PC(PD_REQ_MEM), re = PC(PD_REQ_EPH);
        const i64 nzc = PC(PD_NZ_CPU), nzm = PC(PD_NZ_MEM);

        // ---- phase A: Filter on my nodes ---------------------------------------------------------------------
        uint32_t feas_mask = 0, cnt = 0, fl = 0;
        i64 vflex = kMin64, vhi = kMin64, vnlo = kMin64;
        const bool flex_pod = use_flex_filter && kind != 0;
        const bool fit_req = use_fit_filter && !req_none;
        const int Kc = drain ? 0 : K; // the drain iteration only carries flags
        for (int j = 0; j < Kc; ++j) {
            const i64 i = (i64)j * kSeqThreads + ct;
            if (node0 + i >= n_nodes) break;
            auto C = [&](int c) { return st[(i64)c * st_stride + i]; };
            // every column the verdict can need, loaded before any of them is used (one shared-memory latency)
            const uint32_t nf = (uint32_t)C(ND_FLAGS);
            const i64 thr = C(kind == 1 ? ND_THR_GPU : ND_THR_MEM);
            const i64 fc = C(ND_FREE_CPU), fm_ = C(ND_FREE_MEM), fe = C(ND_FREE_EPH);
            const i64 sflex = C(kind == 1 ? ND_GPU_SCORE : ND_MEM_SCORE);
            const i64 ar = C(ND_ALLOC_RAW);
            // FlexGPU.Filter, flex_gpu.go:41-119, folded into one threshold per pod kind (derive.cuh)
            const bool flex_ok = !flex_pod || (kind != 3 && lim <= thr);
            // Fit.Filter, fit.go:256-329
            bool fit_ok = !use_fit_filter || (nf & NDF_PODS_OK);
            if (fit_req) {
                fit_ok = fit_ok & (rc <= fc) & (rm <= fm_) & (re <= fe);
                uint32_t m = smask;
                while (m) {
                    const int s = __ffs(m) - 1;
                    m &= m - 1;
                    if (PC(PD_FIXED_COLS + s) > C(nd_free_scalar(s))) fit_ok = false;
                }
            }
            const bool valid = nf & FGS_NF_VALID;
            // the reference panics inside NewGPUNode (flex_gpu.go:101) when FlexGPU.Filter gets that far
            const bool reach = flex_pod && kind != 3 && flex_ok && (nf & FGS_NF_REF_PANIC) && (prm.flex_filter_first || fit_ok) && valid;
            if (reach) fl |= 1u;
            if (flex_ok && fit_ok && valid && !reach) {
                feas_mask |= 1u << j;
                cnt++;
                const i64 s = (kind == 1 || kind == 2) ? sflex : 0;
                vflex = s > vflex ? s : vflex;
                vhi = ar > vhi ? ar : vhi;
                vnlo = -ar > vnlo ? -ar : vnlo;
            }
        }
        cnt = __reduce_add_sync(0xffffffffu, cnt);
        fl = __reduce_or_sync(0xffffffffu, fl);
        if (sc_flex) vflex = warp_max_i64(vflex);
        if (sc_alloc) {
            vhi = warp_max_i64(vhi);
            vnlo = warp_max_i64(vnlo);
        }
        if (lane == 0) {
            s_part_a[cw * 4 + 0] = (i64)(cnt | (fl << 24));
            s_part_a[cw * 4 + 1] = vflex;
            s_part_a[cw * 4 + 2] = vhi;
            s_part_a[cw * 4 + 3] = vnlo;
        }
        tick(0); // pod fetch + Filter on my nodes
        __syncthreads(); // (1)

        // ---- while the communication warp exchanges the phase-A results: the part of Score that needs none of
        // them — Fit's Least/MostAllocated (least_allocated.go:29-57, most_allocated.go:29-61,
        // resource_allocation.go:45-117) on my feasible nodes -----------------------------------------------
        if (sc_fit) {
            uint32_t fm = feas_mask;
            while (fm) {
                const int j = __ffs(fm) - 1;
                fm &= fm - 1;
                const i64 i = (i64)j * kSeqThreads + ct;
                auto C = [&](int c) { return st[(i64)c * st_stride + i]; };
                i64 node_score = 0, weight_sum = 0;
                for (int k = 0; k < prm.fit_n_res; ++k) {
                    const int res = prm.fit_res[k];
                    i64 alloc = 0, req = 0;
                    if (res == FGS_RES_CPU) {
                        alloc = C(ND_CAP_CPU);
                        req = C(ND_NZR_CPU) + nzc;
                    } else if (res == FGS_RES_MEM) {
                        alloc = C(ND_CAP_MEM);
                        req = C(ND_NZR_MEM) + nzm;
                    } else if (res == FGS_RES_EPH) {
                        alloc = C(ND_ALLOC_EPH);
                        req = (alloc - C(ND_FREE_EPH)) + re;
                    } else if (res >= FGS_RES_SCALAR0 && res < FGS_RES_SCALAR0 + S) {
                        const int s = res - FGS_RES_SCALAR0;
                        const i64 pr = PC(PD_FIXED_COLS + s);
                        if (pr != 0) { // resource_allocation.go:93-97
                            alloc = C(nd_alloc_scalar(S, s));
                            req = (alloc - C(nd_free_scalar(s))) + pr;
                        }
                    }
                    if (alloc == 0) continue; // resource_allocation.go:64-67
                    i64 rs;
                    if (prm.fit_strategy == 0) rs = req > alloc ? 0 : go_div(wmul(alloc - req, 100), alloc);
                    else rs = go_div(wmul(req > alloc ? alloc : req, 100), alloc);
                    node_score = wadd(node_score, wmul(rs, prm.fit_w[k]));
                    weight_sum = wadd(weight_sum, prm.fit_w[k]);
                }
                s_fit[j * kSeqThreads + ct] = weight_sum == 0 ? 0 : go_div(node_score, weight_sum);
            }
        }
        tick(1); // Fit score, overlapped with exchange A
        __syncthreads(); // (2)
        const i64 count = (i64)(uint32_t)s_fin_a[0];
        if (prof) consume(count);
        tick(2); // waiting for exchange A
        const uint32_t qcode = (uint32_t)(s_fin_a[0] >> 40) & 3u; // CapacityScheduling.PreFilter: 0 pass, 1 over Max, 2 over Σmin
        const bool fpanic = ((s_fin_a[0] >> 32) & 1) && qcode == 0;
        const i64 g_flex = s_fin_a[1], g_hi = s_fin_a[2], g_nlo = s_fin_a[3];

        // ---- phase B: NormalizeScore / weights / selectHost key on my feasible nodes -------------------------
        u64 key = 0;     // 0 = no candidate
        uint32_t eb = 0; // FGS_PODERR_* raised while scoring
        if (count > 0) { // uniform over the grid
            const i64 fmax = g_flex > 0 ? g_flex : 0; // maxCount starts at 0 (normalize_score.go:27)
            const i64 alo = -g_nlo;
            const i64 arange = wsub(g_hi, alo);
            uint32_t ph = (uint32_t)(prm.seed ^ (prm.seed >> 32)) ^ ((uint32_t)p * 0x9E3779B1u);
            ph ^= ph >> 16;
            ph *= 0x85EBCA6Bu;
            ph ^= ph >> 13;
            uint32_t fm = feas_mask;
            while (fm) {
                const int j = __ffs(fm) - 1;
                fm &= fm - 1;
                const i64 i = (i64)j * kSeqThreads + ct;
                auto C = [&](int c) { return st[(i64)c * st_stride + i]; };
                i64 t = 0;
                if (sc_flex) {
                    // FlexGPU.Score (flex_gpu.go:142-166) calls NewGPUNode first, for every pod kind (:148)
                    if ((uint32_t)C(ND_FLAGS) & FGS_NF_REF_PANIC) eb |= FGS_PODERR_REF_PANIC;
                    const i64 s = kind == 1 ? C(ND_GPU_SCORE) : (kind == 2 ? C(ND_MEM_SCORE) : 0);
                    // DefaultNormalizeScore, reverse (normalize_score.go:26-54)
                    const i64 v = fmax == 0 ? 100 : 100 - go_div(wmul(100, s), fmax);
                    if (v > 100 || v < 0) eb |= FGS_PODERR_SCORE_RANGE; // runtime/framework.go:920-923
                    t = wadd(t, wmul(v, prm.w_flex));
                }
                if (sc_fit) {
                    const i64 v = s_fit[j * kSeqThreads + ct];
                    if (v > 100 || v < 0) eb |= FGS_PODERR_SCORE_RANGE;
                    t = wadd(t, wmul(v, prm.w_fit));
                }
                if (sc_alloc) { // allocatable.go:61-72,146-171
                    const i64 v = arange == 0 ? 0 : go_div(wmul(wsub(C(ND_ALLOC_RAW), alo), 100), arange);
                    if (v > 100 || v < 0) eb |= FGS_PODERR_SCORE_RANGE;
                    t = wadd(t, wmul(v, prm.w_alloc));
                }
                // selectHost candidate (generic_scheduler.go:146-167): same key as the batched kernels
                const uint32_t ng = (uint32_t)(node0 + i);
                u64 k;
                if (prm.key_shift == 40) {
                    uint32_t h = (ph ^ ng) * 0x9E3779B1u;
                    h ^= h >> 15;
                    k = ((u64)t << 40) | ((u64)(h & 0xffffu) << 24) | ng;
                } else {
                    k = ((u64)t << 24) | ng;
                }
                k += 1;
                key = k > key ? k : key;
            }
            key = (u64)warp_max_i64((i64)key);
            eb = __reduce_or_sync(0xffffffffu, eb);
            if (lane == 0) {
                s_part_b[cw * 2 + 0] = key;
                s_part_b[cw * 2 + 1] = eb;
            }
            tick(3); // NormalizeScore + key on my feasible nodes
            __syncthreads(); // (3)
            __syncthreads(); // (4)
            key = s_fin_b[0];
            if (prof) consume((i64)key);
            tick(4); // waiting for exchange B
            eb = (uint32_t)s_fin_b[1];
        }

        // ---- verdict of the cycle (every compute thread computes the same) ------------------------------------
        uint32_t err = fpanic ? FGS_PODERR_REF_PANIC : 0u;
        if (count >= 2) err |= eb; // scoring only runs with >= 2 feasible nodes (generic_scheduler.go:112-127)
        i64 sel = -1, score = 0;
        if (key > 0) {
            const u64 k = key - 1;
            sel = (i64)(k & 0xffffffu);
            score = (i64)(k >> prm.key_shift);
        }
        const bool mine = sel >= node0 && sel < node0 + cta_nodes && (int)((sel - node0) % kSeqThreads) == ct;
        if (!drain && (sel < 0 ? (cta == 0 && ct == 0) : mine)) {
            int dev = FGS_RESERVE_SKIPPED;
            bool commit = sel >= 0 && err == 0;
            if (commit && prm.reserve_flex) {
                dev = reserve_on_row(rw + (sel - node0), st_stride, st[(i64)ND_MEM_EACH * st_stride + (sel - node0)], S, D, kind, mem_lim);
                if (dev == FGS_RESERVE_PANIC) err |= FGS_PODERR_REF_PANIC;
                commit = dev >= 0 || dev == FGS_RESERVE_NONE;
            }
            if (commit) {
                // assume: NodeInfo.AddPod (framework/types.go:583-609) on the raw row, then re-derive what a
                // placement can change. All loads first, then the sums, then the stores.
                const i64 i = sel - node0;
                i64* row = rw + i;
                auto A = [&](int c) -> i64& { return row[(i64)c * st_stride]; };
                const i64 o_rc = A(NR_REQ_MCPU), o_rm = A(NR_REQ_MEM), o_re = A(NR_REQ_EPH), o_nzc = A(NR_NZ_MCPU), o_nzm = A(NR_NZ_MEM);
                const i64 o_np = A(NR_NUM_PODS), o_gs = A(NR_FLEX_GPU_SUM), o_ms = A(NR_FLEX_MEM_SUM);
                u64 meta = (u64)A(NR_META);
                const i64 gpuCnt = ((uint32_t)meta & FGS_NF_HAS_GPU_RES) ? A(nr_alloc_scalar(FGS_SCALAR_FLEX_GPU)) : 0;
                // the annotation constructGPUs (gpu_node.go:66-120) will read on the next cycle
                const i64 annot = kind == 0 ? (i64)FGS_ANNOT_NONE : (dev >= 0 ? (i64)dev : PC(pd_cols(S) + SX_ANNOT));
                const bool annot_ok = annot != (i64)FGS_ANNOT_NONE && annot >= 0 && annot < gpuCnt;
                const i64 o_used = (annot_ok && (kind & 2u)) ? A(nr_used(S, (int)annot)) : 0;
                i64 o_sc[FGS_MAX_SCALARS];
#pragma unroll
                for (int s = 0; s < FGS_MAX_SCALARS; ++s) o_sc[s] = s < S ? A(nr_req_scalar(S, s)) : 0;
                bool in_range = true;
                auto put = [&](int c, i64 v) {
                    A(c) = v;
                    in_range = in_range && v > -kMaxQuantity && v < kMaxQuantity;
                };
                put(NR_REQ_MCPU, o_rc + rc);
                put(NR_REQ_MEM, o_rm + rm);
                put(NR_REQ_EPH, o_re + re);
#pragma unroll
                for (int s = 0; s < FGS_MAX_SCALARS; ++s)
                    if (s < S) put(nr_req_scalar(S, s), o_sc[s] + PC(PD_FIXED_COLS + s));
                put(NR_NZ_MCPU, o_nzc + PC(pd_cols(S) + SX_ACCT_CPU));
                put(NR_NZ_MEM, o_nzm + PC(pd_cols(S) + SX_ACCT_MEM));
                put(NR_NUM_PODS, o_np + 1);
                // what nodeResourceLimitSum (flex_gpu.go:133-140) will see
                if (kind & 1u) put(NR_FLEX_GPU_SUM, o_gs + gpu_lim);
                if (kind & 2u) put(NR_FLEX_MEM_SUM, o_ms + mem_lim);
                if (annot != (i64)FGS_ANNOT_NONE) {
                    if (!annot_ok) {
                        meta |= FGS_NF_REF_PANIC; // gpus[index] out of range (gpu_node.go:110,115)
                    } else {
                        if (kind & 1u) meta |= 1ull << (32 + (int)annot);
                        if (kind & 2u) put(nr_used(S, (int)annot), o_used + mem_lim);
                    }
                    A(NR_META) = (i64)meta;
                }
                if (!in_range) atomicMin(prm.status, (int)(p < 0x7fffffff ? p : 0x7ffffffe));
                derive_row<true>(prm.dv, row, st_stride, st + i, st_stride);
                if (SMEM) *s_dirty = (int)i; // the communication warp copies the row to the global tables
                if (quota_on) *s_commit = 1; // → CapacityScheduling.Reserve of this pod, applied by the quota warp
            }
            prm.out[(i64)SO_COUNT * prm.p_pad + p] = count;
            prm.out[(i64)SO_NODE * prm.p_pad + p] = sel;
            prm.out[(i64)SO_SCORE * prm.p_pad + p] = score;
            prm.out[(i64)SO_ERR * prm.p_pad + p] = (i64)err;
            prm.out[(i64)SO_DEVICE * prm.p_pad + p] = (i64)dev;
            prm.out[(i64)SO_COMMITTED * prm.p_pad + p] = commit ? 1 : 0;
            prm.out[(i64)SO_QUOTA * prm.p_pad + p] = (i64)qcode;
        }
        tick(5); // verdict + commit (this thread's share)
    }
    __syncthreads();
    if (SMEM && comm) flush_dirty();
    if (prof)
        for (int k = 0; k < 6; ++k) prm.prof[(size_t)cta * 16 + k] = (u64)pc[k];
    if (cprof)
        for (int k = 0; k < 4; ++k) prm.prof[(size_t)cta * 16 + 8 + k] = (u64)cc[k];
    if (CLUSTER) cluster_sync_all(); // no CTA leaves while a peer could still address its shared memory
}

// FlexGPU.Reserve for a batch of (pod, node) pairs, no commit
__global__ void reserve_kernel(const i64* __restrict__ raw, const i64* __restrict__ nd, i64 np, i64 n_nodes, int S, int D, const i64* __restrict__ pods, i64 p_pad,
                               i64 n_pods, const int* __restrict__ node, int* __restrict__ dev) {
    const i64 p = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n_pods) return;
    const int n = node[p];
    if (n < 0 || n >= n_nodes) {
        dev[p] = FGS_RESERVE_SKIPPED;
        return;
    }
    const uint32_t kind = (uint32_t)pods[(i64)PD_META * p_pad + p] & 3u;
    dev[p] = reserve_on_row(raw + n, np, nd[(i64)ND_MEM_EACH * np + n], S, D, kind, pods[(i64)PD_MEM_LIM * p_pad + p]);
}

} // namespace

// =================================================================================================
// host side
// =================================================================================================

struct SeqState {
    void* d_pods = nullptr; // [(pd_cols(S) + SX_COLS)][cap]
    void* h_pods = nullptr;
    void* d_out = nullptr;  // [SO_COLS][cap]
    void* h_out = nullptr;
    int64_t cap = 0;
    void* d_quota = nullptr; // has_eq | min | max | used | used_present
    void* h_quota = nullptr;
    size_t quota_bytes = 0;
    void* d_mail = nullptr; // mail_a | mail_b | status
    size_t mail_bytes = 0;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    int sm_count = 0, smem_optin = 0;
};

void Engine::seq_free() {
    if (!seq) return;
    if (seq->d_pods) cudaFree(seq->d_pods);
    if (seq->h_pods) cudaFreeHost(seq->h_pods);
    if (seq->d_out) cudaFree(seq->d_out);
    if (seq->h_out) cudaFreeHost(seq->h_out);
    if (seq->d_mail) cudaFree(seq->d_mail);
    if (seq->d_quota) cudaFree(seq->d_quota);
    if (seq->h_quota) cudaFreeHost(seq->h_quota);
    if (seq->ev0) cudaEventDestroy(seq->ev0);
    if (seq->ev1) cudaEventDestroy(seq->ev1);
    delete seq;
    seq = nullptr;
}

static int64_t round_up64(int64_t v, int64_t m) { return (v + m - 1) / m * m; }

static int seq_reserve_buffers(Engine* e, int64_t pp) {
    if (!e->seq) {
        e->seq = new SeqState();
        FGS_CUDA(e, cudaDeviceGetAttribute(&e->seq->sm_count, cudaDevAttrMultiProcessorCount, e->device));
        FGS_CUDA(e, cudaDeviceGetAttribute(&e->seq->smem_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, e->device));
        FGS_CUDA(e, cudaEventCreate(&e->seq->ev0));
        FGS_CUDA(e, cudaEventCreate(&e->seq->ev1));
    }
    SeqState* q = e->seq;
    if (pp > q->cap) {
        if (q->d_pods) cudaFree(q->d_pods);
        if (q->h_pods) cudaFreeHost(q->h_pods);
        if (q->d_out) cudaFree(q->d_out);
        if (q->h_out) cudaFreeHost(q->h_out);
        q->d_pods = q->h_pods = q->d_out = q->h_out = nullptr;
        q->cap = 0;
        const size_t in_bytes = sizeof(int64_t) * (size_t)(pd_cols(e->S) + SX_COLS) * (size_t)pp;
        const size_t out_bytes = sizeof(int64_t) * (size_t)SO_COLS * (size_t)pp;
        FGS_CUDA(e, cudaMalloc(&q->d_pods, in_bytes));
        FGS_CUDA(e, cudaMallocHost(&q->h_pods, in_bytes));
        FGS_CUDA(e, cudaMalloc(&q->d_out, out_bytes));
        FGS_CUDA(e, cudaMallocHost(&q->h_out, out_bytes));
        q->cap = pp;
    }
    return FGS_OK;
}

// pod columns → pinned staging → device, in the layout the kernels read ([column][pp])
static int seq_stage_pods(Engine* e, const fgs_pod_table* t, int64_t pp) {
    SeqState* q = e->seq;
    const int64_t P = t->n_pods;
    const int S = e->S;
    int64_t* h = (int64_t*)q->h_pods;
    auto col = [&](int c, const int64_t* src) {
        std::memcpy(h + (int64_t)c * pp, src, sizeof(int64_t) * (size_t)P);
        std::memset(h + (int64_t)c * pp + P, 0, sizeof(int64_t) * (size_t)(pp - P));
    };
    col(PD_GPU_LIM, t->flex_gpu_limit);
    col(PD_MEM_LIM, t->flex_mem_limit);
    col(PD_REQ_CPU, t->req_mcpu);
    col(PD_REQ_MEM, t->req_mem);
    col(PD_REQ_EPH, t->req_eph);
    col(PD_NZ_CPU, t->nz_mcpu);
    col(PD_NZ_MEM, t->nz_mem);
    for (int s = 0; s < S; ++s) col(PD_FIXED_COLS + s, t->req_scalar + (int64_t)s * P);
    const int x0 = pd_cols(S);
    col(x0 + SX_ACCT_CPU, t->acct_nz_mcpu);
    col(x0 + SX_ACCT_MEM, t->acct_nz_mem);
    int64_t* meta = h + (int64_t)PD_META * pp;
    int64_t* ann = h + (int64_t)(x0 + SX_ANNOT) * pp;
    int64_t* nsc = h + (int64_t)(x0 + SX_NS) * pp;
    for (int64_t i = 0; i < P; ++i) {
        meta[i] = (int64_t)(((uint64_t)t->req_scalar_present[i] << 32) | (uint64_t)t->flags[i]);
        ann[i] = (int64_t)t->index_annot[i];
        nsc[i] = t->ns ? (int64_t)t->ns[i] : -1;
    }
    std::memset(meta + P, 0, sizeof(int64_t) * (size_t)(pp - P));
    std::memset(ann + P, 0, sizeof(int64_t) * (size_t)(pp - P));
    std::memset(nsc + P, 0, sizeof(int64_t) * (size_t)(pp - P));
    FGS_CUDA(e, cudaMemcpyAsync(q->d_pods, q->h_pods, sizeof(int64_t) * (size_t)(x0 + SX_COLS) * (size_t)pp, cudaMemcpyHostToDevice, e->stream));
    return FGS_OK;
}

static int seq_check_inputs(Engine* e, const fgs_pod_table* t, const char* who) {
    if (!e->have_nodes) return e->fail(FGS_E_STATE, std::string(who) + " before fgs_upload_nodes");
    if (!t || t->n_pods < 0) return e->fail(FGS_E_INVALID, std::string(who) + ": null pod table or negative size");
    if (t->n_scalars != e->S) return e->fail(FGS_E_INVALID, std::string(who) + ": pod table n_scalars differs from the engine's");
    if (t->n_pods >= ((int64_t)1 << 31) - 64) return e->fail(FGS_E_TOO_LARGE, std::string(who) + ": at most 2^31 pods per call");
    return FGS_OK;
}

} // namespace fgs

using namespace fgs;

extern "C" int fgs_schedule_sequential(fgs_engine* eh, const fgs_pod_table* t, const fgs_seq_opts* opts, fgs_seq_results* out) {
    if (!eh || !out) return FGS_E_INVALID;
    Engine* e = (Engine*)eh;
    int rc = seq_check_inputs(e, t, "fgs_schedule_sequential");
    if (rc) return rc;
    if (!t->acct_nz_mcpu || !t->acct_nz_mem || !t->index_annot)
        return e->fail(FGS_E_INVALID, "fgs_schedule_sequential: the pod table lacks acct_nz_mcpu / acct_nz_mem / index_annot");
    if (e->node_base != 0 || (e->total_nodes != 0 && e->total_nodes != e->n_nodes))
        return e->fail(FGS_E_STATE, "fgs_schedule_sequential: the engine must hold the whole cluster (node_index_base 0)");
    cudaSetDevice(e->device);
    FGS_CUDA(e, cudaStreamSynchronize(e->stream));
    if (!e->nodes_derived) {
        rc = e->derive_now();
        if (rc) return rc;
    }
    const int64_t P = t->n_pods;
    if (P == 0) return FGS_OK;
    const int64_t pp = round_up64(P, kSeqTileP);
    rc = seq_reserve_buffers(e, pp);
    if (rc) return rc;
    SeqState* q = e->seq;
    rc = seq_stage_pods(e, t, pp);
    if (rc) return rc;

    // ElasticQuota tables (CapacityScheduling in the profile)
    const fgs_quotas* qt = opts ? opts->quotas : nullptr;
    size_t q_off[5] = {0, 0, 0, 0, 0}, q_total = 0;
    if (qt) {
        if (qt->n_scalars != e->S) return e->fail(FGS_E_INVALID, "fgs_schedule_sequential: quotas n_scalars differs from the engine's");
        if (qt->n_ns < 0 || (qt->n_ns > 0 && (!qt->has_eq || !qt->min || !qt->max || !qt->used || !qt->used_scalar_present)))
            return e->fail(FGS_E_INVALID, "fgs_schedule_sequential: incomplete quota tables");
        const size_t NS = (size_t)qt->n_ns, W = (size_t)(2 + e->S);
        auto take = [&](size_t bytes) {
            const size_t o = q_total;
            q_total += (bytes + 255) & ~(size_t)255;
            return o;
        };
        q_off[0] = take(NS);
        q_off[1] = take(sizeof(int64_t) * NS * W);
        q_off[2] = take(sizeof(int64_t) * NS * W);
        q_off[3] = take(sizeof(int64_t) * NS * W);
        q_off[4] = take(sizeof(uint32_t) * NS);
        q_total += 256;
        if (q_total > q->quota_bytes) {
            if (q->d_quota) cudaFree(q->d_quota);
            if (q->h_quota) cudaFreeHost(q->h_quota);
            q->d_quota = q->h_quota = nullptr;
            q->quota_bytes = 0;
            FGS_CUDA(e, cudaMalloc(&q->d_quota, q_total));
            FGS_CUDA(e, cudaMallocHost(&q->h_quota, q_total));
            q->quota_bytes = q_total;
        }
        char* hq = (char*)q->h_quota;
        std::memcpy(hq + q_off[0], qt->has_eq, NS);
        std::memcpy(hq + q_off[1], qt->min, sizeof(int64_t) * NS * W);
        std::memcpy(hq + q_off[2], qt->max, sizeof(int64_t) * NS * W);
        std::memcpy(hq + q_off[3], qt->used, sizeof(int64_t) * NS * W);
        std::memcpy(hq + q_off[4], qt->used_scalar_present, sizeof(uint32_t) * NS);
        FGS_CUDA(e, cudaMemcpyAsync(q->d_quota, q->h_quota, q_total, cudaMemcpyHostToDevice, e->stream));
    }

    // geometry. Small clusters of nodes: ONE thread-block cluster (<= 16 CTAs) whose CTAs exchange through
    // distributed shared memory; otherwise up to one CTA per SM exchanging through global mailboxes. As few nodes
    // per thread as possible either way; node state in shared memory when it fits.
    const int64_t N = std::max<int64_t>(e->n_nodes, 1);
    int k_smem_max = 0;
    while (k_smem_max < kMaxK && seq_smem(e->S, e->D, k_smem_max + 1, 1).total <= q->smem_optin) k_smem_max++;
    const char* env_ctas = getenv("FGS_SEQ_CTAS");
    const char* env_cluster = getenv("FGS_SEQ_CLUSTER");
    const bool cluster_off = env_cluster && env_cluster[0] == '0';
    const bool state_global = getenv("FGS_SEQ_STATE_GLOBAL") && getenv("FGS_SEQ_STATE_GLOBAL")[0] == '1';
    int G = 0, K = 0, cluster = 0;
    if (!cluster_off && !state_global && !env_ctas) {
        for (int k = 1; k <= k_smem_max && !G; ++k)
            for (int g = 1; g <= kMaxCluster; g *= 2)
                if ((int64_t)g * kSeqThreads * k >= N) {
                    G = g;
                    K = k;
                    cluster = g > 1;
                    break;
                }
    }
    if (cluster) { // can a cluster of that size be resident at all? (16 is above the portable limit of 8)
        const SeqSmem Lc = seq_smem(e->S, e->D, K, 1);
        const void* kc = (const void*)seq_kernel<true, true>;
        cudaError_t ce = cudaFuncSetAttribute(kc, cudaFuncAttributeMaxDynamicSharedMemorySize, Lc.total);
        if (ce == cudaSuccess && G > 8) ce = cudaFuncSetAttribute(kc, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
        int n_clusters = 0;
        if (ce == cudaSuccess) {
            cudaLaunchConfig_t lc;
            std::memset(&lc, 0, sizeof(lc));
            lc.gridDim = dim3((unsigned)G);
            lc.blockDim = dim3(kSeqBlock);
            lc.dynamicSmemBytes = (size_t)Lc.total;
            cudaLaunchAttribute at[1];
            at[0].id = cudaLaunchAttributeClusterDimension;
            at[0].val.clusterDim.x = (unsigned)G;
            at[0].val.clusterDim.y = 1;
            at[0].val.clusterDim.z = 1;
            lc.attrs = at;
            lc.numAttrs = 1;
            ce = cudaOccupancyMaxActiveClusters(&n_clusters, kc, &lc);
        }
        if (ce != cudaSuccess || n_clusters < 1) {
            cudaGetLastError();
            G = 0;
            cluster = 0;
        }
    }
    if (!G) {
        G = (int)std::min<int64_t>(q->sm_count, (N + kSeqThreads - 1) / kSeqThreads);
        if (env_ctas) {
            const int v = atoi(env_ctas);
            if (v >= 1) G = std::min(v, q->sm_count);
        }
        const int64_t K64 = (N + (int64_t)G * kSeqThreads - 1) / ((int64_t)G * kSeqThreads);
        if (K64 > kMaxK) return e->fail(FGS_E_TOO_LARGE, "fgs_schedule_sequential: more than 32 nodes per thread (" + std::to_string(N) + " nodes)");
        K = (int)K64;
    }
    const int state_smem = (!state_global && K <= k_smem_max) ? 1 : 0;
    const SeqSmem L = seq_smem(e->S, e->D, K, state_smem);
    const void* kfn = cluster ? (const void*)seq_kernel<true, true>
                              : (state_smem ? (const void*)seq_kernel<true, false> : (const void*)seq_kernel<false, false>);
    FGS_CUDA(e, cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, L.total));

    const size_t a_bytes = sizeof(u64) * 2 * (size_t)G * kWordsA, b_bytes = sizeof(u64) * 2 * (size_t)G * kWordsB;
    const bool want_prof = getenv("FGS_SEQ_PROFILE") && getenv("FGS_SEQ_PROFILE")[0] == '1';
    const size_t prof_bytes = sizeof(u64) * 16 * (size_t)G;
    const size_t need = a_bytes + b_bytes + 64 + prof_bytes;
    if (need > q->mail_bytes) {
        if (q->d_mail) cudaFree(q->d_mail);
        q->d_mail = nullptr;
        q->mail_bytes = 0;
        FGS_CUDA(e, cudaMalloc(&q->d_mail, need));
        q->mail_bytes = need;
    }
    FGS_CUDA(e, cudaMemsetAsync(q->d_mail, 0, need, e->stream));
    int* d_status = (int*)((char*)q->d_mail + a_bytes + b_bytes);
    const int none = 0x7fffffff;
    FGS_CUDA(e, cudaMemcpyAsync(d_status, &none, sizeof(int), cudaMemcpyHostToDevice, e->stream));

    EvalParams ep;
    e->fill_eval_params(&ep, nullptr);
    SeqParams sp;
    std::memset(&sp, 0, sizeof(sp));
    e->fill_derive_params(&sp.dv);
    sp.pods = (const i64*)q->d_pods;
    sp.extra = (const i64*)q->d_pods + (int64_t)pd_cols(e->S) * pp;
    sp.out = (i64*)q->d_out;
    sp.mail_a = (u64*)q->d_mail;
    sp.mail_b = (u64*)((char*)q->d_mail + a_bytes);
    sp.status = d_status;
    sp.prof = want_prof ? (u64*)((char*)q->d_mail + a_bytes + b_bytes + 64) : nullptr;
    sp.n_pods = P;
    sp.p_pad = pp;
    sp.K = K;
    sp.state_smem = state_smem;
    sp.filter_mask = ep.filter_mask;
    sp.score_mask = ep.score_mask;
    sp.flex_filter_first = ep.flex_filter_first;
    sp.w_flex = ep.w_flex;
    sp.w_fit = ep.w_fit;
    sp.w_alloc = ep.w_alloc;
    sp.fit_strategy = ep.fit_strategy;
    sp.fit_n_res = ep.fit_n_res;
    for (int k = 0; k < FGS_MAX_SCORE_RES; ++k) {
        sp.fit_res[k] = ep.fit_res[k];
        sp.fit_w[k] = ep.fit_w[k];
    }
    sp.fit_ignored = ep.fit_ignored;
    sp.seed = e->seed;
    sp.key_shift = ep.key_shift;
    sp.reserve_flex = (opts && opts->reserve_flex) ? 1 : 0;
    if (qt) {
        char* dq = (char*)q->d_quota;
        sp.q_has_eq = (const uint8_t*)(dq + q_off[0]);
        sp.q_min = (const i64*)(dq + q_off[1]);
        sp.q_max = (const i64*)(dq + q_off[2]);
        sp.q_used = (i64*)(dq + q_off[3]);
        sp.q_used_present = (uint32_t*)(dq + q_off[4]);
        sp.n_ns = qt->n_ns;
    }

    FGS_CUDA(e, cudaEventRecord(q->ev0, e->stream));
    void* args[] = {(void*)&sp};
    if (cluster) {
        // one cluster = the whole grid: its CTAs are co-scheduled by construction
        cudaLaunchConfig_t lc;
        std::memset(&lc, 0, sizeof(lc));
        lc.gridDim = dim3((unsigned)G);
        lc.blockDim = dim3(kSeqBlock);
        lc.dynamicSmemBytes = (size_t)L.total;
        lc.stream = e->stream;
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeClusterDimension;
        at[0].val.clusterDim.x = (unsigned)G;
        at[0].val.clusterDim.y = 1;
        at[0].val.clusterDim.z = 1;
        lc.attrs = at;
        lc.numAttrs = 1;
        FGS_CUDA(e, cudaLaunchKernelExC(&lc, kfn, args));
    } else {
        // cooperative launch: all G CTAs must be resident at once (they wait on each other's mailbox words)
        FGS_CUDA(e, cudaLaunchCooperativeKernel(kfn, dim3((unsigned)G), dim3(kSeqBlock), args, (size_t)L.total, e->stream));
    }
    e->launches++;
    FGS_CUDA(e, cudaEventRecord(q->ev1, e->stream));
    FGS_CUDA(e, cudaMemcpyAsync(q->h_out, q->d_out, sizeof(int64_t) * SO_COLS * (size_t)pp, cudaMemcpyDeviceToHost, e->stream));
    int status = none;
    FGS_CUDA(e, cudaMemcpyAsync(&status, d_status, sizeof(int), cudaMemcpyDeviceToHost, e->stream));
    if (qt && q_total) // ElasticQuotaInfo.Used after every CapacityScheduling.Reserve of the queue
        FGS_CUDA(e, cudaMemcpyAsync((char*)q->h_quota + q_off[3], (char*)q->d_quota + q_off[3], q_total - q_off[3], cudaMemcpyDeviceToHost, e->stream));
    FGS_CUDA(e, cudaStreamSynchronize(e->stream));
    if (qt) {
        const size_t NS = (size_t)qt->n_ns, W = (size_t)(2 + e->S);
        if (out->quota_used) std::memcpy(out->quota_used, (char*)q->h_quota + q_off[3], sizeof(int64_t) * NS * W);
        if (out->quota_used_present) std::memcpy(out->quota_used_present, (char*)q->h_quota + q_off[4], sizeof(uint32_t) * NS);
    }
    cudaEventElapsedTime(&e->tm.seq_ms, q->ev0, q->ev1);
    e->tm.seq_ctas = cluster ? -G : G; // negative: one thread-block cluster of that many CTAs
    if (want_prof) { // clock cycles per scheduling cycle and section, CTA 0 and the mean over CTAs
        std::vector<u64> hp((size_t)G * 16);
        FGS_CUDA(e, cudaMemcpy(hp.data(), sp.prof, prof_bytes, cudaMemcpyDeviceToHost));
        static const char* names[16] = {"filter", "fit score", "wait A", "normalize", "wait B", "commit", "", "",
                                        "comm: slowest A", "comm: exchange A", "comm: slowest B", "comm: exchange B", "", "", "", ""};
        for (int k = 0; k < 12; ++k) {
            if (!names[k][0]) continue;
            double sum = 0;
            for (int c = 0; c < G; ++c) sum += (double)hp[(size_t)c * 16 + k];
            fprintf(stderr, "[fgs seq profile] %-17s cta0 %8.1f  mean %8.1f clocks/cycle\n", names[k], (double)hp[k] / (double)P, sum / G / (double)P);
        }
    }
    const int64_t* h = (const int64_t*)q->h_out;
    for (int64_t p = 0; p < P; ++p) {
        if (out->feasible_count) out->feasible_count[p] = (int32_t)h[(int64_t)SO_COUNT * pp + p];
        if (out->selected_node) out->selected_node[p] = (int32_t)h[(int64_t)SO_NODE * pp + p];
        if (out->selected_score) out->selected_score[p] = h[(int64_t)SO_SCORE * pp + p];
        if (out->pod_err) out->pod_err[p] = (uint32_t)h[(int64_t)SO_ERR * pp + p];
        if (out->device_index) out->device_index[p] = (int32_t)h[(int64_t)SO_DEVICE * pp + p];
        if (out->committed) out->committed[p] = (uint8_t)h[(int64_t)SO_COMMITTED * pp + p];
        if (out->quota_code) out->quota_code[p] = (uint8_t)h[(int64_t)SO_QUOTA * pp + p];
    }
    // the derived table changed under the batched path's feet: redo its node-side analysis
    rc = e->analyze_nodes_now();
    if (rc) return rc;
    if (status != none)
        return e->fail(FGS_E_RANGE, "fgs_schedule_sequential: committing pod " + std::to_string(status) +
                                        " took a node aggregate outside |v| < 2^53; results from that pod on are not the reference's");
    return FGS_OK;
}

extern "C" int fgs_reserve(fgs_engine* eh, const fgs_pod_table* t, const int32_t* node, int32_t* device_index) {
    if (!eh) return FGS_E_INVALID;
    Engine* e = (Engine*)eh;
    int rc = seq_check_inputs(e, t, "fgs_reserve");
    if (rc) return rc;
    if (!node || !device_index) return e->fail(FGS_E_INVALID, "fgs_reserve: null node / device_index array");
    cudaSetDevice(e->device);
    FGS_CUDA(e, cudaStreamSynchronize(e->stream));
    const int64_t P = t->n_pods;
    if (P == 0) return FGS_OK;
    const int64_t pp = round_up64(P, kSeqTileP);
    rc = seq_reserve_buffers(e, pp);
    if (rc) return rc;
    SeqState* q = e->seq;
    // only the flags and the fractional limit are read
    int64_t* h = (int64_t*)q->h_pods;
    for (int64_t i = 0; i < P; ++i) {
        h[(int64_t)PD_META * pp + i] = (int64_t)(uint64_t)t->flags[i];
        h[(int64_t)PD_MEM_LIM * pp + i] = t->flex_mem_limit[i];
    }
    int32_t* hn = (int32_t*)q->h_out;
    std::memcpy(hn, node, sizeof(int32_t) * (size_t)P);
    FGS_CUDA(e, cudaMemcpyAsync((int64_t*)q->d_pods + (int64_t)PD_META * pp, h + (int64_t)PD_META * pp, sizeof(int64_t) * (size_t)P,
                                cudaMemcpyHostToDevice, e->stream));
    FGS_CUDA(e, cudaMemcpyAsync((int64_t*)q->d_pods + (int64_t)PD_MEM_LIM * pp, h + (int64_t)PD_MEM_LIM * pp, sizeof(int64_t) * (size_t)P,
                                cudaMemcpyHostToDevice, e->stream));
    int* d_node = (int*)q->d_out;
    int* d_dev = d_node + pp;
    FGS_CUDA(e, cudaMemcpyAsync(d_node, hn, sizeof(int32_t) * (size_t)P, cudaMemcpyHostToDevice, e->stream));
    reserve_kernel<<<(unsigned)((P + 255) / 256), 256, 0, e->stream>>>((const i64*)e->d_raw, (const i64*)e->d_nd, e->n_pad, e->n_nodes, e->S, e->D, (const i64*)q->d_pods, pp, P,
                                                                      d_node, d_dev);
    e->launches++;
    FGS_CUDA(e, cudaGetLastError());
    FGS_CUDA(e, cudaMemcpyAsync(hn + pp, d_dev, sizeof(int32_t) * (size_t)P, cudaMemcpyDeviceToHost, e->stream));
    FGS_CUDA(e, cudaStreamSynchronize(e->stream));
    std::memcpy(device_index, hn + pp, sizeof(int32_t) * (size_t)P);
    return FGS_OK;
}
