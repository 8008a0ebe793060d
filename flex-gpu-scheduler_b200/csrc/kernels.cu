// kernels.cu — sm_100a kernels of the batched Filter+Score path.
//
//   derive_node_columns   pod-independent per-node work the reference redoes per (pod,node):
//                         NewGPUNode/constructGPUs (pkg/flexgpu/gpu_node.go:30-120), GPUScore/MemScore
//                         (:179-199), the step-4/5/7 thresholds of FlexGPU.Filter (flex_gpu.go:85-116),
//                         Allocatable raw score (pkg/noderesources/allocatable.go:120-143).
//   eval_kernel<REDUCE>   pass A: Filter for every (pod,node) + feasibility bitmap (warp ballot) +
//                         per-pod reductions NormalizeScore needs (max flex score, hi/lo allocatable).
//   eval_kernel<EMIT>     pass B: recompute Filter, Score, NormalizeScore, ×weight, Σ plugins, write
//                         the int64 score rows, reduce the per-pod selectHost key.
//
// Mapping: one CTA = 512 consecutive nodes (two per thread, held in registers) × a contiguous range
// of pods; pods stream through shared memory in tiles of 64, fetched with cp.async.bulk (TMA bulk
// copy) into a two-stage mbarrier ring. All arithmetic is integer; the only floats are quotient
// ESTIMATES that an exact int64 multiply-compare then corrects (see qfix).
//
// Integer code, HBM-write bound: no tensor cores on purpose (north_star).
#include <cuda_runtime.h>

#include <cstdint>

#include "kernels.cuh"
#include "derive.cuh"

namespace fgs {

namespace {

typedef long long i64;
typedef unsigned long long u64;

constexpr i64 kI64Min = (i64)0x8000000000000000ull;
constexpr i64 kI64Max = (i64)0x7fffffffffffffffull;

// ---- small PTX wrappers -------------------------------------------------------------------------

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}\n"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    while (!mbar_try_wait(bar, parity)) {
    }
}
// TMA bulk copy global → shared, completion counted on an mbarrier (SASS: UBLKCP)
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(dst)),
                 "l"(__cvta_generic_to_global(src)), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void fence_mbar_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }

// streaming 64-bit store: the score matrix is written once and never re-read by these kernels
__device__ __forceinline__ void st_stream(i64* p, i64 v) { __stcs(p, v); }

// ---- exact small-quotient division --------------------------------------------------------------
// round-to-nearest of a float in (-2^22, 2^22) without the conversion pipe
__device__ __forceinline__ int rn_small(float x) { return __float_as_int(x + 12582912.0f) - 0x4B400000; }

// floor(x / d) for exact integers x >= 0, d > 0, given a float estimate within 0.5 of x/d and a
// quotient below 2^22. k = round(est) is either the quotient or one above it; the sign of
// x − k·d (exact, 64-bit) decides. This is what keeps the scores bit-exact with Go's int64 '/'.
__device__ __forceinline__ i64 qfix(i64 x, i64 d, float est) {
    const int k = rn_small(est);
    const i64 r = x - (i64)((u64)(uint32_t)k * (u64)d);
    return (i64)k + (r >> 63);
}

__device__ __forceinline__ float u64_to_float(u64 a) {
    return fmaf(__uint2float_rn((uint32_t)(a >> 32)), 4294967296.0f, __uint2float_rn((uint32_t)a));
}

// 64-bit signed max over the warp with two 32-bit REDUX ops
__device__ __forceinline__ i64 warp_max_i64(i64 v) {
    const int hi = (int)(v >> 32);
    const uint32_t lo = (uint32_t)v;
    const int mhi = __reduce_max_sync(0xffffffffu, hi);
    const uint32_t mlo = __reduce_max_sync(0xffffffffu, hi == mhi ? lo : 0u);
    return (i64)(((u64)(uint32_t)mhi << 32) | mlo);
}

__device__ __forceinline__ i64 wmul(i64 a, i64 b) { return (i64)((u64)a * (u64)b); }
__device__ __forceinline__ i64 wadd(i64 a, i64 b) { return (i64)((u64)a + (u64)b); }
__device__ __forceinline__ i64 wsub(i64 a, i64 b) { return (i64)((u64)a - (u64)b); }

// =================================================================================================
// derive_node_columns
// =================================================================================================

__global__ void __launch_bounds__(256) derive_node_columns(const __grid_constant__ DeriveParams prm) {
    const i64 n = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= prm.n_pad) return;
    derive_one_node(prm, n);
}

// =================================================================================================
// eval_kernel
// =================================================================================================

// per-pod values every lane reads (warp-uniform): 128 bytes, read with broadcast LDS.128
struct __align__(16) PodU {
    i64 lim;        // the flex limit that applies to the pod's kind
    i64 rc, rm, re; // Fit.Filter request (fit.go:162-180)
    i64 nzc, nzm;   // scoring request, non-zero defaults applied (resource_allocation.go:119-143)
    i64 fmax;       // EMIT: max(0, max feasible FlexGPU score) (normalize_score.go:27-32)
    i64 alo;        // EMIT: lowest Allocatable raw score over feasible nodes
    i64 arange;     // EMIT: highest − lowest (allocatable.go:160)
    float fnzc, fnzm;
    float finv;     // EMIT: 1/fmax
    float ainv;     // EMIT: 100/arange
    uint32_t flags; // bits 0-1: kind (0 none, 1 whole gpu, 2 fractional, 3 both); bit 2: request is empty
    uint32_t smask; // scalar keys Fit.Filter checks (present and not ignored)
    int32_t count;  // EMIT: number of feasible nodes over all shards
    uint32_t hash;  // EMIT: per-pod seed of the tie-break hash
    uint32_t valid; // pod index < n_pods
    uint32_t _pad[3];
};
static_assert(sizeof(PodU) == 128, "PodU layout");

struct SmemLayout {
    // offsets in bytes from the dynamic shared memory base
    int bars, node_cols, pod_raw, podu, pod_scalars, stage_a, stage_b, total;
    int n_node_cols, n_raw_cols;
};

__host__ __device__ inline SmemLayout smem_layout(int S) {
    SmemLayout L;
    L.n_node_cols = 2 * S + 1;                   // free_scalar[S], alloc_scalar[S], alloc_eph
    L.n_raw_cols = pd_cols(S) + 4;               // + flex max, alloc hi, alloc -lo, count (EMIT)
    int off = 0;
    L.bars = off;
    off += 64;
    L.node_cols = off;
    off += L.n_node_cols * kTileN * 8;
    L.pod_raw = off;
    off += 2 * L.n_raw_cols * kTileP * 8;
    L.podu = off;
    off += kTileP * (int)sizeof(PodU);
    L.pod_scalars = off;
    off += S * kTileP * 8;
    L.stage_a = off;                             // REDUCE: int64 wred[3][kTileP][kWarps]; EMIT: u64 wkey[kTileP][kWarps]
    off += 3 * kTileP * kWarps * 8;
    L.stage_b = off;                             // REDUCE: u32 wcnt[kTileP][kWarps] + u32 bm[kTileP][16]; EMIT: u32 werr[kTileP][kWarps]
    off += kTileP * kWarps * 4 + kTileP * 16 * 4;
    L.total = off;
    return L;
}

enum { MODE_REDUCE = 0, MODE_EMIT = 1 };

struct NodeRegs {
    i64 thr_gpu, thr_mem;
    i64 free_cpu, free_mem, free_eph;
    i64 capc, xc, x100c; // cpu: capacity; Least: x = cap − nonzeroRequested, Most: x = nonzeroRequested; 100·x
    i64 capm, xm, x100m;
    i64 gscore, mscore, araw;
    float invc, xqc, invm, xqm; // 100/cap, 100·x/cap
    float gs100f, ms100f;       // 100·score as float
    uint32_t flags;
};

template <int MODE>
__global__ void __launch_bounds__(kThreads, 2) eval_kernel(const __grid_constant__ EvalParams prm) {
    extern __shared__ __align__(128) unsigned char smem[];
    const int S = prm.S;
    const SmemLayout L = smem_layout(S);
    uint64_t* bars = (uint64_t*)(smem + L.bars); // [0],[1]: pod stages, [2]: node columns
    i64* s_node = (i64*)(smem + L.node_cols);
    i64* s_raw = (i64*)(smem + L.pod_raw);
    PodU* s_podu = (PodU*)(smem + L.podu);
    i64* s_pscal = (i64*)(smem + L.pod_scalars);

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const i64 node_tile0 = (i64)blockIdx.x * kTileN;
    const i64 pod_begin = (i64)blockIdx.y * prm.pods_per_cta;
    i64 pod_end = pod_begin + prm.pods_per_cta;
    if (pod_end > prm.p_pad) pod_end = prm.p_pad;
    const int n_tiles = (int)((pod_end - pod_begin + kTileP - 1) / kTileP);
    if (n_tiles <= 0) return;

    const int n_raw_cols = (MODE == MODE_EMIT) ? L.n_raw_cols : pd_cols(S);
    const uint32_t raw_bytes = (uint32_t)(n_raw_cols * kTileP * 8);

    auto issue_pod_tile = [&](int it) { // one elected thread
        const int stage = it & 1;
        const i64 p0 = pod_begin + (i64)it * kTileP;
        i64* dst = s_raw + (size_t)stage * L.n_raw_cols * kTileP;
        mbar_expect_tx(&bars[stage], raw_bytes);
        for (int c = 0; c < pd_cols(S); ++c) bulk_g2s(dst + c * kTileP, prm.pods + (i64)c * prm.p_pad + p0, kTileP * 8, &bars[stage]);
        if (MODE == MODE_EMIT) {
            i64* d2 = dst + pd_cols(S) * kTileP;
            bulk_g2s(d2 + 0 * kTileP, prm.reduce_max + (i64)RR_FLEX_MAX * prm.p_pad + p0, kTileP * 8, &bars[stage]);
            bulk_g2s(d2 + 1 * kTileP, prm.reduce_max + (i64)RR_ALLOC_HI * prm.p_pad + p0, kTileP * 8, &bars[stage]);
            bulk_g2s(d2 + 2 * kTileP, prm.reduce_max + (i64)RR_ALLOC_NEGLO * prm.p_pad + p0, kTileP * 8, &bars[stage]);
            bulk_g2s(d2 + 3 * kTileP, prm.reduce_sum + p0, kTileP * 8, &bars[stage]);
        }
    };

    if (tid == 0) {
        mbar_init(&bars[0], 1);
        mbar_init(&bars[1], 1);
        mbar_init(&bars[2], 1);
        fence_mbar_init();
    }
    __syncthreads();
    if (tid == 0) {
        // node columns that stay in shared memory for the whole CTA
        mbar_expect_tx(&bars[2], (uint32_t)(L.n_node_cols * kTileN * 8));
        for (int s = 0; s < S; ++s) {
            bulk_g2s(s_node + (size_t)s * kTileN, prm.nd + (i64)nd_free_scalar(s) * prm.n_pad + node_tile0, kTileN * 8, &bars[2]);
            bulk_g2s(s_node + (size_t)(S + s) * kTileN, prm.nd + (i64)nd_alloc_scalar(S, s) * prm.n_pad + node_tile0, kTileN * 8, &bars[2]);
        }
        bulk_g2s(s_node + (size_t)(2 * S) * kTileN, prm.nd + (i64)ND_ALLOC_EPH * prm.n_pad + node_tile0, kTileN * 8, &bars[2]);
        issue_pod_tile(0);
        if (n_tiles > 1) issue_pod_tile(1);
    }

    // ---- node tile → registers: thread owns nodes (64·warp + lane) and (+32) of the tile ---------
    NodeRegs nr[kNodesPerThread];
    int nloc[kNodesPerThread];
#pragma unroll
    for (int j = 0; j < kNodesPerThread; ++j) {
        nloc[j] = warp * 64 + j * 32 + lane;
        const i64 n = node_tile0 + nloc[j];
        auto C = [&](int c) { return __ldg(prm.nd + (i64)c * prm.n_pad + n); };
        NodeRegs& r = nr[j];
        r.thr_gpu = C(ND_THR_GPU);
        r.thr_mem = C(ND_THR_MEM);
        r.free_cpu = C(ND_FREE_CPU);
        r.free_mem = C(ND_FREE_MEM);
        r.free_eph = C(ND_FREE_EPH);
        r.flags = (uint32_t)C(ND_FLAGS);
        r.capc = C(ND_CAP_CPU);
        r.capm = C(ND_CAP_MEM);
        const i64 nzrc = C(ND_NZR_CPU), nzrm = C(ND_NZR_MEM);
        r.xc = prm.fit_strategy == 0 ? r.capc - nzrc : nzrc;
        r.xm = prm.fit_strategy == 0 ? r.capm - nzrm : nzrm;
        r.x100c = 100 * r.xc;
        r.x100m = 100 * r.xm;
        // estimates only: double → float once per node per CTA
        r.invc = r.capc > 0 ? (float)(100.0 / (double)r.capc) : 0.f;
        r.invm = r.capm > 0 ? (float)(100.0 / (double)r.capm) : 0.f;
        r.xqc = r.capc > 0 ? (float)(100.0 * (double)r.xc / (double)r.capc) : 0.f;
        r.xqm = r.capm > 0 ? (float)(100.0 * (double)r.xm / (double)r.capm) : 0.f;
        r.gscore = C(ND_GPU_SCORE);
        r.mscore = C(ND_MEM_SCORE);
        r.gs100f = (float)(100.0 * (double)r.gscore);
        r.ms100f = (float)(100.0 * (double)r.mscore);
        r.araw = C(ND_ALLOC_RAW);
    }
    mbar_wait(&bars[2], 0);

    const bool use_flex_filter = prm.filter_mask & FGS_PL_FLEXGPU;
    const bool use_fit_filter = prm.filter_mask & FGS_PL_FIT;
    const bool sc_flex = prm.score_mask & FGS_PL_FLEXGPU;
    const bool sc_fit = prm.score_mask & FGS_PL_FIT;
    const bool sc_alloc = prm.score_mask & FGS_PL_ALLOCATABLE;

    for (int it = 0; it < n_tiles; ++it) {
        const int stage = it & 1;
        const uint32_t parity = (uint32_t)((it >> 1) & 1);
        const i64 tile_p0 = pod_begin + (i64)it * kTileP;
        mbar_wait(&bars[stage], parity);

        // ---- per-pod uniform values ----------------------------------------------------------
        if (tid < kTileP) {
            const i64* raw = s_raw + (size_t)stage * L.n_raw_cols * kTileP;
            auto RC = [&](int c) { return raw[c * kTileP + tid]; };
            PodU u;
            const u64 meta = (u64)RC(PD_META);
            const uint32_t pf = (uint32_t)meta;
            const uint32_t kind = pf & 3u;
            u.flags = kind | ((pf & FGS_PF_REQ_NONE) ? 4u : 0u);
            u.smask = (uint32_t)(meta >> 32) & ~prm.fit_ignored;
            u.lim = kind == 1 ? RC(PD_GPU_LIM) : RC(PD_MEM_LIM);
            u.rc = RC(PD_REQ_CPU);
            u.rm = RC(PD_REQ_MEM);
            u.re = RC(PD_REQ_EPH);
            u.nzc = RC(PD_NZ_CPU);
            u.nzm = RC(PD_NZ_MEM);
            u.fnzc = (float)u.nzc;
            u.fnzm = (float)u.nzm;
            u.valid = (tile_p0 + tid) < prm.n_pods ? 1u : 0u;
            u.fmax = 0;
            u.alo = 0;
            u.arange = 0;
            u.finv = 0.f;
            u.ainv = 0.f;
            u.count = 0;
            u.hash = 0;
            if (MODE == MODE_EMIT) {
                const i64 fm = RC(pd_cols(S) + 0);
                u.fmax = fm > 0 ? fm : 0; // maxCount starts at 0 (normalize_score.go:27)
                u.finv = u.fmax > 0 ? (float)(1.0 / (double)u.fmax) : 0.f;
                const i64 hi = RC(pd_cols(S) + 1), lo = -RC(pd_cols(S) + 2);
                u.count = (int32_t)RC(pd_cols(S) + 3);
                if (u.count > 0) {
                    u.alo = lo;
                    u.arange = wsub(hi, lo);
                    u.ainv = u.arange > 0 ? (float)(100.0 / (double)u.arange) : 0.f;
                }
                uint32_t h = (uint32_t)(prm.seed ^ (prm.seed >> 32)) ^ ((uint32_t)(tile_p0 + tid) * 0x9E3779B1u);
                h ^= h >> 16;
                h *= 0x85EBCA6Bu;
                h ^= h >> 13;
                u.hash = h;
            }
            s_podu[tid] = u;
            for (int s = 0; s < S; ++s) s_pscal[s * kTileP + tid] = RC(PD_FIXED_COLS + s);
        }
        __syncthreads(); // PodU visible; raw stage is free again
        if (tid == 0 && it + 2 < n_tiles) {
            fence_proxy_async();
            issue_pod_tile(it + 2);
        }

        i64* wred = (i64*)(smem + L.stage_a);           // REDUCE [3][kTileP][kWarps]
        u64* wkey = (u64*)(smem + L.stage_a);           // EMIT   [kTileP][kWarps]
        uint32_t* wcnt = (uint32_t*)(smem + L.stage_b); // REDUCE [kTileP][kWarps] (bit31: filter panic); EMIT: error bits
        uint32_t* bm = wcnt + kTileP * kWarps;          // REDUCE [kTileP][16]

        // ---- the pods × nodes loop ----------------------------------------------------------------
        for (int jp = 0; jp < kTileP; ++jp) {
            const PodU u = s_podu[jp];
            if (!u.valid) break; // padding pods sit at the end of the last tile
            const uint32_t kind = u.flags & 3u;
            const bool req_none = u.flags & 4u;
            const i64 p = tile_p0 + jp;

            bool feas[kNodesPerThread];
            bool fpanic = false;
#pragma unroll
            for (int j = 0; j < kNodesPerThread; ++j) {
                const NodeRegs& r = nr[j];
                // FlexGPU.Filter, flex_gpu.go:41-119: kind none → Success; both → unresolvable;
                // else steps 3-5,7 folded into one threshold per kind (derive_node_columns).
                bool flex_ok = true;
                if (use_flex_filter) {
                    if (kind == 3) flex_ok = false;
                    else if (kind == 1) flex_ok = u.lim <= r.thr_gpu;
                    else if (kind == 2) flex_ok = u.lim <= r.thr_mem;
                }
                // Fit.Filter, fit.go:256-329
                bool fit_ok = true;
                if (use_fit_filter) {
                    fit_ok = r.flags & NDF_PODS_OK;
                    if (!req_none) {
                        fit_ok = fit_ok && u.rc <= r.free_cpu && u.rm <= r.free_mem && u.re <= r.free_eph;
                        uint32_t m = u.smask;
                        while (m) {
                            const int s = __ffs(m) - 1;
                            m &= m - 1;
                            if (s_pscal[s * kTileP + jp] > s_node[(size_t)s * kTileN + nloc[j]]) fit_ok = false;
                        }
                    }
                }
                const bool valid = r.flags & FGS_NF_VALID;
                // the reference panics inside NewGPUNode (flex_gpu.go:101) when FlexGPU.Filter gets
                // that far on such a node: report instead of inventing a verdict
                const bool reach = use_flex_filter && (kind == 1 || kind == 2) && flex_ok && (r.flags & FGS_NF_REF_PANIC) &&
                                   (prm.flex_filter_first || fit_ok) && valid;
                fpanic = fpanic || reach;
                feas[j] = flex_ok && fit_ok && valid && !reach;
                if (MODE == MODE_EMIT && prm.feas_from_bitmap) { // node sampling kept a subset of the feasible nodes
                    const uint32_t word = __ldg(prm.bitmap + p * prm.words_per_row + (node_tile0 >> 5) + warp * kNodesPerThread + j);
                    feas[j] = feas[j] && ((word >> lane) & 1u);
                }
            }

            if (MODE == MODE_REDUCE) {
                const uint32_t b0 = __ballot_sync(0xffffffffu, feas[0]);
                const uint32_t b1 = __ballot_sync(0xffffffffu, feas[1]);
                const uint32_t anyp = __ballot_sync(0xffffffffu, fpanic);
                i64 vflex = kI64Min, vhi = kI64Min, vnlo = kI64Min;
#pragma unroll
                for (int j = 0; j < kNodesPerThread; ++j) {
                    if (feas[j]) {
                        // a gpu limit wins over a memory limit (flex_gpu.go:157-159)
                        const i64 s = (kind & 1u) ? nr[j].gscore : (kind == 2 ? nr[j].mscore : 0);
                        vflex = s > vflex ? s : vflex;
                        vhi = nr[j].araw > vhi ? nr[j].araw : vhi;
                        const i64 nl = -nr[j].araw;
                        vnlo = nl > vnlo ? nl : vnlo;
                    }
                }
                if (sc_flex) vflex = warp_max_i64(vflex);
                if (sc_alloc) {
                    vhi = warp_max_i64(vhi);
                    vnlo = warp_max_i64(vnlo);
                }
                if (lane == 0) {
                    wred[(0 * kTileP + jp) * kWarps + warp] = vflex;
                    wred[(1 * kTileP + jp) * kWarps + warp] = vhi;
                    wred[(2 * kTileP + jp) * kWarps + warp] = vnlo;
                    wcnt[jp * kWarps + warp] = (uint32_t)(__popc(b0) + __popc(b1)) | (anyp ? 0x80000000u : 0u);
                    bm[jp * 16 + warp * 2 + 0] = b0;
                    bm[jp * 16 + warp * 2 + 1] = b1;
                }
            } else {
                // ---- Score / NormalizeScore / weight / sum ----------------------------------------
                i64 total[kNodesPerThread];
                u64 key = 0; // 0 = no candidate (valid keys are > 0: see key packing)
                bool e_range = false, e_panic = false;
                const bool scored = u.count >= 2; // scoring runs only then (generic_scheduler.go:112-127)
#pragma unroll
                for (int j = 0; j < kNodesPerThread; ++j) {
                    const NodeRegs& r = nr[j];
                    i64 t = 0;
                    // FlexGPU.Score (flex_gpu.go:142-166) + DefaultNormalizeScore reverse (normalize_score.go:26-54)
                    if (sc_flex && (prm.emit_mask & FGS_PL_FLEXGPU)) {
                        const i64 s = (kind & 1u) ? r.gscore : (kind == 2 ? r.mscore : 0);
                        i64 v;
                        if (prm.emit_raw) {
                            v = s;
                        } else if (u.fmax == 0) {
                            v = 100;
                        } else if (s >= 0) {
                            const float sf = (kind & 1u) ? r.gs100f : r.ms100f;
                            v = 100 - qfix(100 * s, u.fmax, sf * u.finv);
                        } else {
                            v = 100 - wmul(100, s) / u.fmax; // Go '/' truncates toward zero, as does C++
                        }
                        if (!prm.emit_raw) {
                            if (feas[j] && (v > 100 || v < 0)) e_range = true; // runtime/framework.go:920-923
                            v = wmul(v, prm.w_flex);
                        }
                        t = wadd(t, v);
                        // Score calls NewGPUNode first for every pod kind (flex_gpu.go:148)
                        if (feas[j] && (r.flags & FGS_NF_REF_PANIC)) e_panic = true;
                    } else if (sc_flex && feas[j] && (r.flags & FGS_NF_REF_PANIC)) {
                        e_panic = true;
                    }
                    // Fit.Score: Least/MostAllocated (least_allocated.go:29-57, most_allocated.go:29-61)
                    if (sc_fit && (prm.emit_mask & FGS_PL_FIT)) {
                        i64 v;
                        if (prm.fit_fast && r.capc > 0 && r.capm > 0) {
                            i64 scpu, smem_;
                            if (prm.fit_strategy == 0) {
                                // (cap − (nzr + r))·100/cap = (x − r)·100/cap, 0 when r > x
                                scpu = qfix(r.x100c - 100 * u.nzc, r.capc, fmaf(-u.fnzc, r.invc, r.xqc));
                                smem_ = qfix(r.x100m - 100 * u.nzm, r.capm, fmaf(-u.fnzm, r.invm, r.xqm));
                                if (u.nzc > r.xc) scpu = 0;
                                if (u.nzm > r.xm) smem_ = 0;
                            } else {
                                // min(nzr + r, cap)·100/cap
                                scpu = qfix(r.x100c + 100 * u.nzc, r.capc, fmaf(u.fnzc, r.invc, r.xqc));
                                smem_ = qfix(r.x100m + 100 * u.nzm, r.capm, fmaf(u.fnzm, r.invm, r.xqm));
                                if (u.nzc > r.capc - r.xc) scpu = 100;
                                if (u.nzm > r.capm - r.xm) smem_ = 100;
                            }
                            // Σ score·w / Σ w with w in [1,100]: small enough for an exact float floor
                            const int num = (int)scpu * prm.fit_w_cpu + (int)smem_ * prm.fit_w_mem;
                            const float qf = __fadd_rd(((float)num + 0.5f) * prm.fit_inv_wsum, 8388608.0f);
                            v = (i64)(__float_as_int(qf) & 0x7fffff);
                        } else {
                            // generic path: any resource list, excluded (capacity 0) resources, exact '/'
                            i64 node_score = 0, weight_sum = 0;
                            for (int k = 0; k < prm.fit_n_res; ++k) {
                                const int res = prm.fit_res[k];
                                i64 alloc = 0, req = 0;
                                const i64 nzrc = prm.fit_strategy == 0 ? r.capc - r.xc : r.xc;
                                const i64 nzrm = prm.fit_strategy == 0 ? r.capm - r.xm : r.xm;
                                if (res == FGS_RES_CPU) {
                                    alloc = r.capc;
                                    req = nzrc + u.nzc;
                                } else if (res == FGS_RES_MEM) {
                                    alloc = r.capm;
                                    req = nzrm + u.nzm;
                                } else if (res == FGS_RES_EPH) {
                                    alloc = s_node[(size_t)(2 * S) * kTileN + nloc[j]];
                                    req = (alloc - r.free_eph) + u.re;
                                } else if (res >= FGS_RES_SCALAR0 && res < FGS_RES_SCALAR0 + S) {
                                    const int s = res - FGS_RES_SCALAR0;
                                    const i64 pr = s_pscal[s * kTileP + jp];
                                    if (pr != 0) { // resource_allocation.go:93-97
                                        alloc = s_node[(size_t)(S + s) * kTileN + nloc[j]];
                                        req = (alloc - s_node[(size_t)s * kTileN + nloc[j]]) + pr;
                                    }
                                }
                                if (alloc == 0) continue; // resource_allocation.go:64-67
                                i64 rs;
                                if (prm.fit_strategy == 0) rs = req > alloc ? 0 : wmul(alloc - req, 100) / alloc;
                                else rs = wmul(req > alloc ? alloc : req, 100) / alloc;
                                node_score = wadd(node_score, wmul(rs, prm.fit_w[k]));
                                weight_sum = wadd(weight_sum, prm.fit_w[k]);
                            }
                            v = weight_sum == 0 ? 0 : node_score / weight_sum;
                        }
                        if (!prm.emit_raw) {
                            if (feas[j] && (v > 100 || v < 0)) e_range = true;
                            v = wmul(v, prm.w_fit);
                        }
                        t = wadd(t, v);
                    }
                    // Allocatable.Score + NormalizeScore (allocatable.go:61-72,146-171)
                    if (sc_alloc && (prm.emit_mask & FGS_PL_ALLOCATABLE)) {
                        i64 v;
                        if (prm.emit_raw) {
                            v = r.araw;
                        } else if (u.arange == 0) {
                            v = 0;
                        } else {
                            const i64 a = wsub(r.araw, u.alo);
                            if (a >= 0 && u.arange > 0 && u.arange < ((i64)1 << 56)) v = qfix(100 * a, u.arange, u64_to_float((u64)a) * u.ainv);
                            else v = wmul(a, 100) / u.arange;
                        }
                        if (!prm.emit_raw) {
                            if (feas[j] && (v > 100 || v < 0)) e_range = true;
                            v = wmul(v, prm.w_alloc);
                        }
                        t = wadd(t, v);
                    }
                    total[j] = t;
                    if (feas[j]) { // selectHost candidate (generic_scheduler.go:146-167)
                        const uint32_t ng = (uint32_t)(prm.node_base + node_tile0 + nloc[j]);
                        u64 k;
                        if (prm.key_shift == 40) {
                            uint32_t h = (u.hash ^ ng) * 0x9E3779B1u;
                            h ^= h >> 15;
                            k = ((u64)t << 40) | ((u64)(h & 0xffffu) << 24) | ng;
                        } else {
                            k = ((u64)t << 24) | ng;
                        }
                        k += 1; // keep 0 free for "no candidate"
                        key = k > key ? k : key;
                    }
                }
                // score rows: each warp writes two 256-byte runs per pod
                i64* row = (i64*)prm.scores + p * prm.score_stride + node_tile0;
#pragma unroll
                for (int j = 0; j < kNodesPerThread; ++j)
                    if (node_tile0 + nloc[j] < prm.n_nodes) st_stream(row + nloc[j], total[j]);
                const i64 wk = warp_max_i64((i64)key);
                const uint32_t eb = __ballot_sync(0xffffffffu, e_range) ? 1u : 0u;
                const uint32_t ep = __ballot_sync(0xffffffffu, e_panic) ? 2u : 0u;
                if (lane == 0) {
                    wkey[jp * kWarps + warp] = (u64)wk;
                    wcnt[jp * kWarps + warp] = scored ? (eb | ep) : 0u;
                }
            }
        }
        __syncthreads();

        // ---- flush the tile's per-warp partials ---------------------------------------------------
        if (MODE == MODE_REDUCE) {
            for (int idx = tid; idx < 3 * kTileP; idx += kThreads) {
                const int rrow = idx / kTileP, jp = idx % kTileP;
                const i64 p = tile_p0 + jp;
                if (p >= prm.n_pods) continue;
                i64 v = kI64Min;
                for (int w = 0; w < kWarps; ++w) {
                    const i64 x = wred[(rrow * kTileP + jp) * kWarps + w];
                    v = x > v ? x : v;
                }
                if (v != kI64Min) atomicMax((i64*)prm.reduce_max + (i64)rrow * prm.p_pad + p, v);
            }
            if (tid < kTileP) {
                const i64 p = tile_p0 + tid;
                if (p < prm.n_pods) {
                    uint32_t c = 0, pn = 0;
                    for (int w = 0; w < kWarps; ++w) {
                        const uint32_t x = wcnt[tid * kWarps + w];
                        c += x & 0x7fffffffu;
                        pn |= x >> 31;
                    }
                    if (c) atomicAdd((u64*)prm.reduce_sum + p, (u64)c);
                    if (pn) atomicMax((i64*)prm.reduce_max + (i64)RR_FILTER_PANIC * prm.p_pad + p, (i64)1);
                }
            }
            // bitmap: 16 words (64 bytes) per pod row per CTA, written as whole sectors
            const i64 word0 = node_tile0 / 32;
            for (int idx = tid; idx < kTileP * 16; idx += kThreads) {
                const int jp = idx >> 4, w = idx & 15;
                const i64 p = tile_p0 + jp;
                if (p < prm.n_pods && word0 + w < prm.words_per_row) prm.bitmap[p * prm.words_per_row + word0 + w] = bm[idx];
            }
        } else {
            if (tid < kTileP) {
                const i64 p = tile_p0 + tid;
                if (p < prm.n_pods) {
                    u64 k = 0;
                    uint32_t eb = 0;
                    for (int w = 0; w < kWarps; ++w) {
                        const u64 x = wkey[tid * kWarps + w];
                        k = x > k ? x : k;
                        eb |= wcnt[tid * kWarps + w];
                    }
                    if (k) atomicMax((i64*)prm.select_key + p, (i64)k);
                    if (eb & 1u) atomicMax((i64*)prm.select_key + prm.p_pad + p, (i64)1);
                    if (eb & 2u) atomicMax((i64*)prm.select_key + 2 * prm.p_pad + p, (i64)1);
                }
            }
        }
        __syncthreads(); // staging and PodU are rewritten by the next tile
    }
}

} // namespace

// =================================================================================================
// host launchers
// =================================================================================================

size_t eval_smem_bytes(int S) { return (size_t)smem_layout(S).total; }

void launch_derive(const DeriveParams& p, void* stream) {
    const int threads = 256;
    const int blocks = (int)((p.n_pad + threads - 1) / threads);
    derive_node_columns<<<blocks, threads, 0, (cudaStream_t)stream>>>(p);
}

static void eval_grid(const EvalParams& p, dim3* grid) {
    const int node_tiles = (int)(p.n_pad / kTileN);
    const int pod_chunks = (int)((p.p_pad + p.pods_per_cta - 1) / p.pods_per_cta);
    *grid = dim3((unsigned)node_tiles, (unsigned)pod_chunks, 1);
}

cudaError_t eval_setup() {
    cudaError_t e = cudaFuncSetAttribute(eval_kernel<MODE_REDUCE>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    if (e != cudaSuccess) return e;
    return cudaFuncSetAttribute(eval_kernel<MODE_EMIT>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
}

namespace {
__global__ void init_perpod(i64* perpod, i64 p_pad) {
    const i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= p_pad) return;
    perpod[(i64)RR_FLEX_MAX * p_pad + i] = kI64Min;
    perpod[(i64)RR_ALLOC_HI * p_pad + i] = kI64Min;
    perpod[(i64)RR_ALLOC_NEGLO * p_pad + i] = kI64Min;
    for (int r = RR_FILTER_PANIC; r < FGS_REDUCE_ROWS + 1 + 3; ++r) perpod[(i64)r * p_pad + i] = 0;
}
} // namespace

namespace {
// Per-pod epilogue: decode the selectHost key and derive the cycle-error bits.
//  * REF_PANIC: FlexGPU.Filter reached NewGPUNode on a node where it panics, or FlexGPU.Score would (it calls
//    NewGPUNode on every feasible node, flex_gpu.go:148) — scoring only runs with >= 2 feasible nodes
//    (generic_scheduler.go:112-127).
//  * SCORE_RANGE: a feasible node with a negative raw score s makes 100 − (100·s)/max exceed 100 as soon as
//    100·|s| >= max (Go's '/' truncates toward zero) — runtime/framework.go:920-923 aborts the cycle. It is the
//    only way a normalised score on this path leaves [0,100].
__global__ void finalize_pods(const i64* __restrict__ perpod, i64* __restrict__ compact, i64 n_pods, i64 p_pad, int key_shift, int emitted,
                              int sc_flex) {
    const i64 p = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n_pods) return;
    auto R = [&](int row) { return perpod[(i64)row * p_pad + p]; };
    const i64 count = R(FGS_REDUCE_ROWS);
    const i64 key = R(FGS_REDUCE_ROWS + 1);
    uint32_t err = 0;
    if (R(RR_FILTER_PANIC) > 0) err |= FGS_PODERR_REF_PANIC;
    if (sc_flex && emitted && count >= 2) {
        if (R(RR_SCORE_PANIC) > 0) err |= FGS_PODERR_REF_PANIC;
        const i64 fmax = R(RR_FLEX_MAX), neg = R(RR_NEG_MAG);
        if (fmax > 0 && neg > 0 && 100 * neg >= fmax) err |= FGS_PODERR_SCORE_RANGE;
    }
    if (emitted) { // flags raised inside the int64 emit kernel (same conditions, kept as a cross-check)
        if (R(FGS_REDUCE_ROWS + 2) > 0) err |= FGS_PODERR_SCORE_RANGE;
        if (R(FGS_REDUCE_ROWS + 3) > 0) err |= FGS_PODERR_REF_PANIC;
    }
    i64 sel = -1, score = 0;
    if (emitted && key > 0) {
        const u64 k = (u64)key - 1;
        sel = (i64)(k & 0xffffffu);
        score = (i64)(k >> key_shift);
    }
    // 16 bytes per pod for the D2H: {count (28 bits) | error bits << 28 | selected node << 32, score}
    compact[0 * p_pad + p] = (i64)((u64)(uint32_t)((uint32_t)count | (err << 28)) | ((u64)(uint32_t)(int32_t)sel << 32));
    compact[1 * p_pad + p] = score;
}
} // namespace

void launch_finalize_pods(const int64_t* perpod, int64_t* compact, int64_t n_pods, int64_t p_pad, int key_shift, int emitted, int sc_flex,
                          void* stream) {
    if (n_pods <= 0) return;
    finalize_pods<<<(unsigned)((n_pods + 255) / 256), 256, 0, (cudaStream_t)stream>>>((const i64*)perpod, (i64*)compact, n_pods, p_pad, key_shift,
                                                                                     emitted, sc_flex);
}

namespace {
// PD_META of a pod table that arrived column by column (DMA straight from the caller's table): flags | present << 32
__global__ void make_pod_meta(const uint32_t* __restrict__ flags, const uint32_t* __restrict__ present, i64* __restrict__ meta, i64 n_pods, i64 p_pad) {
    const i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= p_pad) return;
    meta[i] = i < n_pods ? (i64)(((u64)present[i] << 32) | (u64)flags[i]) : 0;
}
} // namespace

void launch_make_pod_meta(const uint32_t* flags, const uint32_t* present, int64_t* meta, int64_t n_pods, int64_t p_pad, void* stream) {
    make_pod_meta<<<(unsigned)((p_pad + 255) / 256), 256, 0, (cudaStream_t)stream>>>(flags, present, (i64*)meta, n_pods, p_pad);
}

void launch_init_perpod(int64_t* perpod, int64_t p_pad, void* stream) {
    init_perpod<<<(unsigned)((p_pad + 255) / 256), 256, 0, (cudaStream_t)stream>>>((i64*)perpod, (i64)p_pad);
}

void launch_eval_reduce(const EvalParams& p, void* stream) {
    dim3 grid;
    eval_grid(p, &grid);
    eval_kernel<MODE_REDUCE><<<grid, kThreads, eval_smem_bytes(p.S), (cudaStream_t)stream>>>(p);
}

void launch_eval_emit(const EvalParams& p, void* stream) {
    dim3 grid;
    eval_grid(p, &grid);
    eval_kernel<MODE_EMIT><<<grid, kThreads, eval_smem_bytes(p.S), (cudaStream_t)stream>>>(p);
}

} // namespace fgs
