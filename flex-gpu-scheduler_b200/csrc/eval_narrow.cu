// eval_narrow.cu — int32 evaluation kernels (see kernels_narrow.cu for when they apply and why they are
// exact).
//
//   reduce_narrow<SPEC>  pass A: Filter for every (pod,node) → feasibility bitmap (warp ballot, written as
//                        whole 32-byte sectors) + per-pod feasible count, max FlexGPU score, Allocatable hi/lo.
//   emit_narrow<SPEC>    pass B: Score, NormalizeScore, ×weight, Σ plugins → int64 score rows (streaming
//                        256-byte stores per warp) + the per-pod selectHost key. Feasibility is READ from the
//                        bitmap (staged per pod tile by TMA), not recomputed; an external mask can be
//                        substituted (the feasible list the framework hands to PreScore).
//
//   emit_narrow<SPEC, true> + flexmax_prepass: both passes in one (opt-in, FGS_FUSED=1).
//
// One CTA = 256 threads × a tile of nodes in registers (reduce: 2 per thread, emit: 4) × a range of pods; pod tiles
// of 64 arrive through a two-stage cp.async.bulk + mbarrier ring (loads from this GPU's own tables). SPEC (plugin
// set, strategy, weights equal) is a template parameter, so the pod loops have no configuration branches. The emit
// pass is HBM-write bound (91.7 % of the copy peak on C2, ~36 lane-instructions per (pod,node)), the reduce pass
// integer-issue bound (~20).
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdlib>
#include <type_traits>

#include "kernels.cuh"
#include "kernels_narrow.cuh"

namespace fgs {

namespace {

typedef long long i64;
typedef unsigned long long u64;
constexpr i64 kI64Min = (i64)0x8000000000000000ull;
constexpr int kNone = (int)0x80000000; // node side: below every pod value, sentinels included; also 'no feasible node'
constexpr int kNever = 0x7fffffff;
constexpr int kPass = (int)0x80000001; // pod side: "no limit / no request" — below every real node value, above kNone

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
// TMA bulk copy global → shared (SASS: UBLKCP), completion counted in bytes on an mbarrier
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(dst)),
                 "l"(__cvta_generic_to_global(src)), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void fence_mbar_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }

// floor(x / d) for 0 <= x < 2^31 and the divisor's magic pair (narrow_magic, kernels_narrow.cuh), given 2·x:
// one IMAD.HI and one shift — the quotient is exact, no estimate to repair.
__device__ __forceinline__ int qmagic(uint32_t x2, uint32_t m, int l) { return (int)(__umulhi(x2, m) >> l); }

// Nodes per thread. Each pod iteration pays a fixed cost (broadcast loads of the pod record, loop and pointer
// bookkeeping, one REDUX, one staged store); four nodes per thread spread it over 128 evaluations per warp.
// Measured on C2 (profiles/): four per thread is faster for the emit pass (76.7 % vs 73.5 % of HBM peak), two for
// the reduce pass (its per-node work is nine compares: more nodes per thread only costs occupancy).
constexpr int kNPTe = 4, kNPTr = 2;
constexpr int kTileNe = kThreads * kNPTe, kTileNr = kThreads * kNPTr; // nodes per CTA: 1024 / 512
constexpr int kWordsTe = kTileNe / 32, kWordsTr = kTileNr / 32;       // bitmap words per pod row per CTA
static_assert(kNodePad % kTileNe == 0 && kNodePad % kTileNr == 0, "node padding must cover whole CTA tiles");

constexpr uint32_t SP_FLEX_F = 1, SP_FIT_F = 2, SP_SC_FLEX = 4, SP_SC_FIT = 8, SP_FIT_MOST = 16, SP_SC_ALLOC = 32, SP_W_EQ = 64;

// rare paths kept out of line so that they add no loop-carried address arithmetic to the pod loops
__device__ __noinline__ void raise_flag(i64* row, i64 idx) { atomicMax(row + idx, (i64)1); }

__device__ __forceinline__ i64 warp_max_i64(i64 v) {
    const int hi = (int)(v >> 32);
    const int mhi = __reduce_max_sync(0xffffffffu, hi);
    const uint32_t mlo = __reduce_max_sync(0xffffffffu, hi == mhi ? (uint32_t)v : 0u);
    return (i64)(((u64)(uint32_t)mhi << 32) | mlo);
}

// Both kernels run the same software pipeline over the CTA's pod tiles, ONE __syncthreads per tile:
//   TMA (cp.async.bulk → mbarrier) brings tile k+2's pod columns while tile k is evaluated;
//   warps 0-1 turn tile k+1's columns into the per-pod broadcast records right after finishing tile k;
//   per-warp results of tile k are flushed after the barrier while tile k+1 is already running.
// Every shared buffer is double-buffered by tile parity.

// =================================================================================================
// pass A
// =================================================================================================

struct __align__(16) PodR { // 32 bytes: two broadcast LDS.128 per pod
    int lim_g, lim_m, rc, rm;
    int re, rs0, rs1, kind;
};

struct SmemR {
    int bars, raw, pod, wred, whi, wnlo, bm, total, raw_stage_bytes;
};
__host__ __device__ inline SmemR smem_reduce(bool alloc) {
    SmemR L;
    int off = 0;
    L.bars = off;
    off += 64;
    L.raw = off;
    L.raw_stage_bytes = PN_COLS * kTileP * 4;
    off += 2 * L.raw_stage_bytes;
    L.pod = off;
    off += 2 * kTileP * (int)sizeof(PodR);
    L.wred = off; // int2 {max flex score, count | panic<<31} [2][kTileP][kWarps]
    off += 2 * kTileP * kWarps * 8;
    L.whi = off;
    off += alloc ? 2 * kTileP * kWarps * 8 : 0;
    L.wnlo = off;
    off += alloc ? 2 * kTileP * kWarps * 8 : 0;
    L.bm = off; // uint32 [2][kTileP][kWordsTr]
    off += 2 * kTileP * kWordsTr * 4;
    L.total = off;
    return L;
}

template <uint32_t SPEC>
__global__ void __launch_bounds__(kThreads, 4) reduce_narrow(const __grid_constant__ NarrowParams prm) {
    extern __shared__ __align__(128) unsigned char smem[];
    constexpr bool kAlloc = SPEC & SP_SC_ALLOC;
    const SmemR L = smem_reduce(kAlloc);
    uint64_t* bars = (uint64_t*)(smem + L.bars);

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const i64 node_tile0 = (i64)blockIdx.x * kTileNr;
    const i64 pod_begin = (i64)blockIdx.y * prm.pods_per_cta;
    i64 pod_end = pod_begin + prm.pods_per_cta;
    if (pod_end > prm.p_pad) pod_end = prm.p_pad;
    const int n_tiles = (int)((pod_end - pod_begin + kTileP - 1) / kTileP);
    if (n_tiles <= 0) return;

    auto issue = [&](int it) { // one thread
        const int stage = it & 1;
        const i64 p0 = pod_begin + (i64)it * kTileP;
        int* dst = (int*)(smem + L.raw + stage * L.raw_stage_bytes);
        mbar_expect_tx(&bars[stage], (uint32_t)L.raw_stage_bytes);
        for (int c = 0; c < PN_COLS; ++c) bulk_g2s(dst + c * kTileP, prm.pn + (i64)c * prm.p_pad + p0, kTileP * 4, &bars[stage]);
    };
    auto derive = [&](int it) { // threads 0..63
        const int stage = it & 1;
        mbar_wait(&bars[stage], (uint32_t)((it >> 1) & 1));
        const int* raw = (const int*)(smem + L.raw + stage * L.raw_stage_bytes);
        PodR u;
        u.lim_g = raw[PN_LIM_G * kTileP + tid];
        u.lim_m = raw[PN_LIM_M * kTileP + tid];
        u.rc = raw[PN_RC * kTileP + tid];
        u.rm = raw[PN_RM * kTileP + tid];
        u.re = raw[PN_RE * kTileP + tid];
        u.rs0 = raw[PN_RS0 * kTileP + tid];
        u.rs1 = raw[PN_RS1 * kTileP + tid];
        u.kind = raw[PN_KIND * kTileP + tid];
        // both limits: the Filter side never passes (its limit columns hold kNever) and Score takes GPUScore — a gpu
        // limit wins over a memory limit (flex_gpu.go:157-159) — so from here on the pod is a whole-GPU pod
        if (u.kind == 3) u.kind = 1;
        ((PodR*)(smem + L.pod))[stage * kTileP + tid] = u;
    };
    if (tid == 0) {
        mbar_init(&bars[0], 1);
        mbar_init(&bars[1], 1);
        fence_mbar_init();
    }
    __syncthreads();
    if (tid == 0) {
        issue(0);
        if (n_tiles > 1) issue(1);
    }

    // node tile → registers (thread owns nodes 128·warp + 32·j + lane, so each ballot is one bitmap word)
    int thr_g[kNPTr], thr_m[kNPTr], fc[kNPTr], fm[kNPTr], fe[kNPTr], fs0[kNPTr], fs1[kNPTr], gsc[kNPTr], msc[kNPTr];
    int node_panic[kNPTr];
    i64 araw[kNPTr];
    uint32_t flags_or = 0;
#pragma unroll
    for (int j = 0; j < kNPTr; ++j) {
        const i64 n = node_tile0 + warp * (32 * kNPTr) + j * 32 + lane;
        auto C = [&](int c) { return __ldg(prm.nn + (i64)c * prm.n_pad + n); };
        const uint32_t fl = (uint32_t)C(NN_FLAGS);
        flags_or |= fl;
        thr_g[j] = (SPEC & SP_FLEX_F) ? C(NN_THR_GPU) : kNever;
        thr_m[j] = (SPEC & SP_FLEX_F) ? C(NN_THR_MEM) : kNever;
        fc[j] = (SPEC & SP_FIT_F) ? C(NN_FREE_CPU) : kNever;
        fm[j] = (SPEC & SP_FIT_F) ? C(NN_FREE_MEM) : kNever;
        fe[j] = (SPEC & SP_FIT_F) ? C(NN_FREE_EPH) : kNever;
        fs0[j] = (SPEC & SP_FIT_F) ? C(NN_FS0) : kNever;
        fs1[j] = (SPEC & SP_FIT_F) ? C(NN_FS1) : kNever;
        gsc[j] = C(NN_GSCORE);
        msc[j] = C(NN_MSCORE);
        node_panic[j] = (fl & FGS_NF_REF_PANIC) ? 1 : 0;
        araw[j] = kAlloc ? __ldg((const i64*)prm.nd + (i64)ND_ALLOC_RAW * prm.n_pad + n) : 0;
        // node-only verdicts folded into the compares (every pod value, sentinels included, is > kNone):
        // a row that is not a node fails FlexGPU's side, "too many pods" (fit.go:259) fails Fit's side
        // a node on which NewGPUNode would panic (flex_gpu.go:101) takes no pod that carries a flex limit — their limits
        // are >= 0 — but every pod without one (kPass): the threshold kPass says exactly that. WHICH pods reach the panic
        // (pod_err) is found outside the loop, over the few such nodes (filter_panic_pass).
        if ((SPEC & SP_FLEX_F) && (fl & FGS_NF_REF_PANIC)) thr_g[j] = thr_m[j] = kPass;
        if (!(fl & FGS_NF_VALID)) thr_g[j] = thr_m[j] = kNone;
        if ((SPEC & SP_FIT_F) && !(fl & NDF_PODS_OK)) fc[j] = kNone;
    }
    (void)flags_or;
    (void)node_panic;
    const bool tile_has_panic = false;
    if (tid < kTileP) derive(0);
    __syncthreads();
    if (tid == 0 && n_tiles > 2) {
        fence_proxy_async();
        issue(2);
    }
    const i64 word0 = node_tile0 / 32; // kWordsTr words per pod row per CTA

    for (int it = 0; it < n_tiles; ++it) {
        const int stage = it & 1;
        const i64 tile_p0 = pod_begin + (i64)it * kTileP;
        const i64 left = prm.n_pods - tile_p0;
        const int n_valid = left >= kTileP ? kTileP : (left > 0 ? (int)left : 0);
        const PodR* s_pod = (const PodR*)(smem + L.pod) + stage * kTileP;
        int2* wred = (int2*)(smem + L.wred) + stage * kTileP * kWarps;
        i64* whi = (i64*)(smem + L.whi) + stage * kTileP * kWarps;
        i64* wnlo = (i64*)(smem + L.wnlo) + stage * kTileP * kWarps;
        uint32_t* bm = (uint32_t*)(smem + L.bm) + stage * kTileP * kWordsTr;

        // The pod loop used to exist twice (tiles with / without a node on which NewGPUNode would panic); panic nodes now go
        // through the thresholds and only the plain instance runs. The generic-lambda form is kept on purpose: written as
        // a bare loop the same body compiles to a slower schedule (0.432 vs 0.403 ms on C2, measured).
        auto pod_loop = [&](auto rare_tag) {
        constexpr bool kRare = decltype(rare_tag)::value;
#pragma unroll 8 // measured on C2: 0.403 ms without, 0.387 (4), 0.376 (8), 0.373 (16)
        for (int jp = 0; jp < n_valid; ++jp) {
            const int4 a = *(const int4*)&s_pod[jp].lim_g;
            const int4 b = *(const int4*)&s_pod[jp].re;
            const int kind = b.w;
            bool feas[kNPTr];
            bool fpanic = false;
#pragma unroll
            for (int j = 0; j < kNPTr; ++j) {
                // FlexGPU.Filter: kind none passes, "both" never, else limit <= per-kind threshold (flex_gpu.go:54-116)
                const bool flex_ok = a.x <= thr_g[j] && a.y <= thr_m[j];
                // Fit.Filter: request <= allocatable − requested per resource (fit.go:277-326)
                const bool fit_ok = a.z <= fc[j] && a.w <= fm[j] && b.x <= fe[j] && b.y <= fs0[j] && b.z <= fs1[j];
                bool ok = flex_ok && fit_ok;
                if (kRare) { // NewGPUNode would panic inside FlexGPU.Filter (flex_gpu.go:101)
                    const bool reach = (kind == 1 || kind == 2) && flex_ok && node_panic[j] && (prm.flex_filter_first || fit_ok);
                    fpanic = fpanic || reach;
                    ok = ok && !reach;
                }
                feas[j] = ok;
            }
            uint32_t bal[kNPTr];
            uint32_t cnt = 0;
#pragma unroll
            for (int j = 0; j < kNPTr; ++j) {
                bal[j] = __ballot_sync(0xffffffffu, feas[j]);
                cnt += (uint32_t)__popc(bal[j]);
            }
            int vflex = kNone;
            if (SPEC & SP_SC_FLEX) { // raw FlexGPU score by pod kind (flex_gpu.go:153-165), max over the feasible nodes
                int sj[kNPTr];
#pragma unroll
                for (int j = 0; j < kNPTr; ++j) sj[j] = kind == 1 ? gsc[j] : (kind == 2 ? msc[j] : 0);
#pragma unroll
                for (int j = 0; j < kNPTr; ++j) vflex = max(vflex, feas[j] ? sj[j] : kNone);
                vflex = __reduce_max_sync(0xffffffffu, vflex);
            }
            if (kRare && __ballot_sync(0xffffffffu, fpanic)) cnt |= 0x80000000u;
            if (kAlloc) {
                i64 vhi = kI64Min, vnlo = kI64Min;
#pragma unroll
                for (int j = 0; j < kNPTr; ++j)
                    if (feas[j]) {
                        vhi = araw[j] > vhi ? araw[j] : vhi;
                        vnlo = -araw[j] > vnlo ? -araw[j] : vnlo;
                    }
                vhi = warp_max_i64(vhi);
                vnlo = warp_max_i64(vnlo);
                if (lane == 0) {
                    whi[jp * kWarps + warp] = vhi;
                    wnlo[jp * kWarps + warp] = vnlo;
                }
            }
            if (lane == 0) {
                wred[jp * kWarps + warp] = make_int2(vflex, (int)cnt);
                static_assert(kNPTr == 2, "one 8-byte store per warp and pod");
                *(uint2*)&bm[jp * kWordsTr + warp * kNPTr] = make_uint2(bal[0], bal[1]);
            }
        }
        };
        if (tile_has_panic) pod_loop(std::true_type{});
        else pod_loop(std::false_type{});
        if (tid < kTileP && it + 1 < n_tiles) derive(it + 1);
        __syncthreads();
        if (tid == 0 && it + 3 < n_tiles) { // tile it+1's columns are consumed: that stage takes tile it+3
            fence_proxy_async();
            issue(it + 3);
        }
        // ---- flush tile `it` (overlaps tile it+1 of the faster warps) ------------------------------
        if (tid < n_valid) {
            const i64 p = tile_p0 + tid;
            uint32_t c = 0, pn = 0;
            int vf = kNone;
            i64 vh = kI64Min, vl = kI64Min;
#pragma unroll
            for (int w = 0; w < kWarps; ++w) {
                const int2 x = wred[tid * kWarps + w];
                c += (uint32_t)x.y & 0x7fffffffu;
                pn |= (uint32_t)x.y >> 31;
                vf = x.x > vf ? x.x : vf;
                if (kAlloc) {
                    const i64 h = whi[tid * kWarps + w], l = wnlo[tid * kWarps + w];
                    vh = h > vh ? h : vh;
                    vl = l > vl ? l : vl;
                }
            }
            if (c) atomicAdd((u64*)prm.reduce_sum + p, (u64)c);
            if (pn) atomicMax((i64*)prm.reduce_max + (i64)RR_FILTER_PANIC * prm.p_pad + p, (i64)1);
            if ((SPEC & SP_SC_FLEX) && vf != kNone) {
                // back to the reference's units: merged across shards and read by either emit kernel
                const i64 v = s_pod[tid].kind == 2 ? (i64)vf * ((i64)1 << prm.sh.s[DIM_FM]) : (i64)vf;
                atomicMax((i64*)prm.reduce_max + (i64)RR_FLEX_MAX * prm.p_pad + p, v);
            }
            if (kAlloc) {
                if (vh != kI64Min) atomicMax((i64*)prm.reduce_max + (i64)RR_ALLOC_HI * prm.p_pad + p, vh);
                if (vl != kI64Min) atomicMax((i64*)prm.reduce_max + (i64)RR_ALLOC_NEGLO * prm.p_pad + p, vl);
            }
        }
        for (int idx = tid; idx < n_valid * kWordsTr; idx += kThreads) { // bitmap rows as whole 32-byte sectors
            const int jp = idx / kWordsTr, w = idx % kWordsTr;
            prm.bitmap[(tile_p0 + jp) * prm.words_per_row + word0 + w] = bm[idx];
        }
    }
}

// =================================================================================================
// pass B
// =================================================================================================

struct __align__(16) PodE { // 64 bytes: four broadcast LDS.128 per pod
    int nzc, nzm, nz200c, nz200m;      // non-zero requests and 200 × them (the dividends are kept doubled, see qmagic)
    uint32_t fmag, amag;               // magic multipliers of the pod's two divisors: flex max, Allocatable range
    int fsh, ash;                      // … and their shifts
    int fmax, count, arange32, kind; // arange32 < 0: exact 64-bit path for Allocatable
    uint32_t hmul; // per-pod hash (cross-tile tie mixing in the flush)
    int hpos;      // 8 of its bits at positions 10..17: XORed with the node's bits in the key
    i64 alo;
};

struct SmemE {
    int bars, raw, pod, wkey, bm, total, raw_stage_bytes;
};
// fused: the pass also runs the filters (one pass over pod × node instead of two). Its tiles carry the pod's filter
// columns too, and the 'bm' area holds the per-warp feasible counts instead of staged bitmap rows.
__host__ __device__ inline SmemE smem_emit(bool fused) {
    SmemE L;
    int off = 0;
    L.bars = off;
    off += 64;
    L.raw = off;
    L.raw_stage_bytes = 3 * kTileP * 4 + 4 * kTileP * 8; // kind, nzc, nzm + flex max, alloc hi, alloc −lo, count
    if (fused) L.raw_stage_bytes += 7 * kTileP * 4;     // + limits and requests (PN_LIM_G … PN_RS1)
    off += 2 * L.raw_stage_bytes;
    L.pod = off;
    off += 2 * kTileP * (int)(sizeof(PodE) + (fused ? sizeof(PodR) : 0));
    L.wkey = off;
    off += 2 * kTileP * kWarps * 4;
    L.bm = off; // uint32 [2][kTileP][kWordsTe]: the feasibility words of this node tile, two pod tiles deep
                // (fused: uint32 [2][kTileP][kWarps] feasible nodes per warp, bit 31 = "NewGPUNode would panic")
    off += 2 * kTileP * (fused ? kWarps : kWordsTe) * 4;
    L.total = off;
    return L;
}

// TAIL: the last node tile of a table that fills at most half of it (a 6 250-node shard of C5 fills 106 of 1 024 columns).
// A warp owns 128 nodes and walks every pod of the CTA's chunk, so such a CTA would take as long as a full one with one
// warp in eight working. Here the warps whose nodes lie beyond the table take a share of the pods instead: with A = 1, 2
// or 4 node groups in range, warp w works on node group w mod A and on every (8/A)-th pod.
template <uint32_t SPEC, bool FUSED, bool TAIL = false>
__global__ void __launch_bounds__(kThreads, (TAIL || ((SPEC & SP_SC_ALLOC) && (SPEC & SP_SC_FIT))) ? 2 : 3) emit_narrow(const __grid_constant__ NarrowParams prm) {
    extern __shared__ __align__(128) unsigned char smem[];
    constexpr bool kFlex = SPEC & SP_SC_FLEX, kFit = SPEC & SP_SC_FIT, kAlloc = SPEC & SP_SC_ALLOC, kMost = SPEC & SP_FIT_MOST;
    static_assert(!(FUSED && kAlloc), "the fused pass has no Allocatable hi/lo (they would need the feasible set first)");
    const SmemE L = smem_emit(FUSED);
    PodR* const s_podf_base = (PodR*)(smem + L.pod + 2 * kTileP * (int)sizeof(PodE)); // fused: the pods' filter records
    uint64_t* bars = (uint64_t*)(smem + L.bars); // [0,1] pod columns, [2,3] bitmap rows

    static_assert(!(FUSED && TAIL), "the fused pass writes bitmap words per warp: no pod split there");
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t tile_idx = TAIL ? (uint32_t)prm.tile_first + blockIdx.x : blockIdx.x; // (the main launch starts at tile 0)
    const i64 node_tile0 = TAIL ? (i64)tile_idx * kTileNe : (i64)blockIdx.x * kTileNe;
    const i64 pod_begin = (i64)blockIdx.y * prm.pods_per_cta;
    i64 pod_end = pod_begin + prm.pods_per_cta;
    if (pod_end > prm.p_pad) pod_end = prm.p_pad;
    const int n_tiles = (int)((pod_end - pod_begin + kTileP - 1) / kTileP);
    if (n_tiles <= 0) return;
    const i64 word0 = node_tile0 / 32;
    // TAIL only (the other instantiation keeps the expressions it was tuned with: a different but equivalent form cost it
    // three registers and 1.5 % — 677 vs 667 µs on C2):
    const int ngrp = TAIL ? (warp & (prm.tail_groups - 1)) : 0;   // which 128 nodes of the tile
    const int pod_step = TAIL ? kWarps / prm.tail_groups : 1;      // this warp takes every pod_step-th pod …
    const int pod_first = TAIL ? warp / prm.tail_groups : 0;       // … from this one on

    auto issue_pods = [&](int it) { // one thread
        const int stage = it & 1;
        const i64 p0 = pod_begin + (i64)it * kTileP;
        constexpr int tn = kTileP;
        int* dst = (int*)(smem + L.raw + stage * L.raw_stage_bytes);
        mbar_expect_tx(&bars[stage], (uint32_t)L.raw_stage_bytes);
        bulk_g2s(dst + 0 * kTileP, prm.pn + (i64)PN_KIND * prm.p_pad + p0, tn * 4, &bars[stage]);
        bulk_g2s(dst + 1 * kTileP, prm.pn + (i64)PN_NZC * prm.p_pad + p0, tn * 4, &bars[stage]);
        bulk_g2s(dst + 2 * kTileP, prm.pn + (i64)PN_NZM * prm.p_pad + p0, tn * 4, &bars[stage]);
        i64* d2 = (i64*)(dst + 3 * kTileP);
        bulk_g2s(d2 + 0 * kTileP, prm.reduce_max + (i64)RR_FLEX_MAX * prm.p_pad + p0, tn * 8, &bars[stage]);
        bulk_g2s(d2 + 1 * kTileP, prm.reduce_max + (i64)RR_ALLOC_HI * prm.p_pad + p0, tn * 8, &bars[stage]);
        bulk_g2s(d2 + 2 * kTileP, prm.reduce_max + (i64)RR_ALLOC_NEGLO * prm.p_pad + p0, tn * 8, &bars[stage]);
        bulk_g2s(d2 + 3 * kTileP, prm.reduce_sum + p0, tn * 8, &bars[stage]);
        if (FUSED) {
            int* d3 = (int*)(d2 + 4 * kTileP);
            for (int c = 0; c < 7; ++c) bulk_g2s(d3 + c * kTileP, prm.pn + (i64)(PN_LIM_G + c) * prm.p_pad + p0, kTileP * 4, &bars[stage]);
        }
    };
    // threads 0..63 fetch one 128-byte bitmap row each (rows are words_per_row apart in global memory)
    auto issue_bitmap = [&](int it) {
        const int stage = it & 1;
        const i64 p0 = pod_begin + (i64)it * kTileP;
        uint32_t* s_bm = (uint32_t*)(smem + L.bm);
        if (tid == 0) mbar_expect_tx(&bars[2 + stage], (uint32_t)(kTileP * kWordsTe * 4));
        if (tid < kTileP)
            bulk_g2s(s_bm + (stage * kTileP + tid) * kWordsTe, prm.bitmap + (p0 + tid) * prm.words_per_row + word0, kWordsTe * 4, &bars[2 + stage]);
    };
    auto derive = [&](int it) { // threads 0..63
        const int stage = it & 1;
        const i64 tile_p0 = pod_begin + (i64)it * kTileP;
        mbar_wait(&bars[stage], (uint32_t)((it >> 1) & 1));
        const int* raw = (const int*)(smem + L.raw + stage * L.raw_stage_bytes);
        const i64* raw64 = (const i64*)(raw + 3 * kTileP);
        PodE u;
        u.kind = raw[0 * kTileP + tid];
        u.nzc = raw[1 * kTileP + tid];
        u.nzm = raw[2 * kTileP + tid];
        u.nz200c = (int)(200u * (uint32_t)u.nzc);
        u.nz200m = (int)(200u * (uint32_t)u.nzm);
        i64 fmx = raw64[0 * kTileP + tid];
        if (fmx < 0) fmx = 0; // maxCount starts at 0 (normalize_score.go:27)
        if (u.kind == 2) fmx >>= prm.sh.s[DIM_FM];
        u.fmax = (int)fmx;
        u.fmag = narrow_magic_fast((uint32_t)u.fmax, &u.fsh);
        u.count = (int)raw64[3 * kTileP + tid];
        u.alo = 0;
        u.arange32 = 0;
        u.amag = 0;
        u.ash = 0;
        if (kAlloc && u.count > 0) {
            const i64 hi = raw64[1 * kTileP + tid], lo = -raw64[2 * kTileP + tid];
            const i64 range = (i64)((u64)hi - (u64)lo);
            u.alo = lo;
            if (range >= 0 && range < (1 << 24)) {
                u.arange32 = (int)range;
                u.amag = narrow_magic_fast((uint32_t)range, &u.ash);
            } else { // the exact 64-bit path: the two fields carry the reciprocal of the range (alloc_div100)
                u.arange32 = -1;
                const u64 m = alloc_recip(range);
                u.amag = (uint32_t)m;
                u.ash = (int)(uint32_t)(m >> 32);
            }
        }
        uint32_t h = (uint32_t)(prm.seed ^ (prm.seed >> 32)) ^ ((uint32_t)(tile_p0 + tid) * 0x85EBCA6Bu);
        h ^= h >> 15;
        h *= 0x9E3779B1u;
        u.hmul = h | 1u;
        u.hpos = (int)(((h >> 20) & 0xffu) << 10);
        ((PodE*)(smem + L.pod))[stage * kTileP + tid] = u;
        if (FUSED) {
            const int* f = (const int*)(raw64 + 4 * kTileP);
            PodR r;
            r.lim_g = f[0 * kTileP + tid];
            r.lim_m = f[1 * kTileP + tid];
            r.rc = f[2 * kTileP + tid];
            r.rm = f[3 * kTileP + tid];
            r.re = f[4 * kTileP + tid];
            r.rs0 = f[5 * kTileP + tid];
            r.rs1 = f[6 * kTileP + tid];
            r.kind = u.kind;
            s_podf_base[stage * kTileP + tid] = r;
        }
    };
    if (tid == 0) {
        for (int b = 0; b < 4; ++b) mbar_init(&bars[b], 1);
        static_assert(PN_LIM_G + 6 == PN_RS1, "the fused pass copies seven consecutive pod columns");
        fence_mbar_init();
    }
    __syncthreads();
    if (tid == 0) {
        issue_pods(0);
        if (n_tiles > 1) issue_pods(1);
    }
    if (!FUSED) {
        issue_bitmap(0);
        if (n_tiles > 1) issue_bitmap(1);
    }

    // node tile → registers
    int capc[kNPTe], xc[kNPTe], capm[kNPTe], xm[kNPTe], nloc[kNPTe], inr[kNPTe], mneg[kNPTe], shc[kNPTe], shm[kNPTe];
    uint32_t x200c[kNPTe], x200m[kNPTe], g200[kNPTe], m200[kNPTe], magc[kNPTe], magm[kNPTe];
    uint32_t nbase[kNPTe], nh[kNPTe];
    i64 araw[kNPTe];
    // fused: the node side of the two filters, as in reduce_narrow (which filters run is a run-time choice here: a
    // disabled one compares against kNever)
    int thr_g[kNPTe], thr_m[kNPTe], fc[kNPTe], fm[kNPTe], fe[kNPTe], fs0[kNPTe], fs1[kNPTe], node_panic[kNPTe];
    uint32_t flags_or = 0;
#pragma unroll
    for (int j = 0; j < kNPTe; ++j) {
        nloc[j] = (TAIL ? ngrp : warp) * (32 * kNPTe) + j * 32 + lane;
        if (FUSED) {
            const i64 n = node_tile0 + nloc[j];
            auto C = [&](int c) { return __ldg(prm.nn + (i64)c * prm.n_pad + n); };
            const uint32_t fl = (uint32_t)C(NN_FLAGS);
            const bool flex_f = prm.fused_spec & SP_FLEX_F, fit_f = prm.fused_spec & SP_FIT_F;
            flags_or |= fl;
            thr_g[j] = flex_f ? C(NN_THR_GPU) : kNever;
            thr_m[j] = flex_f ? C(NN_THR_MEM) : kNever;
            fc[j] = fit_f ? C(NN_FREE_CPU) : kNever;
            fm[j] = fit_f ? C(NN_FREE_MEM) : kNever;
            fe[j] = fit_f ? C(NN_FREE_EPH) : kNever;
            fs0[j] = fit_f ? C(NN_FS0) : kNever;
            fs1[j] = fit_f ? C(NN_FS1) : kNever;
            node_panic[j] = (fl & FGS_NF_REF_PANIC) ? 1 : 0;
            if (flex_f && (fl & FGS_NF_REF_PANIC)) thr_g[j] = thr_m[j] = kPass; // see reduce_narrow
            if (!(fl & FGS_NF_VALID)) thr_g[j] = thr_m[j] = kNone; // a row that is not a node fails FlexGPU's side
            if (fit_f && !(fl & NDF_PODS_OK)) fc[j] = kNone;       // "Too many pods" (fit.go:259) fails Fit's side
        }
        nbase[j] = (1u << 18) | (uint32_t)nloc[j];
        nh[j] = (((uint32_t)nloc[j] * 0x9E3779B1u) >> 24) << 10; // the node's 8 tie-break bits, at bits 10..17
        nh[j] |= nbase[j];                                        // … with the marker bit 18 and the node-in-tile: the pod's
                                                                  // hash bits (10..17 only) are XORed into this word
        const i64 n = node_tile0 + nloc[j];
        inr[j] = n < prm.n_nodes ? 1 : 0;
        auto C = [&](int c) { return __ldg(prm.nn + (i64)c * prm.n_pad + n); };
        if (kFit) {
            capc[j] = C(NN_CAP_C);
            xc[j] = C(NN_X_C);
            capm[j] = C(NN_CAP_M);
            xm[j] = C(NN_X_M);
            // Least: a request above what is left scores 0 (least_allocated.go:48-55): the dividend is 200·x − min(200·x, 200·nz), and
            // x < 0 ("everything exceeds it") is kept as x = 0 for that purpose
            x200c[j] = (!kMost && xc[j] < 0) ? 0u : 200u * (uint32_t)xc[j];
            x200m[j] = (!kMost && xm[j] < 0) ? 0u : 200u * (uint32_t)xm[j];
            magc[j] = (uint32_t)C(NN_MAG_C);
            magm[j] = (uint32_t)C(NN_MAG_M);
            const int shs = C(NN_MAG_SH);
            shc[j] = shs & 0xff;
            shm[j] = (shs >> 8) & 0xff;
        }
        if (kFlex) {
            // a negative MemScore (over-committed device) is kept as magnitude + sign: Go's '/' truncates toward
            // zero, so 100 − (100·s)/max = 100 + floor(100·|s|/max). Whether that leaves [0,100] and aborts the
            // cycle is decided per pod by special_pass / finalize_pods, not here.
            const int ms = C(NN_MSCORE);
            g200[j] = 200u * (uint32_t)C(NN_GSCORE);
            m200[j] = 200u * (uint32_t)(ms < 0 ? -ms : ms);
            mneg[j] = ms < 0 ? 1 : 0;
        }
        araw[j] = kAlloc ? __ldg((const i64*)prm.nd + (i64)ND_ALLOC_RAW * prm.n_pad + n) : 0;
    }
    // keep the tail-tile predicates in registers (the compiler would otherwise recompute the 64-bit compare per pod)
    asm volatile("" : "+r"(inr[0]), "+r"(inr[1]), "+r"(inr[2]), "+r"(inr[3]));
    asm volatile("" : "+r"(nh[0]), "+r"(nh[1]), "+r"(nh[2]), "+r"(nh[3])); // opaque: one register per node instead of rebuilding it per pod

    const int c100flex = 100 * prm.w_flex;
    if (kFlex && !FUSED) { // the doubled dividends stay in registers (the compiler would rather rebuild them from the scores per pod)
        asm volatile("" : "+r"(m200[0]), "+r"(m200[1]), "+r"(m200[2]), "+r"(m200[3]));
        asm volatile("" : "+r"(g200[0]), "+r"(g200[1]), "+r"(g200[2]), "+r"(g200[3]));
    }
    uint32_t lanebit = 1u << lane;
    asm volatile("" : "+r"(lanebit)); // opaque: keeps the feasibility test a single LOP3 with predicate output
    (void)flags_or;
    (void)node_panic;
    constexpr bool tile_has_panic = false; // panic nodes are handled through the thresholds (see reduce_narrow)
    // A warp whose 128 nodes all lie beyond the table (the tail of the last node tile: a 6 250-node shard fills 106 of its
    // last 1 024 columns) has no score to write and no key to offer: it leaves zeros in its column of the per-warp keys once
    // and sits the pod loops out, so that the tail tile costs an eighth of the issue slots of a full one.
    const bool warp_active = FUSED || __any_sync(0xffffffffu, inr[0] | inr[1] | inr[2] | inr[3]);
    // a negative MemScore (over-committed device) is rare: warps without one run the MemScore body without the sign handling
    bool warp_neg = false;
    if (kFlex) warp_neg = __any_sync(0xffffffffu, mneg[0] | mneg[1] | mneg[2] | mneg[3]);
    if (!warp_active || TAIL) // (TAIL: the pods this warp skips keep the zero)
        for (int i = lane; i < 2 * kTileP; i += 32) ((uint32_t*)(smem + L.wkey))[i * kWarps + warp] = 0u;
    if (tid < kTileP) derive(0);
    __syncthreads();
    if (tid == 0 && n_tiles > 2) {
        fence_proxy_async();
        issue_pods(2);
    }

    for (int it = 0; it < n_tiles; ++it) {
        const int stage = it & 1;
        const i64 tile_p0 = pod_begin + (i64)it * kTileP;
        const i64 left = prm.n_pods - tile_p0;
        const int n_valid = left >= kTileP ? kTileP : (left > 0 ? (int)left : 0);
        const PodE* s_pod = (const PodE*)(smem + L.pod) + stage * kTileP;
        uint32_t* wkey = (uint32_t*)(smem + L.wkey) + stage * kTileP * kWarps;
        const uint32_t* bmrow = (const uint32_t*)(smem + L.bm) + stage * kTileP * kWordsTe + (TAIL ? ngrp : warp) * kNPTe;
        const PodR* s_podf = s_podf_base + stage * kTileP;
        uint32_t* wcnt = (uint32_t*)(smem + L.bm) + stage * kTileP * kWarps; // fused only
        uint32_t* bm_out = prm.bitmap + tile_p0 * prm.words_per_row + word0 + warp * kNPTe; // fused: this warp's four words
        if (!FUSED) mbar_wait(&bars[2 + stage], (uint32_t)((it >> 1) & 1));

        i64* rowp = (i64*)prm.scores + (TAIL ? tile_p0 + pod_first : tile_p0) * prm.score_stride + node_tile0 + nloc[0];
        // (unrolling this loop costs registers and with them the third CTA per SM: 0.699 vs 0.665 ms)
        const int n_loop = warp_active ? n_valid : 0;
        for (int jp = TAIL ? pod_first : 0; jp < n_loop;
             jp += TAIL ? pod_step : 1, rowp += TAIL ? prm.score_stride * pod_step : prm.score_stride, bm_out += prm.words_per_row) {
            const int4 q2 = *(const int4*)&s_pod[jp].nzc;      // nzc nzm nz200c nz200m
            const uint4 q3 = *(const uint4*)&s_pod[jp].fmag;   // fmag amag fsh ash
            const int4 q4 = *(const int4*)&s_pod[jp].fmax;     // fmax count arange32 kind
            const int kind = q4.w;
            bool feas[kNPTe] = {false, false, false, false};
            if (!FUSED) {
                const uint4 fw = *(const uint4*)(bmrow + jp * kWordsTe);
                feas[0] = (fw.x & lanebit) != 0, feas[1] = (fw.y & lanebit) != 0, feas[2] = (fw.z & lanebit) != 0, feas[3] = (fw.w & lanebit) != 0;
            }
            // fused: the two filters, exactly as reduce_narrow runs them (flex_gpu.go:54-116, fit.go:277-326). As there, the
            // code exists twice: tiles without a node on which NewGPUNode would panic (the normal case) carry none of the
            // bookkeeping for it.
            auto run_filters = [&](auto rare_tag) {
                constexpr bool kRare = decltype(rare_tag)::value;
                const int4 a = *(const int4*)&s_podf[jp].lim_g; // lim_g lim_m rc rm
                const int4 b = *(const int4*)&s_podf[jp].re;    // re rs0 rs1 kind
                bool fpanic = false;
#pragma unroll
                for (int j = 0; j < kNPTe; ++j) {
                    const bool flex_ok = a.x <= thr_g[j] && a.y <= thr_m[j];
                    const bool fit_ok = a.z <= fc[j] && a.w <= fm[j] && b.x <= fe[j] && b.y <= fs0[j] && b.z <= fs1[j];
                    bool ok = flex_ok && fit_ok;
                    if (kRare) { // NewGPUNode would panic inside FlexGPU.Filter (flex_gpu.go:101)
                        const bool reach = kind != 0 && flex_ok && node_panic[j] && (prm.flex_filter_first || fit_ok);
                        fpanic = fpanic || reach;
                        ok = ok && !reach;
                    }
                    feas[j] = ok;
                }
                uint32_t bal[kNPTe], cnt = 0;
#pragma unroll
                for (int j = 0; j < kNPTe; ++j) {
                    bal[j] = __ballot_sync(0xffffffffu, feas[j]);
                    cnt += (uint32_t)__popc(bal[j]);
                }
                if (kRare && __ballot_sync(0xffffffffu, fpanic)) cnt |= 0x80000000u;
                if (lane == 0) {
                    static_assert(kNPTe == 4, "one 16-byte store per warp and pod");
                    *(uint4*)bm_out = make_uint4(bal[0], bal[1], bal[2], bal[3]); // the bitmap row's words of this warp's 128 nodes
                    wcnt[jp * kWarps + warp] = cnt;
                }
            };
            if (FUSED) {
                if (tile_has_panic) run_filters(std::true_type{});
                else run_filters(std::false_type{});
            }
            const uint32_t ph = (uint32_t)s_pod[jp].hpos; // the pod's 8 tie-break bits, already at bits 10..17
            // The pod kind and "max flex score == 0" are uniform over the tile: one copy of the node loop per case
            // instead of per-node selects. KS = 0: the FlexGPU term is the constant 100 (no limit, or max == 0);
            // 1: GPUScore (a gpu limit wins over a memory limit, flex_gpu.go:157-159); 2: MemScore.
            auto pod_body = [&](auto KSEL) {
                constexpr int KS = decltype(KSEL)::value & 3;
                constexpr bool kSigned = (decltype(KSEL)::value & 4) != 0; // KS == 2 only: some node of this warp has a negative MemScore
                int tot[kNPTe];
                uint32_t key = 0;
#pragma unroll
                for (int j = 0; j < kNPTe; ++j) {
                    int t = 0;
                    if (kFlex) { // FlexGPU.Score + DefaultNormalizeScore(reverse) (flex_gpu.go:142-176, normalize_score.go:26-54)
                        // (100 ∓ q) · w as 100·w ∓ q·w: one multiply-add per node
                        t = c100flex;
                        if (KS == 1) {
                            t -= qmagic(g200[j], q3.x, (int)q3.z) * prm.w_flex;
                        } else if (KS == 2) { // magnitude and sign: Go's '/' truncates toward zero
                            const int qw = qmagic(m200[j], q3.x, (int)q3.z) * prm.w_flex;
                            t = (kSigned && mneg[j]) ? t + qw : t - qw;
                        }
                    }
                    if (kFit) { // Least/MostAllocated over cpu and memory, non-zero requests (least_allocated.go:29-57, most_allocated.go:29-61)
                        int sc, sm;
                        if (!kMost) {
                            sc = qmagic(x200c[j] - min(x200c[j], (uint32_t)q2.z), magc[j], shc[j]); // requested > capacity → 0
                            sm = qmagic(x200m[j] - min(x200m[j], (uint32_t)q2.w), magm[j], shm[j]);
                        } else {
                            sc = qmagic(x200c[j] + (uint32_t)q2.z, magc[j], shc[j]);
                            sm = qmagic(x200m[j] + (uint32_t)q2.w, magm[j], shm[j]);
                            sc = q2.x > capc[j] - xc[j] ? 100 : sc; // requested clamps to capacity
                            sm = q2.y > capm[j] - xm[j] ? 100 : sm;
                        }
                        int v;
                        if (SPEC & SP_W_EQ) {
                            v = (sc + sm) >> 1;
                        } else { // Σ score·w / Σ w, w in [1,100]: exact through a float floor
                            const int num = sc * prm.fit_w_cpu + sm * prm.fit_w_mem;
                            v = __float_as_int(__fadd_rd(((float)num + 0.5f) * prm.fit_inv_wsum, 8388608.0f)) & 0x7fffff;
                        }
                        t += v * prm.w_fit;
                    }
                    if (kAlloc) { // Allocatable.NormalizeScore (allocatable.go:146-171)
                        int v = 0;
                        if (q4.z > 0) {
                            const int a = (int)(uint32_t)((u64)araw[j] - (u64)s_pod[jp].alo);
                            v = qmagic(200u * (uint32_t)a, q3.y, (int)q3.w);
                        } else if (q4.z < 0) {
                            const i64 alo = s_pod[jp].alo;
                            const i64 range = (i64)((u64)__ldg((const i64*)prm.reduce_max + (i64)RR_ALLOC_HI * prm.p_pad + tile_p0 + jp) - (u64)alo);
                            v = alloc_div100((i64)((u64)araw[j] - (u64)alo), range, ((u64)q3.w << 32) | (u64)q3.y);
                        }
                        t += v * prm.w_alloc;
                    }
                    tot[j] = t;
                    // selectHost candidate: score+1 (14 bits) | 8 tie-break bits | node-in-tile (10 bits); 0 = none
                    // (generic_scheduler.go:146-167). nbase = 1<<18 | node-in-tile, nh = the node's 8 hash bits in place:
                    // one LOP3 and one IMAD per node
                    const uint32_t k = (uint32_t)t * 262144u + (nh[j] ^ ph);
                    key = max(key, feas[j] ? k : 0u);
                }
#pragma unroll
                for (int j = 0; j < kNPTe; ++j)
                    if (inr[j]) __stcs(rowp + 32 * j, (i64)(u64)(uint32_t)tot[j]);
                key = __reduce_max_sync(0xffffffffu, key);
                if (lane == 0) wkey[jp * kWarps + warp] = key;
            };
            const int ks = (kFlex && q4.x != 0) ? ((kind & 1) ? 1 : (kind == 2 ? 2 : 0)) : 0;
            if (ks == 2) {
                if (warp_neg) pod_body(std::integral_constant<int, 6>{});
                else pod_body(std::integral_constant<int, 2>{});
            } else if (ks == 1) pod_body(std::integral_constant<int, 1>{});
            else pod_body(std::integral_constant<int, 0>{});
        }
        if (tid < kTileP && it + 1 < n_tiles) derive(it + 1);
        __syncthreads();
        if (!FUSED && it + 2 < n_tiles) { // this stage's bitmap rows are consumed: fetch the rows of tile it+2
            if (tid < kTileP) fence_proxy_async();
            issue_bitmap(it + 2);
        }
        if (tid == 0 && it + 3 < n_tiles) issue_pods(it + 3); // tile it+1's columns were consumed by derive(it+1)
        if (tid < n_valid) { // flush tile `it` (overlaps tile it+1 of the other warps)
            const i64 p = tile_p0 + tid;
            uint32_t k = 0;
#pragma unroll
            for (int w = 0; w < kWarps; ++w) {
                const uint32_t x = wkey[tid * kWarps + w];
                k = x > k ? x : k;
            }
            if (k) {
                const u64 t = (u64)(k >> 18) - 1, h8 = (k >> 10) & 0xffu, nl = k & 0x3ffu;
                const u64 ng = (u64)(prm.node_base + node_tile0) + nl;
                // cross-tile ties: mix the tile index into the low hash byte so that no tile is favoured
                const uint32_t h16 = ((uint32_t)h8 << 8) | (((TAIL ? tile_idx : (uint32_t)blockIdx.x) * 0x9E3779B1u ^ s_pod[tid].hmul) >> 24);
                const u64 key64 = ((t << 40) | ((u64)h16 << 24) | ng) + 1; // layout shared with the int64 kernels
                atomicMax((i64*)prm.select_key + p, (i64)key64);
            }
            if (FUSED) { // what reduce_narrow's flush does for the count and the panic flag
                uint32_t c = 0, pn = 0;
#pragma unroll
                for (int w = 0; w < kWarps; ++w) {
                    const uint32_t x = wcnt[tid * kWarps + w];
                    c += x & 0x7fffffffu;
                    pn |= x >> 31;
                }
                if (c) atomicAdd((u64*)prm.reduce_sum + p, (u64)c);
                if (pn) raise_flag((i64*)prm.reduce_max + (i64)RR_FILTER_PANIC * prm.p_pad, p);
            }
        }
    }
}

// ---- launch geometry: one wave of CTAs, each with a long run of pod tiles ------------------------------------
int g_num_sms = 0;
thread_local int g_last_emit_launches = 1; // kernels the last launch_emit_narrow of this thread issued (the tail tile may go separately)

// CTAs per resident slot: 1 = a single wave of long CTAs; more = shorter CTAs the hardware scheduler can
// rebalance. Measured on C2 (profiles/): see kWavesReduce / kWavesEmit. FGS_WAVES_REDUCE / FGS_WAVES_EMIT override.
int env_waves(const char* name, int dflt) {
    const char* v = getenv(name);
    const int w = v ? atoi(v) : 0;
    return w > 0 ? w : dflt;
}

template <typename K>
int pods_per_cta_for(K kernel, int smem, const NarrowParams& p, int waves, int tile_n) {
    if (!g_num_sms) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&g_num_sms, cudaDevAttrMultiProcessorCount, dev);
    }
    int occ = 1;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kernel, kThreads, smem);
    if (occ < 1) occ = 1;
    const i64 slots = (i64)occ * g_num_sms * waves;
    const i64 node_tiles = p.n_pad / tile_n;
    const i64 chunks = slots / node_tiles > 0 ? slots / node_tiles : 1;
    i64 ppc = (p.p_pad + chunks - 1) / chunks;
    ppc = (ppc + kTileP - 1) / kTileP * kTileP;
    if (ppc < kTileP) ppc = kTileP;
    return (int)ppc;
}

template <uint32_t SPEC>
void launch_reduce_spec(NarrowParams p, cudaStream_t st) {
    const int sm = smem_reduce(SPEC & SP_SC_ALLOC).total;
    static int ready = 0;
    if (!ready) {
        cudaFuncSetAttribute(reduce_narrow<SPEC>, cudaFuncAttributeMaxDynamicSharedMemorySize, sm);
        ready = 1;
    }
    static const int waves = env_waves("FGS_WAVES_REDUCE", 12);
    p.pods_per_cta = pods_per_cta_for(reduce_narrow<SPEC>, sm, p, waves, kTileNr);
    const dim3 grid((unsigned)(p.n_pad / kTileNr), (unsigned)((p.p_pad + p.pods_per_cta - 1) / p.pods_per_cta), 1);
    reduce_narrow<SPEC><<<grid, kThreads, sm, st>>>(p);
}
template <uint32_t SPEC, bool FUSED>
void launch_emit_spec(NarrowParams p, cudaStream_t st) {
    const int sm = smem_emit(FUSED).total;
    static int ready = 0;
    if (!ready) {
        cudaFuncSetAttribute(emit_narrow<SPEC, FUSED>, cudaFuncAttributeMaxDynamicSharedMemorySize, sm);
        if (!FUSED) cudaFuncSetAttribute(emit_narrow<SPEC, false, !FUSED>, cudaFuncAttributeMaxDynamicSharedMemorySize, sm);
        ready = 1;
    }
    static const int waves = env_waves(FUSED ? "FGS_WAVES_FUSED" : "FGS_WAVES_EMIT", FUSED ? 6 : 24);
    static const int tail_on = []() { const char* v = getenv("FGS_EMIT_TAIL"); return !(v && v[0] == '0') ? 1 : 0; }();
    const i64 node_tiles = p.n_pad / kTileNe;
    const i64 tail_nodes = p.n_nodes - (node_tiles - 1) * kTileNe; // nodes in the last tile, 1 … 1024
    const int a = (int)((tail_nodes + 32 * kNPTe - 1) / (32 * kNPTe));
    const bool split_tail = !FUSED && tail_on && a <= 4;
    p.tile_first = 0;
    p.tail_groups = kWarps;
    g_last_emit_launches = (split_tail && node_tiles > 1) ? 2 : 1;
    if (!split_tail || node_tiles > 1) {
        NarrowParams q = p;
        if (split_tail) q.n_pad = p.n_pad - kTileNe; // geometry only: the grid below and the chunking skip the tail tile
        p.pods_per_cta = pods_per_cta_for(emit_narrow<SPEC, FUSED>, sm, q, waves, kTileNe);
        const dim3 grid((unsigned)(q.n_pad / kTileNe), (unsigned)((p.p_pad + p.pods_per_cta - 1) / p.pods_per_cta), 1);
        emit_narrow<SPEC, FUSED><<<grid, kThreads, sm, st>>>(p);
    }
    if constexpr (!FUSED) {
        if (split_tail) { // the tail tile: all eight warps on its one, two or four node groups
            NarrowParams q = p;
            q.n_pad = kTileNe;
            p.tile_first = (int)(node_tiles - 1);
            p.tail_groups = a <= 1 ? 1 : (a <= 2 ? 2 : 4);
            p.pods_per_cta = pods_per_cta_for(emit_narrow<SPEC, false, true>, sm, q, 4, kTileNe);
            const dim3 grid(1u, (unsigned)((p.p_pad + p.pods_per_cta - 1) / p.pods_per_cta), 1);
            emit_narrow<SPEC, false, true><<<grid, kThreads, sm, st>>>(p);
        }
    }
}

// ---- the maxima the fused pass needs up front --------------------------------------------------------------------
// DefaultNormalizeScore divides by the highest raw FlexGPU score among the pod's feasible nodes (normalize_score.go:27-35).
// That raw score does not depend on the pod beyond its kind — GPUScore(node) for a whole-GPU pod, MemScore(node) for a
// fractional one (gpu_node.go:179-199) — so with the nodes sorted by it, the maximum is the score of the FIRST feasible
// node in that order: one warp per pod walks the order 32 nodes at a time and almost always stops after the first step.
__global__ void __launch_bounds__(256) flexmax_prepass(const __grid_constant__ NarrowParams prm, const int* __restrict__ ord_g,
                                                       const int* __restrict__ ord_m) {
    const i64 p = ((i64)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (p >= prm.n_pods) return;
    auto P = [&](int c) { return __ldg(prm.pn + (i64)c * prm.p_pad + p); };
    int kind = P(PN_KIND);
    const bool flex_f = prm.fused_spec & SP_FLEX_F, fit_f = prm.fused_spec & SP_FIT_F;
    if (kind == 3) {
        if (flex_f) return; // both limits: FlexGPU.Filter rejects every node (flex_gpu.go:58-61)
        kind = 1;           // a gpu limit wins over a memory limit (flex_gpu.go:157-159)
    }
    if (kind != 1 && kind != 2) return; // no limit: the raw score is 0 on every node
    const int lim_g = P(PN_LIM_G), lim_m = P(PN_LIM_M), rc = P(PN_RC), rm = P(PN_RM), re = P(PN_RE), rs0 = P(PN_RS0), rs1 = P(PN_RS1);
    const int* ord = kind == 2 ? ord_m : ord_g;
    for (i64 base = 0; base < prm.n_nodes; base += 32) {
        const i64 k = base + lane;
        bool ok = false;
        int score = 0;
        if (k < prm.n_nodes) {
            const i64 n = ord[k];
            auto C = [&](int c) { return __ldg(prm.nn + (i64)c * prm.n_pad + n); };
            const uint32_t fl = (uint32_t)C(NN_FLAGS);
            ok = (fl & FGS_NF_VALID) != 0;
            if (flex_f) ok = ok && lim_g <= C(NN_THR_GPU) && lim_m <= C(NN_THR_MEM) && !(fl & FGS_NF_REF_PANIC);
            if (fit_f) ok = ok && (fl & NDF_PODS_OK) && rc <= C(NN_FREE_CPU) && rm <= C(NN_FREE_MEM) && re <= C(NN_FREE_EPH) && rs0 <= C(NN_FS0) && rs1 <= C(NN_FS1);
            score = C(kind == 2 ? NN_MSCORE : NN_GSCORE);
        }
        const uint32_t bal = __ballot_sync(0xffffffffu, ok);
        if (bal) {
            const int v = __shfl_sync(0xffffffffu, score, __ffs(bal) - 1);
            if (lane == 0) prm.reduce_max[(i64)RR_FLEX_MAX * prm.p_pad + p] = kind == 2 ? (i64)v * ((i64)1 << prm.sh.s[DIM_FM]) : (i64)v;
            return;
        }
    }
}

} // namespace

// reduce: the SPEC bits that matter are the two filters, "needs the flex max", "needs Allocatable hi/lo"
bool launch_reduce_narrow(const NarrowParams& p, uint32_t spec, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    switch (spec & (SP_FLEX_F | SP_FIT_F | SP_SC_FLEX | SP_SC_ALLOC)) {
#define RCASE(v)                      \
    case v:                           \
        launch_reduce_spec<v>(p, st); \
        return true;
        RCASE(SP_FLEX_F | SP_FIT_F | SP_SC_FLEX)
        RCASE(SP_FLEX_F | SP_FIT_F | SP_SC_FLEX | SP_SC_ALLOC)
        RCASE(SP_FLEX_F | SP_FIT_F)
        RCASE(SP_FLEX_F | SP_FIT_F | SP_SC_ALLOC)
        RCASE(SP_FLEX_F | SP_SC_FLEX)
        RCASE(SP_FIT_F)
        RCASE(SP_FIT_F | SP_SC_FLEX)
        RCASE(SP_FLEX_F)
#undef RCASE
        default:
            return false;
    }
}

bool narrow_fused_available(uint32_t spec) {
    switch (spec & (SP_SC_FLEX | SP_SC_FIT | SP_FIT_MOST | SP_SC_ALLOC | SP_W_EQ)) {
        case SP_SC_FLEX | SP_SC_FIT | SP_W_EQ:
        case SP_SC_FLEX | SP_SC_FIT:
        case SP_SC_FLEX | SP_SC_FIT | SP_FIT_MOST | SP_W_EQ:
        case SP_SC_FLEX | SP_SC_FIT | SP_FIT_MOST:
        case SP_SC_FLEX:
        case SP_SC_FIT | SP_W_EQ:
        case SP_SC_FIT:
            return true;
        default:
            return false;
    }
}

// One pass over pod × node: flexmax_prepass, then emit_narrow<SPEC, true> (filters + bitmap + count + scores + key).
// ord_g / ord_m: the node indices sorted by GPUScore / MemScore, highest first (launch_narrow_orders).
bool launch_fused_narrow(const NarrowParams& p0, uint32_t spec, const int* ord_g, const int* ord_m, void* stream, void* ev_between, int part) {
    cudaStream_t st = (cudaStream_t)stream;
    NarrowParams p = p0;
    p.fused_spec = spec;
    if (!narrow_fused_available(spec)) return false;
    if (part & 1) {
        if (spec & SP_SC_FLEX) flexmax_prepass<<<(unsigned)((p.n_pods * 32 + 255) / 256), 256, 0, st>>>(p, ord_g, ord_m);
        if (ev_between) cudaEventRecord((cudaEvent_t)ev_between, st);
    }
    if (!(part & 2)) return true;
    switch (spec & (SP_SC_FLEX | SP_SC_FIT | SP_FIT_MOST | SP_SC_ALLOC | SP_W_EQ)) {
#define FCASE(v)                          \
    case v:                               \
        launch_emit_spec<v, true>(p, st); \
        return true;
        FCASE(SP_SC_FLEX | SP_SC_FIT | SP_W_EQ)
        FCASE(SP_SC_FLEX | SP_SC_FIT)
        FCASE(SP_SC_FLEX | SP_SC_FIT | SP_FIT_MOST | SP_W_EQ)
        FCASE(SP_SC_FLEX | SP_SC_FIT | SP_FIT_MOST)
        FCASE(SP_SC_FLEX)
        FCASE(SP_SC_FIT | SP_W_EQ)
        FCASE(SP_SC_FIT)
#undef FCASE
        default:
            return false;
    }
}

int narrow_last_emit_launches() { return g_last_emit_launches; }

bool launch_emit_narrow(const NarrowParams& p, uint32_t spec, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    switch (spec & (SP_SC_FLEX | SP_SC_FIT | SP_FIT_MOST | SP_SC_ALLOC | SP_W_EQ)) {
#define ECASE(v)                           \
    case v:                                \
        launch_emit_spec<v, false>(p, st); \
        return true;
        ECASE(SP_SC_FLEX | SP_SC_FIT | SP_W_EQ) // C2 / C5: FlexGPU + LeastAllocated cpu:1 memory:1
        ECASE(SP_SC_FLEX | SP_SC_FIT)
        ECASE(SP_SC_FLEX | SP_SC_FIT | SP_FIT_MOST | SP_W_EQ)
        ECASE(SP_SC_FLEX | SP_SC_FIT | SP_FIT_MOST)
        ECASE(SP_SC_FLEX | SP_SC_ALLOC) // C1: FlexGPU + Allocatable
        ECASE(SP_SC_FLEX)               // the reference's own profile: FlexGPU only
        ECASE(SP_SC_FIT | SP_W_EQ)
        ECASE(SP_SC_FIT)
        ECASE(SP_SC_FLEX | SP_SC_FIT | SP_SC_ALLOC | SP_W_EQ)
        ECASE(SP_SC_FLEX | SP_SC_FIT | SP_SC_ALLOC)
        ECASE(SP_SC_ALLOC)
#undef ECASE
        default:
            return false;
    }
}

uint32_t narrow_spec_bits(bool flex_f, bool fit_f, bool sc_flex, bool sc_fit, bool fit_most, bool sc_alloc, bool w_eq) {
    return (flex_f ? SP_FLEX_F : 0) | (fit_f ? SP_FIT_F : 0) | (sc_flex ? SP_SC_FLEX : 0) | (sc_fit ? SP_SC_FIT : 0) |
           ((sc_fit && fit_most) ? SP_FIT_MOST : 0) | (sc_alloc ? SP_SC_ALLOC : 0) | ((sc_fit && w_eq) ? SP_W_EQ : 0);
}

} // namespace fgs
