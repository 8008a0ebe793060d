// kernels_narrow.cu — the 32-bit specialisation of the evaluation kernels.
//
// The reference computes in int64, but on real clusters every quantity on this path is a multiple of
// a power of two that leaves fewer than 24 significant bits (memory in MiB/GiB, cpu in milli-cores,
// device memory in MiB …). When the engine can PROVE that for the uploaded node table and the staged
// pod batch (analyze_nodes + the host-side OR/MAX of the pod columns pick one exact shift per
// resource dimension; see Engine::decide_narrow), the tables are rewritten as int32 columns and
// these kernels run: same mapping as kernels.cu (512 nodes per CTA in registers, pod tiles through a
// TMA/mbarrier ring), a third of the instructions. Dividing numerator and denominator of
// floor(100·x/cap) by the same power of two, and both sides of r <= free, changes nothing, so the
// results stay bit-exact; any batch that does not qualify takes the int64 kernels instead.
//
// Plugin configuration is a template parameter (SPEC bits) so that the inner loop carries no
// configuration branches.
#include <cuda_runtime.h>

#include <cstdint>

#include "kernels.cuh"
#include "kernels_narrow.cuh"

namespace fgs {

namespace {

typedef long long i64;
typedef unsigned long long u64;
constexpr i64 kI64Min = (i64)0x8000000000000000ull;
constexpr int kPass = (int)0x80000000;  // pod-side sentinel: "<= anything"
constexpr int kNever = 0x7fffffff;      // pod-side sentinel: "<= nothing"

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
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(dst)),
                 "l"(__cvta_generic_to_global(src)), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void fence_mbar_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }

__device__ __forceinline__ int rn_small(float x) { return __float_as_int(x + 12582912.0f) - 0x4B400000; }

// floor(x / d), 0 <= x < 2^31, 0 < d < 2^24, est within 0.5 of x/d, quotient < 2^22: see qfix in kernels.cu
__device__ __forceinline__ int qfix32(int x, int d, float est) {
    const int k = rn_small(est);
    const int r = (int)((uint32_t)x - (uint32_t)k * (uint32_t)d); // wraps harmlessly on don't-care lanes
    return k + (r >> 31);
}

// ---- narrowing kernels --------------------------------------------------------------------------

// OR and MAX of the derived node columns per resource dimension (valid nodes only)
__global__ void __launch_bounds__(256) analyze_nodes(const i64* __restrict__ nd, i64 n_nodes, i64 n_pad, int S, NodeAnalysis* out) {
    const i64 n = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    u64 o[ND_DIMS];
    i64 mx[ND_DIMS];
    for (int d = 0; d < ND_DIMS; ++d) {
        o[d] = 0;
        mx[d] = 0;
    }
    i64 amin = 0x7fffffffffffffffll, amax = kI64Min;
    int capzero = 0, negscore = 0, anyvalid = 0;
    if (n < n_nodes) {
        auto C = [&](int c) { return nd[(i64)c * n_pad + n]; };
        const uint32_t fl = (uint32_t)C(ND_FLAGS);
        if (fl & FGS_NF_VALID) {
            anyvalid = 1;
            auto acc = [&](int d, i64 v) { // negative values are clamped to -1 by the narrowing, so only v >= 0 matters
                if (v > 0) {
                    o[d] |= (u64)v;
                    if (v > mx[d]) mx[d] = v;
                }
            };
            acc(DIM_CPU, C(ND_FREE_CPU));
            acc(DIM_CPU, C(ND_CAP_CPU));
            acc(DIM_CPU, C(ND_NZR_CPU));
            acc(DIM_MEM, C(ND_FREE_MEM));
            acc(DIM_MEM, C(ND_CAP_MEM));
            acc(DIM_MEM, C(ND_NZR_MEM));
            acc(DIM_EPH, C(ND_FREE_EPH));
            acc(DIM_S0, C(nd_free_scalar(0)));
            acc(DIM_S1, C(nd_free_scalar(1)));
            const i64 tg = C(ND_THR_GPU), tm = C(ND_THR_MEM);
            if (tg != kI64Min) acc(DIM_FG, tg);
            if (tm != kI64Min && tm != 0x7fffffffffffffffll) acc(DIM_FM, tm);
            acc(DIM_FG, C(ND_GPU_SCORE));
            const i64 ms = C(ND_MEM_SCORE);
            acc(DIM_FM, ms < 0 ? -ms : ms); // a negative MemScore (over-committed device) keeps its sign when narrowed
            if (ms < 0) negscore = 1;
            if (C(ND_CAP_CPU) <= 0 || C(ND_CAP_MEM) <= 0) capzero = 1;
            const i64 ar = C(ND_ALLOC_RAW);
            amin = ar;
            amax = ar;
        }
    }
    // block reduce through shared memory atomics, then one global atomic per block and value
    __shared__ u64 s_or[ND_DIMS];
    __shared__ i64 s_mx[ND_DIMS];
    __shared__ i64 s_amin, s_amax;
    __shared__ int s_flags;
    if (threadIdx.x == 0) {
        for (int d = 0; d < ND_DIMS; ++d) {
            s_or[d] = 0;
            s_mx[d] = 0;
        }
        s_amin = 0x7fffffffffffffffll;
        s_amax = kI64Min;
        s_flags = 0;
    }
    __syncthreads();
    for (int d = 0; d < ND_DIMS; ++d) {
        if (o[d]) atomicOr(&s_or[d], o[d]);
        if (mx[d]) atomicMax(&s_mx[d], mx[d]);
    }
    if (anyvalid) {
        atomicMin(&s_amin, amin);
        atomicMax(&s_amax, amax);
    }
    if (capzero | negscore) atomicOr(&s_flags, capzero | (negscore << 1));
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int d = 0; d < ND_DIMS; ++d) {
            if (s_or[d]) atomicOr((u64*)&out->or_bits[d], s_or[d]);
            if (s_mx[d]) atomicMax((i64*)&out->max_val[d], s_mx[d]);
        }
        atomicMin((i64*)&out->araw_min, s_amin);
        atomicMax((i64*)&out->araw_max, s_amax);
        if (s_flags) atomicOr(&out->flags, s_flags);
    }
}

__device__ __forceinline__ int narrow_free(i64 v, int sh) { return v < 0 ? -1 : (int)(v >> sh); }

__global__ void __launch_bounds__(256) narrow_nodes(const i64* __restrict__ nd, i64 n_pad, NarrowShifts sh, int fit_most, int* __restrict__ nn) {
    const i64 n = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= n_pad) return;
    auto C = [&](int c) { return nd[(i64)c * n_pad + n]; };
    auto W = [&](int c, int v) { nn[(i64)c * n_pad + n] = v; };
    auto WF = [&](int c, float v) { nn[(i64)c * n_pad + n] = __float_as_int(v); };
    const i64 tg = C(ND_THR_GPU), tm = C(ND_THR_MEM);
    W(NN_THR_GPU, tg == kI64Min ? -1 : (tg < 0 ? -1 : (int)(tg > 0x3fffffff ? 0x3fffffff : tg)));
    // thr_mem == INT64_MAX cannot happen (min with a finite headroom), negative → -1
    W(NN_THR_MEM, tm == kI64Min ? -1 : narrow_free(tm, sh.s[DIM_FM]));
    W(NN_FREE_CPU, narrow_free(C(ND_FREE_CPU), sh.s[DIM_CPU]));
    W(NN_FREE_MEM, narrow_free(C(ND_FREE_MEM), sh.s[DIM_MEM]));
    W(NN_FREE_EPH, narrow_free(C(ND_FREE_EPH), sh.s[DIM_EPH]));
    W(NN_FS0, narrow_free(C(nd_free_scalar(0)), sh.s[DIM_S0]));
    W(NN_FS1, narrow_free(C(nd_free_scalar(1)), sh.s[DIM_S1]));
    W(NN_FLAGS, (int)(uint32_t)C(ND_FLAGS));
    const int capc = (int)(C(ND_CAP_CPU) >> sh.s[DIM_CPU]), capm = (int)(C(ND_CAP_MEM) >> sh.s[DIM_MEM]);
    const i64 nzrc = C(ND_NZR_CPU) >> sh.s[DIM_CPU], nzrm = C(ND_NZR_MEM) >> sh.s[DIM_MEM];
    // Least: x = cap − nonzeroRequested (negative → -1: every request exceeds it); Most: x = nonzeroRequested
    i64 xc = fit_most ? nzrc : (i64)capc - nzrc, xm = fit_most ? nzrm : (i64)capm - nzrm;
    if (xc < 0) xc = -1;
    if (xm < 0) xm = -1;
    if (xc > 0x3fffffff) xc = 0x3fffffff; // Most on an over-committed node: any request saturates at 100
    if (xm > 0x3fffffff) xm = 0x3fffffff;
    W(NN_CAP_C, capc);
    W(NN_X_C, (int)xc);
    W(NN_CAP_M, capm);
    W(NN_X_M, (int)xm);
    const i64 ms64 = C(ND_MEM_SCORE);
    // exact in both signs: |MemScore| is a multiple of 2^shift (analyze_nodes ORs the magnitude)
    const int gs = (int)C(ND_GPU_SCORE), ms = ms64 < 0 ? -(int)((-ms64) >> sh.s[DIM_FM]) : (int)(ms64 >> sh.s[DIM_FM]);
    W(NN_GSCORE, gs);
    W(NN_MSCORE, ms);
    WF(NN_INV_C, capc > 0 ? (float)(100.0 / (double)capc) : 0.f);
    WF(NN_XQ_C, capc > 0 ? (float)(100.0 * (double)xc / (double)capc) : 0.f);
    WF(NN_INV_M, capm > 0 ? (float)(100.0 / (double)capm) : 0.f);
    WF(NN_XQ_M, capm > 0 ? (float)(100.0 * (double)xm / (double)capm) : 0.f);
    WF(NN_GS100F, (float)(100.0 * (double)gs));
    WF(NN_MS100F, (float)(100.0 * (double)ms));
}

__global__ void __launch_bounds__(256) narrow_pods(const i64* __restrict__ pods, i64 p_pad, int S, NarrowShifts sh, uint32_t fit_ignored,
                                                   int* __restrict__ pn) {
    const i64 p = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= p_pad) return;
    auto C = [&](int c) { return pods[(i64)c * p_pad + p]; };
    auto W = [&](int c, int v) { pn[(i64)c * p_pad + p] = v; };
    const u64 meta = (u64)C(PD_META);
    const uint32_t pf = (uint32_t)meta, kind = pf & 3u;
    const uint32_t smask = (uint32_t)(meta >> 32) & ~fit_ignored;
    const bool none = pf & FGS_PF_REQ_NONE;
    W(PN_LIM_G, kind == 1 ? (int)C(PD_GPU_LIM) : (kind == 3 ? kNever : kPass));
    W(PN_LIM_M, kind == 2 ? (int)(C(PD_MEM_LIM) >> sh.s[DIM_FM]) : (kind == 3 ? kNever : kPass));
    W(PN_RC, none ? kPass : (int)(C(PD_REQ_CPU) >> sh.s[DIM_CPU]));
    W(PN_RM, none ? kPass : (int)(C(PD_REQ_MEM) >> sh.s[DIM_MEM]));
    W(PN_RE, none ? kPass : (int)(C(PD_REQ_EPH) >> sh.s[DIM_EPH]));
    W(PN_RS0, (none || !(smask & 1u)) ? kPass : (int)(C(PD_FIXED_COLS + 0) >> sh.s[DIM_S0]));
    W(PN_RS1, (none || !(smask & 2u)) ? kPass : (int)(C(PD_FIXED_COLS + 1) >> sh.s[DIM_S1]));
    W(PN_NZC, (int)(C(PD_NZ_CPU) >> sh.s[DIM_CPU]));
    W(PN_NZM, (int)(C(PD_NZ_MEM) >> sh.s[DIM_MEM]));
    W(PN_KIND, (int)kind);
}

// ---- evaluation -----------------------------------------------------------------------------------

struct __align__(16) PodN { // 96 bytes, broadcast LDS.128 × 6
    int lim_g, lim_m, rc, rm;
    int re, rs0, rs1, kind;
    int nzc, nzm, nz100c, nz100m;
    float fnzc, fnzm, finv, ainv;
    int fmax, count, arange32, valid; // arange32 < 0: take the exact 64-bit path for Allocatable
    uint32_t hmul, _p0;
    i64 alo;
};
static_assert(sizeof(PodN) == 96, "PodN layout");

struct SmemN {
    int bars, raw, podn, stage_a, stage_b, total;
    int raw_stage_bytes;
};
__host__ __device__ inline SmemN smem_narrow(bool emit) {
    SmemN L;
    int off = 0;
    L.bars = off;
    off += 64;
    L.raw = off;
    L.raw_stage_bytes = PN_COLS * kTileP * 4 + (emit ? 4 * kTileP * 8 : 0);
    off += 2 * L.raw_stage_bytes;
    L.podn = off;
    off += kTileP * (int)sizeof(PodN);
    L.stage_a = off; // REDUCE: int wflex[kTileP][kWarps] + i64 whi[kTileP][kWarps] + i64 wnlo[kTileP][kWarps]; EMIT: u32 wkey
    off += kTileP * kWarps * 4 + 2 * kTileP * kWarps * 8;
    L.stage_b = off; // REDUCE: uint32 wcnt[kTileP][kWarps] + bm[kTileP][16]
    off += kTileP * kWarps * 4 + kTileP * 16 * 4;
    L.total = off;
    return L;
}

// SPEC bits
constexpr uint32_t SP_FLEX_F = 1, SP_FIT_F = 2, SP_SC_FLEX = 4, SP_SC_FIT = 8, SP_FIT_MOST = 16, SP_SC_ALLOC = 32, SP_W_EQ = 64;

template <uint32_t SPEC>
__global__ void __launch_bounds__(kThreads, 3) reduce_narrow(const __grid_constant__ NarrowParams prm) {
    extern __shared__ __align__(128) unsigned char smem[];
    const SmemN L = smem_narrow(false);
    uint64_t* bars = (uint64_t*)(smem + L.bars);
    PodN* s_pod = (PodN*)(smem + L.podn);
    int* wflex = (int*)(smem + L.stage_a);
    i64* whi = (i64*)(smem + L.stage_a + kTileP * kWarps * 4);
    i64* wnlo = whi + kTileP * kWarps;
    uint32_t* wcnt = (uint32_t*)(smem + L.stage_b);
    uint32_t* bm = wcnt + kTileP * kWarps;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const i64 node_tile0 = (i64)blockIdx.x * kTileN;
    const i64 pod_begin = (i64)blockIdx.y * prm.pods_per_cta;
    i64 pod_end = pod_begin + prm.pods_per_cta;
    if (pod_end > prm.p_pad) pod_end = prm.p_pad;
    const int n_tiles = (int)((pod_end - pod_begin + kTileP - 1) / kTileP);
    if (n_tiles <= 0) return;

    auto issue = [&](int it) {
        const int stage = it & 1;
        const i64 p0 = pod_begin + (i64)it * kTileP;
        int* dst = (int*)(smem + L.raw + stage * L.raw_stage_bytes);
        mbar_expect_tx(&bars[stage], (uint32_t)(PN_COLS * kTileP * 4));
        for (int c = 0; c < PN_COLS; ++c) bulk_g2s(dst + c * kTileP, prm.pn + (i64)c * prm.p_pad + p0, kTileP * 4, &bars[stage]);
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

    int thr_g[2], thr_m[2], fc[2], fm[2], fe[2], fs0[2], fs1[2], gsc[2], msc[2];
    uint32_t fl[2];
    i64 araw[2];
    int nloc[2];
#pragma unroll
    for (int j = 0; j < 2; ++j) {
        nloc[j] = warp * 64 + j * 32 + lane;
        const i64 n = node_tile0 + nloc[j];
        auto C = [&](int c) { return __ldg(prm.nn + (i64)c * prm.n_pad + n); };
        thr_g[j] = C(NN_THR_GPU);
        thr_m[j] = C(NN_THR_MEM);
        fc[j] = C(NN_FREE_CPU);
        fm[j] = C(NN_FREE_MEM);
        fe[j] = C(NN_FREE_EPH);
        fs0[j] = C(NN_FS0);
        fs1[j] = C(NN_FS1);
        fl[j] = (uint32_t)C(NN_FLAGS);
        gsc[j] = C(NN_GSCORE);
        msc[j] = C(NN_MSCORE);
        araw[j] = (SPEC & SP_SC_ALLOC) ? __ldg((const i64*)prm.nd + (i64)ND_ALLOC_RAW * prm.n_pad + n) : 0;
        if (!(SPEC & SP_FLEX_F)) thr_g[j] = thr_m[j] = kNever; // every limit passes
        if (!(SPEC & SP_FIT_F)) {
            fc[j] = fm[j] = fe[j] = fs0[j] = fs1[j] = kNever;
            fl[j] |= NDF_PODS_OK;
        }
    }
    const bool tile_has_panic = __syncthreads_or(((fl[0] | fl[1]) & FGS_NF_REF_PANIC) != 0);

    for (int it = 0; it < n_tiles; ++it) {
        const int stage = it & 1;
        const i64 tile_p0 = pod_begin + (i64)it * kTileP;
        mbar_wait(&bars[stage], (uint32_t)((it >> 1) & 1));
        if (tid < kTileP) {
            const int* raw = (const int*)(smem + L.raw + stage * L.raw_stage_bytes);
            auto RC = [&](int c) { return raw[c * kTileP + tid]; };
            PodN u;
            u.lim_g = RC(PN_LIM_G);
            u.lim_m = RC(PN_LIM_M);
            u.rc = RC(PN_RC);
            u.rm = RC(PN_RM);
            u.re = RC(PN_RE);
            u.rs0 = RC(PN_RS0);
            u.rs1 = RC(PN_RS1);
            u.kind = RC(PN_KIND);
            u.valid = (tile_p0 + tid) < prm.n_pods;
            u.nzc = u.nzm = u.nz100c = u.nz100m = u.fmax = u.count = u.arange32 = 0;
            u.fnzc = u.fnzm = u.finv = u.ainv = 0.f;
            u.hmul = 0;
            u._p0 = 0;
            u.alo = 0;
            s_pod[tid] = u;
        }
        __syncthreads();
        if (tid == 0 && it + 2 < n_tiles) {
            fence_proxy_async();
            issue(it + 2);
        }
        for (int jp = 0; jp < kTileP; ++jp) {
            const int4 a = *(const int4*)&s_pod[jp].lim_g;
            const int4 b = *(const int4*)&s_pod[jp].re;
            if (!s_pod[jp].valid) break;
            const int kind = b.w;
            bool feas[2];
            bool fpanic = false;
#pragma unroll
            for (int j = 0; j < 2; ++j) {
                // FlexGPU.Filter folded into two thresholds (flex_gpu.go:85-116); Fit.Filter (fit.go:256-329)
                const bool flex_ok = a.x <= thr_g[j] && a.y <= thr_m[j];
                const bool fit_ok = (fl[j] & NDF_PODS_OK) && a.z <= fc[j] && a.w <= fm[j] && b.x <= fe[j] && b.y <= fs0[j] && b.z <= fs1[j];
                bool ok = flex_ok && fit_ok && (fl[j] & FGS_NF_VALID);
                if (tile_has_panic) { // NewGPUNode would panic inside FlexGPU.Filter (flex_gpu.go:101)
                    const bool reach = (SPEC & SP_FLEX_F) && (kind == 1 || kind == 2) && flex_ok && (fl[j] & FGS_NF_REF_PANIC) &&
                                       (prm.flex_filter_first || fit_ok) && (fl[j] & FGS_NF_VALID);
                    fpanic = fpanic || reach;
                    ok = ok && !reach;
                }
                feas[j] = ok;
            }
            const uint32_t b0 = __ballot_sync(0xffffffffu, feas[0]);
            const uint32_t b1 = __ballot_sync(0xffffffffu, feas[1]);
            int vflex = kPass;
            if (SPEC & SP_SC_FLEX) {
                const int s0 = kind == 1 ? gsc[0] : (kind == 2 ? msc[0] : 0);
                const int s1 = kind == 1 ? gsc[1] : (kind == 2 ? msc[1] : 0);
                vflex = max(feas[0] ? s0 : kPass, feas[1] ? s1 : kPass);
                vflex = __reduce_max_sync(0xffffffffu, vflex);
            }
            i64 vhi = kI64Min, vnlo = kI64Min;
            if (SPEC & SP_SC_ALLOC) {
#pragma unroll
                for (int j = 0; j < 2; ++j)
                    if (feas[j]) {
                        vhi = araw[j] > vhi ? araw[j] : vhi;
                        vnlo = -araw[j] > vnlo ? -araw[j] : vnlo;
                    }
                // 64-bit max with two 32-bit REDUX each
                {
                    const int hi = (int)(vhi >> 32);
                    const int mhi = __reduce_max_sync(0xffffffffu, hi);
                    const uint32_t mlo = __reduce_max_sync(0xffffffffu, hi == mhi ? (uint32_t)vhi : 0u);
                    vhi = (i64)(((u64)(uint32_t)mhi << 32) | mlo);
                }
                {
                    const int hi = (int)(vnlo >> 32);
                    const int mhi = __reduce_max_sync(0xffffffffu, hi);
                    const uint32_t mlo = __reduce_max_sync(0xffffffffu, hi == mhi ? (uint32_t)vnlo : 0u);
                    vnlo = (i64)(((u64)(uint32_t)mhi << 32) | mlo);
                }
            }
            uint32_t anyp = 0;
            if (tile_has_panic) anyp = __ballot_sync(0xffffffffu, fpanic);
            if (lane == 0) {
                wflex[jp * kWarps + warp] = vflex;
                if (SPEC & SP_SC_ALLOC) {
                    whi[jp * kWarps + warp] = vhi;
                    wnlo[jp * kWarps + warp] = vnlo;
                }
                wcnt[jp * kWarps + warp] = (uint32_t)(__popc(b0) + __popc(b1)) | (anyp ? 0x80000000u : 0u);
                bm[jp * 16 + warp * 2 + 0] = b0;
                bm[jp * 16 + warp * 2 + 1] = b1;
            }
        }
        __syncthreads();
        // ---- flush ----
        if (tid < kTileP) {
            const i64 p = tile_p0 + tid;
            if (p < prm.n_pods) {
                uint32_t c = 0, pn = 0;
                int vf = kPass;
                i64 vh = kI64Min, vl = kI64Min;
                for (int w = 0; w < kWarps; ++w) {
                    const uint32_t x = wcnt[tid * kWarps + w];
                    c += x & 0x7fffffffu;
                    pn |= x >> 31;
                    const int f = wflex[tid * kWarps + w];
                    vf = f > vf ? f : vf;
                    if (SPEC & SP_SC_ALLOC) {
                        const i64 h = whi[tid * kWarps + w];
                        vh = h > vh ? h : vh;
                        const i64 l = wnlo[tid * kWarps + w];
                        vl = l > vl ? l : vl;
                    }
                }
                if (c) atomicAdd((u64*)prm.reduce_sum + p, (u64)c);
                if (pn) atomicMax((i64*)prm.reduce_max + (i64)RR_FILTER_PANIC * prm.p_pad + p, (i64)1);
                if ((SPEC & SP_SC_FLEX) && vf != kPass) {
                    // back to the reference's units: the reductions are merged across shards / read by either emit kernel
                    const int kind = s_pod[tid].kind;
                    const i64 v = kind == 2 ? (i64)vf * ((i64)1 << prm.sh.s[DIM_FM]) : (i64)vf;
                    atomicMax((i64*)prm.reduce_max + (i64)RR_FLEX_MAX * prm.p_pad + p, v);
                }
                if (SPEC & SP_SC_ALLOC) {
                    if (vh != kI64Min) atomicMax((i64*)prm.reduce_max + (i64)RR_ALLOC_HI * prm.p_pad + p, vh);
                    if (vl != kI64Min) atomicMax((i64*)prm.reduce_max + (i64)RR_ALLOC_NEGLO * prm.p_pad + p, vl);
                }
            }
        }
        const i64 word0 = node_tile0 / 32;
        for (int idx = tid; idx < kTileP * 16; idx += kThreads) {
            const int jp = idx >> 4, w = idx & 15;
            const i64 p = tile_p0 + jp;
            if (p < prm.n_pods && word0 + w < prm.words_per_row) prm.bitmap[p * prm.words_per_row + word0 + w] = bm[idx];
        }
        __syncthreads();
    }
}

template <uint32_t SPEC>
__global__ void __launch_bounds__(kThreads, 3) emit_narrow(const __grid_constant__ NarrowParams prm) {
    extern __shared__ __align__(128) unsigned char smem[];
    const SmemN L = smem_narrow(true);
    uint64_t* bars = (uint64_t*)(smem + L.bars);
    PodN* s_pod = (PodN*)(smem + L.podn);
    uint32_t* wkey = (uint32_t*)(smem + L.stage_a);

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const i64 node_tile0 = (i64)blockIdx.x * kTileN;
    const i64 pod_begin = (i64)blockIdx.y * prm.pods_per_cta;
    i64 pod_end = pod_begin + prm.pods_per_cta;
    if (pod_end > prm.p_pad) pod_end = prm.p_pad;
    const int n_tiles = (int)((pod_end - pod_begin + kTileP - 1) / kTileP);
    if (n_tiles <= 0) return;

    auto issue = [&](int it) {
        const int stage = it & 1;
        const i64 p0 = pod_begin + (i64)it * kTileP;
        int* dst = (int*)(smem + L.raw + stage * L.raw_stage_bytes);
        mbar_expect_tx(&bars[stage], (uint32_t)L.raw_stage_bytes);
        for (int c = 0; c < PN_COLS; ++c) bulk_g2s(dst + c * kTileP, prm.pn + (i64)c * prm.p_pad + p0, kTileP * 4, &bars[stage]);
        i64* d2 = (i64*)(dst + PN_COLS * kTileP);
        bulk_g2s(d2 + 0 * kTileP, prm.reduce_max + (i64)RR_FLEX_MAX * prm.p_pad + p0, kTileP * 8, &bars[stage]);
        bulk_g2s(d2 + 1 * kTileP, prm.reduce_max + (i64)RR_ALLOC_HI * prm.p_pad + p0, kTileP * 8, &bars[stage]);
        bulk_g2s(d2 + 2 * kTileP, prm.reduce_max + (i64)RR_ALLOC_NEGLO * prm.p_pad + p0, kTileP * 8, &bars[stage]);
        bulk_g2s(d2 + 3 * kTileP, prm.reduce_sum + p0, kTileP * 8, &bars[stage]);
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

    // node tile → registers
    int capc[2], xc[2], x100c[2], capm[2], xm[2], x100m[2], g100[2], m100[2];
    float invc[2], xqc[2], invm[2], xqm[2], g100f[2], m100f[2];
    uint32_t fl[2], nhash[2];
    i64 araw[2];
    int nloc[2];
    bool inrange[2];
#pragma unroll
    for (int j = 0; j < 2; ++j) {
        nloc[j] = warp * 64 + j * 32 + lane;
        const i64 n = node_tile0 + nloc[j];
        inrange[j] = n < prm.n_nodes;
        auto C = [&](int c) { return __ldg(prm.nn + (i64)c * prm.n_pad + n); };
        capc[j] = C(NN_CAP_C);
        xc[j] = C(NN_X_C);
        capm[j] = C(NN_CAP_M);
        xm[j] = C(NN_X_M);
        x100c[j] = 100 * xc[j];
        x100m[j] = 100 * xm[j];
        invc[j] = __int_as_float(C(NN_INV_C));
        xqc[j] = __int_as_float(C(NN_XQ_C));
        invm[j] = __int_as_float(C(NN_INV_M));
        xqm[j] = __int_as_float(C(NN_XQ_M));
        g100[j] = 100 * C(NN_GSCORE);
        m100[j] = 100 * C(NN_MSCORE);
        g100f[j] = __int_as_float(C(NN_GS100F));
        m100f[j] = __int_as_float(C(NN_MS100F));
        fl[j] = (uint32_t)C(NN_FLAGS);
        nhash[j] = (uint32_t)(prm.node_base + n) * 0x9E3779B1u;
        araw[j] = (SPEC & SP_SC_ALLOC) ? __ldg((const i64*)prm.nd + (i64)ND_ALLOC_RAW * prm.n_pad + n) : 0;
    }
    const bool tile_has_panic = __syncthreads_or(((fl[0] | fl[1]) & FGS_NF_REF_PANIC) != 0);
    const bool tile_has_neg = __syncthreads_or(((fl[0] | fl[1]) & NDF_MEMSCORE_NEG) != 0);
    const uint32_t* bm_base = prm.bitmap + node_tile0 / 32 + warp * 2;

    for (int it = 0; it < n_tiles; ++it) {
        const int stage = it & 1;
        const i64 tile_p0 = pod_begin + (i64)it * kTileP;
        mbar_wait(&bars[stage], (uint32_t)((it >> 1) & 1));
        if (tid < kTileP) {
            const int* raw = (const int*)(smem + L.raw + stage * L.raw_stage_bytes);
            const i64* raw64 = (const i64*)(raw + PN_COLS * kTileP);
            auto RC = [&](int c) { return raw[c * kTileP + tid]; };
            PodN u;
            u.lim_g = u.lim_m = u.rc = u.rm = u.re = u.rs0 = u.rs1 = 0;
            u.kind = RC(PN_KIND);
            u.nzc = RC(PN_NZC);
            u.nzm = RC(PN_NZM);
            u.nz100c = (int)(100u * (uint32_t)u.nzc);
            u.nz100m = (int)(100u * (uint32_t)u.nzm);
            u.fnzc = (float)u.nzc;
            u.fnzm = (float)u.nzm;
            u.valid = (tile_p0 + tid) < prm.n_pods;
            i64 fmx = raw64[0 * kTileP + tid];
            if (fmx < 0) fmx = 0; // maxCount starts at 0 (normalize_score.go:27)
            if (u.kind == 2) fmx >>= prm.sh.s[DIM_FM];
            u.fmax = (int)fmx;
            u.finv = fmx > 0 ? (float)(1.0 / (double)fmx) : 0.f;
            u.count = (int)raw64[3 * kTileP + tid];
            u.alo = 0;
            u.arange32 = 0;
            u.ainv = 0.f;
            if ((SPEC & SP_SC_ALLOC) && u.count > 0) {
                const i64 hi = raw64[1 * kTileP + tid], lo = -raw64[2 * kTileP + tid];
                const i64 range = (i64)((u64)hi - (u64)lo);
                u.alo = lo;
                if (range >= 0 && range < (1 << 24)) {
                    u.arange32 = (int)range;
                    u.ainv = range > 0 ? (float)(100.0 / (double)range) : 0.f;
                } else {
                    u.arange32 = -1;
                }
            }
            uint32_t h = (uint32_t)(prm.seed ^ (prm.seed >> 32)) ^ ((uint32_t)(tile_p0 + tid) * 0x85EBCA6Bu);
            h ^= h >> 15;
            u.hmul = h | 1u;
            u._p0 = 0;
            s_pod[tid] = u;
        }
        __syncthreads();
        if (tid == 0 && it + 2 < n_tiles) {
            fence_proxy_async();
            issue(it + 2);
        }

        // feasibility comes from the bitmap the reduce pass wrote (or an external PreScore mask): prefetch one pod ahead
        uint2 wnext = make_uint2(0, 0);
        if (tile_p0 < prm.n_pods) wnext = __ldg((const uint2*)(bm_base + tile_p0 * prm.words_per_row));
        for (int jp = 0; jp < kTileP; ++jp) {
            const i64 p = tile_p0 + jp;
            if (p >= prm.n_pods) break;
            const uint2 wcur = wnext;
            if (p + 1 < prm.n_pods && jp + 1 < kTileP) wnext = __ldg((const uint2*)(bm_base + (p + 1) * prm.words_per_row));
            const int4 q2 = *(const int4*)&s_pod[jp].nzc;    // nzc nzm nz100c nz100m
            const float4 q3 = *(const float4*)&s_pod[jp].fnzc; // fnzc fnzm finv ainv
            const int4 q4 = *(const int4*)&s_pod[jp].fmax;   // fmax count arange32 valid
            const int kind = s_pod[jp].kind;
            const uint32_t hmul = s_pod[jp].hmul;
            const bool feas0 = (wcur.x >> lane) & 1u, feas1 = (wcur.y >> lane) & 1u;
            uint32_t key = 0;
            int tot[2];
#pragma unroll
            for (int j = 0; j < 2; ++j) {
                int t = 0;
                const bool f = j == 0 ? feas0 : feas1;
                if (SPEC & SP_SC_FLEX) { // FlexGPU.Score + DefaultNormalizeScore(reverse), flex_gpu.go:142-176
                    int v = 100;
                    if (q4.x != 0 && kind != 0) {
                        const int s100 = kind == 1 ? g100[j] : m100[j];
                        const float sf = kind == 1 ? g100f[j] : m100f[j];
                        v = 100 - qfix32(s100, q4.x, sf * q3.z);
                        if (tile_has_neg && s100 < 0) {
                            // over-committed device: Go's '/' truncates toward zero like C++; the score leaves
                            // [0,100] and the framework aborts the cycle (runtime/framework.go:920-923)
                            v = 100 - s100 / q4.x;
                            if (f && q4.y >= 2 && (v > 100 || v < 0)) atomicMax((i64*)prm.select_key + prm.p_pad + p, (i64)1);
                        }
                    }
                    t += v * prm.w_flex;
                }
                if (SPEC & SP_SC_FIT) { // Least/MostAllocated over cpu and memory with non-zero requests
                    int sc, sm;
                    if (!(SPEC & SP_FIT_MOST)) {
                        sc = qfix32(x100c[j] - q2.z, capc[j], fmaf(-q3.x, invc[j], xqc[j]));
                        sm = qfix32(x100m[j] - q2.w, capm[j], fmaf(-q3.y, invm[j], xqm[j]));
                        sc = q2.x > xc[j] ? 0 : sc;
                        sm = q2.y > xm[j] ? 0 : sm;
                    } else {
                        sc = qfix32(x100c[j] + q2.z, capc[j], fmaf(q3.x, invc[j], xqc[j]));
                        sm = qfix32(x100m[j] + q2.w, capm[j], fmaf(q3.y, invm[j], xqm[j]));
                        sc = q2.x > capc[j] - xc[j] ? 100 : sc;
                        sm = q2.y > capm[j] - xm[j] ? 100 : sm;
                    }
                    int v;
                    if (SPEC & SP_W_EQ) {
                        v = (sc + sm) >> 1;
                    } else {
                        const int num = sc * prm.fit_w_cpu + sm * prm.fit_w_mem;
                        v = __float_as_int(__fadd_rd(((float)num + 0.5f) * prm.fit_inv_wsum, 8388608.0f)) & 0x7fffff;
                    }
                    t += v * prm.w_fit;
                }
                if (SPEC & SP_SC_ALLOC) { // Allocatable.NormalizeScore, allocatable.go:146-171
                    int v = 0;
                    if (q4.z > 0) {
                        const int a = (int)(uint32_t)((u64)araw[j] - (u64)s_pod[jp].alo);
                        v = qfix32((int)(100u * (uint32_t)a), q4.z, (float)a * q3.w);
                    } else if (q4.z < 0) {
                        const i64 alo = s_pod[jp].alo;
                        const i64 range = (i64)((u64)__ldg((const i64*)prm.reduce_max + (i64)RR_ALLOC_HI * prm.p_pad + p) - (u64)alo);
                        v = (int)((i64)((u64)((i64)((u64)araw[j] - (u64)alo)) * 100ull) / range);
                    }
                    t += v * prm.w_alloc;
                }
                tot[j] = t;
                const uint32_t h8 = (nhash[j] * hmul) >> 24;
                const uint32_t k = ((uint32_t)t << 17) | (h8 << 9) | (uint32_t)nloc[j];
                key = max(key, f ? k + 1u : 0u);
            }
            i64* row = (i64*)prm.scores + p * prm.score_stride + node_tile0;
            if (inrange[0]) __stcs(row + nloc[0], (i64)tot[0]);
            if (inrange[1]) __stcs(row + nloc[1], (i64)tot[1]);
            key = __reduce_max_sync(0xffffffffu, key);
            if (lane == 0) wkey[jp * kWarps + warp] = key;
            if (tile_has_panic && (SPEC & SP_SC_FLEX) && q4.y >= 2) { // FlexGPU.Score → NewGPUNode panics (flex_gpu.go:148)
                const bool pz = (feas0 && (fl[0] & FGS_NF_REF_PANIC)) || (feas1 && (fl[1] & FGS_NF_REF_PANIC));
                if (__ballot_sync(0xffffffffu, pz) && lane == 0) atomicMax((i64*)prm.select_key + 2 * prm.p_pad + p, (i64)1);
            }
        }
        __syncthreads();
        if (tid < kTileP) {
            const i64 p = tile_p0 + tid;
            if (p < prm.n_pods) {
                uint32_t k = 0;
                for (int w = 0; w < kWarps; ++w) {
                    const uint32_t x = wkey[tid * kWarps + w];
                    k = x > k ? x : k;
                }
                if (k) {
                    k -= 1;
                    const u64 t = k >> 17, h8 = (k >> 9) & 0xffu, nl = k & 0x1ffu;
                    const u64 ng = (u64)(prm.node_base + node_tile0) + nl;
                    const u64 key64 = ((t << 40) | (h8 << 32) | ng) + 1; // same layout as the int64 kernels: score | 16-bit hash | node
                    atomicMax((i64*)prm.select_key + p, (i64)key64);
                }
            }
        }
        __syncthreads();
    }
}

template <uint32_t SPEC>
void launch_reduce_spec(const NarrowParams& p, dim3 grid, cudaStream_t st) {
    static bool attr = false;
    if (!attr) {
        cudaFuncSetAttribute(reduce_narrow<SPEC>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_narrow(false).total);
        attr = true;
    }
    reduce_narrow<SPEC><<<grid, kThreads, smem_narrow(false).total, st>>>(p);
}
template <uint32_t SPEC>
void launch_emit_spec(const NarrowParams& p, dim3 grid, cudaStream_t st) {
    static bool attr = false;
    if (!attr) {
        cudaFuncSetAttribute(emit_narrow<SPEC>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_narrow(true).total);
        attr = true;
    }
    emit_narrow<SPEC><<<grid, kThreads, smem_narrow(true).total, st>>>(p);
}

} // namespace

// =================================================================================================

void launch_analyze_nodes(const int64_t* nd, int64_t n_nodes, int64_t n_pad, int S, NodeAnalysis* out, void* stream) {
    if (n_nodes <= 0) return;
    analyze_nodes<<<(unsigned)((n_nodes + 255) / 256), 256, 0, (cudaStream_t)stream>>>((const i64*)nd, n_nodes, n_pad, S, out);
}

void launch_narrow_nodes(const int64_t* nd, int64_t n_pad, const NarrowShifts& sh, int fit_most, int* nn, void* stream) {
    narrow_nodes<<<(unsigned)((n_pad + 255) / 256), 256, 0, (cudaStream_t)stream>>>((const i64*)nd, n_pad, sh, fit_most, nn);
}

void launch_narrow_pods(const int64_t* pods, int64_t p_pad, int S, const NarrowShifts& sh, uint32_t fit_ignored, int* pn, void* stream) {
    narrow_pods<<<(unsigned)((p_pad + 255) / 256), 256, 0, (cudaStream_t)stream>>>((const i64*)pods, p_pad, S, sh, fit_ignored, pn);
}

static dim3 narrow_grid(const NarrowParams& p) {
    return dim3((unsigned)(p.n_pad / kTileN), (unsigned)((p.p_pad + p.pods_per_cta - 1) / p.pods_per_cta), 1);
}

// reduce: the SPEC bits that matter are the two filters, "needs flex max", "needs allocatable hi/lo"
bool launch_reduce_narrow(const NarrowParams& p, uint32_t spec, void* stream) {
    const dim3 g = narrow_grid(p);
    cudaStream_t st = (cudaStream_t)stream;
    const uint32_t key = spec & (SP_FLEX_F | SP_FIT_F | SP_SC_FLEX | SP_SC_ALLOC);
    switch (key) {
#define RCASE(v) \
    case v:      \
        launch_reduce_spec<v>(p, g, st); \
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

bool launch_emit_narrow(const NarrowParams& p, uint32_t spec, void* stream) {
    const dim3 g = narrow_grid(p);
    cudaStream_t st = (cudaStream_t)stream;
    const uint32_t key = spec & (SP_SC_FLEX | SP_SC_FIT | SP_FIT_MOST | SP_SC_ALLOC | SP_W_EQ);
    switch (key) {
#define ECASE(v) \
    case v:      \
        launch_emit_spec<v>(p, g, st); \
        return true;
        ECASE(SP_SC_FLEX | SP_SC_FIT | SP_W_EQ)                               // C2 / C5: FlexGPU + LeastAllocated cpu:1 mem:1
        ECASE(SP_SC_FLEX | SP_SC_FIT)                                         // … other weights
        ECASE(SP_SC_FLEX | SP_SC_FIT | SP_FIT_MOST | SP_W_EQ)
        ECASE(SP_SC_FLEX | SP_SC_FIT | SP_FIT_MOST)
        ECASE(SP_SC_FLEX | SP_SC_ALLOC)                                       // C1: FlexGPU + Allocatable
        ECASE(SP_SC_FLEX)                                                     // the reference's own profile: FlexGPU only
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
