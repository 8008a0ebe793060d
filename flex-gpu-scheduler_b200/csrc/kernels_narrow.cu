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
#include <cub/device/device_radix_sort.cuh>
#include <cuda_runtime.h>

#include <cstdint>

#include "kernels.cuh"
#include "kernels_narrow.cuh"
#include "node_patch.cuh"
#include "derive.cuh"

namespace fgs {

namespace {

typedef long long i64;
typedef unsigned long long u64;
constexpr i64 kI64Min = (i64)0x8000000000000000ull;
constexpr int kPass = (int)0x80000001;  // pod-side sentinel: "<= any node value" (INT_MIN itself is the node-side "never")
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

// What one derived node row adds to the analysis (valid nodes only): OR and MAX per resource dimension, the two flags,
// whether the per-pod error bits depend on the node (special_pass) and with which magnitude
struct RowAnalysis {
    u64 o[ND_DIMS];
    i64 mx[ND_DIMS];
    i64 araw, mag;
    int capzero, negscore, valid, special, panic;
};

__device__ __forceinline__ void analyze_row(const i64* nd, i64 n_pad, i64 n, RowAnalysis& a) {
    for (int d = 0; d < ND_DIMS; ++d) {
        a.o[d] = 0;
        a.mx[d] = 0;
    }
    a.araw = 0, a.mag = 0;
    a.capzero = a.negscore = a.valid = a.special = a.panic = 0;
    auto C = [&](int c) { return nd[(i64)c * n_pad + n]; };
    const uint32_t fl = (uint32_t)C(ND_FLAGS);
    if (!(fl & FGS_NF_VALID)) return;
    a.valid = 1;
    auto acc = [&](int d, i64 v) { // negative values are clamped to -1 by the narrowing, so only v >= 0 matters
        if (v > 0) {
            a.o[d] |= (u64)v;
            if (v > a.mx[d]) a.mx[d] = v;
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
    if (ms < 0) a.negscore = 1;
    if (ms < 0 || (fl & FGS_NF_REF_PANIC)) { // nodes the per-pod error bits depend on (special_pass)
        a.special = 1;
        a.panic = (fl & FGS_NF_REF_PANIC) ? 1 : 0;
        a.mag = ms < 0 ? -ms : 0;
    }
    if (C(ND_CAP_CPU) <= 0 || C(ND_CAP_MEM) <= 0) a.capzero = 1;
    a.araw = C(ND_ALLOC_RAW);
}

__global__ void __launch_bounds__(256) analyze_nodes(const i64* __restrict__ nd, i64 n_nodes, i64 n_pad, int S, NodeAnalysis* out,
                                                     i64* __restrict__ special) {
    const i64 n = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    (void)S;
    RowAnalysis ra;
    for (int d = 0; d < ND_DIMS; ++d) {
        ra.o[d] = 0;
        ra.mx[d] = 0;
    }
    ra.capzero = ra.negscore = ra.valid = 0;
    i64 amin = 0x7fffffffffffffffll, amax = kI64Min;
    if (n < n_nodes) {
        analyze_row(nd, n_pad, n, ra);
        if (ra.valid) {
            if (ra.special) {
                const int k = atomicAdd(&out->n_special, 1);
                special[k] = n | (ra.panic ? ((i64)1 << 40) : 0);
                special[n_pad + k] = ra.mag;
            }
            amin = ra.araw;
            amax = ra.araw;
        }
    }
    u64(&o)[ND_DIMS] = ra.o;
    i64(&mx)[ND_DIMS] = ra.mx;
    const int capzero = ra.capzero, negscore = ra.negscore, anyvalid = ra.valid;
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

// The few nodes that can put a pod's cycle into an error state — a negative MemScore (over-committed device) or
// a node NewGPUNode panics on — are handled here, outside the evaluation loops: one thread per pod looks up the
// feasibility bit of each such node in the bitmap the reduce pass wrote (finalize_pods turns the two rows into
// FGS_PODERR_* once the feasible count and the flex maximum are merged).
__global__ void __launch_bounds__(128) special_pass(const uint32_t* __restrict__ bitmap, i64 words_per_row, const i64* __restrict__ special,
                                                    i64 n_pad, int n_special, const i64* __restrict__ pods, i64* __restrict__ reduce_max,
                                                    i64 n_pods, i64 p_pad) {
    const i64 p = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n_pods) return;
    const uint32_t kind = (uint32_t)pods[(i64)PD_META * p_pad + p] & 3u;
    const uint32_t* row = bitmap + p * words_per_row;
    i64 neg = 0, panic = 0;
    for (int k = 0; k < n_special; ++k) {
        const i64 e = special[k];
        const i64 n = e & 0xffffffffll;
        if (!((row[n >> 5] >> (n & 31)) & 1u)) continue;
        const i64 mag = special[n_pad + k];
        if (kind == 2 && mag > neg) neg = mag; // MemScore is the raw score of fractional pods only (flex_gpu.go:160-162)
        if (e >> 40) panic = 1;
    }
    if (neg) reduce_max[(i64)RR_NEG_MAG * p_pad + p] = neg;
    if (panic) reduce_max[(i64)RR_SCORE_PANIC * p_pad + p] = 1;
}

__device__ __forceinline__ int narrow_free(i64 v, int sh) { return v < 0 ? -1 : (int)(v >> sh); }

template <typename NdPtr>
__device__ __forceinline__ void narrow_one_node(NdPtr nd, i64 n_pad, const NarrowShifts& sh, int fit_most, int* __restrict__ nn, i64 n) {
    auto C = [&](int c) { return nd[(i64)c * n_pad + n]; };
    auto W = [&](int c, int v) { nn[(i64)c * n_pad + n] = v; };
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
    int lc = 0, lm = 0;
    W(NN_MAG_C, (int)narrow_magic(capc > 0 ? (uint32_t)capc : 0u, &lc));
    W(NN_MAG_M, (int)narrow_magic(capm > 0 ? (uint32_t)capm : 0u, &lm));
    W(NN_MAG_SH, lc | (lm << 8));
}


__global__ void __launch_bounds__(256) narrow_nodes(const i64* __restrict__ nd, i64 n_pad, NarrowShifts sh, int fit_most, int* __restrict__ nn) {
    const i64 n = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= n_pad) return;
    narrow_one_node(nd, n_pad, sh, fit_most, nn, n);
}

// fgs_update_nodes for a few rows, one thread per row and no host round trip: the raw columns are stored, the row is derived
// again (derive_row, from the values handed in), narrowed under the shifts in force, and checked against the analysis the
// host decides by. A row that stays inside it — no bit below a dimension's lowest set bit, no value above its maximum, no
// new flag, and the same membership of the special list — changes nothing the host knows; any other row raises the
// flag, and the next call that needs the analysis redoes it for the whole table (Engine::resolve_patch). The small-batch
// kernel reads the flag itself and declines to run, so a patch and the cycle that follows need no synchronisation.
__global__ void __launch_bounds__(128) patch_nodes(const __grid_constant__ PatchParams prm) {
    const int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= prm.n_dirty) return;
    const i64 n = prm.rows ? prm.idx[r] : prm.inline_idx[r];
    const i64* row = prm.rows ? prm.rows + (i64)r * prm.ncols : &prm.inline_rows[r][0];
    const i64 np = prm.dp.n_pad;
    i64* nd = (i64*)prm.dp.nd;
    i64* raw = (i64*)prm.dp.raw;
    const uint32_t fl0 = (uint32_t)nd[(i64)ND_FLAGS * np + n];
    const bool special0 = (fl0 & FGS_NF_VALID) && (nd[(i64)ND_MEM_SCORE * np + n] < 0 || (fl0 & FGS_NF_REF_PANIC));
    for (int c = 0; c < prm.ncols; ++c) raw[(i64)c * np + n] = row[c];
    derive_row(prm.dp, row, 1, nd + n, np);
    RowAnalysis a;
    analyze_row(nd, np, n, a);
    bool inside = special0 == (bool)a.special;
    for (int d = 0; d < ND_DIMS; ++d) {
        const u64 ro = prm.ref.or_bits[d];
        const u64 below = ro ? ((ro & (0 - ro)) - 1) : ~0ull; // the bits under the lowest one any node had set
        if ((a.o[d] & below) || a.mx[d] > prm.ref.max_val[d]) inside = false;
    }
    if ((a.capzero | (a.negscore << 1)) & ~prm.ref.flags) inside = false;
    if (!inside) {
        *prm.d_flag = 1;
        *prm.h_flag = 1;
        return;
    }
    if (a.special) { // its entry of the special list, in place
        for (int k = 0; k < prm.ref.n_special; ++k)
            if ((prm.special[k] & 0xffffffffll) == n) {
                prm.special[k] = n | (a.panic ? ((i64)1 << 40) : 0);
                prm.special[np + k] = a.mag;
            }
    }
    if (prm.narrow_valid) narrow_one_node((const i64*)nd, np, prm.sh, prm.fit_most, prm.nn, n);
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

} // namespace

void launch_analyze_nodes(const int64_t* nd, int64_t n_nodes, int64_t n_pad, int S, NodeAnalysis* out, int64_t* special, void* stream) {
    if (n_nodes <= 0) return;
    analyze_nodes<<<(unsigned)((n_nodes + 255) / 256), 256, 0, (cudaStream_t)stream>>>((const i64*)nd, n_nodes, n_pad, S, out, (i64*)special);
}

void launch_special_pass(const uint32_t* bitmap, int64_t words_per_row, const int64_t* special, int64_t n_pad, int n_special,
                         const int64_t* pods, int64_t* reduce_max, int64_t n_pods, int64_t p_pad, void* stream) {
    if (n_special <= 0 || n_pods <= 0) return;
    special_pass<<<(unsigned)((n_pods + 127) / 128), 128, 0, (cudaStream_t)stream>>>(bitmap, words_per_row, (const i64*)special, n_pad, n_special,
                                                                                  (const i64*)pods, (i64*)reduce_max, n_pods, p_pad);
}

cudaError_t launch_patch_nodes(const PatchParams& p, void* stream) {
    patch_nodes<<<(unsigned)((p.n_dirty + 127) / 128), 128, 0, (cudaStream_t)stream>>>(p);
    return cudaGetLastError();
}

void launch_narrow_nodes(const int64_t* nd, int64_t n_pad, const NarrowShifts& sh, int fit_most, int* nn, void* stream) {
    narrow_nodes<<<(unsigned)((n_pad + 255) / 256), 256, 0, (cudaStream_t)stream>>>((const i64*)nd, n_pad, sh, fit_most, nn);
}

// Which pods would make the reference panic inside FlexGPU.Filter: a pod with a flex limit that gets as far as
// NewGPUNode (flex_gpu.go:101) on a node whose device state is inconsistent (gpu_node.go:40-58,109,115) — it passed the
// limit-sum checks (:85-98) and, when NodeResourcesFit runs first, Fit as well (RunFilterPlugins stops at the first
// failure, runtime/framework.go:667-693). The evaluation loops keep such nodes infeasible for these pods through their
// thresholds; this pass, over the few such nodes only, raises the pods' error flag.
__global__ void __launch_bounds__(128) filter_panic_pass(const int* __restrict__ nn, const int* __restrict__ pn, const i64* __restrict__ special,
                                                         i64 n_pad, int n_special, i64 n_pods, i64 p_pad, int fit_f, int flex_first,
                                                         i64* __restrict__ reduce_max) {
    const i64 p = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n_pods) return;
    auto P = [&](int c) { return pn[(i64)c * p_pad + p]; };
    const int kind = P(PN_KIND);
    if (kind != 1 && kind != 2) return; // no limit: Success before NewGPUNode (:54-56); both: rejected before it (:58-61)
    const int lim_g = P(PN_LIM_G), lim_m = P(PN_LIM_M), rc = P(PN_RC), rm = P(PN_RM), re = P(PN_RE), rs0 = P(PN_RS0), rs1 = P(PN_RS1);
    for (int k = 0; k < n_special; ++k) {
        const i64 e = special[k];
        if (!(e >> 40)) continue;
        const i64 n = e & 0xffffffffll;
        auto C = [&](int c) { return nn[(i64)c * n_pad + n]; };
        const uint32_t fl = (uint32_t)C(NN_FLAGS);
        if (!(fl & FGS_NF_VALID)) continue;
        if (!(lim_g <= C(NN_THR_GPU) && lim_m <= C(NN_THR_MEM))) continue;
        if (fit_f && !flex_first) {
            const bool fit_ok = (fl & NDF_PODS_OK) && rc <= C(NN_FREE_CPU) && rm <= C(NN_FREE_MEM) && re <= C(NN_FREE_EPH) && rs0 <= C(NN_FS0) && rs1 <= C(NN_FS1);
            if (!fit_ok) continue;
        }
        reduce_max[(i64)RR_FILTER_PANIC * p_pad + p] = 1;
        return;
    }
}

void launch_filter_panic_pass(const int* nn, const int* pn, const int64_t* special, int64_t n_pad, int n_special, int64_t n_pods, int64_t p_pad,
                              int fit_f, int flex_first, int64_t* reduce_max, void* stream) {
    if (n_special <= 0 || n_pods <= 0) return;
    filter_panic_pass<<<(unsigned)((n_pods + 127) / 128), 128, 0, (cudaStream_t)stream>>>(nn, pn, (const i64*)special, n_pad, n_special, n_pods, p_pad, fit_f,
                                                                                       flex_first, (i64*)reduce_max);
}

// ---- node orders for the fused pass: indices sorted by GPUScore / MemScore, highest first ---------------------------------
__global__ void iota_kernel(int* v, int n) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) v[i] = i;
}

size_t narrow_orders_scratch_bytes(int64_t n_nodes) {
    size_t tmp = 0;
    cub::DeviceRadixSort::SortPairsDescending(nullptr, tmp, (const int*)nullptr, (int*)nullptr, (const int*)nullptr, (int*)nullptr, (int)n_nodes);
    return ((tmp + 255) / 256) * 256 + 2 * sizeof(int) * (size_t)((n_nodes + 63) / 64 * 64);
}

// ord[0][n_pad] ← by NN_GSCORE, ord[1][n_pad] ← by NN_MSCORE (ties in any order: only the score of the first feasible
// node is used). scratch: narrow_orders_scratch_bytes(n_nodes).
void launch_narrow_orders(const int* nn, int64_t n_nodes, int64_t n_pad, int* ord, void* scratch, size_t scratch_bytes, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    const size_t nn64 = (size_t)((n_nodes + 63) / 64 * 64);
    int* iota = (int*)scratch;
    int* keys_out = iota + nn64;
    void* tmp = (void*)(keys_out + nn64);
    size_t tmp_bytes = scratch_bytes - 2 * sizeof(int) * nn64;
    if (n_nodes <= 0) return;
    iota_kernel<<<(unsigned)((n_nodes + 255) / 256), 256, 0, st>>>(iota, (int)n_nodes);
    cub::DeviceRadixSort::SortPairsDescending(tmp, tmp_bytes, nn + (i64)NN_GSCORE * n_pad, keys_out, iota, ord, (int)n_nodes, 0, 32, st);
    cub::DeviceRadixSort::SortPairsDescending(tmp, tmp_bytes, nn + (i64)NN_MSCORE * n_pad, keys_out, iota, ord + n_pad, (int)n_nodes, 0, 32, st);
}

void launch_narrow_pods(const int64_t* pods, int64_t p_pad, int S, const NarrowShifts& sh, uint32_t fit_ignored, int* pn, void* stream) {
    narrow_pods<<<(unsigned)((p_pad + 255) / 256), 256, 0, (cudaStream_t)stream>>>((const i64*)pods, p_pad, S, sh, fit_ignored, pn);
}

} // namespace fgs
