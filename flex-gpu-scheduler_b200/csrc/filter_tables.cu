// filter_tables.cu — pass A by table look-up (see filter_tables.cuh for the idea).
//
//   ft_build            node side, once per change of the node table: effective thresholds per dimension →
//                       radix sort → distinct levels + the rank of every node → bit rows T_d[r]; node orders by
//                       GPUScore / MemScore / Allocatable raw score.
//   filter_rows         pod side, per batch, ONE WARP PER POD: rank of the pod's value in every dimension (binary search
//                       in the levels), AND of the 5+S rows → the pod's bitmap row, popcount → len(feasibleNodes);
//                       DefaultNormalizeScore's maximum (normalize_score.go:27-35) and Allocatable's hi / lo
//                       (allocatable.go:146-159) are the score of the FIRST feasible node in score order — a raw score
//                       depends on the node and the pod's kind only (flex_gpu.go:153-165, gpu_node.go:179-199).
//
// The verdicts are those of eval_kernel<REDUCE> (kernels.cu) bit for bit: the same int64 thresholds, compared through
// their ranks. Reads of the tables are this GPU's own memory only.
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>
#include <cooperative_groups.h>
#include <cuda_runtime.h>

#include <algorithm>
#include <cstdint>
#include <cstdlib>

#include "filter_tables.cuh"
#include "kernels.cuh"
#include "kernels_narrow.cuh"

namespace cg = cooperative_groups;

namespace fgs {

namespace {

typedef int64_t i64;
typedef uint64_t u64;
constexpr i64 kMin = (i64)0x8000000000000000ull;
constexpr i64 kMax = (i64)0x7fffffffffffffffull;
// sentinels below every real quantity (|v| < 2^53, fgs_types.h):
constexpr i64 kLow0 = kMin;     // node side: the row takes no pod at all
constexpr i64 kLow1 = kMin + 1; // pod side: "no limit / no request" — passes wherever the node takes anything
constexpr i64 kLow2 = kMin + 2; // node side: takes only the pods without a value in this dimension

// ---- node side ---------------------------------------------------------------------------------------------------

// thr[d][n]: what a pod's value in dimension d must not exceed on node n; keys of the three score orders behind them
__global__ void __launch_bounds__(256) ft_thresholds(const i64* __restrict__ nd, i64 n_nodes, i64 n_pad, int S, uint32_t filter_mask,
                                                     i64* __restrict__ thr, i64 stride) {
    const i64 n = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= n_pad) return;
    const int n_dims = 5 + S;
    auto C = [&](int c) { return nd[(i64)c * n_pad + n]; };
    const uint32_t fl = (uint32_t)C(ND_FLAGS);
    const bool valid = n < n_nodes && (fl & FGS_NF_VALID);
    const bool flex_f = filter_mask & FGS_PL_FLEXGPU, fit_f = filter_mask & FGS_PL_FIT;
    // FlexGPU.Filter: limit <= per-kind threshold; a node on which NewGPUNode would panic takes no pod that carries a
    // flex limit (which pods REACH the panic is ft_filter_panic's business) but every pod without one
    auto flex_thr = [&](i64 t) -> i64 {
        if (!valid) return kLow0; // a row that is not a node takes nothing, whichever filters run
        if (!flex_f) return kMax;
        if ((fl & FGS_NF_REF_PANIC) || t == kMin) return kLow2;
        return t;
    };
    thr[(i64)FT_GPU * stride + n] = flex_thr(C(ND_THR_GPU));
    thr[(i64)FT_FMEM * stride + n] = flex_thr(C(ND_THR_MEM));
    // Fit.Filter: "Too many pods" (fit.go:259) rejects every pod, then request <= Allocatable − Requested per resource
    thr[(i64)FT_CPU * stride + n] = !fit_f ? kMax : (!(fl & NDF_PODS_OK) ? kLow0 : C(ND_FREE_CPU));
    thr[(i64)FT_MEM * stride + n] = !fit_f ? kMax : C(ND_FREE_MEM);
    thr[(i64)FT_EPH * stride + n] = !fit_f ? kMax : C(ND_FREE_EPH);
    for (int s = 0; s < S; ++s) thr[(i64)(FT_SCALAR0 + s) * stride + n] = !fit_f ? kMax : C(nd_free_scalar(s));
    thr[(i64)(n_dims + FT_ORD_GSCORE) * stride + n] = C(ND_GPU_SCORE);
    thr[(i64)(n_dims + FT_ORD_MSCORE) * stride + n] = C(ND_MEM_SCORE);
    thr[(i64)(n_dims + FT_ORD_ARAW) * stride + n] = C(ND_ALLOC_RAW);
}

__global__ void __launch_bounds__(256) ft_iota(int* v, i64 n) {
    const i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) v[i] = (int)i;
}

__global__ void __launch_bounds__(256) ft_heads(const i64* __restrict__ ks, i64 n, int* __restrict__ head) {
    const i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) head[i] = (i == 0 || ks[i] != ks[i - 1]) ? 1 : 0;
}

// sorted position i → rank = (number of distinct keys up to i) − 1; levels[rank] = the key; node_rank[node] = rank
__global__ void __launch_bounds__(256) ft_scatter(const i64* __restrict__ ks, const int* __restrict__ perm, const int* __restrict__ head,
                                                  const int* __restrict__ scan, i64 n, i64* __restrict__ levels, int* __restrict__ node_rank,
                                                  int* __restrict__ count) {
    const i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int r = scan[i] - 1;
    node_rank[perm[i]] = r;
    if (head[i]) levels[r] = ks[i];
    if (i == n - 1) *count = scan[i];
}

// Node tables of up to 16 384 (padded) rows: every sort of the build — the 5+S threshold columns ascending, the three score
// orders descending — in ONE launch, a CTA per column, bitonic in shared memory on (key, node index) pairs (the order a
// stable radix sort gives, so both forms of the build produce the same tables bit for bit), followed for the threshold columns
// by what ft_heads / InclusiveSum / ft_scatter do: rank = number of distinct keys before, levels[rank], node_rank[node], count.
constexpr int kSortThreads = 1024;
constexpr int kSortMax = 16384; // 16 384 x 12 bytes = 192 KB of the 227 KB a CTA may have
struct FtSortParams {
    const i64* thr; // [n_dims + FT_ORDERS][stride]
    i64 stride, n_pad, n_nodes;
    int n_dims;
    i64* levels;    // [n_dims][stride + 1]
    int* node_rank; // [n_dims][stride]
    int* counts;    // [n_dims]
    int* ord;       // [FT_ORDERS][stride]
    i64* ord_val;   // [FT_ORDERS][stride]
};
__global__ void __launch_bounds__(kSortThreads) ft_sort_small(const __grid_constant__ FtSortParams prm) {
    extern __shared__ __align__(16) unsigned char sort_smem[];
    const int col = blockIdx.x;
    const bool is_order = col >= prm.n_dims;
    const int n = (int)(is_order ? prm.n_nodes : prm.n_pad);
    int M = 1024;
    while (M < n) M <<= 1;
    i64* key = (i64*)sort_smem;
    int* val = (int*)(sort_smem + (size_t)M * 8);
    __shared__ int s_warp[kSortThreads / 32];
    const int tid = threadIdx.x;
    const i64* src = prm.thr + (i64)col * prm.stride;
    const i64 pad = is_order ? kMin : kMax;
    for (int i = tid; i < M; i += kSortThreads) {
        key[i] = i < n ? src[i] : pad;
        val[i] = i;
    }
    __syncthreads();
    // a before b: ascending by (key, index); the orders descending by key, ascending by index
    auto before = [&](i64 ka, int va, i64 kb, int vb) { return ka != kb ? (is_order ? ka > kb : ka < kb) : va < vb; };
    for (int k = 2; k <= M; k <<= 1) {
        for (int j = k >> 1; j > 0; j >>= 1) {
            for (int t = tid; t < (M >> 1); t += kSortThreads) {
                const int lo = ((t & ~(j - 1)) << 1) | (t & (j - 1)), hi = lo | j; // the pair (lo, lo + j)
                const bool up = (lo & k) == 0;
                const i64 ka = key[lo], kb = key[hi];
                const int va = val[lo], vb = val[hi];
                if (before(kb, vb, ka, va) == up) {
                    key[lo] = kb, key[hi] = ka;
                    val[lo] = vb, val[hi] = va;
                }
            }
            __syncthreads();
        }
    }
    if (is_order) {
        const int o = col - prm.n_dims;
        for (int i = tid; i < n; i += kSortThreads) {
            prm.ord[(i64)o * prm.stride + i] = val[i];
            prm.ord_val[(i64)o * prm.stride + i] = key[i];
        }
        return;
    }
    // heads and their inclusive scan: a thread owns M / 1024 consecutive positions
    const int per = M / kSortThreads, i0 = tid * per;
    int local = 0;
    for (int u = 0; u < per; ++u) {
        const int i = i0 + u;
        if (i < n && (i == 0 || key[i] != key[i - 1])) ++local;
    }
    int incl = local;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int v = __shfl_up_sync(0xffffffffu, incl, o);
        if ((tid & 31) >= o) incl += v;
    }
    if ((tid & 31) == 31) s_warp[tid >> 5] = incl;
    __syncthreads();
    if (tid < 32) {
        int w = s_warp[tid];
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int v = __shfl_up_sync(0xffffffffu, w, o);
            if (tid >= o) w += v;
        }
        s_warp[tid] = w;
    }
    __syncthreads();
    int run = incl - local + ((tid >> 5) ? s_warp[(tid >> 5) - 1] : 0); // distinct keys before this thread's first position
    i64* levels = prm.levels + (i64)col * (prm.stride + 1);
    int* node_rank = prm.node_rank + (i64)col * prm.stride;
    for (int u = 0; u < per; ++u) {
        const int i = i0 + u;
        if (i >= n) break;
        const bool head = i == 0 || key[i] != key[i - 1];
        if (head) levels[run] = key[i];
        run += head ? 1 : 0;
        node_rank[val[i]] = run - 1;
        if (i == n - 1) prm.counts[col] = run;
    }
}

// T_d[r][w] = bits of the nodes 32w..32w+31 whose rank in dimension d is >= r, for r = 0..R_d (row R_d is empty).
// One CTA = 32 consecutive words × 8 rows at a time; a thread keeps the 32 ranks of its word in registers.
constexpr int kFtRowsPerCta = 64;
struct FtTabParams {
    uint32_t* table;
    const int* node_rank; // [n_dims][stride]
    i64 stride, W;
    i64 row0[kFtMaxDims];
    int R[kFtMaxDims];
};
__global__ void __launch_bounds__(256) ft_fill_rows(const __grid_constant__ FtTabParams prm) {
    const int d = blockIdx.z;
    const int R = prm.R[d];
    const int r_begin = blockIdx.y * kFtRowsPerCta;
    if (r_begin > R) return;
    const i64 w = (i64)blockIdx.x * 32 + (threadIdx.x & 31);
    if (w >= prm.W) return;
    int rk[32];
    const int4* src = (const int4*)(prm.node_rank + (i64)d * prm.stride + w * 32);
#pragma unroll
    for (int q = 0; q < 8; ++q) {
        const int4 v = __ldg(src + q);
        rk[4 * q] = v.x, rk[4 * q + 1] = v.y, rk[4 * q + 2] = v.z, rk[4 * q + 3] = v.w;
    }
    const int r_end = min(R, r_begin + kFtRowsPerCta - 1);
    for (int r = r_begin + (threadIdx.x >> 5); r <= r_end; r += 8) {
        uint32_t word = 0;
#pragma unroll
        for (int b = 0; b < 32; ++b) word |= (rk[b] >= r ? 1u : 0u) << b;
        prm.table[(prm.row0[d] + r) * prm.W + w] = word;
    }
}

// ---- pod side ----------------------------------------------------------------------------------------------------

__device__ __forceinline__ i64 warp_max_i64(i64 v) {
    const int hi = (int)(v >> 32);
    const int mhi = __reduce_max_sync(0xffffffffu, hi);
    const uint32_t mlo = __reduce_max_sync(0xffffffffu, hi == mhi ? (uint32_t)v : 0u);
    return (i64)(((u64)(uint32_t)mhi << 32) | mlo);
}

constexpr int kWalkChunks = 24; // order walk: 32 nodes per step; after this many empty steps the set bits are scanned instead

// value of the first feasible node in an order (front) or of the last one (back); `row` = the pod's bitmap row in shared memory
__device__ __noinline__ bool walk_order(const int* __restrict__ ord, const i64* __restrict__ val, i64 n_nodes, const uint32_t* row, int lane,
                                           bool from_back, i64* out) {
    for (int step = 0; step < kWalkChunks; ++step) {
        const i64 k0 = (i64)step * 32 + lane;
        if ((i64)step * 32 >= n_nodes) return false;
        const i64 k = from_back ? n_nodes - 1 - k0 : k0;
        bool ok = false;
        if (k0 < n_nodes) {
            const int n = __ldg(ord + k);
            ok = (row[n >> 5] >> (n & 31)) & 1u;
        }
        const uint32_t bal = __ballot_sync(0xffffffffu, ok);
        if (bal) {
            const i64 kf = (i64)step * 32 + (__ffs(bal) - 1);
            *out = __ldg(val + (from_back ? n_nodes - 1 - kf : kf));
            return true;
        }
    }
    return false;
}

// the slow way round, bounded by the number of feasible nodes: max of col[n] (or of −col[n]) over the set bits
__device__ __noinline__ i64 scan_bits(const uint32_t* row, i64 W, const i64* __restrict__ col, int lane, bool negate) {
    i64 best = kMin;
    for (i64 w = lane; w < W; w += 32) {
        uint32_t x = row[w];
        while (x) {
            const int bit = __ffs(x) - 1;
            x &= x - 1;
            const i64 v = __ldg(col + w * 32 + bit);
            const i64 c = negate ? -v : v;
            best = c > best ? c : best;
        }
    }
    return warp_max_i64(best);
}

constexpr int kRowsWarps = 8;

// The row of every (pod, dimension): thread per pair — the pod's value in that dimension, its rank among the levels
// (lower bound: the number of levels below it), → row index in the table, or the all-ones row when the row cannot
// clear a bit: rank 0 (every node's threshold is at least the value), or the flex dimension the pod's kind does not
// use (its row is "every valid node", which the other flex row already implies).
__global__ void __launch_bounds__(256) ft_pod_rows(const __grid_constant__ FtRowsParams prm, int* __restrict__ rowidx) {
    const i64 p = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    const int d = blockIdx.y;
    if (p >= prm.p_pad) return;
    const int ones = (int)prm.v.ones_row;
    if (p >= prm.n_pods) {
        rowidx[(i64)d * prm.p_pad + p] = ones;
        return;
    }
    auto P = [&](int c) { return __ldg(prm.pods + (i64)c * prm.p_pad + p); };
    const u64 meta = (u64)P(PD_META);
    const uint32_t pf = (uint32_t)meta, kind = pf & 3u;
    const bool none = pf & FGS_PF_REQ_NONE;
    const uint32_t smask = (uint32_t)(meta >> 32) & ~prm.fit_ignored;
    i64 v;
    bool skip = false;
    if (d == FT_GPU) {
        v = kind == 1 ? P(PD_GPU_LIM) : (kind == 3 ? kMax : kLow1); // both limits: flex_gpu.go:58-61
        skip = kind == 2;
    } else if (d == FT_FMEM) {
        v = kind == 2 ? P(PD_MEM_LIM) : kLow1;
        skip = kind != 2;
    } else if (d == FT_CPU) v = none ? kLow1 : P(PD_REQ_CPU); // fit.go:270-275
    else if (d == FT_MEM) v = none ? kLow1 : P(PD_REQ_MEM);
    else if (d == FT_EPH) v = none ? kLow1 : P(PD_REQ_EPH);
    else v = (none || !((smask >> (d - FT_SCALAR0)) & 1u)) ? kLow1 : P(PD_FIXED_COLS + (d - FT_SCALAR0));
    int out = ones;
    if (!skip) {
        const i64* lev = prm.v.levels + (i64)d * prm.v.lev_stride;
        int lo = 0, hi = prm.v.R[d];
        while (lo < hi) {
            const int mid = (lo + hi) >> 1;
            if (__ldg(lev + mid) < v) lo = mid + 1;
            else hi = mid;
        }
        const bool flex_dim = d == FT_GPU || d == FT_FMEM; // never dropped: it also keeps rows that are not nodes out
        if (lo > 0 || flex_dim) out = (int)(prm.v.row0[d] + lo);
    }
    rowidx[(i64)d * prm.p_pad + p] = out;
}

// ONE WARP PER POD: AND of the pod's rows → bitmap row + count; first feasible node in score order → the maxima.
template <int NDIMS> // 0 = any number of dimensions (S > 2)
__global__ void __launch_bounds__(kRowsWarps * 32, NDIMS ? 5 : 4) filter_rows(const __grid_constant__ FtRowsParams prm, const int* __restrict__ rowidx) {
    extern __shared__ __align__(16) uint32_t s_rows[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const i64 p = (i64)blockIdx.x * kRowsWarps + warp;
    if (p >= prm.n_pods) return;
    const int n_dims = NDIMS ? NDIMS : prm.v.n_dims;
    const i64 W = prm.v.W;
    uint32_t* row = s_rows + (size_t)warp * W;

    // row offset of lane d's dimension in 16-byte units (rows are multiples of 128 bytes)
    const uint32_t my_off = lane < n_dims ? (uint32_t)(((i64)__ldg(rowidx + (i64)lane * prm.p_pad + p) * W) >> 2) : 0u;
    const uint32_t kind = (uint32_t)__ldg(prm.pods + (i64)PD_META * prm.p_pad + p) & 3u;
    // the first step of the order walks does not depend on the row: have its loads in flight now
    const int o_flex = (kind & 1u) ? FT_ORD_GSCORE : FT_ORD_MSCORE; // a gpu limit wins over a memory limit (flex_gpu.go:157-159)
    int n_f = 0, n_a = 0, n_b = 0;
    i64 v_f = 0, v_a = 0, v_b = 0;
    const bool want_flex = prm.sc_flex && kind != 0;
    if (lane < prm.n_nodes) {
        if (want_flex) {
            n_f = __ldg(prm.v.ord + (i64)o_flex * prm.n_pad + lane);
            v_f = __ldg(prm.v.ord_val + (i64)o_flex * prm.n_pad + lane);
        }
        if (prm.sc_alloc) {
            n_a = __ldg(prm.v.ord + (i64)FT_ORD_ARAW * prm.n_pad + lane);
            v_a = __ldg(prm.v.ord_val + (i64)FT_ORD_ARAW * prm.n_pad + lane);
            n_b = __ldg(prm.v.ord + (i64)FT_ORD_ARAW * prm.n_pad + prm.n_nodes - 1 - lane);
            v_b = __ldg(prm.v.ord_val + (i64)FT_ORD_ARAW * prm.n_pad + prm.n_nodes - 1 - lane);
        }
    }

    // ---- the row: AND over the dimensions, 4 words per lane and step (one 16-byte load per dimension) ---------------
    uint32_t cnt = 0;
    const uint4* T = (const uint4*)prm.v.table;
    uint4* out4 = (uint4*)(prm.bitmap + p * W);
    uint4* row4 = (uint4*)row;
    const int W4 = (int)(W >> 2);
    uint32_t off[NDIMS ? NDIMS : 1];
    if (NDIMS) {
#pragma unroll
        for (int d = 0; d < NDIMS; ++d) off[d] = __shfl_sync(0xffffffffu, my_off, d);
    }
    for (int c0 = 0; c0 < W4; c0 += 32) {
        const int c = c0 + lane;
        uint4 a = make_uint4(0xffffffffu, 0xffffffffu, 0xffffffffu, 0xffffffffu);
        if (NDIMS) {
            if (c < W4) {
                uint4 x[NDIMS ? NDIMS : 1];
#pragma unroll
                for (int d = 0; d < NDIMS; ++d) x[d] = __ldg(T + (size_t)off[d] + c);
#pragma unroll
                for (int d = 0; d < NDIMS; ++d) a.x &= x[d].x, a.y &= x[d].y, a.z &= x[d].z, a.w &= x[d].w;
            }
        } else {
            for (int d = 0; d < n_dims; ++d) {
                const uint32_t o = __shfl_sync(0xffffffffu, my_off, d);
                if (c < W4) {
                    const uint4 x = __ldg(T + (size_t)o + c);
                    a.x &= x.x, a.y &= x.y, a.z &= x.z, a.w &= x.w;
                }
            }
        }
        if (c < W4) {
            row4[c] = a;
            out4[c] = a;
            cnt += (uint32_t)(__popc(a.x) + __popc(a.y) + __popc(a.z) + __popc(a.w));
        }
    }
    cnt = __reduce_add_sync(0xffffffffu, cnt);
    __syncwarp();

    // ---- what NormalizeScore needs: the best raw scores among the feasible nodes ------------------------------------
    i64 vflex = kMin, vhi = kMin, vnlo = kMin;
    if (cnt > 0) {
        auto bit = [&](int n) { return (row[n >> 5] >> (n & 31)) & 1u; };
        auto first_of = [&](int n, i64 v, bool* found) -> i64 { // step 0 of a walk from the values loaded above
            const uint32_t bal = __ballot_sync(0xffffffffu, lane < prm.n_nodes && bit(n));
            *found = bal != 0;
            return __shfl_sync(0xffffffffu, v, bal ? __ffs(bal) - 1 : 0);
        };
        if (prm.sc_flex) { // raw FlexGPU score by pod kind (flex_gpu.go:153-165)
            if (kind == 0) {
                vflex = 0;
            } else {
                bool found;
                vflex = first_of(n_f, v_f, &found);
                if (!found && !walk_order(prm.v.ord + (i64)o_flex * prm.n_pad, prm.v.ord_val + (i64)o_flex * prm.n_pad, prm.n_nodes, row, lane, false, &vflex))
                    vflex = scan_bits(row, W, prm.nd + (i64)((kind & 1u) ? ND_GPU_SCORE : ND_MEM_SCORE) * prm.n_pad, lane, false);
            }
        }
        if (prm.sc_alloc) {
            const int* ord = prm.v.ord + (i64)FT_ORD_ARAW * prm.n_pad;
            const i64* val = prm.v.ord_val + (i64)FT_ORD_ARAW * prm.n_pad;
            const i64* col = prm.nd + (i64)ND_ALLOC_RAW * prm.n_pad;
            bool found;
            vhi = first_of(n_a, v_a, &found);
            if (!found && !walk_order(ord, val, prm.n_nodes, row, lane, false, &vhi)) vhi = scan_bits(row, W, col, lane, false);
            i64 lo = first_of(n_b, v_b, &found);
            if (found || walk_order(ord, val, prm.n_nodes, row, lane, true, &lo)) vnlo = -lo;
            else vnlo = scan_bits(row, W, col, lane, true);
        }
    }
    // ---- the nodes that feed the pod's error bits: lane k looks at special node k, k + 32, … -----------------------------
    // RR_FILTER_PANIC: the pod reaches NewGPUNode's panic inside FlexGPU.Filter on such a node (flex_gpu.go:101) — it has ONE
    // flex limit, passes the limit-sum checks (:85-98) and, when NodeResourcesFit runs first, Fit as well;
    // RR_NEG_MAG / RR_SCORE_PANIC: a FEASIBLE node with a negative MemScore (fractional pods only) / on which Score's
    // NewGPUNode panics (:148) — finalize_pods turns them into FGS_PODERR_* once counts and maxima are merged.
    i64 neg = 0;
    int spanic = 0, fpanic = 0;
    if (prm.n_special > 0) {
        const bool flex_f = prm.filter_mask & FGS_PL_FLEXGPU;
        const bool one_limit = kind == 1 || kind == 2;
        i64 lim = 0, rc = 0, rm = 0, re = 0;
        uint32_t smask = 0;
        bool none = false;
        const bool fit_first = (prm.filter_mask & FGS_PL_FIT) && !prm.flex_first;
        if (flex_f && one_limit) {
            auto P = [&](int c) { return __ldg(prm.pods + (i64)c * prm.p_pad + p); };
            const u64 meta = (u64)P(PD_META);
            lim = kind == 1 ? P(PD_GPU_LIM) : P(PD_MEM_LIM);
            none = (uint32_t)meta & FGS_PF_REQ_NONE;
            smask = (uint32_t)(meta >> 32) & ~prm.fit_ignored;
            if (fit_first) rc = P(PD_REQ_CPU), rm = P(PD_REQ_MEM), re = P(PD_REQ_EPH);
        }
        for (int k = lane; k < prm.n_special; k += 32) {
            const i64 e = __ldg(prm.special + k);
            const i64 n = e & 0xffffffffll;
            const bool is_panic = (e >> 40) != 0;
            if ((row[n >> 5] >> (n & 31)) & 1u) {
                const i64 mag = __ldg(prm.special + prm.n_pad + k);
                if (kind == 2 && mag > neg) neg = mag; // MemScore is the raw score of fractional pods only (flex_gpu.go:160-162)
                if (is_panic) spanic = 1;
            }
            if (is_panic && flex_f && one_limit && !fpanic) {
                auto C = [&](int c) { return __ldg(prm.nd + (i64)c * prm.n_pad + n); };
                const uint32_t fl = (uint32_t)C(ND_FLAGS);
                bool reach = (fl & FGS_NF_VALID) && lim <= C(kind == 1 ? ND_THR_GPU : ND_THR_MEM);
                if (reach && fit_first) {
                    bool fit_ok = fl & NDF_PODS_OK;
                    if (fit_ok && !none) {
                        fit_ok = rc <= C(ND_FREE_CPU) && rm <= C(ND_FREE_MEM) && re <= C(ND_FREE_EPH);
                        for (int sidx = 0; sidx < prm.S && fit_ok; ++sidx)
                            if (((smask >> sidx) & 1u) && __ldg(prm.pods + (i64)(PD_FIXED_COLS + sidx) * prm.p_pad + p) > C(nd_free_scalar(sidx))) fit_ok = false;
                    }
                    reach = fit_ok;
                }
                if (reach) fpanic = 1;
            }
        }
        neg = warp_max_i64(neg);
        spanic = __any_sync(0xffffffffu, spanic);
        fpanic = __any_sync(0xffffffffu, fpanic);
    }
    if (lane == 0) {
        prm.reduce_sum[p] = (i64)cnt;
        prm.reduce_max[(i64)RR_FLEX_MAX * prm.p_pad + p] = vflex;
        prm.reduce_max[(i64)RR_ALLOC_HI * prm.p_pad + p] = vhi;
        prm.reduce_max[(i64)RR_ALLOC_NEGLO * prm.p_pad + p] = vnlo;
        prm.reduce_max[(i64)RR_FILTER_PANIC * prm.p_pad + p] = fpanic ? 1 : 0;
        prm.reduce_max[(i64)RR_NEG_MAG * prm.p_pad + p] = neg;
        prm.reduce_max[(i64)RR_SCORE_PANIC * prm.p_pad + p] = spanic ? 1 : 0;
        for (int r = 0; r < 3; ++r) prm.select_key[(i64)r * prm.p_pad + p] = 0; // the emit pass max-accumulates into these
    }
}

// The bitmap ANDed with the caller's mask — the feasible list the framework hands to RunScorePlugins when other filter
// plugins of the profile rejected some of the engine's nodes — and everything NormalizeScore needs over what is left: one
// warp per pod, as filter_rows. With the node orders at hand the maxima are first-feasible walks, else the set bits are
// scanned.
__global__ void __launch_bounds__(kRowsWarps * 32) masked_rows(const __grid_constant__ FtMaskParams prm) {
    extern __shared__ __align__(16) uint32_t s_rows[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const i64 p = (i64)blockIdx.x * kRowsWarps + warp;
    if (p >= prm.n_pods) return;
    const i64 W = prm.W;
    uint32_t* row = s_rows + (size_t)warp * W;
    const uint32_t kind = (uint32_t)__ldg(prm.pods + (i64)PD_META * prm.p_pad + p) & 3u;
    uint4* bm4 = (uint4*)(prm.bitmap + p * W);
    const uint4* mk4 = (const uint4*)(prm.mask + p * W);
    uint4* row4 = (uint4*)row;
    const int W4 = (int)(W >> 2);
    uint32_t cnt = 0;
    for (int c = lane; c < W4; c += 32) {
        uint4 a = bm4[c];
        const uint4 m = __ldg(mk4 + c);
        a.x &= m.x, a.y &= m.y, a.z &= m.z, a.w &= m.w;
        row4[c] = a;
        bm4[c] = a;
        cnt += (uint32_t)(__popc(a.x) + __popc(a.y) + __popc(a.z) + __popc(a.w));
    }
    __syncwarp();
    cnt = __reduce_add_sync(0xffffffffu, cnt);
    i64 vflex = kMin, vhi = kMin, vnlo = kMin;
    if (cnt > 0) {
        if (prm.sc_flex) {
            if (kind == 0) {
                vflex = 0;
            } else {
                const int o = (kind & 1u) ? FT_ORD_GSCORE : FT_ORD_MSCORE;
                if (!(prm.have_orders && walk_order(prm.v.ord + (i64)o * prm.n_pad, prm.v.ord_val + (i64)o * prm.n_pad, prm.n_nodes, row, lane, false, &vflex)))
                    vflex = scan_bits(row, W, prm.nd + (i64)((kind & 1u) ? ND_GPU_SCORE : ND_MEM_SCORE) * prm.n_pad, lane, false);
            }
        }
        if (prm.sc_alloc) {
            const int* ord = prm.v.ord + (i64)FT_ORD_ARAW * prm.n_pad;
            const i64* val = prm.v.ord_val + (i64)FT_ORD_ARAW * prm.n_pad;
            const i64* col = prm.nd + (i64)ND_ALLOC_RAW * prm.n_pad;
            if (!(prm.have_orders && walk_order(ord, val, prm.n_nodes, row, lane, false, &vhi))) vhi = scan_bits(row, W, col, lane, false);
            i64 lo;
            if (prm.have_orders && walk_order(ord, val, prm.n_nodes, row, lane, true, &lo)) vnlo = -lo;
            else vnlo = scan_bits(row, W, col, lane, true);
        }
    }
    if (lane == 0) {
        prm.reduce_sum[p] = (i64)cnt;
        prm.reduce_max[(i64)RR_FLEX_MAX * prm.p_pad + p] = vflex;
        prm.reduce_max[(i64)RR_ALLOC_HI * prm.p_pad + p] = vhi;
        prm.reduce_max[(i64)RR_ALLOC_NEGLO * prm.p_pad + p] = vnlo;
        prm.reduce_max[(i64)RR_NEG_MAG * prm.p_pad + p] = 0; // rewritten by special_pass over the nodes that are left
        prm.reduce_max[(i64)RR_SCORE_PANIC * prm.p_pad + p] = 0;
        for (int r = 0; r < 3; ++r) prm.select_key[(i64)r * prm.p_pad + p] = 0; // the emit pass max-accumulates into these
    }
}

// ---- the whole cycle of one pod in one CTA (small batches) -------------------------------------------------------------------
constexpr int kSmallThreads = 1024; // one CTA per pod and few pods: the SM is the CTA's alone, latency is what counts
constexpr int kNoneI = (int)0x80000000, kPassI = (int)0x80000001, kNeverI = 0x7fffffff; // as in kernels_narrow.cu

__device__ __forceinline__ int qmagic32(uint32_t x2, uint32_t m, int l) { return (int)(__umulhi(x2, m) >> l); }

// lower bound (number of elements below v) in a sorted array by one warp: 32 probes per step instead of one — three dependent
// loads where a binary search over 5 000 levels needs thirteen
__device__ __forceinline__ int warp_lower_bound(const i64* __restrict__ lev, int n, i64 v, int lane) {
    int lo = 0, hi = n; // everything before lo is < v, everything from hi on is >= v
    while (hi - lo > 32) {
        const int span = hi - lo, stride = (span + 31) >> 5;
        int idx = lo + (lane + 1) * stride - 1;
        idx = idx < hi ? idx : hi - 1;
        const uint32_t below = __ballot_sync(0xffffffffu, __ldg(lev + idx) < v); // a prefix of the lanes
        const int c = __popc(below);
        const int last_true = lo + c * stride - 1, first_false = lo + (c + 1) * stride - 1;
        if (c > 0) lo = (last_true < hi ? last_true : hi - 1) + 1;
        if (c < 32) hi = first_false < hi ? first_false : hi - 1;
        if (lo > hi) lo = hi;
    }
    const int idx = lo + lane;
    const uint32_t below = __ballot_sync(0xffffffffu, idx < hi && __ldg(lev + idx) < v);
    return lo + __popc(below);
}

template <int CL>
__global__ void __launch_bounds__(kSmallThreads) small_cycle(const __grid_constant__ SmallParams prm) {
    extern __shared__ __align__(16) uint32_t s_row[]; // the pod's bitmap row, then reduction scratch
    __shared__ i64 s_pod[PD_FIXED_COLS + FGS_MAX_SCALARS];
    __shared__ uint32_t s_off[kFtMaxDims];
    __shared__ i64 s_red[4];      // flex max, alloc hi, alloc -lo, negative-score magnitude
    __shared__ int s_flags[2];    // score-time panic, filter-time panic
    __shared__ uint32_t s_cnt[kSmallThreads / 32];
    __shared__ u64 s_key[kSmallThreads / 32];
    __shared__ i64 s_xall[CL][8]; // CL > 1: row r = what CTA r's slice contributes (count, maxima, flags), stored here by CTA r itself
    __shared__ u64 s_keys[CL];    // CL > 1, rank 0: every slice's best selectHost key
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const i64 p = blockIdx.x / CL;
    const int rank = CL > 1 ? (int)cg::this_cluster().block_rank() : 0;
    const i64 W = prm.W, pp = prm.p_pad;
    const int S = prm.S, ncols = pd_cols(S), n_dims = 5 + S;
    // this CTA's slice of the node list, in whole 1024-node units
    i64 n_lo = 0, n_hi = prm.n_pad;
    if (CL > 1) {
        const i64 units = prm.n_pad >> 10, per = (units + CL - 1) / CL;
        n_lo = (i64)rank * per << 10;
        n_hi = n_lo + (per << 10);
        if (n_lo > prm.n_pad) n_lo = prm.n_pad;
        if (n_hi > prm.n_pad) n_hi = prm.n_pad;
    }
    (void)n_dims;

    if (prm.spec_flag && *(const volatile int*)prm.spec_flag) { // a patched row left the node analysis: the host redoes it and calls again
        if (tid == 0 && rank == 0 && atomicAdd(prm.d_done, 1u) == (unsigned)(prm.n_pods - 1)) {
            *prm.d_done = 0;
            __threadfence_system();
            *prm.h_flag = prm.seq | kSmallDeclined;
        }
        return; // every CTA of a cluster reads the same word: none is left waiting at a cluster barrier
    }
    // ---- the pod's columns: mapped host memory → shared memory and the device table ----------------------------------------
    if (tid < ncols) {
        const i64 v = prm.n_inline ? prm.pods_inline[p][tid] : prm.h_pods[(i64)tid * kSmallMaxPods + p];
        s_pod[tid] = v;
        if (rank == 0) prm.d_pods[(i64)tid * pp + p] = v;
    }
    __syncthreads();
    const u64 meta = (u64)s_pod[PD_META];
    const uint32_t pf = (uint32_t)meta, kind = pf & 3u;
    const bool none = pf & FGS_PF_REQ_NONE;
    const uint32_t smask = (uint32_t)(meta >> 32) & ~prm.fit_ignored;
    // int32 pod values, as narrow_pods derives them
    const int nzc = (int)(s_pod[PD_NZ_CPU] >> prm.sh[DIM_CPU]), nzm = (int)(s_pod[PD_NZ_MEM] >> prm.sh[DIM_MEM]);
    if (tid == 0 && rank == 0) {
        auto Wn = [&](int c, int v) { prm.d_pn[(i64)c * pp + p] = v; };
        Wn(PN_LIM_G, kind == 1 ? (int)s_pod[PD_GPU_LIM] : (kind == 3 ? kNeverI : kPassI));
        Wn(PN_LIM_M, kind == 2 ? (int)(s_pod[PD_MEM_LIM] >> prm.sh[DIM_FM]) : (kind == 3 ? kNeverI : kPassI));
        Wn(PN_RC, none ? kPassI : (int)(s_pod[PD_REQ_CPU] >> prm.sh[DIM_CPU]));
        Wn(PN_RM, none ? kPassI : (int)(s_pod[PD_REQ_MEM] >> prm.sh[DIM_MEM]));
        Wn(PN_RE, none ? kPassI : (int)(s_pod[PD_REQ_EPH] >> prm.sh[DIM_EPH]));
        Wn(PN_RS0, (none || !(smask & 1u)) ? kPassI : (int)(s_pod[PD_FIXED_COLS + 0] >> prm.sh[DIM_S0]));
        Wn(PN_RS1, (none || !(smask & 2u)) ? kPassI : (int)(s_pod[PD_FIXED_COLS + 1] >> prm.sh[DIM_S1]));
        Wn(PN_NZC, nzc);
        Wn(PN_NZM, nzm);
        Wn(PN_KIND, (int)kind);
    }
    const bool sc_flex = prm.score_mask & FGS_PL_FLEXGPU, sc_fit = prm.score_mask & FGS_PL_FIT, sc_alloc = prm.score_mask & FGS_PL_ALLOCATABLE;
    uint32_t count = 0;
    if (CL > 1 || prm.direct) {
        // ---- no level tables (the node table changed since they were built, and a handful of pods does not pay for a rebuild):
        // Filter by comparing, on the int32 node table as reduce_narrow does (flex_gpu.go:54-116, fit.go:256-329). A thread
        // takes four neighbouring nodes per step with 16-byte loads, eight lanes make one bitmap word.
        __shared__ i64 s_w[4][kSmallThreads / 32];
        __shared__ int s_wf[2][kSmallThreads / 32];
        const bool flex_f = prm.filter_mask & FGS_PL_FLEXGPU, fit_f = prm.filter_mask & FGS_PL_FIT;
        const int lim_g = kind == 1 ? (int)s_pod[PD_GPU_LIM] : (kind == 3 ? kNeverI : kPassI);
        const int lim_m = kind == 2 ? (int)(s_pod[PD_MEM_LIM] >> prm.sh[DIM_FM]) : (kind == 3 ? kNeverI : kPassI);
        const int rc = none ? kPassI : (int)(s_pod[PD_REQ_CPU] >> prm.sh[DIM_CPU]);
        const int rm = none ? kPassI : (int)(s_pod[PD_REQ_MEM] >> prm.sh[DIM_MEM]);
        const int re = none ? kPassI : (int)(s_pod[PD_REQ_EPH] >> prm.sh[DIM_EPH]);
        const int rs0 = (none || !(smask & 1u)) ? kPassI : (int)(s_pod[PD_FIXED_COLS + 0] >> prm.sh[DIM_S0]);
        const int rs1 = (none || !(smask & 2u)) ? kPassI : (int)(s_pod[PD_FIXED_COLS + 1] >> prm.sh[DIM_S1]);
        const bool has_limit = kind == 1 || kind == 2;
        i64 vflex = kMin, vhi = kMin, vnlo = kMin, neg = 0;
        int spanic = 0, fpanic = 0;
        uint32_t cnt = 0;
        for (i64 n0 = n_lo + (i64)tid * 4; n0 < n_hi; n0 += (i64)kSmallThreads * 4) { // warp-uniform trip count: whole 1024-node units
            auto C4 = [&](int c) { return __ldg((const int4*)(prm.nn + (i64)c * prm.n_pad + n0)); };
            const int4 fl4 = C4(NN_FLAGS), tg4 = C4(NN_THR_GPU), tm4 = C4(NN_THR_MEM), fc4 = C4(NN_FREE_CPU), fm4 = C4(NN_FREE_MEM);
            const int4 fe4 = C4(NN_FREE_EPH), f04 = C4(NN_FS0), f14 = C4(NN_FS1), gs4 = C4(NN_GSCORE), ms4 = C4(NN_MSCORE);
            auto at = [](const int4& v, int k) { return k == 0 ? v.x : (k == 1 ? v.y : (k == 2 ? v.z : v.w)); };
            uint32_t bits = 0;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const uint32_t fl = (uint32_t)at(fl4, k);
                const bool valid = fl & FGS_NF_VALID;
                const bool flex_ok = !flex_f || (lim_g <= at(tg4, k) && lim_m <= at(tm4, k));
                const bool fit_ok = !fit_f || ((fl & NDF_PODS_OK) && rc <= at(fc4, k) && rm <= at(fm4, k) && re <= at(fe4, k) && rs0 <= at(f04, k) &&
                                               rs1 <= at(f14, k));
                // NewGPUNode would panic inside FlexGPU.Filter (flex_gpu.go:101) when the filter gets that far on such a node
                const bool on_panic = flex_f && has_limit && (fl & FGS_NF_REF_PANIC);
                if (on_panic && valid && flex_ok && (prm.flex_first || fit_ok)) fpanic = 1;
                const bool feas = valid && flex_ok && fit_ok && !on_panic;
                if (feas) {
                    bits |= 1u << k;
                    const int ms = at(ms4, k);
                    const i64 sc = (kind & 1u) ? (i64)at(gs4, k) : (kind == 2 ? ((i64)ms << prm.sh[DIM_FM]) : 0); // exact: see narrow_nodes
                    vflex = sc > vflex ? sc : vflex;
                    if (sc_alloc) {
                        const i64 ar = __ldg(prm.nd + (i64)ND_ALLOC_RAW * prm.n_pad + n0 + k);
                        vhi = ar > vhi ? ar : vhi;
                        vnlo = -ar > vnlo ? -ar : vnlo;
                    }
                    if (kind == 2 && ms < 0) {
                        const i64 mag = (i64)(-(i64)ms) << prm.sh[DIM_FM];
                        neg = mag > neg ? mag : neg;
                    }
                    if (fl & FGS_NF_REF_PANIC) spanic = 1;
                }
            }
            uint32_t word = bits << (4 * (lane & 7));
            word |= __shfl_xor_sync(0xffffffffu, word, 1);
            word |= __shfl_xor_sync(0xffffffffu, word, 2);
            word |= __shfl_xor_sync(0xffffffffu, word, 4);
            if ((lane & 7) == 0) {
                s_row[n0 >> 5] = word;
                prm.bitmap[p * W + (n0 >> 5)] = word;
                cnt += (uint32_t)__popc(word);
            }
        }
        cnt += __shfl_xor_sync(0xffffffffu, cnt, 8);
        cnt += __shfl_xor_sync(0xffffffffu, cnt, 16);
        vflex = warp_max_i64(vflex), vhi = warp_max_i64(vhi), vnlo = warp_max_i64(vnlo), neg = warp_max_i64(neg);
        spanic = __any_sync(0xffffffffu, spanic);
        fpanic = __any_sync(0xffffffffu, fpanic);
        if (lane == 0) {
            s_cnt[warp] = cnt;
            s_w[0][warp] = vflex, s_w[1][warp] = vhi, s_w[2][warp] = vnlo, s_w[3][warp] = neg;
            s_wf[0][warp] = spanic ? 1 : 0, s_wf[1][warp] = fpanic ? 1 : 0;
        }
        __syncthreads();
#pragma unroll
        for (int w = 0; w < kSmallThreads / 32; ++w) count += s_cnt[w];
        if (warp == 0) {
            i64 r0 = warp_max_i64(s_w[0][lane]), r1 = warp_max_i64(s_w[1][lane]), r2 = warp_max_i64(s_w[2][lane]), r3 = warp_max_i64(s_w[3][lane]);
            const int f0 = __any_sync(0xffffffffu, s_wf[0][lane]), f1 = __any_sync(0xffffffffu, s_wf[1][lane]);
            if (lane == 0) {
                s_red[0] = sc_flex ? r0 : kMin, s_red[1] = sc_alloc ? r1 : kMin, s_red[2] = sc_alloc ? r2 : kMin, s_red[3] = r3;
                s_flags[0] = f0 ? 1 : 0, s_flags[1] = f1 ? 1 : 0;
            }
        }
        __syncthreads();
        if (CL > 1) { // the other slices' share, through distributed shared memory: every CTA stores its seven numbers into row
                      // `rank` of every CTA's table (lane r → CTA r), one cluster barrier, and reads its own copy
            cg::cluster_group cluster = cg::this_cluster();
            if (warp == 0 && lane < CL) {
                i64* dst = cluster.map_shared_rank(&s_xall[rank][0], lane);
                dst[0] = (i64)count, dst[1] = s_red[0], dst[2] = s_red[1], dst[3] = s_red[2], dst[4] = s_red[3];
                dst[5] = s_flags[0], dst[6] = s_flags[1];
            }
            cluster.sync();
            if (warp == 0) {
                const bool on = lane < CL;
                i64 c = on ? s_xall[on ? lane : 0][0] : 0;
                const i64 r0 = warp_max_i64(on ? s_xall[on ? lane : 0][1] : kMin), r1 = warp_max_i64(on ? s_xall[on ? lane : 0][2] : kMin);
                const i64 r2 = warp_max_i64(on ? s_xall[on ? lane : 0][3] : kMin), r3 = warp_max_i64(on ? s_xall[on ? lane : 0][4] : 0);
                const int f0 = __any_sync(0xffffffffu, on && s_xall[on ? lane : 0][5] != 0);
                const int f1 = __any_sync(0xffffffffu, on && s_xall[on ? lane : 0][6] != 0);
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
                if (lane == 0) {
                    s_red[0] = r0, s_red[1] = r1, s_red[2] = r2, s_red[3] = r3;
                    s_flags[0] = f0 ? 1 : 0, s_flags[1] = f1 ? 1 : 0;
                    s_cnt[0] = (uint32_t)c;
                }
            }
            __syncthreads();
            count = s_cnt[0];
        }
    } else {
    // ---- the row of every dimension (ft_pod_rows): warp d searches dimension d's levels ------------------------------------
    if (warp < n_dims) {
        const int d = warp;
        i64 v;
        bool skip = false;
        if (d == FT_GPU) {
            v = kind == 1 ? s_pod[PD_GPU_LIM] : (kind == 3 ? kMax : kLow1);
            skip = kind == 2;
        } else if (d == FT_FMEM) {
            v = kind == 2 ? s_pod[PD_MEM_LIM] : kLow1;
            skip = kind != 2;
        } else if (d == FT_CPU) v = none ? kLow1 : s_pod[PD_REQ_CPU];
        else if (d == FT_MEM) v = none ? kLow1 : s_pod[PD_REQ_MEM];
        else if (d == FT_EPH) v = none ? kLow1 : s_pod[PD_REQ_EPH];
        else v = (none || !((smask >> (d - FT_SCALAR0)) & 1u)) ? kLow1 : s_pod[PD_FIXED_COLS + (d - FT_SCALAR0)];
        i64 rowi = prm.v.ones_row;
        if (!skip) {
            const int lo = warp_lower_bound(prm.v.levels + (i64)d * prm.v.lev_stride, prm.v.R[d], v, lane);
            if (lo > 0 || d == FT_GPU || d == FT_FMEM) rowi = prm.v.row0[d] + lo;
        }
        if (lane == 0) s_off[d] = (uint32_t)((rowi * W) >> 2);
    }
    __syncthreads();
    // ---- the bitmap row and the feasible count ----------------------------------------------------------------------------
    {
        const uint4* T = (const uint4*)prm.v.table;
        uint4* out4 = (uint4*)(prm.bitmap + p * W);
        uint4* row4 = (uint4*)s_row;
        const int W4 = (int)(W >> 2);
        uint32_t cnt = 0;
        for (int c = tid; c < W4; c += kSmallThreads) {
            uint4 a = make_uint4(0xffffffffu, 0xffffffffu, 0xffffffffu, 0xffffffffu);
            for (int d = 0; d < n_dims; ++d) {
                const uint4 x = __ldg(T + (size_t)s_off[d] + c);
                a.x &= x.x, a.y &= x.y, a.z &= x.z, a.w &= x.w;
            }
            row4[c] = a;
            out4[c] = a;
            cnt += (uint32_t)(__popc(a.x) + __popc(a.y) + __popc(a.z) + __popc(a.w));
        }
        cnt = __reduce_add_sync(0xffffffffu, cnt);
        if (lane == 0) s_cnt[warp] = cnt;
    }
    __syncthreads();
#pragma unroll
    for (int w = 0; w < kSmallThreads / 32; ++w) count += s_cnt[w];
    // ---- NormalizeScore's maxima (warp 0) and the special nodes (warp 1): as filter_rows -----------------------------------
    if (warp == 0) {
        i64 vflex = kMin, vhi = kMin, vnlo = kMin;
        if (count > 0) {
            if (sc_flex) {
                if (kind == 0) {
                    vflex = 0;
                } else {
                    const int o = (kind & 1u) ? FT_ORD_GSCORE : FT_ORD_MSCORE;
                    if (!walk_order(prm.v.ord + (i64)o * prm.n_pad, prm.v.ord_val + (i64)o * prm.n_pad, prm.n_nodes, s_row, lane, false, &vflex))
                        vflex = scan_bits(s_row, W, prm.nd + (i64)((kind & 1u) ? ND_GPU_SCORE : ND_MEM_SCORE) * prm.n_pad, lane, false);
                }
            }
            if (sc_alloc) {
                const int* ord = prm.v.ord + (i64)FT_ORD_ARAW * prm.n_pad;
                const i64* val = prm.v.ord_val + (i64)FT_ORD_ARAW * prm.n_pad;
                const i64* col = prm.nd + (i64)ND_ALLOC_RAW * prm.n_pad;
                if (!walk_order(ord, val, prm.n_nodes, s_row, lane, false, &vhi)) vhi = scan_bits(s_row, W, col, lane, false);
                i64 lo;
                if (walk_order(ord, val, prm.n_nodes, s_row, lane, true, &lo)) vnlo = -lo;
                else vnlo = scan_bits(s_row, W, col, lane, true);
            }
        }
        if (lane == 0) s_red[0] = vflex, s_red[1] = vhi, s_red[2] = vnlo;
    } else if (warp == 1) {
        i64 neg = 0;
        int spanic = 0, fpanic = 0;
        if (prm.n_special > 0) {
            const bool flex_f = prm.filter_mask & FGS_PL_FLEXGPU;
            const bool one_limit = kind == 1 || kind == 2;
            const i64 lim = kind == 1 ? s_pod[PD_GPU_LIM] : s_pod[PD_MEM_LIM];
            const bool fit_first = (prm.filter_mask & FGS_PL_FIT) && !prm.flex_first;
            for (int k = lane; k < prm.n_special; k += 32) {
                const i64 e = __ldg(prm.special + k);
                const i64 n = e & 0xffffffffll;
                const bool is_panic = (e >> 40) != 0;
                if ((s_row[n >> 5] >> (n & 31)) & 1u) {
                    const i64 mag = __ldg(prm.special + prm.n_pad + k);
                    if (kind == 2 && mag > neg) neg = mag;
                    if (is_panic) spanic = 1;
                }
                if (is_panic && flex_f && one_limit && !fpanic) {
                    auto C = [&](int c) { return __ldg(prm.nd + (i64)c * prm.n_pad + n); };
                    const uint32_t fl = (uint32_t)C(ND_FLAGS);
                    bool reach = (fl & FGS_NF_VALID) && lim <= C(kind == 1 ? ND_THR_GPU : ND_THR_MEM);
                    if (reach && fit_first) {
                        bool fit_ok = fl & NDF_PODS_OK;
                        if (fit_ok && !none) {
                            fit_ok = s_pod[PD_REQ_CPU] <= C(ND_FREE_CPU) && s_pod[PD_REQ_MEM] <= C(ND_FREE_MEM) && s_pod[PD_REQ_EPH] <= C(ND_FREE_EPH);
                            for (int sidx = 0; sidx < S && fit_ok; ++sidx)
                                if (((smask >> sidx) & 1u) && s_pod[PD_FIXED_COLS + sidx] > C(nd_free_scalar(sidx))) fit_ok = false;
                        }
                        reach = fit_ok;
                    }
                    if (reach) fpanic = 1;
                }
            }
            neg = warp_max_i64(neg);
            spanic = __any_sync(0xffffffffu, spanic);
            fpanic = __any_sync(0xffffffffu, fpanic);
        }
        if (lane == 0) s_red[3] = neg, s_flags[0] = spanic ? 1 : 0, s_flags[1] = fpanic ? 1 : 0;
    }
    __syncthreads();
    }
    // ---- Score, NormalizeScore, weights, sum, selectHost key: emit_narrow's arithmetic, one node at a time --------------------
    const i64 raw_fmax = s_red[0];
    i64 fmx = raw_fmax < 0 ? 0 : raw_fmax; // maxCount starts at 0 (normalize_score.go:27)
    if (kind == 2) fmx >>= prm.sh[DIM_FM];
    const int fmax = (int)fmx;
    int fsh = 0, ash = 0;
    const uint32_t fmag = narrow_magic_fast((uint32_t)fmax, &fsh);
    i64 alo = 0, arange64 = 0;
    int arange32 = 0;
    uint32_t amag = 0;
    u64 arecip = 0;
    if (sc_alloc && count > 0) {
        const i64 hi = s_red[1], lo = -s_red[2];
        arange64 = (i64)((u64)hi - (u64)lo);
        alo = lo;
        if (arange64 >= 0 && arange64 < (1 << 24)) {
            arange32 = (int)arange64;
            amag = narrow_magic_fast((uint32_t)arange64, &ash);
        } else {
            arange32 = -1;
            arecip = alloc_recip(arange64);
        }
    }
    uint32_t h = (uint32_t)(prm.seed ^ (prm.seed >> 32)) ^ ((uint32_t)p * 0x85EBCA6Bu);
    h ^= h >> 15;
    h *= 0x9E3779B1u;
    const uint32_t hmul = h | 1u, ph8 = (h >> 20) & 0xffu;
    const int ks = (sc_flex && fmax != 0) ? ((kind & 1u) ? 1 : (kind == 2 ? 2 : 0)) : 0;
    const uint32_t nz200c = 200u * (uint32_t)nzc, nz200m = 200u * (uint32_t)nzm;
    u64 best = 0;
    i64* srow = prm.scores + p * prm.score_stride;
    // (unrolled: the column loads of several nodes in flight together — the loop is a chain of L2 round trips otherwise)
    const i64 n_end = n_hi < prm.n_nodes ? n_hi : prm.n_nodes;
#pragma unroll 4
    for (i64 n = n_lo + tid; n < n_end; n += kSmallThreads) {
        auto C = [&](int c) { return __ldg(prm.nn + (i64)c * prm.n_pad + n); };
        int t = 0;
        if (sc_flex) {
            int v = 100;
            if (ks == 1) {
                v = 100 - qmagic32(200u * (uint32_t)C(NN_GSCORE), fmag, fsh);
            } else if (ks == 2) {
                const int ms = C(NN_MSCORE);
                const int q = qmagic32(200u * (uint32_t)(ms < 0 ? -ms : ms), fmag, fsh);
                v = ms < 0 ? 100 + q : 100 - q;
            }
            t += v * prm.w_flex;
        }
        if (sc_fit) {
            const int capc = C(NN_CAP_C), xc = C(NN_X_C), capm = C(NN_CAP_M), xm = C(NN_X_M);
            const uint32_t magc = (uint32_t)C(NN_MAG_C), magm = (uint32_t)C(NN_MAG_M);
            const int shs = C(NN_MAG_SH), shc = shs & 0xff, shm = (shs >> 8) & 0xff;
            int sc, sm;
            if (!prm.fit_most) {
                sc = qmagic32(200u * (uint32_t)xc - nz200c, magc, shc);
                sm = qmagic32(200u * (uint32_t)xm - nz200m, magm, shm);
                sc = nzc > xc ? 0 : sc;
                sm = nzm > xm ? 0 : sm;
            } else {
                sc = qmagic32(200u * (uint32_t)xc + nz200c, magc, shc);
                sm = qmagic32(200u * (uint32_t)xm + nz200m, magm, shm);
                sc = nzc > capc - xc ? 100 : sc;
                sm = nzm > capm - xm ? 100 : sm;
            }
            int v;
            if (prm.w_eq) {
                v = (sc + sm) >> 1;
            } else {
                const int num = sc * prm.fit_w_cpu + sm * prm.fit_w_mem;
                v = __float_as_int(__fadd_rd(((float)num + 0.5f) * prm.fit_inv_wsum, 8388608.0f)) & 0x7fffff;
            }
            t += v * prm.w_fit;
        }
        if (sc_alloc) {
            int v = 0;
            const i64 araw = __ldg(prm.nd + (i64)ND_ALLOC_RAW * prm.n_pad + n);
            if (arange32 > 0) {
                const int a = (int)(uint32_t)((u64)araw - (u64)alo);
                v = qmagic32(200u * (uint32_t)a, amag, ash);
            } else if (arange32 < 0) {
                v = alloc_div100((i64)((u64)araw - (u64)alo), arange64, arecip);
            }
            t += v * prm.w_alloc;
        }
        srow[n] = (i64)(u64)(uint32_t)t;
        if ((s_row[n >> 5] >> (n & 31)) & 1u) { // selectHost candidate: the key emit_narrow's tiles arrive at
            const uint32_t nloc = (uint32_t)(n & 1023), tile = (uint32_t)(n >> 10);
            const uint32_t h8 = ((nloc * 0x9E3779B1u) >> 24) ^ ph8;
            const uint32_t h16 = (h8 << 8) | ((tile * 0x9E3779B1u ^ hmul) >> 24);
            const u64 key = (((u64)(uint32_t)t << 40) | ((u64)h16 << 24) | (u64)(prm.node_base + n)) + 1;
            best = key > best ? key : best;
        }
    }
    best = (u64)warp_max_i64((i64)best);
    if (lane == 0) s_key[warp] = best;
    __syncthreads();
    // ---- the per-pod rows and the result record (finalize_pods) ---------------------------------------------------------------
    if (CL > 1) { // the cluster's best key: every slice's goes into rank 0's table; after the barrier only rank 0 has work left
        if (tid == 0) {
            u64 key = 0;
#pragma unroll
            for (int w = 0; w < kSmallThreads / 32; ++w) key = s_key[w] > key ? s_key[w] : key;
            *cg::this_cluster().map_shared_rank(&s_keys[rank], 0) = key;
        }
        cg::this_cluster().sync();
        if (rank != 0) return; // (nobody reads a finished CTA's shared memory: the stores above went the other way)
    }
    if (tid == 0) {
        u64 key = 0;
        if (CL > 1) {
#pragma unroll
            for (int r = 0; r < CL; ++r) key = s_keys[r] > key ? s_keys[r] : key;
        } else {
#pragma unroll
            for (int w = 0; w < kSmallThreads / 32; ++w) key = s_key[w] > key ? s_key[w] : key;
        }
        i64* rm = prm.perpod;
        rm[(i64)RR_FLEX_MAX * pp + p] = raw_fmax;
        rm[(i64)RR_ALLOC_HI * pp + p] = s_red[1];
        rm[(i64)RR_ALLOC_NEGLO * pp + p] = s_red[2];
        rm[(i64)RR_FILTER_PANIC * pp + p] = s_flags[1];
        rm[(i64)RR_NEG_MAG * pp + p] = s_red[3];
        rm[(i64)RR_SCORE_PANIC * pp + p] = s_flags[0];
        rm[(i64)FGS_REDUCE_ROWS * pp + p] = (i64)count;
        rm[(i64)(FGS_REDUCE_ROWS + 1) * pp + p] = (i64)key;
        rm[(i64)(FGS_REDUCE_ROWS + 2) * pp + p] = 0;
        rm[(i64)(FGS_REDUCE_ROWS + 3) * pp + p] = 0;
        uint32_t err = 0;
        if (s_flags[1]) err |= FGS_PODERR_REF_PANIC;
        if (sc_flex && count >= 2) {
            if (s_flags[0]) err |= FGS_PODERR_REF_PANIC;
            if (raw_fmax > 0 && s_red[3] > 0 && 100 * s_red[3] >= raw_fmax) err |= FGS_PODERR_SCORE_RANGE;
        }
        i64 sel = -1, score = 0;
        if (key > 0) {
            const u64 k = key - 1;
            sel = (i64)(k & 0xffffffu);
            score = (i64)(k >> 40);
        }
        const i64 w0 = (i64)((u64)(uint32_t)(count | (err << 28)) | ((u64)(uint32_t)(int32_t)sel << 32));
        rm[(i64)(FGS_REDUCE_ROWS + 4) * pp + p] = w0;
        rm[(i64)(FGS_REDUCE_ROWS + 5) * pp + p] = score;
        prm.h_compact[p] = w0;
        prm.h_compact[kSmallMaxPods + p] = score;
        __threadfence_system();
        if (prm.n_pods == 1) { // the record above is the only one
            *prm.h_flag = prm.seq;
        } else if (atomicAdd(prm.d_done, 1u) == (unsigned)(prm.n_pods - 1)) { // every other pod's record is out: tell the host
            *prm.d_done = 0;
            __threadfence_system();
            *prm.h_flag = prm.seq;
        }
    }
}

// Which pods would make the reference panic inside FlexGPU.Filter (see filter_panic_pass in kernels_narrow.cu): a pod
// with ONE flex limit that passes the limit-sum checks (flex_gpu.go:85-98) on a node whose device state makes NewGPUNode
// panic (:101, gpu_node.go:40-58,109,115) and, when NodeResourcesFit runs first, passes Fit too (runtime/framework.go:667-693).
__global__ void __launch_bounds__(128) ft_filter_panic(const i64* __restrict__ nd, const i64* __restrict__ pods, const i64* __restrict__ special,
                                                       i64 n_pad, int n_special, i64 n_pods, i64 p_pad, int S, uint32_t filter_mask,
                                                       uint32_t fit_ignored, int flex_first, i64* __restrict__ reduce_max) {
    const i64 p = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n_pods) return;
    auto P = [&](int c) { return pods[(i64)c * p_pad + p]; };
    const u64 meta = (u64)P(PD_META);
    const uint32_t pf = (uint32_t)meta, kind = pf & 3u;
    if (kind != 1 && kind != 2) return;
    const i64 lim = kind == 1 ? P(PD_GPU_LIM) : P(PD_MEM_LIM);
    const bool none = pf & FGS_PF_REQ_NONE;
    const uint32_t smask = (uint32_t)(meta >> 32) & ~fit_ignored;
    const bool fit_first = (filter_mask & FGS_PL_FIT) && !flex_first;
    for (int k = 0; k < n_special; ++k) {
        const i64 e = special[k];
        if (!(e >> 40)) continue;
        const i64 n = e & 0xffffffffll;
        auto C = [&](int c) { return nd[(i64)c * n_pad + n]; };
        const uint32_t fl = (uint32_t)C(ND_FLAGS);
        if (!(fl & FGS_NF_VALID)) continue;
        if (!(lim <= C(kind == 1 ? ND_THR_GPU : ND_THR_MEM))) continue;
        if (fit_first) {
            bool fit_ok = fl & NDF_PODS_OK;
            if (fit_ok && !none) {
                fit_ok = P(PD_REQ_CPU) <= C(ND_FREE_CPU) && P(PD_REQ_MEM) <= C(ND_FREE_MEM) && P(PD_REQ_EPH) <= C(ND_FREE_EPH);
                for (int s = 0; s < S && fit_ok; ++s)
                    if (((smask >> s) & 1u) && P(PD_FIXED_COLS + s) > C(nd_free_scalar(s))) fit_ok = false;
            }
            if (!fit_ok) continue;
        }
        reduce_max[(i64)RR_FILTER_PANIC * p_pad + p] = 1;
        return;
    }
}

cudaError_t grow(void** p, size_t bytes) {
    if (*p) cudaFree(*p);
    *p = nullptr;
    return cudaMalloc(p, bytes);
}

} // namespace

void FilterTables::free_all() {
    for (void** p : {&d_table, &d_levels, &d_node_rank, &d_ord, &d_ord_val, &d_thr, &d_sort, &d_counts, &d_rowidx})
        if (*p) {
            cudaFree(*p);
            *p = nullptr;
        }
    table_bytes = sort_bytes = 0;
    rowidx_cap = 0;
    n_cap = 0;
    dims_cap = 0;
    valid = false;
}

#define FT_TRY(call)                      \
    do {                                  \
        cudaError_t _e = (call);          \
        if (_e != cudaSuccess) return _e; \
    } while (0)

cudaError_t ft_build(FilterTables* ft, const FtBuildParams& bp, cudaStream_t st, int64_t* launches) {
    ft->valid = false;
    ft->refused = false;
    const i64 n = bp.n_pad, n_nodes = bp.n_nodes;
    const int n_dims = 5 + bp.S;
    if (n <= 0 || n_nodes <= 0) return cudaSuccess;
    if (n > ((i64)1 << 22)) { // (R+1) · n/8 bytes per dimension: beyond any sensible table long before this
        ft->refused = true;
        return cudaSuccess;
    }
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    cudaEventCreate(&ev0);
    cudaEventCreate(&ev1);
    cudaEventRecord(ev0, st);
    // ---- buffers --------------------------------------------------------------------------------------------------
    size_t tmp_sort = 0, tmp_scan = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, tmp_sort, (const i64*)nullptr, (i64*)nullptr, (const int*)nullptr, (int*)nullptr, (int)n);
    cub::DeviceScan::InclusiveSum(nullptr, tmp_scan, (const int*)nullptr, (int*)nullptr, (int)n);
    const size_t tmp_bytes = (std::max(tmp_sort, tmp_scan) + 255) / 256 * 256;
    const size_t n8 = (size_t)n * 8, n4 = ((size_t)n * 4 + 255) / 256 * 256;
    const size_t need_sort = n8 /*keys out*/ + 4 * n4 /*iota, perm, head, scan*/ + tmp_bytes;
    if (n > ft->n_cap || n_dims > ft->dims_cap) {
        ft->n_cap = std::max(n, ft->n_cap);
        ft->dims_cap = std::max(n_dims, ft->dims_cap);
        const size_t nc = (size_t)ft->n_cap, dc = (size_t)ft->dims_cap;
        FT_TRY(grow(&ft->d_levels, sizeof(i64) * dc * (nc + 1)));
        FT_TRY(grow(&ft->d_node_rank, sizeof(int) * dc * nc));
        FT_TRY(grow(&ft->d_ord, sizeof(int) * FT_ORDERS * nc));
        FT_TRY(grow(&ft->d_ord_val, sizeof(i64) * FT_ORDERS * nc));
        FT_TRY(grow(&ft->d_thr, sizeof(i64) * (dc + FT_ORDERS) * nc));
        if (!ft->d_counts) FT_TRY(cudaMalloc(&ft->d_counts, sizeof(int) * kFtMaxDims));
    }
    if (need_sort > ft->sort_bytes) {
        FT_TRY(grow(&ft->d_sort, need_sort));
        ft->sort_bytes = need_sort;
    }
    const i64 stride = ft->n_cap;
    i64* thr = (i64*)ft->d_thr;
    char* sp = (char*)ft->d_sort;
    i64* keys_out = (i64*)sp;
    int* iota = (int*)(sp + n8);
    int* perm = (int*)(sp + n8 + n4);
    int* head = (int*)(sp + n8 + 2 * n4);
    int* scan = (int*)(sp + n8 + 3 * n4);
    void* tmp = sp + n8 + 4 * n4;
    size_t tb = tmp_bytes;
    const unsigned nb = (unsigned)((n + 255) / 256);

    ft_thresholds<<<nb, 256, 0, st>>>(bp.nd, n_nodes, n, bp.S, bp.filter_mask, thr, stride);
    *launches += 1;
    const char* sort_env = getenv("FGS_FT_SORT"); // "cub": the library sorts whatever the size (read per build: the tests switch it)
    const bool small_sort_on = !(sort_env && sort_env[0] == 'c');
    const bool small_sort = small_sort_on && n <= kSortMax;
    if (small_sort) {
        FtSortParams sp;
        sp.thr = thr, sp.stride = stride, sp.n_pad = n, sp.n_nodes = n_nodes, sp.n_dims = n_dims;
        sp.levels = (i64*)ft->d_levels, sp.node_rank = (int*)ft->d_node_rank, sp.counts = (int*)ft->d_counts;
        sp.ord = (int*)ft->d_ord, sp.ord_val = (i64*)ft->d_ord_val;
        int M = 1024;
        while (M < n) M <<= 1;
        const size_t sm = (size_t)M * 12;
        static bool attr_set = false;
        if (!attr_set) {
            FT_TRY(cudaFuncSetAttribute(ft_sort_small, cudaFuncAttributeMaxDynamicSharedMemorySize, kSortMax * 12));
            attr_set = true;
        }
        ft_sort_small<<<(unsigned)(n_dims + FT_ORDERS), kSortThreads, sm, st>>>(sp);
        *launches += 1;
        FT_TRY(cudaGetLastError());
    } else {
    ft_iota<<<nb, 256, 0, st>>>(iota, n);
    *launches += 1;
    for (int d = 0; d < n_dims; ++d) {
        FT_TRY(cub::DeviceRadixSort::SortPairs(tmp, tb, (const i64*)(thr + (i64)d * stride), keys_out, (const int*)iota, perm, (int)n, 0, 64, st));
        ft_heads<<<nb, 256, 0, st>>>(keys_out, n, head);
        FT_TRY(cub::DeviceScan::InclusiveSum(tmp, tb, (const int*)head, scan, (int)n, st));
        ft_scatter<<<nb, 256, 0, st>>>(keys_out, perm, head, scan, n, (i64*)ft->d_levels + (i64)d * (stride + 1), (int*)ft->d_node_rank + (i64)d * stride,
                                       (int*)ft->d_counts + d);
        *launches += 2;
    }
    // node orders, highest score first (nodes only: padding rows are never feasible)
    for (int o = 0; o < FT_ORDERS; ++o) {
        FT_TRY(cub::DeviceRadixSort::SortPairsDescending(tmp, tb, (const i64*)(thr + (i64)(n_dims + o) * stride), (i64*)ft->d_ord_val + (i64)o * stride,
                                                         (const int*)iota, (int*)ft->d_ord + (i64)o * stride, (int)n_nodes, 0, 64, st));
    }
    }
    int counts[kFtMaxDims] = {0};
    FT_TRY(cudaMemcpyAsync(counts, ft->d_counts, sizeof(int) * n_dims, cudaMemcpyDeviceToHost, st));
    FT_TRY(cudaStreamSynchronize(st));
    // ---- the table ------------------------------------------------------------------------------------------------------
    FtView& v = ft->v;
    v.n_dims = n_dims;
    v.W = n / 32;
    v.lev_stride = stride + 1;
    i64 rows = 0;
    int rmax = 0;
    for (int d = 0; d < n_dims; ++d) {
        v.R[d] = counts[d];
        v.row0[d] = rows;
        rows += (i64)counts[d] + 1;
        rmax = std::max(rmax, counts[d]);
    }
    ft->rows_total = rows;
    v.ones_row = rows; // one more row, all ones: what a dimension that cannot clear a bit reads
    const size_t need = (size_t)(rows + 1) * (size_t)v.W * 4;
    if (need > ft->limit_bytes) {
        ft->refused = true;
        cudaEventDestroy(ev0);
        cudaEventDestroy(ev1);
        return cudaSuccess;
    }
    if (need > ft->table_bytes) {
        FT_TRY(grow(&ft->d_table, need));
        ft->table_bytes = need;
    }
    FtTabParams tp;
    tp.table = (uint32_t*)ft->d_table;
    tp.node_rank = (const int*)ft->d_node_rank;
    tp.stride = stride;
    tp.W = v.W;
    for (int d = 0; d < kFtMaxDims; ++d) {
        tp.row0[d] = d < n_dims ? v.row0[d] : 0;
        tp.R[d] = d < n_dims ? v.R[d] : -1;
    }
    const dim3 grid((unsigned)((v.W + 31) / 32), (unsigned)(rmax / kFtRowsPerCta + 1), (unsigned)n_dims);
    ft_fill_rows<<<grid, 256, 0, st>>>(tp);
    FT_TRY(cudaMemsetAsync((uint32_t*)ft->d_table + (size_t)rows * (size_t)v.W, 0xff, (size_t)v.W * 4, st));
    *launches += 1;
    FT_TRY(cudaGetLastError());
    v.table = (const uint32_t*)ft->d_table;
    v.levels = (const i64*)ft->d_levels;
    v.ord = (const int*)ft->d_ord;
    v.ord_val = (const i64*)ft->d_ord_val;
    ft->n_pad = n;
    ft->n_nodes = n_nodes;
    cudaEventRecord(ev1, st);
    cudaEventSynchronize(ev1);
    float ms = 0;
    cudaEventElapsedTime(&ms, ev0, ev1);
    ft->build_ms = ms;
    cudaEventDestroy(ev0);
    cudaEventDestroy(ev1);
    ft->valid = true;
    return cudaSuccess;
}

bool ft_rows_fit(int64_t W) { return (size_t)W * 4 * kRowsWarps <= 200 * 1024; }

cudaError_t ft_launch_rows(const FtRowsParams& p, int* rowidx, cudaStream_t st) {
    if (p.n_pods <= 0) return cudaSuccess;
    const dim3 rgrid((unsigned)((p.p_pad + 255) / 256), (unsigned)p.v.n_dims);
    ft_pod_rows<<<rgrid, 256, 0, st>>>(p, rowidx);
    const size_t smem = (size_t)p.v.W * 4 * kRowsWarps;
    const unsigned grid = (unsigned)((p.n_pods + kRowsWarps - 1) / kRowsWarps);
    if (p.v.n_dims == 7) {
        if (smem > 48 * 1024) FT_TRY(cudaFuncSetAttribute(filter_rows<7>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        filter_rows<7><<<grid, kRowsWarps * 32, smem, st>>>(p, rowidx);
    } else {
        if (smem > 48 * 1024) FT_TRY(cudaFuncSetAttribute(filter_rows<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        filter_rows<0><<<grid, kRowsWarps * 32, smem, st>>>(p, rowidx);
    }
    return cudaGetLastError();
}

cudaError_t ft_launch_masked(const FtMaskParams& p, cudaStream_t st) {
    if (p.n_pods <= 0) return cudaSuccess;
    const size_t smem = (size_t)p.W * 4 * kRowsWarps;
    if (smem > 48 * 1024) FT_TRY(cudaFuncSetAttribute(masked_rows, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    masked_rows<<<(unsigned)((p.n_pods + kRowsWarps - 1) / kRowsWarps), kRowsWarps * 32, smem, st>>>(p);
    return cudaGetLastError();
}

cudaError_t ft_launch_small(const SmallParams& p, cudaStream_t st) {
    if (p.n_pods <= 0) return cudaSuccess;
    const size_t smem = (size_t)p.W * 4;
    if (p.cluster > 1) {
        if (smem > 48 * 1024) FT_TRY(cudaFuncSetAttribute(small_cycle<kSmallCluster>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3((unsigned)p.n_pods * kSmallCluster);
        cfg.blockDim = dim3(kSmallThreads);
        cfg.dynamicSmemBytes = smem;
        cfg.stream = st;
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeClusterDimension;
        at[0].val.clusterDim.x = kSmallCluster, at[0].val.clusterDim.y = 1, at[0].val.clusterDim.z = 1;
        cfg.attrs = at;
        cfg.numAttrs = 1;
        return cudaLaunchKernelEx(&cfg, small_cycle<kSmallCluster>, p);
    }
    if (smem > 48 * 1024) FT_TRY(cudaFuncSetAttribute(small_cycle<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    small_cycle<1><<<(unsigned)p.n_pods, kSmallThreads, smem, st>>>(p);
    return cudaGetLastError();
}

void ft_launch_filter_panic(const int64_t* nd, const int64_t* pods, const int64_t* special, int64_t n_pad, int n_special, int64_t n_pods,
                            int64_t p_pad, int S, uint32_t filter_mask, uint32_t fit_ignored, int flex_first, int64_t* reduce_max,
                            cudaStream_t st) {
    if (n_special <= 0 || n_pods <= 0) return;
    ft_filter_panic<<<(unsigned)((n_pods + 127) / 128), 128, 0, st>>>((const i64*)nd, (const i64*)pods, (const i64*)special, n_pad, n_special, n_pods,
                                                                     p_pad, S, filter_mask, fit_ignored, flex_first, (i64*)reduce_max);
}

} // namespace fgs
