// sample.cu — the reference's node sampling on top of the full feasibility bitmap.
//
//   numFeasibleNodesToFind        vendor/k8s.io/kubernetes/pkg/scheduler/generic_scheduler.go:171-191
//   findNodesThatPassFilters      :266-338 — pod p visits nodes[(nextStartNodeIndex + i) % N], i = 0,1,…, keeps the
//                                 first K feasible ones, stops at the node that would be the (K+1)-th, and advances
//                                 nextStartNodeIndex by the nodes it counted (:329-330)
//
// With the reference's 16 workers WHICH K nodes are kept is a race (:302-308); this is the walk as ONE worker does
// it, which is deterministic. The reduce pass has already evaluated Filter on every node, so the walk is a
// rank/select problem on each pod's bitmap row:
//   sample_chain        one warp, pods in order (pod p+1 starts where pod p stopped — a true dependency chain):
//                       popcount prefix over the row, rank of the start position, select of the (K+1)-th set bit
//                       in rotated order; rows stream through shared memory by TMA bulk copy, 4 stages
//   sample_truncate     pods × words in parallel: clear the feasible bits outside the walked interval
//   reduce_from_bitmap  one warp per pod: the maxima NormalizeScore needs, over the KEPT nodes only
// Score / NormalizeScore / selectHost then run on the kept nodes (the emit pass reads feasibility from the bitmap).
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdlib>

#include "kernels.cuh"
#include "sample.cuh"

namespace fgs {
namespace {

typedef long long i64;
typedef unsigned long long u64;
constexpr i64 kMin64 = (i64)0x8000000000000000ull;
constexpr int kStages = 4;

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

// position of the r-th (0-based) set bit of x: five popcount halvings (the library's __fns is a long software loop)
__device__ __forceinline__ int select_bit(uint32_t x, int r) {
    int pos = 0;
#pragma unroll
    for (int w = 16; w >= 1; w >>= 1) {
        const uint32_t low = x & ((1u << w) - 1u);
        const int c = __popc(low);
        if (r >= c) {
            r -= c;
            x >>= w;
            pos += w;
        } else {
            x = low;
        }
    }
    return pos;
}

// ---- the walk: one warp, one pod after the other ------------------------------------------------------------
// block_pods == 0: the whole batch from start0. Otherwise CTA b walks pods [b·block_pods, (b+1)·block_pods) from starts[b]
// (sample_block_starts), all blocks at once.
__global__ void __launch_bounds__(32, 1) sample_chain(const uint32_t* __restrict__ bitmap, i64 W, i64 n_nodes, i64 n_pods_all, i64 K, i64 start0,
                                                      int2* __restrict__ walk, i64* __restrict__ reduce_sum, i64* __restrict__ next_start,
                                                      i64 block_pods, const int* __restrict__ starts) {
    const i64 pod0 = block_pods ? (i64)blockIdx.x * block_pods : 0;
    i64 n_pods = block_pods ? (n_pods_all - pod0 < block_pods ? n_pods_all - pod0 : block_pods) : n_pods_all;
    if (n_pods <= 0) return;
    if (block_pods) {
        start0 = starts[blockIdx.x];
        bitmap += pod0 * W;
        walk += pod0;
        reduce_sum += pod0;
    }
    extern __shared__ __align__(128) unsigned char smem[];
    uint64_t* bars = (uint64_t*)smem;                // [kStages]
    uint32_t* rows = (uint32_t*)(smem + 64);         // [kStages][W]
    const int lane = threadIdx.x;
    const uint32_t row_bytes = (uint32_t)(W * 4);
    if (lane == 0) {
        for (int k = 0; k < kStages; ++k) mbar_init(&bars[k], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();
    auto issue = [&](i64 p) {
        const int st = (int)(p % kStages);
        mbar_expect_tx(&bars[st], row_bytes);
        bulk_g2s(rows + (size_t)st * W, bitmap + p * W, row_bytes, &bars[st]);
    };
    if (lane == 0)
        for (i64 p = 0; p < kStages - 1 && p < n_pods; ++p) issue(p);

    const int cpl = (int)((W + 31) / 32); // words per lane, consecutive
    const i64 N = n_nodes;
    i64 s = ((start0 % N) + N) % N;
    for (i64 p = 0; p < n_pods; ++p) {
        const int st = (int)(p % kStages);
        if (lane == 0 && p + kStages - 1 < n_pods) {
            // the stage being refilled held row p-1: every lane finished with it before the __syncwarp below
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            issue(p + kStages - 1);
        }
        mbar_wait(&bars[st], (uint32_t)((p / kStages) & 1));
        const uint32_t* row = rows + (size_t)st * W;
        const int w0 = lane * cpl;
        const int sw = (int)(s >> 5), Wi = (int)W;
        const uint32_t below = (1u << (s & 31)) - 1u; // bits of word sw that lie before s
        int c = 0, b = 0;
        for (int k = 0; k < cpl; ++k) {
            const int w = w0 + k;
            if (w >= Wi) break;
            const uint32_t x = row[w];
            const int pc = __popc(x);
            c += pc;
            b += w < sw ? pc : (w == sw ? __popc(x & below) : 0);
        }
        // exclusive prefix of the lanes' counts
        int incl = c;
#pragma unroll
        for (int off = 1; off < 32; off <<= 1) {
            const int v = __shfl_up_sync(0xffffffffu, incl, off);
            if (lane >= off) incl += v;
        }
        const int excl = incl - c;
        const int T = __shfl_sync(0xffffffffu, incl, 31);
        const int before = __reduce_add_sync(0xffffffffu, b);
        const int after = T - before;
        i64 processed = N, kept = T, nxt = s;
        if ((i64)T >= K + 1) {
            // 0-based rank, in row order, of the (K+1)-th feasible node of the rotated walk
            const int r = (i64)after >= K + 1 ? before + (int)K : (int)K - after;
            const uint32_t owner = __ballot_sync(0xffffffffu, excl <= r && r < excl + c);
            const int ol = __ffs(owner) - 1;
            int pos = -1;
            if (lane == ol) {
                int rem = r - excl;
                for (int k = 0; k < cpl; ++k) {
                    const uint32_t x = row[w0 + k];
                    const int pc = __popc(x);
                    if (rem < pc) {
                        pos = (int)((w0 + k) * 32) + select_bit(x, rem);
                        break;
                    }
                    rem -= pc;
                }
            }
            pos = __shfl_sync(0xffffffffu, pos, ol);
            processed = (i64)pos - s; // kept + failed; the (K+1)-th itself is dropped
            if (processed < 0) processed += N;
            kept = K;
            nxt = pos;
        }
        if (lane == 0) {
            walk[p] = make_int2((int)s, (int)processed);
            reduce_sum[p] = kept; // len(feasibleNodes)
        }
        s = nxt;
        __syncwarp();
    }
    if (lane == 0 && !block_pods) *next_start = s; // (blocks: sample_block_starts wrote it)
}

// ---- the walk without the chain --------------------------------------------------------------------------------------
// Pod p turns a start index s into the next one: with b_0 < b_1 < … < b_{T-1} the feasible nodes of its row and i(s) the number
// of them before s, next(s) = b_{(i(s) + K) mod T} when T > K, and s itself otherwise. That is a monotone (cyclically
// order-preserving) step function with at most T steps, and so is every composition of such functions — with fewer and fewer
// steps, because two starts between which a row has no feasible node stay together for good. sample_compose (a CTA per
// block of kBlockPods pods) composes its block's functions: it keeps the distinct values reachable after each pod (at first
// all feasible nodes of the block's first pod with T > K), maps them through the next pod's rank/select, and every few
// pods merges equal neighbours; the result is, for every interval i of that first pod, where a walk that enters the block
// in it stands after the block. sample_block_starts (one warp) then hops from block to block — one rank in the first pod's row
// and one table look-up each — and sample_chain walks all blocks at once from the starts it found. The chain of P dependent
// steps becomes P / kBlockPods hops plus kBlockPods steps.
constexpr int kBlockPods = 256;
constexpr int kComposeThreads = 512;
constexpr int kComposeSmallNodes = 16384; // up to here all four uint16 arrays live in shared memory (128 KB + the row ring)
constexpr int kComposeMaxNodes = 65528;   // beyond: the two arrays only the merges touch live in global scratch (values are 16-bit)
constexpr int kComposeStagesMax = 4;
constexpr int kMergeEvery = 8;

struct ComposeParams {
    const uint32_t* bitmap;
    i64 W, n_nodes, n_pods, K;
    unsigned short* out; // [n_blocks][n_nodes]: interval of the block's first walking pod → start index after the block
    int* blk_q;          // [n_blocks]: that pod (index in the batch), or -1: no pod of the block moves the start
    int* blk_T;          // [n_blocks]: its feasible count
    unsigned short* g_idx; // [n_blocks][n_nodes] and
    unsigned short* g_grp; // [n_blocks][n_nodes]: the merge arrays of a long node table (NULL: shared memory)
    int stages;            // depth of the row ring
};

__global__ void __launch_bounds__(kComposeThreads) sample_compose(const __grid_constant__ ComposeParams prm) {
    extern __shared__ __align__(128) unsigned char smem[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int W = (int)prm.W, N = (int)prm.n_nodes, K = (int)prm.K;
    const i64 pod0 = (i64)blockIdx.x * kBlockPods;
    const int n_pods = (int)(prm.n_pods - pod0 < kBlockPods ? prm.n_pods - pod0 : kBlockPods);
    const int kComposeStages = prm.stages;
    uint64_t* bars = (uint64_t*)smem;                                  // [stages]
    uint32_t* rows = (uint32_t*)(smem + 64);                           // [stages][W]
    int* pre = (int*)(rows + (size_t)kComposeStages * W);              // [W + 1] feasible nodes before word w
    unsigned short* val = (unsigned short*)(pre + W + 1 + ((W + 1) & 1)); // [N] the distinct reachable start indices, in cyclic order
    unsigned short* aux = val + N;                                     // [N] new values / merged values
    unsigned short* idx = prm.g_idx ? prm.g_idx + (size_t)blockIdx.x * (size_t)N : aux + N; // [N] slot → merged slot
    unsigned short* grp = prm.g_grp ? prm.g_grp + (size_t)blockIdx.x * (size_t)N : idx + N; // [T_q] initial slot → current slot
    __shared__ int s_warp[kComposeThreads / 32];
    __shared__ int s_m, s_T;
    const uint32_t row_bytes = (uint32_t)(W * 4);
    if (tid == 0) {
        for (int k = 0; k < kComposeStages; ++k) mbar_init(&bars[k], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    auto issue = [&](int p) { // one thread
        const int st = p % kComposeStages;
        mbar_expect_tx(&bars[st], row_bytes);
        bulk_g2s(rows + (size_t)st * W, prm.bitmap + (pod0 + p) * prm.W, row_bytes, &bars[st]);
    };
    if (tid == 0)
        for (int p = 0; p < kComposeStages - 1 && p < n_pods; ++p) issue(p);

    // block-wide inclusive scan of one int per thread (in thread order); returns the inclusive value, *total ← the sum
    auto block_scan = [&](int x, int* total) {
        int incl = x;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int v = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += v;
        }
        __syncthreads(); // s_warp may still be read from the scan before
        if (lane == 31) s_warp[warp] = incl;
        __syncthreads();
        int base = 0, tot = 0;
#pragma unroll
        for (int w = 0; w < kComposeThreads / 32; ++w) {
            const int v = s_warp[w];
            if (w < warp) base += v;
            tot += v;
        }
        *total = tot;
        return incl + base;
    };

    int q = -1, Tq = 0, m = 0, since_merge = 0;
    for (int p = 0; p < n_pods; ++p) {
        const int st = p % kComposeStages;
        if (tid == 0 && p + kComposeStages - 1 < n_pods) {
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); // the stage being refilled held row p-1: all threads passed the barrier below
            issue(p + kComposeStages - 1);
        }
        mbar_wait(&bars[st], (uint32_t)((p / kComposeStages) & 1));
        const uint32_t* row = rows + (size_t)st * W;
        // pre[w] = feasible nodes in words 0..w-1 (W <= 512: one or two words per thread would do; a strided loop keeps it general)
        {
            const int per = (W + kComposeThreads - 1) / kComposeThreads, w0 = tid * per;
            int c = 0;
            for (int u = 0; u < per; ++u)
                if (w0 + u < W) c += __popc(row[w0 + u]);
            int T = 0;
            const int incl = block_scan(c, &T);
            int run = incl - c;
            for (int u = 0; u < per; ++u)
                if (w0 + u < W) {
                    pre[w0 + u] = run;
                    run += __popc(row[w0 + u]);
                }
            if (tid == 0) pre[W] = T, s_T = T;
            __syncthreads();
        }
        const int T = s_T;
        if (T >= K + 1) {
            auto select = [&](int r) { // position of the r-th (0-based) feasible node of the row
                int lo = 0, hi = W; // pre[lo] <= r < pre[hi]
                while (hi - lo > 1) {
                    const int mid = (lo + hi) >> 1;
                    if (pre[mid] <= r) lo = mid;
                    else hi = mid;
                }
                return lo * 32 + select_bit(row[lo], r - pre[lo]);
            };
            if (q < 0) { // the block's first pod that moves the start: every feasible node of its row is reachable
                q = p, Tq = T, m = T;
                for (int j = tid; j < T; j += kComposeThreads) {
                    val[j] = (unsigned short)select(j);
                    grp[j] = (unsigned short)j;
                }
                since_merge = 0;
            } else {
                for (int j = tid; j < m; j += kComposeThreads) {
                    const int v = val[j];
                    const int sw = v >> 5;
                    int r = pre[sw] + __popc(row[sw] & ((1u << (v & 31)) - 1u)) + K;
                    if (r >= T) r -= T;
                    aux[j] = (unsigned short)select(r);
                }
                __syncthreads();
                for (int j = tid; j < m; j += kComposeThreads) val[j] = aux[j];
                ++since_merge;
            }
            __syncthreads();
            if (since_merge >= kMergeEvery && m > 1) { // equal values are neighbours (cyclically): one slot per run
                since_merge = 0;
                const int per = (m + kComposeThreads - 1) / kComposeThreads, j0 = tid * per;
                const int last = val[m - 1];
                int c = 0;
                for (int u = 0; u < per; ++u) {
                    const int j = j0 + u;
                    if (j < m && val[j] != (j ? val[j - 1] : last)) ++c;
                }
                int heads = 0;
                const int incl = block_scan(c, &heads);
                // (a cyclic sequence changes value at no position or at two or more: heads is 0 or >= 2)
                int run = incl - c;
                for (int u = 0; u < per; ++u) {
                    const int j = j0 + u;
                    if (j >= m) break;
                    const bool head = val[j] != (j ? val[j - 1] : last);
                    run += head ? 1 : 0;
                    const int slot = heads == 0 ? 0 : (run > 0 ? run - 1 : heads - 1); // values before the first change continue the LAST run
                    idx[j] = (unsigned short)slot;
                    if (head || (heads == 0 && j == 0)) aux[slot] = (unsigned short)val[j];
                }
                if (heads == 0) heads = 1; // every value equal: one slot
                __syncthreads();
                for (int i = tid; i < Tq; i += kComposeThreads) grp[i] = idx[grp[i]];
                __syncthreads();
                for (int j = tid; j < heads; j += kComposeThreads) val[j] = aux[j];
                if (tid == 0) s_m = heads;
                __syncthreads();
                m = s_m;
            }
        }
        __syncthreads(); // every thread is done with this stage's row before the ring refills it
    }
    if (tid == 0) prm.blk_q[blockIdx.x] = q >= 0 ? (int)(pod0 + q) : -1, prm.blk_T[blockIdx.x] = Tq;
    if (q >= 0) {
        unsigned short* out = prm.out + (size_t)blockIdx.x * (size_t)N;
        for (int i = tid; i < Tq; i += kComposeThreads) {
            int j = i + K;
            if (j >= Tq) j -= Tq;
            out[i] = val[grp[j]];
        }
    }
}

// block b starts where block b-1 left the index: one warp, one hop per block
__global__ void __launch_bounds__(32, 1) sample_block_starts(const uint32_t* __restrict__ bitmap, i64 W, i64 n_nodes, i64 out_stride, int n_blocks, i64 start0,
                                                             const unsigned short* __restrict__ out, const int* __restrict__ blk_q,
                                                             const int* __restrict__ blk_T, int* __restrict__ starts, i64* __restrict__ next_start) {
    const int lane = threadIdx.x;
    const i64 N = n_nodes;
    int s = (int)(((start0 % N) + N) % N);
    for (int b = 0; b < n_blocks; ++b) {
        if (lane == 0) starts[b] = s;
        const int q = blk_q[b];
        if (q < 0) continue;
        const uint32_t* row = bitmap + (i64)q * W;
        const int sw = s >> 5;
        int before = 0;
        for (int w = lane; w <= sw && w < (int)W; w += 32) {
            const uint32_t x = __ldg(row + w);
            before += w < sw ? __popc(x) : __popc(x & ((1u << (s & 31)) - 1u));
        }
        before = __reduce_add_sync(0xffffffffu, before);
        const int T = blk_T[b];
        if (before >= T) before -= T; // behind the last feasible node: the interval of the first one
        s = (int)__ldg(out + (size_t)b * (size_t)out_stride + before);
    }
    if (lane == 0) *next_start = s;
}

// ---- clear what the walk did not keep ----------------------------------------------------------------------------
__global__ void __launch_bounds__(256) sample_truncate(uint32_t* __restrict__ bitmap, i64 W, i64 n_nodes, i64 n_pods, const int2* __restrict__ walk) {
    const i64 idx = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= n_pods * W) return;
    const i64 p = idx / W, w = idx % W;
    const int2 wk = walk[p];
    const i64 s = wk.x, L = wk.y, N = n_nodes;
    if (L >= N) return; // everything was walked
    // kept: rotated position (n − s) mod N < L  ⇔  n in [s, s+L) or, wrapped, n in [0, s+L−N)
    const i64 lo = w * 32;
    auto range_mask = [&](i64 a, i64 b) -> uint32_t { // bits of this word with a <= n < b
        i64 x = a > lo ? a - lo : 0, y = b - lo;
        if (y <= 0 || x >= 32) return 0u;
        if (y > 32) y = 32;
        const uint32_t hi_m = y == 32 ? 0xffffffffu : ((1u << y) - 1u);
        return hi_m & ~((1u << x) - 1u);
    };
    uint32_t m = range_mask(s, s + L < N ? s + L : N);
    if (s + L > N) m |= range_mask(0, s + L - N);
    const uint32_t x = bitmap[idx];
    if (x & ~m) bitmap[idx] = x & m;
}

// ---- NormalizeScore's reductions over the kept nodes ---------------------------------------------------------------
__global__ void __launch_bounds__(256) reduce_from_bitmap(const uint32_t* __restrict__ bitmap, i64 W, i64 n_pods, i64 p_pad, i64 n_pad,
                                                          const i64* __restrict__ nd, const i64* __restrict__ pods, i64* __restrict__ reduce_max,
                                                          const i64* __restrict__ only /* NULL, or only[p] != 0 selects the pods */) {
    const i64 p = ((i64)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (p >= n_pods) return;
    if (only && only[p] == 0) return;
    const uint32_t kind = (uint32_t)pods[(i64)PD_META * p_pad + p] & 3u;
    const i64* sc = nd + (i64)((kind & 1u) ? ND_GPU_SCORE : ND_MEM_SCORE) * n_pad; // a gpu limit wins (flex_gpu.go:157-159)
    const i64* ar = nd + (i64)ND_ALLOC_RAW * n_pad;
    i64 vflex = kMin64, vhi = kMin64, vnlo = kMin64;
    for (i64 w = lane; w < W; w += 32) {
        uint32_t x = bitmap[p * W + w];
        while (x) {
            const int bit = __ffs(x) - 1;
            x &= x - 1;
            const i64 n = w * 32 + bit;
            const i64 s = kind != 0 ? sc[n] : 0; // FlexGPU.Score, flex_gpu.go:150-165
            const i64 a = ar[n];
            vflex = s > vflex ? s : vflex;
            vhi = a > vhi ? a : vhi;
            vnlo = -a > vnlo ? -a : vnlo;
        }
    }
    auto wmax = [&](i64 v) {
        const int hi = (int)(v >> 32);
        const int mhi = __reduce_max_sync(0xffffffffu, hi);
        const uint32_t mlo = __reduce_max_sync(0xffffffffu, hi == mhi ? (uint32_t)v : 0u);
        return (i64)(((u64)(uint32_t)mhi << 32) | mlo);
    };
    vflex = wmax(vflex);
    vhi = wmax(vhi);
    vnlo = wmax(vnlo);
    if (lane == 0) {
        reduce_max[(i64)RR_FLEX_MAX * p_pad + p] = vflex;
        reduce_max[(i64)RR_ALLOC_HI * p_pad + p] = vhi;
        reduce_max[(i64)RR_ALLOC_NEGLO * p_pad + p] = vnlo;
        reduce_max[(i64)RR_NEG_MAG * p_pad + p] = 0;     // rewritten by special_pass over the kept nodes
        reduce_max[(i64)RR_SCORE_PANIC * p_pad + p] = 0;
    }
}

} // namespace

int64_t num_feasible_nodes_to_find(int64_t n_nodes, int32_t percentage) {
    // generic_scheduler.go:43-53,171-191; int32 arithmetic as in Go
    const int32_t n = (int32_t)n_nodes;
    if (n < 100 || percentage >= 100) return n;
    int32_t adaptive = percentage;
    if (adaptive <= 0) {
        adaptive = 50 - n / 125;
        if (adaptive < 5) adaptive = 5;
    }
    const int32_t k = n * adaptive / 100;
    return k < 100 ? 100 : k;
}

static bool compose_applies(i64 n_nodes, i64 n_pods) {
    const char* v = getenv("FGS_SAMPLE_SERIAL"); // 1: always the serial walk
    if (v && v[0] == '1') return false;
    return n_nodes <= kComposeMaxNodes && n_pods >= 4 * kBlockPods;
}

int64_t sample_scratch_bytes(int64_t n_nodes, int64_t n_pods) {
    if (!compose_applies(n_nodes, n_pods)) return 0;
    const i64 nb = (n_pods + kBlockPods - 1) / kBlockPods;
    const i64 n8 = (n_nodes + 7) & ~7ll;
    const i64 tables = n_nodes > kComposeSmallNodes ? 3 : 1; // out, and for a long node table the two merge arrays
    return (i64)sizeof(unsigned short) * nb * n8 * tables + (i64)sizeof(int) * 3 * (nb + 8);
}

int sample_launch_count(int64_t n_nodes, int64_t n_pods, bool have_scratch) { return have_scratch && compose_applies(n_nodes, n_pods) ? 5 : 3; }

cudaError_t launch_sampling(const SampleParams& sp, void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    if (sp.n_pods <= 0 || sp.n_nodes <= 0) return cudaSuccess;
    const size_t smem = 64 + (size_t)kStages * (size_t)sp.words_per_row * 4;
    cudaError_t ce = cudaFuncSetAttribute(sample_chain, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (ce != cudaSuccess) return ce;
    if (sp.scratch && compose_applies(sp.n_nodes, sp.n_pods) && sp.scratch_bytes >= sample_scratch_bytes(sp.n_nodes, sp.n_pods)) {
        const i64 nb = (sp.n_pods + kBlockPods - 1) / kBlockPods;
        const i64 n8 = (sp.n_nodes + 7) & ~7ll;
        ComposeParams cp;
        cp.bitmap = sp.bitmap, cp.W = sp.words_per_row, cp.n_nodes = n8, cp.n_pods = sp.n_pods, cp.K = sp.k;
        cp.out = (unsigned short*)sp.scratch;
        const bool big = sp.n_nodes > kComposeSmallNodes;
        cp.g_idx = big ? cp.out + (size_t)nb * (size_t)n8 : nullptr;
        cp.g_grp = big ? cp.g_idx + (size_t)nb * (size_t)n8 : nullptr;
        cp.stages = big ? 2 : kComposeStagesMax;
        cp.blk_q = (int*)((char*)sp.scratch + sizeof(unsigned short) * nb * n8 * (big ? 3 : 1));
        cp.blk_T = cp.blk_q + (nb + 8);
        int* starts = cp.blk_T + (nb + 8);
        const size_t W = (size_t)sp.words_per_row;
        const size_t csmem = 64 + (size_t)cp.stages * W * 4 + ((W + 2) & ~(size_t)1) * 4 + (big ? 2 : 4) * (size_t)n8 * 2;
        ce = cudaFuncSetAttribute(sample_compose, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)csmem);
        if (ce != cudaSuccess) return ce;
        sample_compose<<<(unsigned)nb, kComposeThreads, csmem, stream>>>(cp);
        sample_block_starts<<<1, 32, 0, stream>>>(sp.bitmap, sp.words_per_row, sp.n_nodes, n8, (int)nb, sp.start, cp.out, cp.blk_q, cp.blk_T, starts, (i64*)sp.next_start);
        sample_chain<<<(unsigned)nb, 32, smem, stream>>>(sp.bitmap, sp.words_per_row, sp.n_nodes, sp.n_pods, sp.k, sp.start, (int2*)sp.walk,
                                                        (i64*)sp.reduce_sum, (i64*)sp.next_start, kBlockPods, starts);
    } else {
        sample_chain<<<1, 32, smem, stream>>>(sp.bitmap, sp.words_per_row, sp.n_nodes, sp.n_pods, sp.k, sp.start, (int2*)sp.walk, (i64*)sp.reduce_sum,
                                               (i64*)sp.next_start, 0, nullptr);
    }
    const i64 total = sp.n_pods * sp.words_per_row;
    sample_truncate<<<(unsigned)((total + 255) / 256), 256, 0, stream>>>(sp.bitmap, sp.words_per_row, sp.n_nodes, sp.n_pods, (const int2*)sp.walk);
    reduce_from_bitmap<<<(unsigned)((sp.n_pods * 32 + 255) / 256), 256, 0, stream>>>(sp.bitmap, sp.words_per_row, sp.n_pods, sp.p_pad, sp.n_pad,
                                                                                    (const i64*)sp.nd, (const i64*)sp.pods, (i64*)sp.reduce_max, nullptr);
    return cudaGetLastError();
}

cudaError_t launch_reduce_from_bitmap(const SampleParams& sp, const int64_t* only, void* stream_) {
    if (sp.n_pods <= 0) return cudaSuccess;
    reduce_from_bitmap<<<(unsigned)((sp.n_pods * 32 + 255) / 256), 256, 0, (cudaStream_t)stream_>>>(
        sp.bitmap, sp.words_per_row, sp.n_pods, sp.p_pad, sp.n_pad, (const i64*)sp.nd, (const i64*)sp.pods, (i64*)sp.reduce_max, (const i64*)only);
    return cudaGetLastError();
}

} // namespace fgs
