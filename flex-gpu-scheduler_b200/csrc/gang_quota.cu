// gang_quota.cu — Coscheduling and CapacityScheduling PreFilter arithmetic, and the per-plugin
// Filter status plane.
//
//   gang_prefix_scan    per resource column: running sum of the nodes' left-over capacity in
//                       snapshot order (CheckClusterResource walks nodes in that order and a
//                       negative left-over can undo an earlier surplus — pkg/coscheduling/core/core.go:322-342)
//   gang_check_groups   per PodGroup: max over prefixes of (prefix + correction for the group's own
//                       pods, getNodeResource :349-382) against MinResources ∪ {pods: MinMember}
//   group_histograms    sibling and assigned-pod counts (:158-167, :301-318)
//   quota_aggregate /   ElasticQuota checks (pkg/capacityscheduling/capacity_scheduling.go:201-275,
//   quota_check_pods    elasticquota.go:40-51,90-100,161-181)
//   status_plane        framework.Code per (pod,node) and plugin, for the Go Filter answers
//
// These are small, L2-resident problems: latency-bound, not HBM-bound.
#include <cuda_runtime.h>

#include <atomic>

#include <algorithm>
#include <cstring>
#include <vector>

#include "engine.h"
#include "kernels.cuh"
#include "podmath.h"

namespace fgs {

typedef long long i64;
typedef unsigned long long u64;
constexpr i64 kI64Min = (i64)0x8000000000000000ull;

#define FGS_CUDA(e, call)                                                    \
    do {                                                                     \
        cudaError_t _err = (call);                                           \
        if (_err != cudaSuccess) {                                           \
            (e)->err = std::string(#call) + ": " + cudaGetErrorString(_err); \
            return FGS_E_CUDA;                                               \
        }                                                                    \
    } while (0)

// growable device + pinned-host arena for the small tables of these checks
struct GangState {
    void* d = nullptr;
    void* h = nullptr;
    size_t cap = 0;
};

void Engine::gang_free() {
    if (!gang) return;
    if (gang->d) cudaFree(gang->d);
    if (gang->h) cudaFreeHost(gang->h);
    delete gang;
    gang = nullptr;
}

static int arena_reserve(Engine* e, size_t bytes) {
    if (!e->gang) e->gang = new GangState();
    GangState* g = e->gang;
    if (bytes <= g->cap) return FGS_OK;
    if (g->d) cudaFree(g->d);
    if (g->h) cudaFreeHost(g->h);
    g->d = g->h = nullptr;
    g->cap = 0;
    const size_t want = bytes + bytes / 4 + 4096;
    FGS_CUDA(e, cudaMalloc(&g->d, want));
    FGS_CUDA(e, cudaMallocHost(&g->h, want));
    g->cap = want;
    return FGS_OK;
}

struct Bump {
    size_t off = 0;
    size_t take(size_t bytes) {
        const size_t o = off;
        off += (bytes + 255) & ~(size_t)255;
        return o;
    }
};

namespace {

// ---- gang ----------------------------------------------------------------------------------------

constexpr int kScanThreads = 1024;

// cand[k][n] = Σ_{valid i<=n} left_k(i) for valid n, INT64_MIN for nodes that are not part of the walk
// (core.go:324-326); colmax[k] = max_n cand[k][n].
__global__ void __launch_bounds__(kScanThreads) gang_prefix_scan(const i64* __restrict__ nd, const i64* __restrict__ raw, i64 n_nodes, i64 n_pad,
                                                                 int S, i64* __restrict__ cand, i64* __restrict__ colmax) {
    const int k = blockIdx.x; // column: 0 cpu, 1 mem, 2 eph, 3 pods, 4+s scalars
    const int tid = threadIdx.x;
    __shared__ i64 s_sum[kScanThreads];
    __shared__ i64 s_max[kScanThreads];
    const i64 chunk = (n_nodes + kScanThreads - 1) / kScanThreads;
    const i64 a = (i64)tid * chunk, b = a + chunk < n_nodes ? a + chunk : n_nodes;
    auto left = [&](i64 n) -> i64 {
        const uint32_t fl = (uint32_t)nd[(i64)ND_FLAGS * n_pad + n];
        if (!(fl & FGS_NF_VALID)) return 0;
        if (k == 0) return nd[(i64)ND_FREE_CPU * n_pad + n];
        if (k == 1) return nd[(i64)ND_FREE_MEM * n_pad + n];
        if (k == 2) return nd[(i64)ND_FREE_EPH * n_pad + n];
        if (k == 3) return nd[(i64)ND_LEFT_PODS * n_pad + n];
        const int s = k - 4;
        // leftResource.ScalarResources only has the node's allocatable keys (core.go:372-379)
        const uint32_t present = (uint32_t)raw[(i64)NR_PRESENT * n_pad + n];
        return ((present >> s) & 1u) ? nd[(i64)nd_free_scalar(s) * n_pad + n] : 0;
    };
    i64 sum = 0;
    for (i64 n = a; n < b; ++n) sum += left(n);
    s_sum[tid] = sum;
    __syncthreads();
    // exclusive scan of the chunk sums (Hillis-Steele on 1024 values)
    for (int off = 1; off < kScanThreads; off <<= 1) {
        i64 v = tid >= off ? s_sum[tid - off] : 0;
        __syncthreads();
        s_sum[tid] += v;
        __syncthreads();
    }
    i64 run = s_sum[tid] - sum;
    i64 mx = kI64Min;
    for (i64 n = a; n < b; ++n) {
        run += left(n);
        const bool valid = ((uint32_t)nd[(i64)ND_FLAGS * n_pad + n]) & FGS_NF_VALID;
        const i64 c = valid ? run : kI64Min;
        cand[(i64)k * n_pad + n] = c;
        mx = c > mx ? c : mx;
    }
    s_max[tid] = mx;
    __syncthreads();
    for (int off = kScanThreads / 2; off > 0; off >>= 1) {
        if (tid < off) s_max[tid] = s_max[tid] > s_max[tid + off] ? s_max[tid] : s_max[tid + off];
        __syncthreads();
    }
    if (tid == 0) colmax[k] = s_max[0];
}

// counts[0][g] = pods carrying the group's label (pods on nodes + pending pods), counts[1][g] = pods on
// nodes with the label, the group's namespace and a node name
__global__ void group_histograms(const int32_t* __restrict__ ex_pg, const int32_t* __restrict__ ex_ns, const uint8_t* __restrict__ ex_named,
                                 i64 n_ex, const int32_t* __restrict__ pend_pg, i64 n_pend, const int32_t* __restrict__ pg_ns, i64 G,
                                 int32_t* __restrict__ counts) {
    const i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n_ex) {
        const int32_t g = ex_pg[i];
        if (g >= 0 && g < G) {
            atomicAdd(&counts[g], 1);
            if (ex_named[i] && ex_ns[i] == pg_ns[g]) atomicAdd(&counts[G + g], 1);
        }
    } else if (i < n_ex + n_pend) {
        const int32_t g = pend_pg[i - n_ex];
        if (g >= 0 && g < G) atomicAdd(&counts[g], 1);
    }
}

struct GangGroupParams {
    const i64* cand;      // [K][n_pad]
    const i64* colmax;    // [K]
    const i64* min_res;   // [G][K]
    const uint32_t* min_present;
    const int32_t* min_member;
    const uint8_t* gflags; // bit0 exists, bit1 has_min, bit2 denied, bit3 permitted
    const int32_t* counts; // [2][G]
    const i64* corr_off;   // [G+1]
    const int32_t* corr_node;
    const i64* corr_delta; // [n_corr][K]
    uint8_t* gcode;        // [G]
    i64 n_nodes, n_pad, G;
    int K;
};

// one CTA per PodGroup
__global__ void __launch_bounds__(256) gang_check_groups(const __grid_constant__ GangGroupParams prm) {
    const i64 g = blockIdx.x;
    const int tid = threadIdx.x;
    const uint8_t fl = prm.gflags[g];
    __shared__ i64 s_red[256];
    __shared__ int s_fail;
    // core.go:149-178 — everything that does not need the cluster walk
    uint8_t code = 0;
    bool walk = false;
    if (!(fl & 1u)) code = 0;
    else if (fl & 4u) code = 1;
    else if (prm.counts[g] < prm.min_member[g]) code = 2;
    else if (!(fl & 2u)) code = 0;
    else if (fl & 8u) code = 0;
    else walk = true;
    if (!walk) {
        if (tid == 0) prm.gcode[g] = code;
        return;
    }
    const int K = prm.K;
    const i64 c0 = prm.corr_off[g], c1 = prm.corr_off[g + 1];
    const int nc = (int)(c1 - c0);
    const uint32_t keys = prm.min_present[g] | (1u << 3); // ∪ {pods} (core.go:185-187)
    if (tid == 0) s_fail = 0;
    __syncthreads();
    for (int k = 0; k < K; ++k) {
        if (!((keys >> k) & 1u)) continue;
        const i64 req = k == 3 ? (i64)prm.min_member[g] : prm.min_res[g * K + k];
        i64 best = kI64Min;
        if (nc == 0) {
            best = prm.colmax[k];
        } else {
            // the group's own pods are removed from their nodes (core.go:349-359): from node
            // corr_node[j] on, every prefix grows by Δ_j
            for (i64 n = tid; n < prm.n_nodes; n += blockDim.x) {
                const i64 c = prm.cand[(i64)k * prm.n_pad + n];
                if (c == kI64Min) continue;
                // entries are sorted by node: add every Δ with node <= n
                int lo = 0, hi = nc;
                while (lo < hi) {
                    const int mid = (lo + hi) >> 1;
                    if ((i64)prm.corr_node[c0 + mid] <= n) lo = mid + 1;
                    else hi = mid;
                }
                // corr_delta holds running sums per group, so one read suffices
                const i64 add = lo > 0 ? prm.corr_delta[(c0 + lo - 1) * K + k] : 0;
                const i64 v = c + add;
                best = v > best ? v : best;
            }
            s_red[tid] = best;
            __syncthreads();
            for (int off = 128; off > 0; off >>= 1) {
                if (tid < off) s_red[tid] = s_red[tid] > s_red[tid + off] ? s_red[tid] : s_red[tid + off];
                __syncthreads();
            }
            best = s_red[0];
            __syncthreads();
        }
        // the key is deleted at the first node where the remainder drops to <= 0 (core.go:331-334):
        // ∃ prefix m >= 1 with Σ left >= req. No valid node → best == INT64_MIN → gap.
        if (tid == 0 && !(best != kI64Min && best >= req)) s_fail = 1;
    }
    __syncthreads();
    if (tid == 0) prm.gcode[g] = s_fail ? 3 : 0;
}

__global__ void gang_pod_codes(const int32_t* __restrict__ pend_pg, i64 n_pend, const uint8_t* __restrict__ gcode,
                               const uint8_t* __restrict__ gflags, const int32_t* __restrict__ counts, const int32_t* __restrict__ min_member,
                               i64 G, uint8_t* __restrict__ prefilter, uint8_t* __restrict__ permit, uint8_t* __restrict__ postfilter) {
    const i64 p = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n_pend) return;
    const int32_t g = pend_pg[p];
    if (g < 0 || g >= G) {
        prefilter[p] = 0;
        permit[p] = 2; // PodGroupNotSpecified (core.go:201-203)
        postfilter[p] = FGS_GANG_POSTFILTER_NO_GROUP;
        return;
    }
    prefilter[p] = gcode[g];
    if (!(gflags[g] & 1u)) permit[p] = 3; // PodGroupNotFound (:204-207)
    else permit[p] = (counts[G + g] + 1 >= min_member[g]) ? 0 : 1; // :209-215
    // Coscheduling.PostFilter (pkg/coscheduling/coscheduling.go:140-176): what happens to the pod's GROUP once the pod
    // turned out unschedulable
    uint8_t pf;
    const int assigned = counts[G + g], mm = min_member[g];
    if (!(gflags[g] & 1u)) {
        pf = FGS_GANG_POSTFILTER_NO_GROUP; // GetPodGroup found no PodGroup object (:143-147)
    } else if (assigned >= mm) {
        pf = FGS_GANG_POSTFILTER_ENOUGH_ASSIGNED; // :151-155
    } else {
        // notAssignedPercentage := float32(int(MinMember)-assigned) / float32(MinMember); <= 0.1 (:159-163): IEEE single
        // precision division, round to nearest, as Go compiles it; 0.1 is the float32 nearest to one tenth
        const float pct = __fdiv_rn((float)(mm - assigned), (float)mm);
        pf = pct <= 0.1f ? FGS_GANG_POSTFILTER_SMALL_GAP : FGS_GANG_POSTFILTER_REJECT_GROUP;
    }
    postfilter[p] = pf;
}

// ---- quota ---------------------------------------------------------------------------------------

struct QuotaParams {
    const uint8_t* has_eq;   // [NS]
    const i64* qmin;         // [NS][W]
    const i64* qmax;
    const i64* qused;
    const uint32_t* used_present;
    i64* agg;                // [2][W]: Σused, Σmin   (written by quota_aggregate)
    uint32_t* agg_present;   // [1]: keys of Σused
    uint8_t* ns_over_min;    // [NS] usedOverMin()
    i64 n_ns;
    int S, W;
    // pods (column-major [cols][P]): req_mcpu, req_mem, req_scalar[S]
    const i64* pod_req;
    const uint32_t* pod_present;
    const int32_t* pod_ns;
    const int32_t* pod_prio;
    const i64* pod_uid;      // may be null
    i64 n_pods;
    const i64* nom_req;      // same layout for nominated pods
    const uint32_t* nom_present;
    const int32_t* nom_ns;
    const int32_t* nom_prio;
    const i64* nom_uid;
    i64 n_nom;
    uint8_t* code;           // [P]
};

// Σ over namespaces of Used and Min (aggregatedUsedOverMinWith, elasticquota.go:40-51) and the
// per-namespace usedOverMin() (:98-100). One CTA; NS is a few hundred.
__global__ void __launch_bounds__(256) quota_aggregate(const __grid_constant__ QuotaParams prm) {
    const int tid = threadIdx.x;
    const int W = prm.W, S = prm.S;
    __shared__ i64 s_acc[256];
    __shared__ uint32_t s_present;
    if (tid == 0) s_present = 0;
    __syncthreads();
    for (int which = 0; which < 2; ++which)
        for (int c = 0; c < W; ++c) {
            i64 acc = 0;
            for (i64 k = tid; k < prm.n_ns; k += blockDim.x) {
                if (!prm.has_eq[k]) continue;
                if (which == 0) {
                    const bool present = c < 2 || ((prm.used_present[k] >> (c - 2)) & 1u);
                    if (present) acc += prm.qused[k * W + c];
                } else {
                    acc += prm.qmin[k * W + c];
                }
            }
            s_acc[tid] = acc;
            __syncthreads();
            for (int off = 128; off > 0; off >>= 1) {
                if (tid < off) s_acc[tid] += s_acc[tid + off];
                __syncthreads();
            }
            if (tid == 0) prm.agg[which * W + c] = s_acc[0];
            __syncthreads();
        }
    uint32_t pr = 0;
    for (i64 k = tid; k < prm.n_ns; k += blockDim.x) {
        if (!prm.has_eq[k]) continue;
        pr |= prm.used_present[k];
        // cmp(Used, Min): scalar loop over Used's keys only (elasticquota.go:161-181)
        bool over = prm.qused[k * W + 0] > prm.qmin[k * W + 0] || prm.qused[k * W + 1] > prm.qmin[k * W + 1];
        for (int s = 0; s < S; ++s)
            if ((prm.used_present[k] >> s) & 1u)
                if (prm.qused[k * W + 2 + s] > prm.qmin[k * W + 2 + s]) over = true;
        prm.ns_over_min[k] = over ? 1 : 0;
    }
    if (pr) atomicOr(&s_present, pr);
    __syncthreads();
    if (tid == 0) *prm.agg_present = s_present;
}

// one thread per pending pod: capacity_scheduling.go:201-275
__global__ void __launch_bounds__(256) quota_check_pods(const __grid_constant__ QuotaParams prm) {
    const i64 p = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= prm.n_pods) return;
    const int S = prm.S, W = prm.W;
    const int32_t ns = prm.pod_ns[p];
    if (ns < 0 || ns >= prm.n_ns || !prm.has_eq[ns]) { // :210-217
        prm.code[p] = 0;
        return;
    }
    i64 in_eq[2 + FGS_MAX_SCALARS], all[2 + FGS_MAX_SCALARS];
    uint32_t in_keys = 0, all_keys = 0;
    for (int c = 0; c < W; ++c) in_eq[c] = all[c] = 0;
    const int32_t prio = prm.pod_prio[p];
    for (i64 k = 0; k < prm.n_nom; ++k) { // :233-255
        if (prm.pod_uid && prm.nom_uid && prm.nom_uid[k] == prm.pod_uid[p]) continue;
        const int32_t pns = prm.nom_ns[k];
        if (pns < 0 || pns >= prm.n_ns || !prm.has_eq[pns]) continue;
        const bool same = pns == ns && prm.nom_prio[k] >= prio;
        const bool other = pns != ns && !prm.ns_over_min[pns];
        if (!same && !other) continue;
        const uint32_t keys = prm.nom_present[k];
        for (int c = 0; c < W; ++c) {
            const i64 v = prm.nom_req[(i64)c * prm.n_nom + k];
            all[c] += v;
            if (same) in_eq[c] += v;
        }
        all_keys |= keys;
        if (same) in_keys |= keys;
    }
    const uint32_t pkeys = prm.pod_present[p];
    for (int c = 0; c < W; ++c) { // :257-258
        const i64 v = prm.pod_req[(i64)c * prm.n_pods + p];
        in_eq[c] += v;
        all[c] += v;
    }
    in_keys |= pkeys;
    all_keys |= pkeys;
    // usedOverMaxWith = cmp2(podRequest, Used, Max) (elasticquota.go:94-96,165-181)
    const i64* used = prm.qused + (i64)ns * W;
    const i64* mx = prm.qmax + (i64)ns * W;
    const uint32_t up = prm.used_present[ns];
    bool over_max = in_eq[0] + used[0] > mx[0] || in_eq[1] + used[1] > mx[1];
    for (int s = 0; s < S; ++s)
        if ((in_keys >> s) & 1u) {
            const i64 u = ((up >> s) & 1u) ? used[2 + s] : 0;
            if (in_eq[2 + s] + u > mx[2 + s]) over_max = true;
        }
    if (over_max) { // :266-268
        prm.code[p] = 1;
        return;
    }
    // aggregatedUsedOverMinWith: x1 = Σused + request, keys = Σused's ∪ the request's
    const uint32_t akeys = *prm.agg_present | all_keys;
    bool over_min = prm.agg[0] + all[0] > prm.agg[W + 0] || prm.agg[1] + all[1] > prm.agg[W + 1];
    for (int s = 0; s < S; ++s)
        if ((akeys >> s) & 1u)
            if (prm.agg[2 + s] + all[2 + s] > prm.agg[W + 2 + s]) over_min = true;
    prm.code[p] = over_min ? 2 : 0; // :270-272
}

// ---- status plane ----------------------------------------------------------------------------------

struct StatusParams {
    const i64* nd;
    const i64* pods;
    i64 n_nodes, n_pad, n_pods, p_pad;
    int S;
    uint32_t plugin; // 0 merged (profile order, early exit), FGS_PL_FLEXGPU, FGS_PL_FIT
    uint32_t filter_mask, fit_ignored;
    int flex_filter_first;
    uint8_t* out; // [P][n_nodes]
    // reason plane only
    const i64* raw;
    int D;
    uint16_t* out16;
};

__device__ __forceinline__ uint8_t flex_code(const StatusParams& prm, i64 p, i64 n, bool* panic) {
    const u64 meta = (u64)prm.pods[(i64)PD_META * prm.p_pad + p];
    const uint32_t kind = (uint32_t)meta & 3u;
    const uint32_t fl = (uint32_t)prm.nd[(i64)ND_FLAGS * prm.n_pad + n];
    *panic = false;
    if (kind == 0) return FGS_CODE_SUCCESS;        // flex_gpu.go:54-56
    if (kind == 3) return FGS_CODE_UNRESOLVABLE;   // :58-61
    if (!(fl & FGS_NF_HAS_FLEX)) return FGS_CODE_UNRESOLVABLE; // :64-74
    const i64 lim = prm.pods[(i64)(kind == 1 ? PD_GPU_LIM : PD_MEM_LIM) * prm.p_pad + p];
    const i64 thr = prm.nd[(i64)(kind == 1 ? ND_THR_GPU : ND_THR_MEM) * prm.n_pad + n];
    if (lim > thr) return FGS_CODE_UNSCHEDULABLE;  // :85-98,109-116
    if (fl & FGS_NF_REF_PANIC) *panic = true;      // :101
    return FGS_CODE_SUCCESS;
}

__device__ __forceinline__ uint8_t fit_code(const StatusParams& prm, i64 p, i64 n) {
    const u64 meta = (u64)prm.pods[(i64)PD_META * prm.p_pad + p];
    const uint32_t pf = (uint32_t)meta;
    const uint32_t fl = (uint32_t)prm.nd[(i64)ND_FLAGS * prm.n_pad + n];
    bool ok = fl & NDF_PODS_OK;
    if (!(pf & FGS_PF_REQ_NONE)) {
        auto P = [&](int c) { return prm.pods[(i64)c * prm.p_pad + p]; };
        auto N = [&](int c) { return prm.nd[(i64)c * prm.n_pad + n]; };
        ok = ok && P(PD_REQ_CPU) <= N(ND_FREE_CPU) && P(PD_REQ_MEM) <= N(ND_FREE_MEM) && P(PD_REQ_EPH) <= N(ND_FREE_EPH);
        uint32_t m = (uint32_t)(meta >> 32) & ~prm.fit_ignored;
        while (m) {
            const int s = __ffs(m) - 1;
            m &= m - 1;
            if (P(PD_FIXED_COLS + s) > N(nd_free_scalar(s))) ok = false;
        }
    }
    return ok ? FGS_CODE_SUCCESS : FGS_CODE_UNSCHEDULABLE;
}

__global__ void __launch_bounds__(256) status_plane(const __grid_constant__ StatusParams prm) {
    const i64 n = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    const i64 p = blockIdx.y;
    if (n >= prm.n_nodes || p >= prm.n_pods) return;
    uint8_t out = FGS_CODE_SUCCESS;
    bool panic = false;
    if (prm.plugin == FGS_PL_FLEXGPU) {
        out = flex_code(prm, p, n, &panic);
        if (panic) out = 0xF0 | FGS_CODE_ERROR;
    } else if (prm.plugin == FGS_PL_FIT) {
        out = fit_code(prm, p, n);
    } else if (!((uint32_t)prm.nd[(i64)ND_FLAGS * prm.n_pad + n] & FGS_NF_VALID)) {
        // a NodeInfo without a Node is never in the list findNodesThatPassFilters walks (the cache keeps such
        // entries out of the snapshot list, internal/cache/cache.go:254,290,314-325): plain Error, no plugin
        out = FGS_CODE_ERROR;
    } else {
        // RunFilterPlugins: profile order, first failure wins (runtime/framework.go:667-693)
        const uint32_t order[2] = {prm.flex_filter_first ? FGS_PL_FLEXGPU : FGS_PL_FIT, prm.flex_filter_first ? FGS_PL_FIT : FGS_PL_FLEXGPU};
        for (int k = 0; k < 2; ++k) {
            if (!(prm.filter_mask & order[k])) continue;
            uint8_t c;
            if (order[k] == FGS_PL_FLEXGPU) {
                c = flex_code(prm, p, n, &panic);
                if (panic) {
                    out = 0xF0 | FGS_CODE_ERROR;
                    break;
                }
            } else {
                c = fit_code(prm, p, n);
            }
            if (c != FGS_CODE_SUCCESS) {
                out = (uint8_t)((order[k] << 4) | c);
                break;
            }
        }
    }
    prm.out[p * prm.n_nodes + n] = out;
}

// WHY a plugin rejected the node — what the reference puts into the Status message (NodeToStatusMap,
// generic_scheduler.go:309-313): FlexGPU's five messages (flex_gpu.go:60,67,73,92,97,114) as FGS_FLEX_REASON_*, and
// Fit's list of InsufficientResource entries (fit.go:256-329: every failing dimension is reported) as a bit mask.
__device__ __forceinline__ uint16_t reason_of(const StatusParams& prm, uint32_t plugin, i64 p, i64 n) {
    auto P = [&](int c) { return prm.pods[(i64)c * prm.p_pad + p]; };
    auto N = [&](int c) { return prm.nd[(i64)c * prm.n_pad + n]; };
    auto R = [&](int c) { return prm.raw[(i64)c * prm.n_pad + n]; };
    const u64 meta = (u64)P(PD_META);
    const uint32_t pf = (uint32_t)meta, kind = pf & 3u;
    const uint32_t fl = (uint32_t)N(ND_FLAGS);
    uint16_t out = 0;
    if (plugin == FGS_PL_FLEXGPU) {
        const int S = prm.S;
        if (kind == 0) out = FGS_FLEX_REASON_NONE;
        else if (kind == 3) out = FGS_FLEX_REASON_CONFLICT;
        else if (!(fl & FGS_NF_HAS_FLEX)) out = FGS_FLEX_REASON_UNKNOWN_RESOURCE;
        else {
            const i64 gl = kind == 1 ? P(PD_GPU_LIM) : 0, ml = kind == 2 ? P(PD_MEM_LIM) : 0;
            const i64 gpu_head = R(nr_alloc_scalar(FGS_SCALAR_FLEX_GPU)) - R(NR_FLEX_GPU_SUM);
            const i64 mem_head = R(nr_alloc_scalar(FGS_SCALAR_FLEX_MEM)) - R(NR_FLEX_MEM_SUM);
            if (gl > gpu_head) out = FGS_FLEX_REASON_INSUFFICIENT_GPU;       // :90-93
            else if (ml > mem_head) out = FGS_FLEX_REASON_INSUFFICIENT_MEM;  // :95-98
            else if (fl & FGS_NF_REF_PANIC) out = FGS_FLEX_REASON_PANIC;     // NewGPUNode, :101
            else if (kind == 2) {                                            // MemAssumeFitIndexes empty, :109-116
                const i64 each = N(ND_MEM_EACH), cnt = R(nr_alloc_scalar(FGS_SCALAR_FLEX_GPU));
                const uint32_t mono = (uint32_t)((u64)R(NR_META) >> 32);
                bool fits = false;
                for (int d = 0; d < prm.D && (i64)d < cnt; ++d)
                    if (!((mono >> d) & 1u) && each >= R(nr_used(S, d)) + ml) fits = true;
                if (!fits) out = FGS_FLEX_REASON_NO_FIT_INDEX;
            }
        }
    } else {
        if (!(fl & NDF_PODS_OK)) out |= FGS_FIT_REASON_PODS; // "Too many pods", fit.go:259-268
        if (!(pf & FGS_PF_REQ_NONE)) {
            if (P(PD_REQ_CPU) > N(ND_FREE_CPU)) out |= FGS_FIT_REASON_CPU;
            if (P(PD_REQ_MEM) > N(ND_FREE_MEM)) out |= FGS_FIT_REASON_MEM;
            if (P(PD_REQ_EPH) > N(ND_FREE_EPH)) out |= FGS_FIT_REASON_EPH;
            uint32_t m = (uint32_t)(meta >> 32) & ~prm.fit_ignored;
            while (m) {
                const int s = __ffs(m) - 1;
                m &= m - 1;
                if (P(PD_FIXED_COLS + s) > N(nd_free_scalar(s))) out |= (uint16_t)(FGS_FIT_REASON_SCALAR0 << s);
            }
        }
    }
    return out;
}

__global__ void __launch_bounds__(256) reason_plane(const __grid_constant__ StatusParams prm) {
    const i64 n = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    const i64 p = blockIdx.y;
    if (n >= prm.n_nodes || p >= prm.n_pods) return;
    prm.out16[p * prm.n_nodes + n] = reason_of(prm, prm.plugin, p, n);
}

// FitError (vendor/k8s.io/kubernetes/pkg/scheduler/framework/types.go:221-250): the histogram of the reasons in the
// NodeToStatusMap of a pod no node took — per pod, how many nodes carry each reason string. A node's status is the
// status of the FIRST filter plugin that rejected it (RunFilterPlugins, profile order, early exit); Fit lists every
// insufficient resource, so one node may count under several Fit reasons. One CTA per (pod, 2048 nodes), counts in
// shared memory, one atomicAdd per non-zero slot. Pods with a feasible node are skipped: there is no FitError for them.
__global__ void __launch_bounds__(256) fit_error_histogram(const __grid_constant__ StatusParams prm, const i64* __restrict__ feasible_count,
                                                           int* __restrict__ counts) {
    __shared__ int sh[FGS_DIAG_SLOTS];
    const i64 p = blockIdx.y;
    if (p >= prm.n_pods || feasible_count[p] > 0) return;
    if (threadIdx.x < FGS_DIAG_SLOTS) sh[threadIdx.x] = 0;
    __syncthreads();
    const uint32_t order[2] = {prm.flex_filter_first ? FGS_PL_FLEXGPU : FGS_PL_FIT, prm.flex_filter_first ? FGS_PL_FIT : FGS_PL_FLEXGPU};
    for (int it = 0; it < 8; ++it) {
        const i64 n = ((i64)blockIdx.x * 8 + it) * blockDim.x + threadIdx.x;
        if (n >= prm.n_nodes) break;
        if (!((uint32_t)prm.nd[(i64)ND_FLAGS * prm.n_pad + n] & FGS_NF_VALID)) {
            atomicAdd(&sh[FGS_DIAG_NOT_A_NODE], 1);
            continue;
        }
        for (int k = 0; k < 2; ++k) {
            if (!(prm.filter_mask & order[k])) continue;
            const uint16_t r = reason_of(prm, order[k], p, n);
            if (!r) continue;
            if (order[k] == FGS_PL_FLEXGPU) {
                atomicAdd(&sh[r == FGS_FLEX_REASON_PANIC ? FGS_DIAG_FLEX_PANIC : FGS_DIAG_FLEX0 + (int)r - 1], 1);
            } else {
                uint32_t m = r;
                while (m) {
                    const int b = __ffs(m) - 1;
                    m &= m - 1;
                    atomicAdd(&sh[FGS_DIAG_FIT0 + b], 1);
                }
            }
            break; // the first rejecting plugin's status is the node's status
        }
    }
    __syncthreads();
    if (threadIdx.x < FGS_DIAG_SLOTS && sh[threadIdx.x]) atomicAdd(&counts[p * FGS_DIAG_SLOTS + threadIdx.x], sh[threadIdx.x]);
}

} // namespace
} // namespace fgs

using namespace fgs;

// =================================================================================================

extern "C" int fgs_reason_plane(fgs_engine* eh, uint32_t plugin, uint16_t* out) {
    if (!eh || !out) return FGS_E_INVALID;
    Engine* e = (Engine*)eh;
    FGS_ENTER(e);
    if (!e->have_pods) return e->fail(FGS_E_STATE, "fgs_reason_plane before an evaluation");
    if (plugin != FGS_PL_FLEXGPU && plugin != FGS_PL_FIT) return e->fail(FGS_E_INVALID, "fgs_reason_plane: plugin must be FlexGPU or Fit");
    cudaSetDevice(e->device);
    const size_t cells = (size_t)e->n_pods * (size_t)e->n_nodes;
    if (cells == 0) return FGS_OK;
    uint16_t* d = nullptr;
    FGS_CUDA(e, cudaMalloc(&d, cells * sizeof(uint16_t)));
    StatusParams sp;
    std::memset(&sp, 0, sizeof(sp));
    sp.nd = (const i64*)e->d_nd;
    sp.raw = (const i64*)e->d_raw;
    sp.pods = (const i64*)e->d_pods;
    sp.n_nodes = e->n_nodes;
    sp.n_pad = e->n_pad;
    sp.p_pad = e->p_pad;
    sp.S = e->S;
    sp.D = e->D;
    sp.plugin = plugin;
    sp.fit_ignored = e->cfg.fit_ignored_scalar_mask;
    for (int64_t p0 = 0; p0 < e->n_pods; p0 += 32768) { // grid.y is limited to 65535
        StatusParams s2 = sp;
        const int64_t cnt = std::min<int64_t>(32768, e->n_pods - p0);
        s2.pods = sp.pods + p0;
        s2.n_pods = cnt;
        s2.out16 = d + p0 * e->n_nodes;
        dim3 grid((unsigned)((e->n_nodes + 255) / 256), (unsigned)cnt);
        reason_plane<<<grid, 256, 0, e->stream>>>(s2);
        e->launches++;
    }
    cudaError_t ce = cudaGetLastError();
    if (ce == cudaSuccess) ce = cudaMemcpyAsync(out, d, cells * sizeof(uint16_t), cudaMemcpyDeviceToHost, e->stream);
    if (ce == cudaSuccess) ce = cudaStreamSynchronize(e->stream);
    cudaFree(d);
    if (ce != cudaSuccess) return e->fail(FGS_E_CUDA, std::string("fgs_reason_plane: ") + cudaGetErrorString(ce));
    return FGS_OK;
}

extern "C" int fgs_status_plane(fgs_engine* eh, uint32_t plugin, uint8_t* out) {
    if (!eh || !out) return FGS_E_INVALID;
    Engine* e = (Engine*)eh;
    FGS_ENTER(e);
    if (!e->have_pods) return e->fail(FGS_E_STATE, "fgs_status_plane before an evaluation");
    if (plugin != 0 && plugin != FGS_PL_FLEXGPU && plugin != FGS_PL_FIT) return e->fail(FGS_E_INVALID, "fgs_status_plane: plugin must be 0, FlexGPU or Fit");
    cudaSetDevice(e->device);
    const size_t bytes = (size_t)e->n_pods * (size_t)e->n_nodes;
    if (bytes == 0) return FGS_OK;
    uint8_t* d = nullptr;
    FGS_CUDA(e, cudaMalloc(&d, bytes));
    EvalParams ep;
    e->fill_eval_params(&ep, nullptr);
    StatusParams sp;
    std::memset(&sp, 0, sizeof(sp));
    sp.nd = (const i64*)e->d_nd;
    sp.pods = (const i64*)e->d_pods;
    sp.n_nodes = e->n_nodes;
    sp.n_pad = e->n_pad;
    sp.n_pods = e->n_pods;
    sp.p_pad = e->p_pad;
    sp.S = e->S;
    sp.plugin = plugin;
    sp.filter_mask = e->cfg.filter_mask;
    sp.fit_ignored = e->cfg.fit_ignored_scalar_mask;
    sp.flex_filter_first = ep.flex_filter_first;
    sp.out = d;
    int rc = FGS_OK;
    // grid.y is limited to 65535: slice the pod range
    for (int64_t p0 = 0; p0 < e->n_pods && rc == FGS_OK; p0 += 32768) {
        StatusParams s2 = sp;
        const int64_t cnt = std::min<int64_t>(32768, e->n_pods - p0);
        s2.pods = sp.pods + p0;
        s2.n_pods = cnt;
        s2.out = d + p0 * e->n_nodes;
        dim3 grid((unsigned)((e->n_nodes + 255) / 256), (unsigned)cnt);
        status_plane<<<grid, 256, 0, e->stream>>>(s2);
        e->launches++;
    }
    cudaError_t ce = cudaGetLastError();
    if (ce == cudaSuccess) ce = cudaMemcpyAsync(out, d, bytes, cudaMemcpyDeviceToHost, e->stream);
    if (ce == cudaSuccess) ce = cudaStreamSynchronize(e->stream);
    cudaFree(d);
    if (ce != cudaSuccess) return e->fail(FGS_E_CUDA, std::string("fgs_status_plane: ") + cudaGetErrorString(ce));
    return rc;
}

static int gang_check_impl(Engine* e, const fgs_snapshot* snap, const fgs_podset* pending, const fgs_podgroups* pgs, uint8_t* prefilter_code,
                           uint8_t* permit_status, int32_t* assigned, uint8_t* postfilter_code);

extern "C" int fgs_gang_check(fgs_engine* eh, const fgs_snapshot* snap, const fgs_podset* pending, const fgs_podgroups* pgs,
                              uint8_t* prefilter_code, uint8_t* permit_status, int32_t* assigned) {
    if (!eh || !snap || !pending || !pgs) return FGS_E_INVALID;
    Engine* e = (Engine*)eh;
    FGS_ENTER(e);
    return gang_check_impl(e, snap, pending, pgs, prefilter_code, permit_status, assigned, nullptr);
}

extern "C" int fgs_gang_postfilter(fgs_engine* eh, const fgs_snapshot* snap, const fgs_podset* pending, const fgs_podgroups* pgs,
                                   uint8_t* postfilter_code, int32_t* assigned) {
    if (!eh || !snap || !pending || !pgs || !postfilter_code) return FGS_E_INVALID;
    Engine* e = (Engine*)eh;
    FGS_ENTER(e);
    return gang_check_impl(e, snap, pending, pgs, nullptr, nullptr, assigned, postfilter_code);
}

static int gang_check_impl(Engine* e, const fgs_snapshot* snap, const fgs_podset* pending, const fgs_podgroups* pgs, uint8_t* prefilter_code,
                           uint8_t* permit_status, int32_t* assigned, uint8_t* postfilter_code) {
    if (!e->have_nodes) return e->fail(FGS_E_STATE, "fgs_gang_check before fgs_upload_nodes");
    if (snap->n_nodes != e->n_nodes) return e->fail(FGS_E_INVALID, "fgs_gang_check: snapshot does not match the uploaded node table");
    if (snap->n_scalars != e->S || pgs->n_scalars != e->S) return e->fail(FGS_E_INVALID, "fgs_gang_check: n_scalars mismatch");
    cudaSetDevice(e->device);
    if (!e->nodes_derived) return e->fail(FGS_E_STATE, "fgs_gang_check: node columns not derived");
    const int S = e->S, K = 4 + S;
    const int64_t G = pgs->n_groups, NE = snap->pods.n_pods, NP = pending->n_pods, N = e->n_nodes;

    // corrections: the group's own pods on nodes, as running sums per group in node order
    // (host-side packing of the sparse per-group term; the dense walk runs on the GPU)
    std::vector<std::vector<std::pair<int32_t, int64_t>>> by_group((size_t)G);
    for (int64_t n = 0; n < N; ++n) {
        if (!snap->node_valid[n]) continue;
        for (int64_t q = snap->pod_off[n]; q < snap->pod_off[n + 1]; ++q) {
            const int32_t g = snap->pods.pg[q];
            if (g >= 0 && g < G) by_group[(size_t)g].push_back({(int32_t)n, q});
        }
    }
    std::vector<int64_t> corr_off((size_t)G + 1, 0);
    for (int64_t g = 0; g < G; ++g) corr_off[(size_t)g + 1] = corr_off[(size_t)g] + (int64_t)by_group[(size_t)g].size();
    const int64_t NC = corr_off[(size_t)G];

    Bump b;
    const size_t o_ex_pg = b.take(sizeof(int32_t) * NE), o_ex_ns = b.take(sizeof(int32_t) * NE), o_ex_named = b.take(NE);
    const size_t o_pend_pg = b.take(sizeof(int32_t) * NP);
    const size_t o_pg_ns = b.take(sizeof(int32_t) * G), o_mm = b.take(sizeof(int32_t) * G), o_gfl = b.take(G);
    const size_t o_minres = b.take(sizeof(int64_t) * G * K), o_minp = b.take(sizeof(uint32_t) * G);
    const size_t o_coff = b.take(sizeof(int64_t) * (G + 1)), o_cnode = b.take(sizeof(int32_t) * NC), o_cdelta = b.take(sizeof(int64_t) * NC * K);
    const size_t in_bytes = b.off;
    const size_t o_cand = b.take(sizeof(int64_t) * K * e->n_pad), o_colmax = b.take(sizeof(int64_t) * K);
    const size_t o_counts = b.take(sizeof(int32_t) * 2 * G), o_gcode = b.take(G);
    const size_t o_pre = b.take(NP), o_permit = b.take(NP), o_post = b.take(NP);
    int rc = arena_reserve(e, b.off);
    if (rc) return rc;
    FGS_CUDA(e, cudaStreamSynchronize(e->stream));
    char* h = (char*)e->gang->h;
    char* d = (char*)e->gang->d;
    if (NE) {
        std::memcpy(h + o_ex_pg, snap->pods.pg, sizeof(int32_t) * NE);
        std::memcpy(h + o_ex_ns, snap->pods.ns, sizeof(int32_t) * NE);
        std::memcpy(h + o_ex_named, snap->pods.has_node_name, NE);
    }
    if (NP) std::memcpy(h + o_pend_pg, pending->pg, sizeof(int32_t) * NP);
    for (int64_t g = 0; g < G; ++g) {
        ((int32_t*)(h + o_pg_ns))[g] = pgs->ns[g];
        ((int32_t*)(h + o_mm))[g] = pgs->min_member[g];
        ((uint8_t*)(h + o_gfl))[g] = (uint8_t)((pgs->exists[g] ? 1 : 0) | (pgs->has_min_resources[g] ? 2 : 0) | (pgs->denied[g] ? 4 : 0) |
                                               (pgs->permitted[g] ? 8 : 0));
        ((uint32_t*)(h + o_minp))[g] = pgs->min_res_present[g];
    }
    if (G) std::memcpy(h + o_minres, pgs->min_res, sizeof(int64_t) * G * K);
    std::memcpy(h + o_coff, corr_off.data(), sizeof(int64_t) * (G + 1));
    {
        int32_t* cn = (int32_t*)(h + o_cnode);
        int64_t* cd = (int64_t*)(h + o_cdelta);
        for (int64_t g = 0; g < G; ++g) {
            int64_t run[4 + FGS_MAX_SCALARS] = {0};
            int64_t j = corr_off[(size_t)g];
            for (auto& pr : by_group[(size_t)g]) { // already in node order
                const PodRequest r = pod_request(&snap->pods, pr.second); // RemovePod's calculateResource (types.go:715-745)
                run[0] += r.v[0];
                run[1] += r.v[1];
                run[2] += r.v[2];
                run[3] += 1;
                for (int s = 0; s < S; ++s)
                    if ((snap->alloc_scalar_present[pr.first] >> s) & 1u) run[4 + s] += r.v[3 + s];
                cn[j] = pr.first;
                for (int k = 0; k < K; ++k) cd[j * K + k] = run[k];
                ++j;
            }
        }
    }
    FGS_CUDA(e, cudaMemcpyAsync(d, h, in_bytes, cudaMemcpyHostToDevice, e->stream));
    FGS_CUDA(e, cudaMemsetAsync(d + o_counts, 0, sizeof(int32_t) * 2 * G, e->stream));
    if (N > 0) {
        gang_prefix_scan<<<K, kScanThreads, 0, e->stream>>>((const i64*)e->d_nd, (const i64*)e->d_raw, N, e->n_pad, S, (i64*)(d + o_cand),
                                                             (i64*)(d + o_colmax));
        e->launches++;
    } else {
        std::vector<int64_t> mn((size_t)K, INT64_MIN);
        FGS_CUDA(e, cudaMemcpyAsync(d + o_colmax, mn.data(), sizeof(int64_t) * K, cudaMemcpyHostToDevice, e->stream));
        FGS_CUDA(e, cudaStreamSynchronize(e->stream));
    }
    if (NE + NP > 0 && G > 0) {
        group_histograms<<<(unsigned)((NE + NP + 255) / 256), 256, 0, e->stream>>>(
            (const int32_t*)(d + o_ex_pg), (const int32_t*)(d + o_ex_ns), (const uint8_t*)(d + o_ex_named), NE, (const int32_t*)(d + o_pend_pg), NP,
            (const int32_t*)(d + o_pg_ns), G, (int32_t*)(d + o_counts));
        e->launches++;
    }
    if (G > 0) {
        GangGroupParams gp;
        std::memset(&gp, 0, sizeof(gp));
        gp.cand = (const i64*)(d + o_cand);
        gp.colmax = (const i64*)(d + o_colmax);
        gp.min_res = (const i64*)(d + o_minres);
        gp.min_present = (const uint32_t*)(d + o_minp);
        gp.min_member = (const int32_t*)(d + o_mm);
        gp.gflags = (const uint8_t*)(d + o_gfl);
        gp.counts = (const int32_t*)(d + o_counts);
        gp.corr_off = (const i64*)(d + o_coff);
        gp.corr_node = (const int32_t*)(d + o_cnode);
        gp.corr_delta = (const i64*)(d + o_cdelta);
        gp.gcode = (uint8_t*)(d + o_gcode);
        gp.n_nodes = N;
        gp.n_pad = e->n_pad;
        gp.G = G;
        gp.K = K;
        gang_check_groups<<<(unsigned)G, 256, 0, e->stream>>>(gp);
        e->launches++;
    }
    if (NP > 0) {
        gang_pod_codes<<<(unsigned)((NP + 255) / 256), 256, 0, e->stream>>>((const int32_t*)(d + o_pend_pg), NP, (const uint8_t*)(d + o_gcode),
                                                                            (const uint8_t*)(d + o_gfl), (const int32_t*)(d + o_counts),
                                                                            (const int32_t*)(d + o_mm), G, (uint8_t*)(d + o_pre),
                                                                            (uint8_t*)(d + o_permit), (uint8_t*)(d + o_post));
        e->launches++;
    }
    FGS_CUDA(e, cudaGetLastError());
    if (NP > 0) {
        FGS_CUDA(e, cudaMemcpyAsync(h + o_pre, d + o_pre, NP, cudaMemcpyDeviceToHost, e->stream));
        FGS_CUDA(e, cudaMemcpyAsync(h + o_permit, d + o_permit, NP, cudaMemcpyDeviceToHost, e->stream));
        FGS_CUDA(e, cudaMemcpyAsync(h + o_post, d + o_post, NP, cudaMemcpyDeviceToHost, e->stream));
    }
    if (G > 0) FGS_CUDA(e, cudaMemcpyAsync(h + o_counts, d + o_counts, sizeof(int32_t) * 2 * G, cudaMemcpyDeviceToHost, e->stream));
    FGS_CUDA(e, cudaStreamSynchronize(e->stream));
    if (prefilter_code && NP) std::memcpy(prefilter_code, h + o_pre, NP);
    if (permit_status && NP) std::memcpy(permit_status, h + o_permit, NP);
    if (postfilter_code && NP) std::memcpy(postfilter_code, h + o_post, NP);
    if (assigned && G) std::memcpy(assigned, h + o_counts + sizeof(int32_t) * G, sizeof(int32_t) * G);
    return FGS_OK;
}

extern "C" int fgs_quota_check(fgs_engine* eh, const fgs_pod_table* pods, const fgs_quotas* q, const fgs_pod_table* nominated,
                               const int64_t* pod_uid, const int64_t* nominated_uid, uint8_t* code) {
    if (!eh || !pods || !q || !code) return FGS_E_INVALID;
    Engine* e = (Engine*)eh;
    FGS_ENTER(e);
    if (pods->n_scalars != e->S || q->n_scalars != e->S) return e->fail(FGS_E_INVALID, "fgs_quota_check: n_scalars mismatch");
    if (nominated && nominated->n_scalars != e->S) return e->fail(FGS_E_INVALID, "fgs_quota_check: nominated n_scalars mismatch");
    cudaSetDevice(e->device);
    const int S = e->S, W = 2 + S;
    const int64_t P = pods->n_pods, NS = q->n_ns, Q = nominated ? nominated->n_pods : 0;
    if (P == 0) return FGS_OK;
    Bump b;
    const size_t o_has = b.take(NS), o_min = b.take(sizeof(int64_t) * NS * W), o_max = b.take(sizeof(int64_t) * NS * W),
                 o_used = b.take(sizeof(int64_t) * NS * W), o_up = b.take(sizeof(uint32_t) * NS);
    const size_t o_preq = b.take(sizeof(int64_t) * W * P), o_ppres = b.take(sizeof(uint32_t) * P), o_pns = b.take(sizeof(int32_t) * P),
                 o_pprio = b.take(sizeof(int32_t) * P), o_puid = b.take(sizeof(int64_t) * P);
    const size_t o_nreq = b.take(sizeof(int64_t) * W * Q), o_npres = b.take(sizeof(uint32_t) * Q), o_nns = b.take(sizeof(int32_t) * Q),
                 o_nprio = b.take(sizeof(int32_t) * Q), o_nuid = b.take(sizeof(int64_t) * Q);
    const size_t in_bytes = b.off;
    const size_t o_agg = b.take(sizeof(int64_t) * 2 * W), o_aggp = b.take(sizeof(uint32_t)), o_over = b.take(NS), o_code = b.take(P);
    int rc = arena_reserve(e, b.off);
    if (rc) return rc;
    FGS_CUDA(e, cudaStreamSynchronize(e->stream));
    char* h = (char*)e->gang->h;
    char* d = (char*)e->gang->d;
    std::memcpy(h + o_has, q->has_eq, NS);
    std::memcpy(h + o_min, q->min, sizeof(int64_t) * NS * W);
    std::memcpy(h + o_max, q->max, sizeof(int64_t) * NS * W);
    std::memcpy(h + o_used, q->used, sizeof(int64_t) * NS * W);
    std::memcpy(h + o_up, q->used_scalar_present, sizeof(uint32_t) * NS);
    auto put_pods = [&](const fgs_pod_table* t, int64_t n, size_t oreq, size_t opres, size_t ons, size_t oprio) {
        if (n == 0) return;
        int64_t* r = (int64_t*)(h + oreq);
        std::memcpy(r + 0 * n, t->req_mcpu, sizeof(int64_t) * n);
        std::memcpy(r + 1 * n, t->req_mem, sizeof(int64_t) * n);
        for (int s = 0; s < S; ++s) std::memcpy(r + (int64_t)(2 + s) * n, t->req_scalar + (int64_t)s * n, sizeof(int64_t) * n);
        std::memcpy(h + opres, t->req_scalar_present, sizeof(uint32_t) * n);
        std::memcpy(h + ons, t->ns, sizeof(int32_t) * n);
        std::memcpy(h + oprio, t->priority, sizeof(int32_t) * n);
    };
    put_pods(pods, P, o_preq, o_ppres, o_pns, o_pprio);
    if (Q) put_pods(nominated, Q, o_nreq, o_npres, o_nns, o_nprio);
    const bool uids = pod_uid && nominated_uid && Q > 0;
    if (uids) {
        std::memcpy(h + o_puid, pod_uid, sizeof(int64_t) * P);
        std::memcpy(h + o_nuid, nominated_uid, sizeof(int64_t) * Q);
    }
    FGS_CUDA(e, cudaMemcpyAsync(d, h, in_bytes, cudaMemcpyHostToDevice, e->stream));
    QuotaParams qp;
    std::memset(&qp, 0, sizeof(qp));
    qp.has_eq = (const uint8_t*)(d + o_has);
    qp.qmin = (const i64*)(d + o_min);
    qp.qmax = (const i64*)(d + o_max);
    qp.qused = (const i64*)(d + o_used);
    qp.used_present = (const uint32_t*)(d + o_up);
    qp.agg = (i64*)(d + o_agg);
    qp.agg_present = (uint32_t*)(d + o_aggp);
    qp.ns_over_min = (uint8_t*)(d + o_over);
    qp.n_ns = NS;
    qp.S = S;
    qp.W = W;
    qp.pod_req = (const i64*)(d + o_preq);
    qp.pod_present = (const uint32_t*)(d + o_ppres);
    qp.pod_ns = (const int32_t*)(d + o_pns);
    qp.pod_prio = (const int32_t*)(d + o_pprio);
    qp.pod_uid = uids ? (const i64*)(d + o_puid) : nullptr;
    qp.n_pods = P;
    qp.nom_req = (const i64*)(d + o_nreq);
    qp.nom_present = (const uint32_t*)(d + o_npres);
    qp.nom_ns = (const int32_t*)(d + o_nns);
    qp.nom_prio = (const int32_t*)(d + o_nprio);
    qp.nom_uid = uids ? (const i64*)(d + o_nuid) : nullptr;
    qp.n_nom = Q;
    qp.code = (uint8_t*)(d + o_code);
    quota_aggregate<<<1, 256, 0, e->stream>>>(qp);
    quota_check_pods<<<(unsigned)((P + 255) / 256), 256, 0, e->stream>>>(qp);
    e->launches += 2;
    FGS_CUDA(e, cudaGetLastError());
    FGS_CUDA(e, cudaMemcpyAsync(h + o_code, d + o_code, P, cudaMemcpyDeviceToHost, e->stream));
    FGS_CUDA(e, cudaStreamSynchronize(e->stream));
    std::memcpy(code, h + o_code, P);
    return FGS_OK;
}

// =================================================================================================
// fgs_quota_check_static — the ElasticQuota check for callers that are not allowed to allocate
// =================================================================================================
//
// Who needs this: code that runs where device allocation is off the table — a scheduler plugin called on a
// latency-critical path that must never block on the CUDA allocator, a host that hands this library a fixed
// arena of device memory and forbids anything beyond it, or a caller that is in the middle of recording work
// and may not touch the allocator. fgs_quota_check sizes and grows an internal arena with cudaMalloc; this
// variant does not: the tables and the result live in device memory the caller already owns, and the only
// working memory the kernels need (the Σused / Σmin aggregate, its key mask, one byte per namespace) sits in
// the fixed-size __device__ array below, which every GPU that loads libfgs carries until the process exits
// (about 4.6 KB). It runs the same two kernels as fgs_quota_check (quota_aggregate, quota_check_pods), without
// nominated pods.
//
// Rules that follow from the shared, fixed scratch:
//   * at most FGS_STATIC_QUOTA_MAX_NS namespaces and FGS_MAX_SCALARS scalar slots — larger inputs are refused;
//   * the FIRST call on each device looks up the device address of the scratch (cudaGetSymbolAddress) and keeps
//     it; make that first call during start-up, before any stream capture begins on that device;
//   * the scratch keeps whatever the previous call left in it; each call clears the part it reads before use;
//   * all calls on one device share the scratch: two calls whose kernels may overlap on the GPU (different
//     streams, no ordering between them) race. Order them on one stream or with events.
#define FGS_STATIC_QUOTA_MAX_NS 4096

namespace {
constexpr size_t kStaticAggOff = 0;        // int64 agg[2][2 + FGS_MAX_SCALARS]
constexpr size_t kStaticPresentOff = 256;  // uint32 key mask of Σused
constexpr size_t kStaticOverOff = 512;     // uint8 ns_over_min[FGS_STATIC_QUOTA_MAX_NS]
constexpr size_t kStaticBytes = kStaticOverOff + FGS_STATIC_QUOTA_MAX_NS;
constexpr int kStaticMaxDevices = 64;

// In host code this name is only a handle for cudaGetSymbolAddress — not an address a kernel could use.
static __device__ __align__(256) unsigned char g_quota_static_scratch[kStaticBytes];
// device address of the array, per device ordinal: resolved by the first call on a device. Two host threads may make that
// first call together: both resolve the same address, the atomic only keeps the publication well defined. The address
// does not outlive the device's primary context (cudaDeviceReset): a process that resets a device must not call again.
static std::atomic<void*> g_quota_static_dev_ptr[kStaticMaxDevices];

void static_err(char* buf, int64_t len, const char* msg) {
    if (!buf || len <= 0) return;
    int64_t i = 0;
    for (; msg[i] && i < len - 1; ++i) buf[i] = msg[i];
    buf[i] = 0;
}
} // namespace

extern "C" int fgs_quota_check_static(void* stream, int32_t n_scalars, int64_t n_ns, const uint8_t* dev_has_eq, const int64_t* dev_min,
                                      const int64_t* dev_max, const int64_t* dev_used, const uint32_t* dev_used_present, int64_t n_pods,
                                      const int64_t* dev_pod_req, const uint32_t* dev_pod_present, const int32_t* dev_pod_ns,
                                      uint8_t* dev_code, char* errbuf, int64_t errbuf_len) {
    if (n_scalars < 2 || n_scalars > FGS_MAX_SCALARS) {
        static_err(errbuf, errbuf_len, "fgs_quota_check_static: n_scalars must be in [2, 8]");
        return FGS_E_INVALID;
    }
    if (n_ns < 0 || n_ns > FGS_STATIC_QUOTA_MAX_NS) {
        static_err(errbuf, errbuf_len, "fgs_quota_check_static: more than 4096 namespaces do not fit the fixed scratch; use fgs_quota_check");
        return FGS_E_TOO_LARGE;
    }
    if (n_pods < 0 || (n_pods > 0 && (!dev_pod_req || !dev_pod_present || !dev_pod_ns || !dev_code)) ||
        (n_ns > 0 && (!dev_has_eq || !dev_min || !dev_max || !dev_used || !dev_used_present))) {
        static_err(errbuf, errbuf_len, "fgs_quota_check_static: null table pointer");
        return FGS_E_INVALID;
    }
    if (n_pods == 0) return FGS_OK;
    int dev = -1;
    cudaError_t ce = cudaGetDevice(&dev);
    if (ce != cudaSuccess || dev < 0 || dev >= kStaticMaxDevices) {
        static_err(errbuf, errbuf_len, "fgs_quota_check_static: no current CUDA device (or ordinal >= 64)");
        return FGS_E_CUDA;
    }
    // once per device, with that device current: resolve the device address of the scratch array
    void* resolved = g_quota_static_dev_ptr[dev].load(std::memory_order_acquire);
    if (!resolved) {
        void* ptr = nullptr;
        ce = cudaGetSymbolAddress((void**)&ptr, g_quota_static_scratch);
        if (ce != cudaSuccess || !ptr) {
            static_err(errbuf, errbuf_len, cudaGetErrorString(ce));
            return FGS_E_CUDA;
        }
        g_quota_static_dev_ptr[dev].store(ptr, std::memory_order_release);
        resolved = ptr;
    }
    unsigned char* scratch = (unsigned char*)resolved;
    cudaStream_t st = (cudaStream_t)stream;
    // the array still holds the previous call's aggregate and flags: start from zero
    ce = cudaMemsetAsync(scratch, 0, kStaticOverOff + (size_t)n_ns, st);
    if (ce != cudaSuccess) {
        static_err(errbuf, errbuf_len, cudaGetErrorString(ce));
        return FGS_E_CUDA;
    }
    QuotaParams qp;
    std::memset(&qp, 0, sizeof(qp));
    qp.has_eq = dev_has_eq;
    qp.qmin = (const i64*)dev_min;
    qp.qmax = (const i64*)dev_max;
    qp.qused = (const i64*)dev_used;
    qp.used_present = dev_used_present;
    qp.agg = (i64*)(scratch + kStaticAggOff);
    qp.agg_present = (uint32_t*)(scratch + kStaticPresentOff);
    qp.ns_over_min = scratch + kStaticOverOff;
    qp.n_ns = n_ns;
    qp.S = n_scalars;
    qp.W = 2 + n_scalars;
    qp.pod_req = (const i64*)dev_pod_req;
    qp.pod_present = dev_pod_present;
    qp.pod_ns = dev_pod_ns;
    qp.pod_prio = dev_pod_ns; // priorities only matter against nominated pods, and this variant has none
    qp.pod_uid = nullptr;
    qp.n_pods = n_pods;
    qp.n_nom = 0;
    qp.code = dev_code;
    quota_aggregate<<<1, 256, 0, st>>>(qp);
    quota_check_pods<<<(unsigned)((n_pods + 255) / 256), 256, 0, st>>>(qp);
    ce = cudaGetLastError();
    if (ce != cudaSuccess) {
        static_err(errbuf, errbuf_len, cudaGetErrorString(ce));
        return FGS_E_CUDA;
    }
    return FGS_OK;
}

extern "C" int fgs_fit_error_histogram(fgs_engine* eh, int32_t* counts) {
    if (!eh || !counts) return FGS_E_INVALID;
    Engine* e = (Engine*)eh;
    FGS_ENTER(e);
    if (!e->have_pods) return e->fail(FGS_E_STATE, "fgs_fit_error_histogram before an evaluation");
    if (e->in_flight) return e->fail(FGS_E_STATE, "fgs_fit_error_histogram while submitted batches are in flight");
    cudaSetDevice(e->device);
    const int64_t P = e->n_pods;
    if (P == 0) return FGS_OK;
    int* d = nullptr;
    const size_t bytes = sizeof(int) * (size_t)P * FGS_DIAG_SLOTS;
    FGS_CUDA(e, cudaMalloc(&d, bytes));
    cudaError_t ce = cudaMemsetAsync(d, 0, bytes, e->stream);
    EvalParams ep;
    e->fill_eval_params(&ep, nullptr);
    StatusParams sp;
    std::memset(&sp, 0, sizeof(sp));
    sp.nd = (const i64*)e->d_nd;
    sp.raw = (const i64*)e->d_raw;
    sp.pods = (const i64*)e->d_pods;
    sp.n_nodes = e->n_nodes;
    sp.n_pad = e->n_pad;
    sp.p_pad = e->p_pad;
    sp.S = e->S;
    sp.D = e->D;
    sp.filter_mask = ep.filter_mask;
    sp.fit_ignored = ep.fit_ignored;
    sp.flex_filter_first = ep.flex_filter_first;
    const i64* d_count = (const i64*)e->d_perpod + (int64_t)FGS_REDUCE_ROWS * e->p_pad; // merged over the shards when there are any
    for (int64_t p0 = 0; p0 < P && ce == cudaSuccess; p0 += 32768) { // grid.y is limited to 65535
        StatusParams s2 = sp;
        const int64_t cnt = std::min<int64_t>(32768, P - p0);
        s2.pods = sp.pods + p0;
        s2.n_pods = cnt;
        dim3 grid((unsigned)((e->n_nodes + 2047) / 2048), (unsigned)cnt);
        fit_error_histogram<<<grid, 256, 0, e->stream>>>(s2, d_count + p0, d + p0 * FGS_DIAG_SLOTS);
        e->launches++;
        ce = cudaGetLastError();
    }
    if (ce == cudaSuccess) ce = cudaMemcpyAsync(counts, d, bytes, cudaMemcpyDeviceToHost, e->stream);
    if (ce == cudaSuccess) ce = cudaStreamSynchronize(e->stream);
    cudaFree(d);
    if (ce != cudaSuccess) return e->fail(FGS_E_CUDA, std::string("fgs_fit_error_histogram: ") + cudaGetErrorString(ce));
    return FGS_OK;
}
