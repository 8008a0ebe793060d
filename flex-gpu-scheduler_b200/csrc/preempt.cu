// preempt.cu — batched dry-run preemption (second half): preemptor.SelectVictimsOnNode
// (pkg/capacityscheduling/capacity_scheduling.go:465-644) for ONE unschedulable pod on EVERY node at once.
//
// preemption.Evaluator.DryRunPreemption (vendor/k8s.io/kubernetes/pkg/scheduler/framework/preemption/
// preemption.go:546-597) clones the NodeInfo and the CycleState per node, so nodes are independent: one thread per
// node walks that node's pods — pick the potential victims by the ElasticQuota rules (:505-557), remove them all,
// run the filters (:573), check the quota bounds (:580-585), then re-add ("reprieve") the victims from the most
// important one down, keeping each one the pod still fits with (:589-634). The node state is the engine's raw row
// plus the deltas of the removed pods; device use (monopoly / used memory per index) is rebuilt from the pods that
// are still there, as constructGPUs does (gpu_node.go:66-120); the ElasticQuota state is the uploaded table minus
// the requests of the removed pods (the RemovePod / AddPod PreFilterExtensions, :277-318).
// No PodDisruptionBudgets, empty PodNominator; sort.Slice = insertion sort (what Go runs below 12 elements; with more
// pods per node Go's quicksort may order EQUALLY important pods differently — distinct (priority, start time) pairs
// are unaffected). A latency-bound side path: N threads, O(pods²) work per node.
#include <cuda_runtime.h>

#include <cstring>
#include <string>
#include <vector>

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

constexpr int kMaxPodsPerNode = 128;

// snapshot-pod columns (int64 [VP_FIXED + S][T])
enum VictimCol : int { VP_NS = 0, VP_PRIO, VP_START, VP_KIND, VP_GPU_LIM, VP_MEM_LIM, VP_ANNOT, VP_KEYS, VP_REQ_CPU, VP_REQ_MEM, VP_REQ_EPH, VP_FIXED };

struct PreemptParams {
    const i64* raw;      // engine raw node table
    const i64* vp;       // snapshot-pod columns
    const i64* pod_off;  // [N+1]
    i64 n_nodes, n_pad, T;
    int S, D;
    // ElasticQuota table (has_eq == NULL: none)
    const uint8_t* has_eq;
    const i64* qmin;
    const i64* qmax;
    const i64* qused;
    const uint32_t* qpresent;
    i64 n_ns;
    i64 agg_used[2 + FGS_MAX_SCALARS], agg_min[2 + FGS_MAX_SCALARS]; // Σ over the namespaces with a quota
    uint32_t agg_present;
    int tracks_pods;
    // the preemptor
    i64 p_req[3 + FGS_MAX_SCALARS]; // cpu, mem, eph, scalars
    uint32_t p_keys, p_flags;       // scalar keys of the request; FGS_PF_*
    i64 p_gpu_lim, p_mem_lim;
    int p_ns, p_prio;
    // filters
    uint32_t filter_mask, fit_ignored;
    int flex_filter_first;
    // out
    uint8_t* status;  // [N]
    uint8_t* victim;  // [T]
    int* n_victims;   // [N]
    i64* cand;        // [4][N] per candidate node: highest victim priority, Σ (priority + 2^31), victims, earliest start
                      // time among the victims of the highest priority — what pickOneNodeForPreemption compares
    int* picked;      // the chosen candidate, -1 if there is none
};

struct Mask128 {
    u64 w[2];
    __device__ bool get(int i) const { return (w[i >> 6] >> (i & 63)) & 1ull; }
    __device__ void set(int i) { w[i >> 6] |= 1ull << (i & 63); }
    __device__ void clr(int i) { w[i >> 6] &= ~(1ull << (i & 63)); }
};

__global__ void __launch_bounds__(64) select_victims_kernel(const __grid_constant__ PreemptParams prm) {
    const i64 n = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= prm.n_nodes) return;
    const int S = prm.S, D = prm.D, W = 2 + S;
    const i64 np = prm.n_pad, T = prm.T;
    auto R = [&](int c) { return prm.raw[(i64)c * np + n]; };
    const i64 q0 = prm.pod_off[n];
    const int cnt = (int)(prm.pod_off[n + 1] - q0);
    auto V = [&](int c, int a) { return prm.vp[(i64)c * T + q0 + a]; };
    const u64 meta = (u64)R(NR_META);
    const uint32_t nflags = (uint32_t)meta;
    prm.n_victims[n] = 0;
    if (!(nflags & FGS_NF_VALID)) {
        prm.status[n] = FGS_PREEMPT_NOT_A_NODE;
        return;
    }
    auto hasEQ = [&](i64 ns) { return prm.has_eq && ns >= 0 && ns < prm.n_ns && prm.has_eq[ns]; };
    // util.MoreImportantPod: higher priority, then the earlier start time
    auto moreImportant = [&](int a, int b) {
        const i64 pa = V(VP_PRIO, a), pb = V(VP_PRIO, b);
        return pa != pb ? pa > pb : V(VP_START, a) < V(VP_START, b);
    };

    Mask128 removed = {{0, 0}}, touched = {{0, 0}}; // removed now / removed at least once (its scalar keys exist in Used)
    // ---- node state = raw row − removed pods ------------------------------------------------------------------
    auto nodeReq = [&](int c) { // c: 0 cpu, 1 mem, 2 eph, 3+s scalar
        i64 v = c == 0 ? R(NR_REQ_MCPU) : (c == 1 ? R(NR_REQ_MEM) : (c == 2 ? R(NR_REQ_EPH) : R(nr_req_scalar(S, c - 3))));
        for (int a = 0; a < cnt; ++a)
            if (removed.get(a)) v -= V(c < 3 ? VP_REQ_CPU + c : VP_FIXED + (c - 3), a);
        return v;
    };
    auto numRemoved = [&]() { return __popcll(removed.w[0]) + __popcll(removed.w[1]); };
    // ---- ElasticQuota state = uploaded table − removed pods (only when the quotas track their pods) ---------------
    auto usedOf = [&](i64 ns, int c, uint32_t* keys) { // Used[ns] column c (0 cpu, 1 mem, 2+s scalar) and its scalar keys
        uint32_t k = prm.qpresent[ns];
        i64 v = (c < 2 || ((k >> (c - 2)) & 1u)) ? prm.qused[ns * W + c] : 0; // a key Used does not have reads 0
        if (prm.tracks_pods)
            for (int a = 0; a < cnt; ++a)
                if (V(VP_NS, a) == ns) {
                    if (touched.get(a)) k |= (uint32_t)V(VP_KEYS, a);
                    if (removed.get(a)) v -= c == 0 ? V(VP_REQ_CPU, a) : (c == 1 ? V(VP_REQ_MEM, a) : V(VP_FIXED + c - 2, a));
                }
        *keys = k;
        if (c >= 2 && !((k >> (c - 2)) & 1u)) v = 0; // a key Used does not have reads 0
        return v;
    };
    // cmp2(x1 = request-like {vals, keys}, x2 = Used[ns], y = bound[ns]) — elasticquota.go:165-181
    auto overBound = [&](const i64* x1, uint32_t x1keys, i64 ns, const i64* bound) {
        uint32_t k;
        if (x1[0] + usedOf(ns, 0, &k) > bound[ns * W + 0]) return true;
        if (x1[1] + usedOf(ns, 1, &k) > bound[ns * W + 1]) return true;
        for (int s = 0; s < S; ++s)
            if ((x1keys >> s) & 1u)
                if (x1[2 + s] + usedOf(ns, 2 + s, &k) > bound[ns * W + 2 + s]) return true;
        return false;
    };
    auto usedOverMin = [&](i64 ns) { // cmp(Used, Min): Used's own scalar keys (:98-100,161-163)
        uint32_t k;
        if (usedOf(ns, 0, &k) > prm.qmin[ns * W + 0]) return true;
        if (usedOf(ns, 1, &k) > prm.qmin[ns * W + 1]) return true;
        for (int s = 0; s < S; ++s) {
            const i64 u = usedOf(ns, 2 + s, &k);
            if (((k >> s) & 1u) && u > prm.qmin[ns * W + 2 + s]) return true;
        }
        return false;
    };
    i64 preq[2 + FGS_MAX_SCALARS]; // the preemptor's request in quota columns (no ephemeral storage)
    preq[0] = prm.p_req[0];
    preq[1] = prm.p_req[1];
    for (int s = 0; s < S; ++s) preq[2 + s] = prm.p_req[3 + s];
    auto overQuota = [&]() { // usedOverMaxWith(podReq) || aggregatedUsedOverMinWith(podReq) — :580-585, :606-607
        if (overBound(preq, prm.p_keys, prm.p_ns, prm.qmax)) return true;
        uint32_t keys = prm.agg_present | prm.p_keys;
        i64 agg[2 + FGS_MAX_SCALARS];
        for (int c = 0; c < W; ++c) agg[c] = prm.agg_used[c] + preq[c];
        if (prm.tracks_pods)
            for (int a = 0; a < cnt; ++a) {
                if (!hasEQ(V(VP_NS, a))) continue;
                if (touched.get(a)) keys |= (uint32_t)V(VP_KEYS, a);
                if (removed.get(a)) {
                    agg[0] -= V(VP_REQ_CPU, a);
                    agg[1] -= V(VP_REQ_MEM, a);
                    for (int s = 0; s < S; ++s) agg[2 + s] -= V(VP_FIXED + s, a);
                }
            }
        if (agg[0] > prm.agg_min[0] || agg[1] > prm.agg_min[1]) return true;
        for (int s = 0; s < S; ++s)
            if (((keys >> s) & 1u) && agg[2 + s] > prm.agg_min[2 + s]) return true;
        return false;
    };

    // ---- RunFilterPlugins on the node as it is now (runtime/framework.go:667-693) -----------------------------------
    const uint32_t kind = prm.p_flags & 3u;
    const i64 gpuAlloc = R(nr_alloc_scalar(FGS_SCALAR_FLEX_GPU)), memAlloc = R(nr_alloc_scalar(FGS_SCALAR_FLEX_MEM));
    enum { OK = 0, FAIL = 1, PANIC = 2 };
    auto flex = [&]() -> int { // FlexGPU.Filter, flex_gpu.go:41-119
        if (kind == 0) return OK;
        if (kind == 3) return FAIL;
        if (!(nflags & FGS_NF_HAS_FLEX)) return FAIL;
        i64 gsum = R(NR_FLEX_GPU_SUM), msum = R(NR_FLEX_MEM_SUM);
        for (int a = 0; a < cnt; ++a)
            if (removed.get(a)) {
                const uint32_t k = (uint32_t)V(VP_KIND, a);
                if (k & 1u) gsum -= V(VP_GPU_LIM, a);
                if (k & 2u) msum -= V(VP_MEM_LIM, a);
            }
        const i64 gl = kind == 1 ? prm.p_gpu_lim : 0, ml = kind == 2 ? prm.p_mem_lim : 0;
        if (gsum + gl > gpuAlloc) return FAIL; // :90-93
        if (msum + ml > memAlloc) return FAIL; // :95-98
        // NewGPUNode / constructGPUs over the pods that are still on the node (gpu_node.go:30-120)
        if (gpuAlloc <= 0) return PANIC; // divide by zero / negative make (a missing key reads 0)
        i64 used[FGS_MAX_DEVICES];
        uint32_t mono = 0;
        for (int d = 0; d < FGS_MAX_DEVICES; ++d) used[d] = 0;
        for (int a = 0; a < cnt; ++a) {
            if (removed.get(a)) continue;
            const uint32_t k = (uint32_t)V(VP_KIND, a);
            const i64 annot = V(VP_ANNOT, a);
            if (k == 0 || annot == (i64)FGS_ANNOT_NONE) continue;
            if (annot < 0 || annot >= gpuAlloc) return PANIC; // gpus[index] out of range (:110,115)
            if (k & 1u) mono |= 1u << annot;
            if (k & 2u) used[annot] += V(VP_MEM_LIM, a);
        }
        if (kind == 2) { // MemAssumeFitIndexes empty, :109-116
            const i64 each = ((nflags & FGS_NF_HAS_MEM_RES) ? memAlloc : 0) / gpuAlloc;
            bool fits = false;
            for (int d = 0; d < D && (i64)d < gpuAlloc; ++d)
                if (!((mono >> d) & 1u) && each >= used[d] + ml) fits = true;
            if (!fits) return FAIL;
        }
        return OK;
    };
    auto fit = [&]() -> int { // Fit.Filter → fitsRequest, fit.go:256-329
        bool bad = (R(NR_NUM_PODS) - numRemoved()) + 1 > R(NR_ALLOC_PODS);
        if (!(prm.p_flags & FGS_PF_REQ_NONE)) {
            bad = bad || prm.p_req[0] > R(NR_ALLOC_MCPU) - nodeReq(0) || prm.p_req[1] > R(NR_ALLOC_MEM) - nodeReq(1) ||
                  prm.p_req[2] > R(NR_ALLOC_EPH) - nodeReq(2);
            uint32_t mk = prm.p_keys & ~prm.fit_ignored;
            while (mk) {
                const int s = __ffs(mk) - 1;
                mk &= mk - 1;
                if (prm.p_req[3 + s] > R(nr_alloc_scalar(s)) - nodeReq(3 + s)) bad = true;
            }
        }
        return bad ? FAIL : OK;
    };
    auto filters = [&]() -> int {
        int verdict = OK;
        for (int k = 0; k < 2 && verdict == OK; ++k) {
            const bool is_flex = (k == 0) == (prm.flex_filter_first != 0);
            if (is_flex && (prm.filter_mask & FGS_PL_FLEXGPU)) verdict = flex();
            if (!is_flex && (prm.filter_mask & FGS_PL_FIT)) verdict = fit();
        }
        return verdict;
    };

    // ---- :505 the node's pods, least important first (insertion sort with less = !MoreImportantPod) -------------------
    if (cnt > kMaxPodsPerNode) { // refused on the host; kept as a guard
        prm.status[n] = FGS_PREEMPT_ERROR;
        return;
    }
    unsigned char order[kMaxPodsPerNode];
    for (int i = 0; i < cnt; ++i) order[i] = (unsigned char)i;
    for (int i = 1; i < cnt; ++i)
        for (int j = i; j > 0 && !moreImportant(order[j], order[j - 1]); --j) {
            const unsigned char t = order[j];
            order[j] = order[j - 1];
            order[j - 1] = t;
        }
    // ---- :506-557 potential victims, removed as they are found ----------------------------------------------------------
    unsigned char potential[kMaxPodsPerNode];
    int n_pot = 0;
    const bool withEQ = hasEQ(prm.p_ns);
    auto take = [&](int a) {
        potential[n_pot++] = (unsigned char)a;
        removed.set(a);
        touched.set(a);
    };
    if (withEQ) {
        const bool moreThanMin = overBound(preq, prm.p_keys, prm.p_ns, prm.qmin); // usedOverMinWith(podReq), before any removal
        for (int i = 0; i < cnt; ++i) {
            const int a = order[i];
            const i64 ns = V(VP_NS, a);
            if (!hasEQ(ns)) continue;
            if (moreThanMin) {
                if (ns == prm.p_ns && V(VP_PRIO, a) < prm.p_prio) take(a);
            } else {
                if (ns != prm.p_ns && usedOverMin(ns)) take(a);
            }
        }
    } else {
        for (int i = 0; i < cnt; ++i) {
            const int a = order[i];
            if (hasEQ(V(VP_NS, a))) continue;
            if (V(VP_PRIO, a) < prm.p_prio) take(a);
        }
    }
    if (n_pot == 0) {
        prm.status[n] = FGS_PREEMPT_NO_VICTIMS;
        return;
    }
    int f = filters(); // :573-575
    if (f != OK) {
        prm.status[n] = f == PANIC ? FGS_PREEMPT_PANIC : FGS_PREEMPT_STILL_UNFIT;
        return;
    }
    if (withEQ && overQuota()) { // :580-585
        prm.status[n] = FGS_PREEMPT_QUOTA;
        return;
    }
    // ---- :589-634 reprieve, most important first -----------------------------------------------------------------------
    for (int i = 1; i < n_pot; ++i)
        for (int j = i; j > 0 && moreImportant(potential[j], potential[j - 1]); --j) {
            const unsigned char t = potential[j];
            potential[j] = potential[j - 1];
            potential[j - 1] = t;
        }
    int n_vic = 0;
    uint8_t st = FGS_PREEMPT_CANDIDATE;
    for (int i = 0; i < n_pot; ++i) {
        const int a = potential[i];
        removed.clr(a); // addPod
        f = filters();
        if (f == PANIC) {
            st = FGS_PREEMPT_PANIC;
            break;
        }
        const bool ok = f == OK;
        bool is_victim = false;
        if (!ok) {
            removed.set(a);
            is_victim = true;
        }
        if (withEQ && overQuota()) {
            if (!ok) { // RemovePod of a pod that has just been removed
                st = FGS_PREEMPT_ERROR;
                break;
            }
            removed.set(a);
            is_victim = true;
        }
        if (is_victim) n_vic++;
    }
    if (st == FGS_PREEMPT_CANDIDATE && n_vic == 0) st = FGS_PREEMPT_ERROR; // preemption.go:578-580
    prm.status[n] = st;
    if (st == FGS_PREEMPT_CANDIDATE) {
        prm.n_victims[n] = n_vic;
        i64 hi = INT64_MIN, sum = 0, earliest = 0;
        for (int a = 0; a < cnt; ++a)
            if (removed.get(a)) {
                prm.victim[q0 + a] = 1; // what is still removed at the end is exactly the victim list
                const i64 pr = V(VP_PRIO, a), t0 = V(VP_START, a);
                sum += pr + 2147483648ll;
                if (pr > hi) hi = pr, earliest = t0; // util.GetEarliestPodStartTime (scheduler/util/utils.go:56-78)
                else if (pr == hi && t0 < earliest) earliest = t0;
            }
        prm.cand[n] = hi;
        prm.cand[prm.n_nodes + n] = sum;
        prm.cand[2 * prm.n_nodes + n] = n_vic;
        prm.cand[3 * prm.n_nodes + n] = earliest;
    }
}

// pickOneNodeForPreemption (vendor/k8s.io/kubernetes/pkg/scheduler/framework/preemption/preemption.go:386-505) over the
// candidates: no PodDisruptionBudgets here, so criterion 1 ties everywhere; then the lowest "highest victim priority",
// the smallest sum of priorities, the fewest victims, the LATEST earliest-start-time of the highest-priority victims.
// The reference breaks what is left by Go map order ("sort of randomly"); here the lowest node index wins.
struct PickKey {
    i64 hi, sum, cnt, start;
    int node;
};
__device__ __forceinline__ bool pick_better(const PickKey& a, const PickKey& b) { // is a a better choice than b?
    if (b.node < 0) return a.node >= 0;
    if (a.node < 0) return false;
    if (a.hi != b.hi) return a.hi < b.hi;
    if (a.sum != b.sum) return a.sum < b.sum;
    if (a.cnt != b.cnt) return a.cnt < b.cnt;
    if (a.start != b.start) return a.start > b.start;
    return a.node < b.node;
}
__global__ void __launch_bounds__(1024) pick_candidate_kernel(const uint8_t* __restrict__ status, const i64* __restrict__ cand, i64 N, int* __restrict__ picked) {
    __shared__ PickKey sh[1024];
    PickKey best = {0, 0, 0, 0, -1};
    for (i64 n = threadIdx.x; n < N; n += blockDim.x) {
        if (status[n] != FGS_PREEMPT_CANDIDATE) continue;
        const PickKey k = {cand[n], cand[N + n], cand[2 * N + n], cand[3 * N + n], (int)n};
        if (pick_better(k, best)) best = k;
    }
    sh[threadIdx.x] = best;
    __syncthreads();
    for (int w = blockDim.x / 2; w > 0; w >>= 1) {
        if ((int)threadIdx.x < w && pick_better(sh[threadIdx.x + w], sh[threadIdx.x])) sh[threadIdx.x] = sh[threadIdx.x + w];
        __syncthreads();
    }
    if (threadIdx.x == 0) *picked = sh[0].node;
}

} // namespace

} // namespace fgs

using namespace fgs;

// The snapshot's pods as device columns, kept between calls: a PostFilter storm asks about many pods against ONE
// snapshot, and packing ~10 pods per node on the host dwarfs the kernel. Valid for the same fgs_snapshot object as long
// as the node table has not been uploaded or patched again (Engine::node_gen).
namespace fgs {
struct PreemptCache {
    const fgs_snapshot* snap = nullptr;
    uint64_t node_gen = 0;
    bool valid = false;
    int64_t N = 0, T = 0;
    int64_t* d_vp = nullptr; // [ncols][Tn] | pod_off [N + 1]
    size_t vp_bytes = 0;
    char* d_work = nullptr;  // quota tables + outputs of one call
    size_t work_bytes = 0;
};
void Engine::preempt_free() {
    if (!pcache) return;
    if (pcache->d_vp) cudaFree(pcache->d_vp);
    if (pcache->d_work) cudaFree(pcache->d_work);
    delete pcache;
    pcache = nullptr;
}
} // namespace fgs

static int preempt_prepare(Engine* e, const fgs_snapshot* snap) {
    if (!e->pcache) e->pcache = new PreemptCache();
    PreemptCache* pc = e->pcache;
    if (pc->valid && pc->snap == snap && pc->node_gen == e->node_gen && pc->N == snap->n_nodes && pc->T == snap->pods.n_pods) return FGS_OK;
    pc->valid = false;
    const int64_t N = snap->n_nodes, T = snap->pods.n_pods;
    const int S = e->S;
    for (int64_t n = 0; n < N; ++n)
        if (snap->pod_off[n + 1] - snap->pod_off[n] > kMaxPodsPerNode)
            return e->fail(FGS_E_TOO_LARGE, "fgs_select_victims: node " + std::to_string(n) + " has more than 128 pods");
    PackedPods vp; // the pods of the snapshot through the packer's arithmetic
    std::string err;
    const int rc = pack_pods(&snap->pods, 0, T, &vp, &err);
    if (rc) return e->fail(rc, err);
    const int ncols = VP_FIXED + S;
    const size_t Tn = (size_t)std::max<int64_t>(T, 1);
    std::vector<int64_t> h((size_t)ncols * Tn + (size_t)N + 1);
    auto col = [&](int c, int64_t a) -> int64_t& { return h[(size_t)c * Tn + (size_t)a]; };
    for (int64_t a = 0; a < T; ++a) {
        col(VP_NS, a) = vp.ns[(size_t)a];
        col(VP_PRIO, a) = vp.priority[(size_t)a];
        col(VP_START, a) = snap->pods.start_time ? snap->pods.start_time[a] : 0;
        col(VP_KIND, a) = vp.flags[(size_t)a] & 3u;
        col(VP_GPU_LIM, a) = vp.flex_gpu_limit[(size_t)a];
        col(VP_MEM_LIM, a) = vp.flex_mem_limit[(size_t)a];
        col(VP_ANNOT, a) = vp.index_annot[(size_t)a];
        col(VP_KEYS, a) = vp.req_scalar_present[(size_t)a];
        col(VP_REQ_CPU, a) = vp.req_mcpu[(size_t)a];
        col(VP_REQ_MEM, a) = vp.req_mem[(size_t)a];
        col(VP_REQ_EPH, a) = vp.req_eph[(size_t)a];
        for (int s = 0; s < S; ++s) col(VP_FIXED + s, a) = vp.req_scalar[(size_t)((int64_t)s * T + a)];
    }
    std::memcpy(h.data() + (size_t)ncols * Tn, snap->pod_off, sizeof(int64_t) * (size_t)(N + 1));
    const size_t bytes = h.size() * sizeof(int64_t);
    FGS_CUDA(e, cudaSetDevice(e->device));
    FGS_CUDA(e, cudaStreamSynchronize(e->stream));
    if (bytes > pc->vp_bytes) {
        if (pc->d_vp) cudaFree(pc->d_vp);
        pc->d_vp = nullptr;
        pc->vp_bytes = 0;
        FGS_CUDA(e, cudaMalloc(&pc->d_vp, bytes));
        pc->vp_bytes = bytes;
    }
    FGS_CUDA(e, cudaMemcpy(pc->d_vp, h.data(), bytes, cudaMemcpyHostToDevice));
    pc->snap = snap;
    pc->node_gen = e->node_gen;
    pc->N = N;
    pc->T = T;
    pc->valid = true;
    return FGS_OK;
}

extern "C" int fgs_select_victims(fgs_engine* eh, const fgs_snapshot* snap, const fgs_podset* pending, int64_t p, const fgs_quotas* q,
                                  int32_t tracks_pods, uint8_t* node_status, uint8_t* victim, int32_t* n_victims, int32_t* picked_node) {
    if (!eh || !snap || !pending || !node_status) return FGS_E_INVALID;
    Engine* e = (Engine*)eh;
    FGS_ENTER(e);
    if (!e->have_nodes) return e->fail(FGS_E_STATE, "fgs_select_victims before fgs_upload_nodes");
    if (snap->n_nodes != e->n_nodes || e->node_base != 0) return e->fail(FGS_E_INVALID, "fgs_select_victims: the snapshot is not the uploaded node table (whole cluster, one engine)");
    if (snap->n_scalars != e->S || pending->n_scalars != e->S || (q && q->n_scalars != e->S)) return e->fail(FGS_E_INVALID, "fgs_select_victims: n_scalars mismatch");
    if (p < 0 || p >= pending->n_pods) return e->fail(FGS_E_INVALID, "fgs_select_victims: pod index out of range");
    const int64_t N = snap->n_nodes, T = snap->pods.n_pods;
    const int S = e->S, W = 2 + S;
    if (picked_node) *picked_node = -1;
    if (N == 0) return FGS_OK;
    int rc = preempt_prepare(e, snap);
    if (rc) return rc;
    PreemptCache* pc = e->pcache;
    PackedPods pp; // the preemptor
    std::string err;
    rc = pack_pods(pending, p, p + 1, &pp, &err);
    if (rc) return e->fail(rc, err);
    const int ncols = VP_FIXED + S;
    const size_t Tn = (size_t)std::max<int64_t>(T, 1), NS = q ? (size_t)q->n_ns : 0;
    std::vector<int64_t> hq(3 * NS * (size_t)W + 1);
    PreemptParams prm;
    std::memset(&prm, 0, sizeof(prm));
    if (q && NS) {
        std::memcpy(hq.data(), q->min, sizeof(int64_t) * NS * W);
        std::memcpy(hq.data() + NS * W, q->max, sizeof(int64_t) * NS * W);
        std::memcpy(hq.data() + 2 * NS * W, q->used, sizeof(int64_t) * NS * W);
        for (size_t k = 0; k < NS; ++k) { // Σ Used and Σ Min over the namespaces with a quota (elasticquota.go:40-51)
            if (!q->has_eq[k]) continue;
            const uint32_t up = q->used_scalar_present[k];
            prm.agg_present |= up;
            for (int c = 0; c < W; ++c) {
                if (c < 2 || ((up >> (c - 2)) & 1u)) prm.agg_used[c] += q->used[k * W + c];
                prm.agg_min[c] += q->min[k * W + c];
            }
        }
    }
    cudaSetDevice(e->device);
    FGS_CUDA(e, cudaStreamSynchronize(e->stream));
    const size_t q_bytes = ((sizeof(int64_t) * 3 * NS * (size_t)W + 63) / 64) * 64;
    const size_t small_bytes = ((NS + sizeof(uint32_t) * NS + 64 + 63) / 64) * 64;      // has_eq, used_scalar_present
    const size_t out_bytes = (size_t)N + Tn + sizeof(int) * (size_t)N + 4 * sizeof(int64_t) * (size_t)N + 256; // status, victim, n_victims, cand, picked
    const size_t need = q_bytes + small_bytes + out_bytes + 256;
    if (need > pc->work_bytes) {
        if (pc->d_work) cudaFree(pc->d_work);
        pc->d_work = nullptr;
        pc->work_bytes = 0;
        FGS_CUDA(e, cudaMalloc(&pc->d_work, need));
        pc->work_bytes = need;
    }
    char* d_q = pc->d_work;
    char* d_small = d_q + q_bytes;
    char* d_out = d_small + small_bytes;
    cudaError_t ce = cudaSuccess;
    if (NS) ce = cudaMemcpyAsync(d_q, hq.data(), sizeof(int64_t) * 3 * NS * (size_t)W, cudaMemcpyHostToDevice, e->stream);
    if (ce == cudaSuccess && NS) ce = cudaMemcpyAsync(d_small, q->has_eq, NS, cudaMemcpyHostToDevice, e->stream);
    char* d_pres = d_small + ((NS + 15) / 16) * 16;
    if (ce == cudaSuccess && NS) ce = cudaMemcpyAsync(d_pres, q->used_scalar_present, sizeof(uint32_t) * NS, cudaMemcpyHostToDevice, e->stream);
    if (ce == cudaSuccess) ce = cudaMemsetAsync(d_out, 0, out_bytes, e->stream);
    prm.raw = (const i64*)e->d_raw;
    prm.vp = (const i64*)pc->d_vp;
    prm.pod_off = prm.vp + (size_t)ncols * Tn;
    prm.n_nodes = N;
    prm.n_pad = e->n_pad;
    prm.T = (i64)Tn;
    prm.S = S;
    prm.D = e->D;
    if (q && NS) {
        prm.has_eq = (const uint8_t*)d_small;
        prm.qmin = (const i64*)d_q;
        prm.qmax = prm.qmin + NS * W;
        prm.qused = prm.qmax + NS * W;
        prm.qpresent = (const uint32_t*)d_pres;
        prm.n_ns = (i64)NS;
    }
    prm.tracks_pods = tracks_pods ? 1 : 0;
    prm.p_req[0] = pp.req_mcpu[0];
    prm.p_req[1] = pp.req_mem[0];
    prm.p_req[2] = pp.req_eph[0];
    for (int s = 0; s < S; ++s) prm.p_req[3 + s] = pp.req_scalar[(size_t)s];
    prm.p_keys = pp.req_scalar_present[0];
    prm.p_flags = pp.flags[0];
    prm.p_gpu_lim = pp.flex_gpu_limit[0];
    prm.p_mem_lim = pp.flex_mem_limit[0];
    prm.p_ns = pp.ns[0];
    prm.p_prio = pp.priority[0];
    EvalParams ep;
    e->fill_eval_params(&ep, nullptr);
    prm.filter_mask = ep.filter_mask;
    prm.fit_ignored = ep.fit_ignored;
    prm.flex_filter_first = ep.flex_filter_first;
    prm.status = (uint8_t*)d_out;
    prm.victim = prm.status + ((N + 15) / 16) * 16;
    prm.n_victims = (int*)(prm.victim + ((Tn + 15) / 16) * 16);
    prm.cand = (i64*)((char*)prm.n_victims + ((sizeof(int) * (size_t)N + 15) / 16) * 16);
    prm.picked = (int*)(prm.cand + 4 * (size_t)N);
    if (ce == cudaSuccess) {
        select_victims_kernel<<<(unsigned)((N + 63) / 64), 64, 0, e->stream>>>(prm);
        e->launches++;
        if (picked_node) {
            pick_candidate_kernel<<<1, 1024, 0, e->stream>>>(prm.status, prm.cand, N, prm.picked);
            e->launches++;
        }
        ce = cudaGetLastError();
    }
    int h_picked = -1;
    if (ce == cudaSuccess && picked_node) ce = cudaMemcpyAsync(&h_picked, prm.picked, sizeof(int), cudaMemcpyDeviceToHost, e->stream);
    std::vector<int> h_cnt((size_t)N);
    if (ce == cudaSuccess) ce = cudaMemcpyAsync(node_status, prm.status, (size_t)N, cudaMemcpyDeviceToHost, e->stream);
    if (ce == cudaSuccess && victim && T) ce = cudaMemcpyAsync(victim, prm.victim, (size_t)T, cudaMemcpyDeviceToHost, e->stream);
    if (ce == cudaSuccess && n_victims) ce = cudaMemcpyAsync(h_cnt.data(), prm.n_victims, sizeof(int) * (size_t)N, cudaMemcpyDeviceToHost, e->stream);
    if (ce == cudaSuccess) ce = cudaStreamSynchronize(e->stream); // (the host copies above come from pageable memory: already staged)
    if (ce != cudaSuccess) return e->fail(FGS_E_CUDA, std::string("fgs_select_victims: ") + cudaGetErrorString(ce));
    if (n_victims)
        for (int64_t n = 0; n < N; ++n) n_victims[n] = h_cnt[(size_t)n];
    if (picked_node) *picked_node = h_picked;
    return FGS_OK;
}
