// nominated.cu — the second filter pass of RunFilterPluginsWithNominatedPods.
//
//   RunFilterPluginsWithNominatedPods   vendor/k8s.io/kubernetes/pkg/scheduler/framework/runtime/framework.go:752-793
//   addNominatedPods                    :795-823 — nodeInfo.Clone() + AddPodInfo for every pod nominated to the node
//                                       whose priority is >= the evaluated pod's (and that is not the pod itself)
//
// A node with such pods is filtered twice: with them added, and — only if that passed — without them; the node is
// feasible when both pass. The reduce pass has produced the "without" verdict for every (pod, node); this pass
// evaluates the "with" verdict for the few nodes the PodNominator names — one thread per (pod, nominated node):
// it adds the yielding-to pods to the node's RAW row (Requested, pod count, flex limit sums, device use by their
// index annotations), runs FlexGPU.Filter and Fit.Filter on that row in profile order, and clears the feasibility
// bit when the augmented node rejects the pod. Pods that lost a node get their NormalizeScore reductions redone
// from the bitmap. Scores are untouched: Score plugins read the snapshot's NodeInfo, never the augmented clone.
#include <cuda_runtime.h>

#include <algorithm>
#include <cstring>
#include <numeric>
#include <string>
#include <vector>

#include "engine.h"
#include "kernels.cuh"
#include "kernels_narrow.cuh"
#include "sample.cuh"

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

// nominated-pod columns (int64 [NQ_FIXED + S][Q], sorted by node)
enum NomCol : int { NQ_PRIO = 0, NQ_UID, NQ_KIND, NQ_GPU_LIM, NQ_MEM_LIM, NQ_ANNOT, NQ_REQ_CPU, NQ_REQ_MEM, NQ_REQ_EPH, NQ_FIXED };
enum AuxCol : int { AX_PRIO = 0, AX_UID, AX_DIRTY, AX_COLS };

struct NomParams {
    const i64* raw;        // raw node table
    const i64* pods;       // device pod table of the batch
    i64* aux;              // [AX_COLS][p_pad]
    const i64* nom;        // nominated-pod columns
    const i64* nom_nodes;  // [M] local node index
    const i64* nom_off;    // [M+1] range of each node in the nominated-pod columns
    uint32_t* bitmap;
    i64* reduce_max;
    i64* reduce_sum;
    i64 n_pad, n_pods, p_pad, words_per_row, nom_q;
    int S, D, has_uid;
    uint32_t filter_mask, fit_ignored;
    int flex_filter_first;
};

__global__ void __launch_bounds__(128) nominated_pass(const __grid_constant__ NomParams prm) {
    const i64 p = (i64)blockIdx.y * blockDim.x + threadIdx.x;
    if (p >= prm.n_pods) return;
    const int m = blockIdx.x;
    const i64 n = prm.nom_nodes[m];
    const int S = prm.S, D = prm.D;
    const i64 np = prm.n_pad, Q = prm.nom_q;
    auto R = [&](int c) { return prm.raw[(i64)c * np + n]; };
    auto P = [&](int c) { return prm.pods[(i64)c * prm.p_pad + p]; };
    auto NQ = [&](int c, i64 q) { return prm.nom[(i64)c * Q + q]; };
    const i64 my_prio = prm.aux[(i64)AX_PRIO * prm.p_pad + p], my_uid = prm.aux[(i64)AX_UID * prm.p_pad + p];
    u64 meta = (u64)R(NR_META);
    if (!((uint32_t)meta & FGS_NF_VALID)) return; // addNominatedPods: nodeInfo.Node() == nil → nothing added (:797-800)

    // ---- addNominatedPods: the node row with the pods this one must yield to ----------------------------------
    i64 req_cpu = R(NR_REQ_MCPU), req_mem = R(NR_REQ_MEM), req_eph = R(NR_REQ_EPH), num_pods = R(NR_NUM_PODS);
    i64 gsum = R(NR_FLEX_GPU_SUM), msum = R(NR_FLEX_MEM_SUM);
    i64 req_sc[FGS_MAX_SCALARS], used[FGS_MAX_DEVICES];
    for (int s = 0; s < FGS_MAX_SCALARS; ++s) req_sc[s] = s < S ? R(nr_req_scalar(S, s)) : 0;
    for (int d = 0; d < FGS_MAX_DEVICES; ++d) used[d] = d < D ? R(nr_used(S, d)) : 0;
    const i64 gpuAlloc = R(nr_alloc_scalar(FGS_SCALAR_FLEX_GPU)), memAlloc = R(nr_alloc_scalar(FGS_SCALAR_FLEX_MEM));
    bool added = false;
    for (i64 q = prm.nom_off[m]; q < prm.nom_off[m + 1]; ++q) {
        if (NQ(NQ_PRIO, q) < my_prio) continue;                        // :811
        if (prm.has_uid && NQ(NQ_UID, q) == my_uid) continue;          // :812 the pod itself
        added = true;
        // NodeInfo.AddPodInfo (framework/types.go:583-609): Requested += calculateResource(pod), one more pod
        req_cpu += NQ(NQ_REQ_CPU, q);
        req_mem += NQ(NQ_REQ_MEM, q);
        req_eph += NQ(NQ_REQ_EPH, q);
        for (int s = 0; s < S; ++s) req_sc[s] += NQ(NQ_FIXED + s, q);
        num_pods++;
        // what nodeResourceLimitSum and constructGPUs see of it (flex_gpu.go:133-140, gpu_node.go:66-120)
        const uint32_t k = (uint32_t)NQ(NQ_KIND, q);
        if (k & 1u) gsum += NQ(NQ_GPU_LIM, q);
        if (k & 2u) msum += NQ(NQ_MEM_LIM, q);
        const i64 annot = NQ(NQ_ANNOT, q);
        if (k != 0 && annot != (i64)FGS_ANNOT_NONE) {
            if (annot < 0 || annot >= gpuAlloc) {
                meta |= FGS_NF_REF_PANIC; // gpus[index] out of range
            } else {
                if (k & 1u) meta |= 1ull << (32 + (int)annot);
                if (k & 2u) used[annot] += NQ(NQ_MEM_LIM, q);
            }
        }
    }
    if (!added) return; // podsAdded == false: one pass only, and the reduce pass ran it

    // ---- RunFilterPlugins on the augmented row (runtime/framework.go:667-693) -------------------------------------
    const u64 pmeta = (u64)P(PD_META);
    const uint32_t pf = (uint32_t)pmeta, kind = pf & 3u, flags = (uint32_t)meta, mono = (uint32_t)(meta >> 32);
    enum { OK = 0, FAIL = 1, PANIC = 2 };
    auto flex = [&]() -> int { // FlexGPU.Filter, flex_gpu.go:41-119
        if (kind == 0) return OK;
        if (kind == 3) return FAIL;
        if (!(flags & FGS_NF_HAS_FLEX)) return FAIL;
        const i64 gl = kind == 1 ? P(PD_GPU_LIM) : 0, ml = kind == 2 ? P(PD_MEM_LIM) : 0;
        if (gsum + gl > gpuAlloc) return FAIL; // :90-93
        if (msum + ml > memAlloc) return FAIL; // :95-98
        if (flags & FGS_NF_REF_PANIC) return PANIC; // NewGPUNode, :101
        if (kind == 2) {                            // MemAssumeFitIndexes empty, :109-116
            const i64 each = ((flags & FGS_NF_HAS_MEM_RES) ? memAlloc : 0) / gpuAlloc; // gpuAlloc > 0: REF_PANIC covers the rest
            bool fits = false;
            for (int d = 0; d < D && (i64)d < gpuAlloc; ++d)
                if (!((mono >> d) & 1u) && each >= used[d] + ml) fits = true;
            if (!fits) return FAIL;
        }
        return OK;
    };
    auto fit = [&]() -> int { // Fit.Filter → fitsRequest, fit.go:256-329
        bool bad = num_pods + 1 > R(NR_ALLOC_PODS);
        if (!(pf & FGS_PF_REQ_NONE)) {
            bad = bad || P(PD_REQ_CPU) > R(NR_ALLOC_MCPU) - req_cpu || P(PD_REQ_MEM) > R(NR_ALLOC_MEM) - req_mem ||
                  P(PD_REQ_EPH) > R(NR_ALLOC_EPH) - req_eph;
            uint32_t mk = (uint32_t)(pmeta >> 32) & ~prm.fit_ignored;
            while (mk) {
                const int s = __ffs(mk) - 1;
                mk &= mk - 1;
                if (P(PD_FIXED_COLS + s) > R(nr_alloc_scalar(s)) - req_sc[s]) bad = true;
            }
        }
        return bad ? FAIL : OK;
    };
    int verdict = OK;
    for (int k = 0; k < 2 && verdict == OK; ++k) {
        const bool is_flex = (k == 0) == (prm.flex_filter_first != 0);
        if (is_flex && (prm.filter_mask & FGS_PL_FLEXGPU)) verdict = flex();
        if (!is_flex && (prm.filter_mask & FGS_PL_FIT)) verdict = fit();
    }
    if (verdict == OK) return; // the second pass (without the nominated pods) decides, and the reduce pass ran it
    if (verdict == PANIC) prm.reduce_max[(i64)RR_FILTER_PANIC * prm.p_pad + p] = 1;
    uint32_t* word = prm.bitmap + p * prm.words_per_row + (n >> 5);
    const uint32_t bit = 1u << (n & 31);
    if (atomicAnd(word, ~bit) & bit) { // the node was feasible without them: it is not any more
        atomicAdd((u64*)prm.reduce_sum + p, (u64)-1ll);
        prm.aux[(i64)AX_DIRTY * prm.p_pad + p] = 1;
    }
}

} // namespace

int Engine::stage_nominated_aux(const fgs_pod_table* t) {
    Engine* e = this;
    if (e->nom_q == 0) return FGS_OK;
    const int64_t P = t->n_pods, pp = e->p_pad;
    if (pp > e->aux_cap) {
        if (e->d_pod_aux) cudaFree(e->d_pod_aux);
        if (e->h_pod_aux) cudaFreeHost(e->h_pod_aux);
        e->d_pod_aux = e->h_pod_aux = nullptr;
        e->aux_cap = 0;
        FGS_CUDA(e, cudaMalloc(&e->d_pod_aux, sizeof(int64_t) * AX_COLS * pp));
        FGS_CUDA(e, cudaMallocHost(&e->h_pod_aux, sizeof(int64_t) * AX_COLS * pp));
        e->aux_cap = pp;
    }
    int64_t* h = (int64_t*)e->h_pod_aux;
    std::memset(h, 0, sizeof(int64_t) * AX_COLS * pp);
    for (int64_t i = 0; i < P; ++i) {
        h[(int64_t)AX_PRIO * pp + i] = t->priority ? t->priority[i] : 0;
        h[(int64_t)AX_UID * pp + i] = t->uid ? t->uid[i] : -1;
    }
    e->aux_has_uid = t->uid != nullptr;
    FGS_CUDA(e, cudaMemcpyAsync(e->d_pod_aux, h, sizeof(int64_t) * AX_COLS * pp, cudaMemcpyHostToDevice, e->stream));
    return FGS_OK;
}

int Engine::run_nominated_pass() {
    Engine* e = this;
    if (e->nom_q == 0 || e->nom_m == 0 || e->n_pods == 0) return FGS_OK;
    if (!e->d_pod_aux || e->aux_cap < e->p_pad) return e->fail(FGS_E_STATE, "nominated pods were set after this batch was staged: stage it again");
    EvalParams ep;
    e->fill_eval_params(&ep, nullptr);
    NomParams np;
    std::memset(&np, 0, sizeof(np));
    const int ncols = NQ_FIXED + e->S;
    np.raw = (const i64*)e->d_raw;
    np.pods = (const i64*)e->d_pods;
    np.aux = (i64*)e->d_pod_aux;
    np.nom = (const i64*)e->d_nom;
    np.nom_nodes = np.nom + (int64_t)ncols * e->nom_q;
    np.nom_off = np.nom_nodes + e->nom_m;
    np.bitmap = (uint32_t*)e->d_bitmap;
    np.reduce_max = (i64*)ep.reduce_max;
    np.reduce_sum = (i64*)ep.reduce_sum;
    np.n_pad = e->n_pad;
    np.n_pods = e->n_pods;
    np.p_pad = e->p_pad;
    np.words_per_row = e->words_per_row;
    np.nom_q = e->nom_q;
    np.S = e->S;
    np.D = e->D;
    np.has_uid = e->aux_has_uid ? 1 : 0;
    np.filter_mask = ep.filter_mask;
    np.fit_ignored = ep.fit_ignored;
    np.flex_filter_first = ep.flex_filter_first;
    const int64_t pp = e->p_pad;
    FGS_CUDA(e, cudaMemsetAsync((int64_t*)e->d_pod_aux + (int64_t)AX_DIRTY * pp, 0, sizeof(int64_t) * pp, e->stream));
    dim3 grid((unsigned)e->nom_m, (unsigned)((e->n_pods + 127) / 128));
    nominated_pass<<<grid, 128, 0, e->stream>>>(np);
    e->launches++;
    FGS_CUDA(e, cudaGetLastError());
    // the pods that lost a node: NormalizeScore's reductions over what is left, then the special-node rows again
    SampleParams sp;
    std::memset(&sp, 0, sizeof(sp));
    sp.bitmap = (uint32_t*)e->d_bitmap;
    sp.nd = (const int64_t*)e->d_nd;
    sp.pods = (const int64_t*)e->d_pods;
    sp.reduce_max = (int64_t*)ep.reduce_max;
    sp.words_per_row = e->words_per_row;
    sp.n_nodes = e->n_nodes;
    sp.n_pad = e->n_pad;
    sp.n_pods = e->n_pods;
    sp.p_pad = e->p_pad;
    FGS_CUDA(e, launch_reduce_from_bitmap(sp, (const int64_t*)e->d_pod_aux + (int64_t)AX_DIRTY * pp, e->stream));
    e->launches++;
    if ((e->cfg.score_mask & FGS_PL_FLEXGPU) && e->node_an_valid && e->node_an.n_special > 0) {
        launch_special_pass((const uint32_t*)e->d_bitmap, e->words_per_row, (const int64_t*)e->d_special, e->n_pad, e->node_an.n_special,
                            (const int64_t*)e->d_pods, (int64_t*)e->d_perpod, e->n_pods, e->p_pad, e->stream);
        e->launches++;
        FGS_CUDA(e, cudaGetLastError());
    }
    return FGS_OK;
}

} // namespace fgs

using namespace fgs;

extern "C" int fgs_set_nominated(fgs_engine* eh, const fgs_pod_table* pods, const int32_t* node, const int64_t* uid) {
    if (!eh) return FGS_E_INVALID;
    Engine* e = (Engine*)eh;
    FGS_ENTER(e);
    if (e->in_flight) return e->fail(FGS_E_STATE, "fgs_set_nominated while submitted batches are in flight: fgs_collect them first");
    cudaSetDevice(e->device);
    FGS_CUDA(e, cudaStreamSynchronize(e->stream));
    e->nom_q = e->nom_m = 0;
    if (!pods || pods->n_pods == 0) return FGS_OK; // cleared
    if (!node) return e->fail(FGS_E_INVALID, "fgs_set_nominated: null node array");
    if (pods->n_scalars != e->S) return e->fail(FGS_E_INVALID, "fgs_set_nominated: n_scalars differs from the engine's");
    if (!e->have_nodes) return e->fail(FGS_E_STATE, "fgs_set_nominated before fgs_upload_nodes");
    const int S = e->S, ncols = NQ_FIXED + S;
    // the pods nominated to a node of THIS shard, grouped by node (stable: the PodNominator's order within a node)
    std::vector<int64_t> idx;
    for (int64_t q = 0; q < pods->n_pods; ++q) {
        const int64_t local = (int64_t)node[q] - e->node_base;
        if (node[q] >= 0 && local >= 0 && local < e->n_nodes) idx.push_back(q);
    }
    std::stable_sort(idx.begin(), idx.end(), [&](int64_t a, int64_t b) { return node[a] < node[b]; });
    const int64_t Q = (int64_t)idx.size();
    if (Q == 0) return FGS_OK;
    std::vector<int64_t> nodes, off;
    for (int64_t k = 0; k < Q; ++k) {
        const int64_t local = (int64_t)node[idx[(size_t)k]] - e->node_base;
        if (nodes.empty() || nodes.back() != local) {
            nodes.push_back(local);
            off.push_back(k);
        }
    }
    off.push_back(Q);
    const int64_t M = (int64_t)nodes.size();
    if (M > 65535) return e->fail(FGS_E_TOO_LARGE, "fgs_set_nominated: more than 65535 nodes with nominated pods");
    std::vector<int64_t> h((size_t)(ncols * Q + M + M + 1));
    auto col = [&](int c, int64_t k) -> int64_t& { return h[(size_t)((int64_t)c * Q + k)]; };
    for (int64_t k = 0; k < Q; ++k) {
        const int64_t q = idx[(size_t)k];
        col(NQ_PRIO, k) = pods->priority ? pods->priority[q] : 0;
        col(NQ_UID, k) = uid ? uid[q] : -2 - k; // never equal to a pending pod's uid (-1 = "none" there)
        col(NQ_KIND, k) = pods->flags[q] & 3u;
        col(NQ_GPU_LIM, k) = pods->flex_gpu_limit[q];
        col(NQ_MEM_LIM, k) = pods->flex_mem_limit[q];
        col(NQ_ANNOT, k) = pods->index_annot ? (int64_t)pods->index_annot[q] : (int64_t)FGS_ANNOT_NONE;
        col(NQ_REQ_CPU, k) = pods->req_mcpu[q];
        col(NQ_REQ_MEM, k) = pods->req_mem[q];
        col(NQ_REQ_EPH, k) = pods->req_eph[q];
        for (int s = 0; s < S; ++s) col(NQ_FIXED + s, k) = pods->req_scalar[(int64_t)s * pods->n_pods + q];
    }
    std::copy(nodes.begin(), nodes.end(), h.begin() + (size_t)(ncols * Q));
    std::copy(off.begin(), off.end(), h.begin() + (size_t)(ncols * Q + M));
    const int64_t bytes = (int64_t)(h.size() * sizeof(int64_t));
    if (bytes > e->nom_cap_bytes) {
        if (e->d_nom) cudaFree(e->d_nom);
        e->d_nom = nullptr;
        e->nom_cap_bytes = 0;
        FGS_CUDA(e, cudaMalloc(&e->d_nom, (size_t)bytes));
        e->nom_cap_bytes = bytes;
    }
    FGS_CUDA(e, cudaMemcpy(e->d_nom, h.data(), (size_t)bytes, cudaMemcpyHostToDevice));
    e->nom_q = Q;
    e->nom_m = M;
    return FGS_OK;
}
