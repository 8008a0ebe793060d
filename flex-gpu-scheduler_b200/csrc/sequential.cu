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

#include "seq_kernel.cuh"


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
    FGS_ENTER(e);
    if (e->in_flight) return e->fail(FGS_E_STATE, "fgs_schedule_sequential while submitted batches are in flight (it rewrites the node table): fgs_collect them first");
    int rc = seq_check_inputs(e, t, "fgs_schedule_sequential");
    if (rc) return rc;
    if (!t->acct_nz_mcpu || !t->acct_nz_mem || !t->index_annot)
        return e->fail(FGS_E_INVALID, "fgs_schedule_sequential: the pod table lacks acct_nz_mcpu / acct_nz_mem / index_annot");
    if (e->node_base != 0 || (e->total_nodes != 0 && e->total_nodes != e->n_nodes))
        return e->fail(FGS_E_STATE, "fgs_schedule_sequential: the engine must hold the whole cluster (node_index_base 0)");
    cudaSetDevice(e->device);
    FGS_CUDA(e, cudaStreamSynchronize(e->stream));
    rc = e->nodes_ready();
    if (rc) return rc;
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
    // speculative commit (two versions of every row in shared memory): one node per thread only
    const bool spec_off = getenv("FGS_SEQ_SPEC") && getenv("FGS_SEQ_SPEC")[0] == '0';
    const bool spec_fits = !spec_off && !state_global && seq_smem(e->S, e->D, 1, 2).total <= q->smem_optin;
    auto pick = [&](bool cl, bool ssm, bool sp) -> const void* {
        if (cl) return sp ? (const void*)seq_kernel<true, true, true> : (const void*)seq_kernel<true, true, false>;
        if (ssm) return sp ? (const void*)seq_kernel<true, false, true> : (const void*)seq_kernel<true, false, false>;
        return (const void*)seq_kernel<false, false, false>;
    };
    int G = 0, K = 0, cluster = 0;
    if (!cluster_off && !state_global && !env_ctas && k_smem_max >= 1) {
        for (int g = 1; g <= kMaxCluster; g *= 2) // one node per thread, or the single-cluster variant does not pay
            if ((int64_t)g * kSeqThreads >= N) {
                G = g;
                K = 1;
                cluster = g > 1;
                break;
            }
    }
    if (cluster) { // can a cluster of that size be resident at all? (16 is above the portable limit of 8)
        const SeqSmem Lc = seq_smem(e->S, e->D, K, spec_fits ? 2 : 1);
        const void* kc = pick(true, true, spec_fits);
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
    const bool spec = state_smem && K == 1 && spec_fits;
    const SeqSmem L = seq_smem(e->S, e->D, K, state_smem ? (spec ? 2 : 1) : 0);
    const void* kfn = pick(cluster != 0, state_smem != 0, spec);
    FGS_CUDA(e, cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, L.total));
    if (cluster && G > 8) FGS_CUDA(e, cudaFuncSetAttribute(kfn, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));

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
    FGS_ENTER(e);
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
