// peer_merge.cu — the two cross-shard merges of the node-sharded path as ONE kernel each over NVLink
// peer memory, in place of the NCCL all-reduces of flex-gpu-scheduler_b200/distributed.py.
//
// What crosses the shards is per-pod only: after the reduce pass the NormalizeScore maxima (DefaultNormalizeScore's
// maxCount, pkg/flexgpu/flex_gpu.go:126 → vendor/…/plugins/helper/normalize_score.go:27-35; the Allocatable hi/lo of
// pkg/noderesources/allocatable.go:121-153) and len(feasibleNodes); after the emit pass the selectHost key
// (schedule_one.go:783-806). Each rank owns one engine and one exchange buffer in symmetric memory; every rank maps all
// of them (and, where the NVSwitch offers it, one multicast address over all).
//
//   multicast (NVLS)      part := my partials            — local stores
//                         barrier
//                         my slice := multimem.ld_reduce  — the switch reduces the R copies on the way in
//                         multimem.st → every rank's out  — the switch replicates on the way out
//                         barrier;  d_perpod := out
//   peer pointers         inbox[me] of rank s := my partials of slice s   (stores over NVLink, for every s)
//                         barrier
//                         my slice := reduce over inbox[0..R); out slice of every rank := it
//                         barrier;  d_perpod := out
//
// The barrier is per CTA: CTA b of every rank works on sub-range b of every slice, so it only has to meet CTA b of
// the peers (flag[b][src] := epoch with st.release.sys, spin with ld.acquire.sys). Epochs only grow, flags are never
// reset. The grid is at most one CTA per SM so that all CTAs are resident (a CTA spins on a peer, never on a CTA of
// its own GPU). A spin that lasts longer than kSpinLimitNs gives up and raises the error word instead of hanging.
#include <cuda_runtime.h>

#include <cstring>
#include <string>

#include "engine.h"

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

constexpr int kMaxRanks = FGS_PEER_MAX_RANKS;
constexpr int kPeerMaxCtas = 128;
constexpr int kPeerThreads = 512;
constexpr u64 kSpinLimitNs = 10ull * 1000 * 1000 * 1000;
constexpr i64 kHeaderBytes = 4096 + (i64)kPeerMaxCtas * kMaxRanks * 8; // error word page + flags

struct PeerParams {
    i64* local;          // the engine's per-pod block (first row of the merge)
    i64 len;             // int64 elements to merge
    i64 sum_begin, sum_end; // elements merged with SUM; the rest with MAX
    i64 per;             // slice length per rank (even)
    i64 region;          // int64 elements per region (inbox / part, out)
    char* buf[kMaxRanks]; // every rank's exchange buffer as mapped here
    char* mc;            // multicast mapping of the same buffer, or NULL
    int rank, world;
    u64 epoch;           // barrier values of this launch: epoch + 1, epoch + 2
};

__device__ __forceinline__ void st_release_sys(u64* p, u64 v) { asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory"); }
__device__ __forceinline__ u64 ld_acquire_sys(const u64* p) {
    u64 v;
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ u64 globaltimer_ns() {
    u64 t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}
__device__ __forceinline__ i64 mc_ld_max(const i64* p) {
    i64 v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.max.s64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ i64 mc_ld_add(const i64* p) {
    u64 v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return (i64)v;
}
__device__ __forceinline__ void mc_st(i64* p, i64 v) { asm volatile("multimem.st.relaxed.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory"); }

__device__ __forceinline__ u64* flag_of(char* base, int cta, int src) { return (u64*)(base + 4096) + (i64)cta * kMaxRanks + src; }
__device__ __forceinline__ i64* region_of(char* base, i64 region, int k) { return (i64*)(base + kHeaderBytes) + (i64)k * region; }

// all writes of this CTA → visible to CTA `blockIdx.x` of every peer once it has seen `value`; and the converse.
// false: a peer did not arrive within kSpinLimitNs (or an earlier merge on this rank already gave up): the error word of
// the own buffer is set and the caller must stop — what the peers would have written is not there.
__device__ __forceinline__ bool cta_barrier(const PeerParams& prm, u64 value) {
    __shared__ int s_failed;
    if (threadIdx.x == 0) s_failed = 0;
    __syncthreads();
    if ((int)threadIdx.x < prm.world) {
        const int peer = threadIdx.x;
        __threadfence_system();
        st_release_sys(flag_of(prm.buf[peer], blockIdx.x, prm.rank), value);
        const u64* mine = flag_of(prm.buf[prm.rank], blockIdx.x, peer);
        u64 t0 = 0;
        for (unsigned polls = 1; ld_acquire_sys(mine) < value; polls++) {
            if ((polls & 4095u) == 0) { // the timer is slow to read: look at it now and then
                const u64 t = globaltimer_ns();
                if (t0 == 0) t0 = t;
                if (t - t0 > kSpinLimitNs || *(volatile int*)prm.buf[prm.rank] != 0) {
                    atomicCAS((int*)prm.buf[prm.rank], 0, 1 + peer); // error word: the (first) peer that never arrived
                    s_failed = 1;
                    break;
                }
            }
        }
    }
    __syncthreads();
    return s_failed == 0;
}

template <bool MC>
__global__ void __launch_bounds__(kPeerThreads) peer_merge_kernel(const PeerParams prm) {
    const int R = prm.world, me = prm.rank;
    const i64 per = prm.per;
    const i64 sub = ((per + gridDim.x - 1) / gridDim.x + 1) & ~1ll; // this CTA's share of every slice (even)
    const i64 s0 = (i64)blockIdx.x * sub;
    const i64 s1 = s0 + sub < per ? s0 + sub : per;
    if (*(volatile int*)prm.buf[me] != 0) return; // an earlier merge timed out: the ranks are out of step, publish nothing more

    // 1. hand the partials over
    if (MC) {
        i64* part = region_of(prm.buf[me], prm.region, 0);
        for (int s = 0; s < R; s++) {
            const i64 base = (i64)s * per;
            for (i64 i = s0 + 2 * threadIdx.x; i < s1; i += 2 * kPeerThreads) {
                const i64 g = base + i;
                if (g + 1 < prm.len) *(longlong2*)(part + g) = *(const longlong2*)(prm.local + g);
                else if (g < prm.len) part[g] = prm.local[g];
            }
        }
    } else {
        for (int k = 1; k < R; k++) {
            const int s = (me + k) % R; // stagger the targets
            const i64 base = (i64)s * per;
            i64* inbox = region_of(prm.buf[s], prm.region, 0) + (i64)me * per;
            for (i64 i = s0 + 2 * threadIdx.x; i < s1; i += 2 * kPeerThreads) {
                const i64 g = base + i;
                if (g + 1 < prm.len) *(longlong2*)(inbox + i) = *(const longlong2*)(prm.local + g);
                else if (g < prm.len) inbox[i] = prm.local[g];
            }
        }
    }
    if (!cta_barrier(prm, prm.epoch + 1)) return;

    // 2. reduce my slice, hand the result to everyone
    {
        const i64 base = (i64)me * per;
        if (MC) {
            const i64* part_mc = region_of(prm.mc, prm.region, 0);
            i64* out_mc = region_of(prm.mc, prm.region, 1);
            // a multimem load is a round trip through the switch: keep kIlp of them in flight per thread
            constexpr int kIlp = 8;
            for (i64 i0 = s0 + threadIdx.x; i0 < s1; i0 += (i64)kIlp * kPeerThreads) {
                i64 v[kIlp];
#pragma unroll
                for (int u = 0; u < kIlp; u++) {
                    const i64 i = i0 + (i64)u * kPeerThreads, g = base + i;
                    if (i < s1 && g < prm.len) v[u] = (g >= prm.sum_begin && g < prm.sum_end) ? mc_ld_add(part_mc + g) : mc_ld_max(part_mc + g);
                }
#pragma unroll
                for (int u = 0; u < kIlp; u++) {
                    const i64 i = i0 + (i64)u * kPeerThreads, g = base + i;
                    if (i < s1 && g < prm.len) mc_st(out_mc + g, v[u]);
                }
            }
        } else {
            const i64* inbox = region_of(prm.buf[me], prm.region, 0);
            for (i64 i = s0 + threadIdx.x; i < s1; i += kPeerThreads) {
                const i64 g = base + i;
                if (g >= prm.len) break;
                const bool sum = g >= prm.sum_begin && g < prm.sum_end;
                i64 v = prm.local[g];
                for (int k = 1; k < R; k++) {
                    const i64 w = inbox[(i64)((me + k) % R) * per + i];
                    v = sum ? (i64)((u64)v + (u64)w) : (w > v ? w : v);
                }
                prm.local[g] = v;
                for (int k = 1; k < R; k++) region_of(prm.buf[(me + k) % R], prm.region, 1)[g] = v;
            }
        }
    }
    if (!cta_barrier(prm, prm.epoch + 2)) return;

    // 3. the merged vector → the engine's block (the own slice is already in place without multicast)
    {
        const i64* out = region_of(prm.buf[me], prm.region, 1);
        for (int s = 0; s < R; s++) {
            if (!MC && s == me) continue;
            const i64 base = (i64)s * per;
            for (i64 i = s0 + 2 * threadIdx.x; i < s1; i += 2 * kPeerThreads) {
                const i64 g = base + i;
                if (g + 1 < prm.len) *(longlong2*)(prm.local + g) = *(const longlong2*)(out + g);
                else if (g < prm.len) prm.local[g] = out[g];
            }
        }
    }
}

} // namespace

struct PeerState {
    fgs_peer_group g;
    u64 epoch = 0;
    int ctas = 128;        // merge alone on the GPU: latency matters, take many CTAs
    int ctas_overlap = 16; // merge beside a pass of another batch (pipeline): it is hidden anyway, leave the SMs to the pass
    bool use_mc = false;
    int* h_err = nullptr;  // pinned: the error word as read back with each batch's per-pod results (one per pipeline slot + 1)
    ~PeerState() {
        if (h_err) cudaFreeHost(h_err);
    }
};

void Engine::peer_free() {
    delete peer;
    peer = nullptr;
}

} // namespace fgs

using namespace fgs;

extern "C" uint64_t fgs_peer_bytes(int64_t max_pods) {
    const i64 p_pad = (max_pods + 63) / 64 * 64 + 64;
    const i64 region = (i64)(FGS_REDUCE_ROWS + 1) * p_pad + 2 * kMaxRanks;
    return (uint64_t)(kHeaderBytes + 2 * region * 8);
}

extern "C" int fgs_peer_attach(fgs_engine* eh, const fgs_peer_group* g) {
    Engine* e = (Engine*)eh;
    if (!e) return FGS_E_INVALID;
    FGS_ENTER(e);
    if (e->in_flight) return e->fail(FGS_E_STATE, "fgs_peer_attach while submitted batches are in flight: fgs_collect them first");
    if (!g) { // detach
        e->peer_free();
        return FGS_OK;
    }
    if (g->world < 2 || g->world > kMaxRanks || g->rank < 0 || g->rank >= g->world) return e->fail(FGS_E_INVALID, "fgs_peer_attach: world must be 2.." + std::to_string(kMaxRanks));
    for (int r = 0; r < g->world; r++)
        if (!g->buffers[r]) return e->fail(FGS_E_INVALID, "fgs_peer_attach: missing buffer of rank " + std::to_string(r));
    if (g->bytes < (uint64_t)kHeaderBytes + 64) return e->fail(FGS_E_INVALID, "fgs_peer_attach: exchange buffer too small");
    FGS_CUDA(e, cudaSetDevice(e->device));
    // header (error word, flags) := 0 once; the caller synchronises the ranks between attach and the first merge
    FGS_CUDA(e, cudaMemsetAsync(g->buffers[g->rank], 0, (size_t)kHeaderBytes, e->stream));
    FGS_CUDA(e, cudaStreamSynchronize(e->stream));
    if (!e->peer) e->peer = new PeerState();
    if (!e->peer->h_err) FGS_CUDA(e, cudaMallocHost((void**)&e->peer->h_err, sizeof(int) * (Engine::kSlots + 1)));
    memset(e->peer->h_err, 0, sizeof(int) * (Engine::kSlots + 1));
    e->peer->g = *g;
    e->peer->epoch = 0;
    e->peer->use_mc = g->multicast != nullptr;
    if (const char* s = getenv("FGS_PEER_NO_MULTICAST"))
        if (*s == '1') e->peer->use_mc = false;
    if (const char* s = getenv("FGS_PEER_CTAS")) {
        int v = atoi(s);
        if (v >= 1 && v <= kPeerMaxCtas) e->peer->ctas = v;
    }
    if (const char* s = getenv("FGS_PEER_CTAS_OVERLAP")) {
        int v = atoi(s);
        if (v >= 1 && v <= kPeerMaxCtas) e->peer->ctas_overlap = v;
    }
    return FGS_OK;
}

// which = 0: after the reduce pass (reduce_max rows: MAX, reduce_sum row: SUM); 1: after the emit pass (select_key: MAX).
// Works on the per-pod block the engine's members alias at the time of the call.
int fgs::Engine::peer_merge(int which, cudaStream_t on, bool overlapped) {
    Engine* e = this;
    const i64 row0 = which == 0 ? 0 : FGS_REDUCE_ROWS + 1, rows = which == 0 ? FGS_REDUCE_ROWS + 1 : 3, sum_row = which == 0 ? FGS_REDUCE_ROWS : -1;
    PeerState* ps = e->peer;
    if (!ps) return e->fail(FGS_E_STATE, "fgs_peer_merge_* before fgs_peer_attach");
    if (!e->have_pods) return e->fail(FGS_E_STATE, "fgs_peer_merge_* before fgs_stage_pods");
    PeerParams prm;
    memset(&prm, 0, sizeof prm);
    prm.local = (i64*)e->d_perpod + row0 * e->p_pad;
    prm.len = rows * e->p_pad;
    prm.sum_begin = sum_row >= 0 ? sum_row * e->p_pad : 0;
    prm.sum_end = sum_row >= 0 ? (sum_row + 1) * e->p_pad : 0;
    prm.world = ps->g.world;
    prm.rank = ps->g.rank;
    prm.per = ((prm.len + prm.world - 1) / prm.world + 1) & ~1ll;
    prm.region = (i64)((ps->g.bytes - (uint64_t)kHeaderBytes) / 16) & ~1ll;
    if (prm.per * prm.world > prm.region) return e->fail(FGS_E_INVALID, "fgs_peer_merge_*: exchange buffer smaller than fgs_peer_bytes(max_pods)");
    for (int r = 0; r < prm.world; r++) prm.buf[r] = (char*)ps->g.buffers[r];
    prm.mc = (char*)ps->g.multicast;
    prm.epoch = ps->epoch;
    ps->epoch += 2;
    FGS_CUDA(e, cudaSetDevice(e->device));
    const int ctas = overlapped ? ps->ctas_overlap : ps->ctas;
    if (ps->use_mc)
        peer_merge_kernel<true><<<ctas, kPeerThreads, 0, on>>>(prm);
    else
        peer_merge_kernel<false><<<ctas, kPeerThreads, 0, on>>>(prm);
    FGS_CUDA(e, cudaGetLastError());
    e->launches++;
    return FGS_OK;
}

extern "C" int fgs_peer_merge_reduce(fgs_engine* eh) {
    Engine* e = (Engine*)eh;
    if (!e) return FGS_E_INVALID;
    FGS_ENTER(e);
    return e->peer_merge(0, e->stream);
}

extern "C" int fgs_peer_merge_select(fgs_engine* eh) {
    Engine* e = (Engine*)eh;
    if (!e) return FGS_E_INVALID;
    FGS_ENTER(e);
    return e->peer_merge(1, e->stream);
}

// the error word travels to the host with every batch's results: enqueue its copy behind the merge on `on` …
int fgs::Engine::peer_fetch_error(int slot, cudaStream_t on) {
    Engine* e = this;
    if (!e->peer) return FGS_OK;
    FGS_CUDA(e, cudaMemcpyAsync(e->peer->h_err + slot, e->peer->g.buffers[e->peer->g.rank], sizeof(int), cudaMemcpyDeviceToHost, on));
    return FGS_OK;
}
// … and look at it once that stream has been waited for
int fgs::Engine::peer_check_error(int slot) {
    if (!peer || !peer->h_err) return FGS_OK;
    const int word = peer->h_err[slot];
    if (word) return fail(FGS_E_STATE, "peer merge: rank " + std::to_string(word - 1) + " never reached the barrier (timed out); this batch's results are not merged");
    return FGS_OK;
}

extern "C" int fgs_peer_status(fgs_engine* eh, int32_t* uses_multicast) {
    Engine* e = (Engine*)eh;
    if (!e) return FGS_E_INVALID;
    FGS_ENTER(e);
    if (!e->peer) return e->fail(FGS_E_STATE, "fgs_peer_status before fgs_peer_attach");
    FGS_CUDA(e, cudaSetDevice(e->device));
    FGS_CUDA(e, cudaStreamSynchronize(e->stream));
    if (e->merge_stream) FGS_CUDA(e, cudaStreamSynchronize(e->merge_stream)); // the pipeline's merges run there
    int word = 0;
    FGS_CUDA(e, cudaMemcpy(&word, e->peer->g.buffers[e->peer->g.rank], sizeof word, cudaMemcpyDeviceToHost));
    if (uses_multicast) *uses_multicast = e->peer->use_mc ? 1 : 0;
    if (word) return e->fail(FGS_E_STATE, "peer merge: rank " + std::to_string(word - 1) + " never reached the barrier (timed out)");
    return FGS_OK;
}
