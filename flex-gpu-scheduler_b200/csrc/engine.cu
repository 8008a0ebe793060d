// engine.cu — host side of libfgs: device memory, H2D/D2H staging, kernel sequencing and the
// extern "C" entry points declared in include/fgs.h. There is no CPU evaluation path in this file:
// every verdict and score comes from the kernels in kernels.cu / gang_quota.cu.
#if defined(__x86_64__)
#include <emmintrin.h>
#endif
#include <omp.h>
#include <cuda_runtime.h>

#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <new>
#include <string>
#include <thread>
#include <vector>

#include "engine.h"
#include "filter_tables.cuh"
#include "kernels.cuh"
#include "node_patch.cuh"
#include "pack.h"
#include "sample.cuh"

namespace fgs {

static std::string g_create_error;

#define FGS_CUDA(e, call)                                                                               \
    do {                                                                                                \
        cudaError_t _err = (call);                                                                      \
        if (_err != cudaSuccess) {                                                                      \
            (e)->err = std::string(#call) + ": " + cudaGetErrorString(_err);                            \
            return FGS_E_CUDA;                                                                          \
        }                                                                                               \
    } while (0)

static int64_t round_up(int64_t v, int64_t m) { return (v + m - 1) / m * m; }

static uint64_t this_thread_tag() {
    static thread_local const char tag = 0; // its address is unique among the live threads and never 0
    return (uint64_t)(uintptr_t)&tag;
}

EngineEntry::EngineEntry(Engine* e_) : e(e_), ok(false) {
    const uint64_t me = this_thread_tag();
    uint64_t cur = e->owner.load(std::memory_order_acquire);
    if (cur == me) {
        e->owner_depth++;
        ok = true;
        return;
    }
    uint64_t expect = 0;
    if (e->owner.compare_exchange_strong(expect, me, std::memory_order_acq_rel)) {
        e->owner_depth = 1;
        ok = true;
    }
}

EngineEntry::~EngineEntry() {
    if (ok && --e->owner_depth == 0) e->owner.store(0, std::memory_order_release);
}

int Engine::fail(int code, const std::string& msg) {
    err = msg;
    return code;
}

int Engine::ensure_nodes(int64_t n, int D_) {
    const int64_t np = round_up(std::max<int64_t>(n, 1), kNodePad);
    if (np <= n_pad_cap && D_ <= D_cap) {
        n_pad = np;
        return FGS_OK;
    }
    if (d_raw) cudaFree(d_raw);
    if (d_nd) cudaFree(d_nd);
    if (h_raw) cudaFreeHost(h_raw);
    d_raw = d_nd = nullptr;
    h_raw = nullptr;
    n_pad_cap = np;
    D_cap = std::max(D_, D_cap);
    n_pad = np;
    FGS_CUDA(this, cudaMalloc(&d_raw, sizeof(int64_t) * nr_cols(S, D_cap) * np));
    FGS_CUDA(this, cudaMalloc(&d_nd, sizeof(int64_t) * nd_cols(S) * np));
    FGS_CUDA(this, cudaMallocHost(&h_raw, sizeof(int64_t) * nr_cols(S, D_cap) * np));
    return FGS_OK;
}

int Engine::ensure_pods(int64_t p) {
    const int64_t pp = round_up(std::max<int64_t>(p, 1), kTileP);
    const int64_t wpr = n_pad / 32;
    const int64_t stride = n_pad;
    const bool planes_ok = (pp <= p_plane_cap) && (wpr == words_per_row) && (stride == score_stride);
    if (pp > p_pad_cap) {
        if (d_pods) cudaFree(d_pods);
        if (d_perpod) cudaFree(d_perpod);
        if (h_pods) cudaFreeHost(h_pods);
        if (h_perpod) cudaFreeHost(h_perpod);
        d_pods = d_perpod = nullptr;
        h_pods = h_perpod = nullptr;
        p_pad_cap = pp;
        FGS_CUDA(this, cudaMalloc(&d_pods, sizeof(int64_t) * (pd_cols(S) + 1) * pp)); // + one column of scratch (flags / present words on the DMA path)
        FGS_CUDA(this, cudaMalloc(&d_perpod, sizeof(int64_t) * kPerPodRows * pp));
        FGS_CUDA(this, cudaMallocHost(&h_pods, sizeof(int64_t) * pd_cols(S) * pp));
        FGS_CUDA(this, cudaMallocHost(&h_perpod, sizeof(int64_t) * 4 * pp));
    }
    if (!planes_ok) {
        if (d_bitmap) cudaFree(d_bitmap);
        if (d_scores) cudaFree(d_scores);
        d_bitmap = nullptr;
        d_scores = nullptr;
        p_plane_cap = std::max(pp, p_plane_cap);
        words_per_row = wpr;
        score_stride = stride;
        FGS_CUDA(this, cudaMalloc(&d_bitmap, sizeof(uint32_t) * p_plane_cap * words_per_row));
        FGS_CUDA(this, cudaMalloc(&d_scores, sizeof(int64_t) * p_plane_cap * score_stride));
    }
    p_pad = pp;
    return FGS_OK;
}

void Engine::free_all() {
    if (d_raw) cudaFree(d_raw);
    if (d_nd) cudaFree(d_nd);
    if (d_pods) cudaFree(d_pods);
    if (d_perpod) cudaFree(d_perpod);
    if (d_bitmap) cudaFree(d_bitmap);
    if (d_scores) cudaFree(d_scores);
    for (auto& en : nn_ring)
        if (en.d) cudaFree(en.d);
    if (d_pn) cudaFree(d_pn);
    if (d_an) cudaFree(d_an);
    if (d_special) cudaFree(d_special);
    if (d_order) cudaFree(d_order);
    if (d_order_scratch) cudaFree(d_order_scratch);
    if (d_mask) cudaFree(d_mask);
    if (d_walk) cudaFree(d_walk);
    if (h_walk) cudaFreeHost(h_walk);
    if (h_small) cudaFreeHost(h_small);
    if (d_small_done) cudaFree(d_small_done);
    if (d_sample_scratch) cudaFree(d_sample_scratch);
    if (d_patch_flag) cudaFree(d_patch_flag);
    if (h_patch_flag) cudaFreeHost((void*)h_patch_flag);
    if (h_patch) cudaFreeHost(h_patch);
    if (d_patch) cudaFree(d_patch);
    if (d_nom) cudaFree(d_nom);
    if (d_pod_aux) cudaFree(d_pod_aux);
    if (h_pod_aux) cudaFreeHost(h_pod_aux);
    { // the pipeline slots the members do not alias right now
        for (int j = 0; j < kSlots; j++) {
            if (j == cur_slot) continue;
            Slot& o = slots[j];
            if (o.d_pods) cudaFree(o.d_pods);
            if (o.d_perpod) cudaFree(o.d_perpod);
            if (o.d_pn) cudaFree(o.d_pn);
            if (o.h_pods) cudaFreeHost(o.h_pods);
            if (o.h_perpod) cudaFreeHost(o.h_perpod);
        }
        for (auto& sl : slots) {
            if (sl.ev_done) cudaEventDestroy(sl.ev_done);
            for (cudaEvent_t ev : {sl.ev_reduce, sl.ev_merge1, sl.ev_emit, sl.ev_merge2, sl.ev_t0, sl.ev_emit0})
                if (ev) cudaEventDestroy(ev);
        }
        if (merge_stream) cudaStreamDestroy(merge_stream);
        if (d_bitmap_alt) cudaFree(d_bitmap_alt);
        if (d_bitmap_alt2) cudaFree(d_bitmap_alt2);
        if (ev_h2d) cudaEventDestroy(ev_h2d);
        if (ev_kernels) cudaEventDestroy(ev_kernels);
        if (front_stream) cudaStreamDestroy(front_stream);
        for (cudaEvent_t ev : {ev_front, ev_back, ev_plane[0], ev_plane[1], ev_plane[2]})
            if (ev) cudaEventDestroy(ev);
        if (copy_stream) cudaStreamDestroy(copy_stream);
        if (d2h_stream) cudaStreamDestroy(d2h_stream);
    }
    if (h_raw) cudaFreeHost(h_raw);
    if (h_pods) cudaFreeHost(h_pods);
    if (h_perpod) cudaFreeHost(h_perpod);
    for (auto& ev : events)
        if (ev) cudaEventDestroy(ev);
    if (own_stream && stream) cudaStreamDestroy(stream);
    ft_free();
    gang_free();
    seq_free();
    peer_free();
    preempt_free();
}

// plugin args → kernel parameters; validation mirrors the reference constructors
int Engine::set_plugins(const fgs_plugin_cfg* c) {
    if (!c) return fail(FGS_E_INVALID, "fgs_set_plugins: null config");
    if (c->filter_mask & ~(FGS_PL_FLEXGPU | FGS_PL_FIT)) return fail(FGS_E_INVALID, "fgs_set_plugins: unknown filter plugin bit");
    if (c->score_mask & ~(FGS_PL_FLEXGPU | FGS_PL_FIT | FGS_PL_ALLOCATABLE))
        return fail(FGS_E_INVALID, "fgs_set_plugins: unknown score plugin bit");
    if (c->score_mask & FGS_PL_ALLOCATABLE) {
        if (c->alloc_mode != 0 && c->alloc_mode != 1) return fail(FGS_E_INVALID, "invalid mode, got " + std::to_string(c->alloc_mode));
        if (c->alloc_n_res < 1 || c->alloc_n_res > FGS_MAX_SCORE_RES) return fail(FGS_E_INVALID, "NodeResourcesAllocatable: 1..8 resources");
        for (int k = 0; k < c->alloc_n_res; ++k)
            if (c->alloc_res_w[k] <= 0) // pkg/noderesources/allocatable.go:51-58
                return fail(FGS_E_INVALID, "resource Weight of slot " + std::to_string(c->alloc_res[k]) + " should be a positive value, got " +
                                               std::to_string(c->alloc_res_w[k]));
    }
    if (c->score_mask & FGS_PL_FIT) {
        if (c->fit_strategy != 0 && c->fit_strategy != 1)
            return fail(FGS_E_INVALID, "scoring strategy " + std::to_string(c->fit_strategy) + " is not supported");
        if (c->fit_n_res < 1 || c->fit_n_res > FGS_MAX_SCORE_RES) return fail(FGS_E_INVALID, "NodeResourcesFit: 1..8 scoring resources");
        for (int k = 0; k < c->fit_n_res; ++k) {
            // ValidateNodeResourcesFitArgs: weight in [1,100]
            if (c->fit_res_w[k] < 1 || c->fit_res_w[k] > 100) return fail(FGS_E_INVALID, "NodeResourcesFit: resource weight not in [1, 100]");
            if (c->fit_res[k] >= FGS_RES_SCALAR0 + S) return fail(FGS_E_INVALID, "NodeResourcesFit: resource slot out of range");
        }
    }
    for (int64_t w : {c->weight_flex, c->weight_fit, c->weight_alloc})
        if (w < 0) return fail(FGS_E_INVALID, "score plugin weight must be >= 0");
    cfg = *c;
    have_cfg = true;
    nodes_derived = false; // Allocatable raw scores depend on the args
    return FGS_OK;
}

void Engine::fill_eval_params(EvalParams* p, const fgs_eval_opts* opts) const {
    std::memset(p, 0, sizeof(*p));
    p->nd = (const int64_t*)d_nd;
    p->pods = (const int64_t*)d_pods;
    p->reduce_max = (int64_t*)d_perpod;
    p->reduce_sum = (int64_t*)d_perpod + (int64_t)FGS_REDUCE_ROWS * p_pad;
    p->select_key = (int64_t*)d_perpod + (int64_t)(FGS_REDUCE_ROWS + 1) * p_pad;
    p->bitmap = (uint32_t*)d_bitmap;
    p->scores = (int64_t*)d_scores;
    p->n_nodes = n_nodes;
    p->n_pad = n_pad;
    p->node_base = node_base;
    p->n_pods = n_pods;
    p->p_pad = p_pad;
    p->words_per_row = words_per_row;
    p->score_stride = score_stride;
    p->S = S;
    p->filter_mask = cfg.filter_mask;
    p->score_mask = cfg.score_mask;
    p->emit_mask = (opts && opts->emit_mask) ? opts->emit_mask : cfg.score_mask;
    p->emit_raw = opts ? opts->emit_raw : 0;
    // profile order of the two filter plugins (runtime/framework.go:674-689 early exit)
    p->flex_filter_first = 1;
    for (int k = 0; k < 2 && cfg.filter_order[k]; ++k) {
        if (cfg.filter_order[k] == (int)FGS_PL_FLEXGPU) break;
        if (cfg.filter_order[k] == (int)FGS_PL_FIT) {
            p->flex_filter_first = 0;
            break;
        }
    }
    p->w_flex = cfg.weight_flex ? cfg.weight_flex : 1; // runtime/framework.go:405-408
    p->w_fit = cfg.weight_fit ? cfg.weight_fit : 1;
    p->w_alloc = cfg.weight_alloc ? cfg.weight_alloc : 1;
    p->fit_strategy = cfg.fit_strategy;
    p->fit_n_res = (cfg.score_mask & FGS_PL_FIT) ? cfg.fit_n_res : 0;
    p->fit_ignored = cfg.fit_ignored_scalar_mask;
    int wc = 0, wm = 0;
    bool fast = true;
    for (int k = 0; k < p->fit_n_res; ++k) {
        p->fit_res[k] = cfg.fit_res[k];
        p->fit_w[k] = cfg.fit_res_w[k];
        if (cfg.fit_res[k] == FGS_RES_CPU && wc == 0) wc = (int)cfg.fit_res_w[k];
        else if (cfg.fit_res[k] == FGS_RES_MEM && wm == 0) wm = (int)cfg.fit_res_w[k];
        else fast = false;
    }
    fast = fast && wc > 0 && wm > 0;
    p->fit_fast = fast ? 1 : 0;
    p->fit_w_cpu = wc;
    p->fit_w_mem = wm;
    p->fit_wsum = wc + wm;
    p->fit_inv_wsum = (wc + wm) > 0 ? 1.0f / (float)(wc + wm) : 0.f;
    p->seed = seed;
    // selectHost key: score field above a 16-bit hash and a 24-bit node index when Σw·100 < 2^23
    int64_t max_total = 0;
    if (cfg.score_mask & FGS_PL_FLEXGPU) max_total += 100 * p->w_flex;
    if (cfg.score_mask & FGS_PL_FIT) max_total += 100 * p->w_fit;
    if (cfg.score_mask & FGS_PL_ALLOCATABLE) max_total += 100 * p->w_alloc;
    p->key_shift = (max_total < ((int64_t)1 << 23) && !p->emit_raw) ? 40 : 24;
    p->feas_from_bitmap = (sampled || masked) ? 1 : 0;
    // pods per CTA: enough CTAs for a few waves of 148 SMs × 2 resident CTAs
    const int64_t node_tiles = n_pad / kTileN;
    const int64_t target_ctas = 148 * 2 * 4;
    int64_t chunks = std::max<int64_t>(1, target_ctas / std::max<int64_t>(1, node_tiles));
    int64_t ppc = round_up(std::max<int64_t>(kTileP, (p_pad + chunks - 1) / chunks), kTileP);
    if (ppc > 65536) ppc = 65536;
    p->pods_per_cta = (int)ppc;
}

} // namespace fgs

using namespace fgs;

// =================================================================================================
// C ABI
// =================================================================================================

extern "C" const char* fgs_last_error(const fgs_engine* e) {
    if (!e) return g_create_error.c_str();
    return ((const Engine*)e)->err.c_str();
}

extern "C" int fgs_create(const fgs_config* c, fgs_engine** out) {
    if (!c || !out) {
        g_create_error = "fgs_create: null argument";
        return FGS_E_INVALID;
    }
    if (c->n_scalars < 2 || c->n_scalars > FGS_MAX_SCALARS) {
        g_create_error = "fgs_create: n_scalars must be in [2, 8]";
        return FGS_E_INVALID;
    }
    if (c->node_index_base < 0 || c->node_index_base > ((int64_t)1 << 24) - 2 || c->total_nodes < 0 || c->total_nodes > ((int64_t)1 << 24) - 2) {
        g_create_error = "fgs_create: node_index_base / total_nodes must be in [0, 2^24-2] (the selectHost key carries the global node index in 24 bits)";
        return FGS_E_TOO_LARGE;
    }
    int ndev = 0;
    cudaError_t ce = cudaGetDeviceCount(&ndev);
    if (ce != cudaSuccess || ndev <= 0 || c->device < 0 || c->device >= ndev) {
        g_create_error = std::string("fgs_create: no usable CUDA device (") + (ce != cudaSuccess ? cudaGetErrorString(ce) : "ordinal out of range") +
                         "); libfgs has no CPU fallback";
        return FGS_E_CUDA;
    }
    ce = cudaSetDevice(c->device);
    if (ce != cudaSuccess) {
        g_create_error = std::string("cudaSetDevice: ") + cudaGetErrorString(ce);
        return FGS_E_CUDA;
    }
    Engine* e = new (std::nothrow) Engine();
    if (!e) return FGS_E_INVALID;
    e->device = c->device;
    e->S = c->n_scalars;
    e->node_base = c->node_index_base;
    e->total_nodes = c->total_nodes;
    e->seed = c->select_seed;
    {
        const char* fw = getenv("FGS_FORCE_INT64");
        e->force_wide = fw && fw[0] == '1';
        // the fused pass is opt-in: on C2 it is slower than the two passes (1.19 vs 1.10 ms per step: 109 registers per
        // thread leave two CTAs per SM, and the loop does not hide its latencies with 16 warps) — DESIGN.md §3
        const char* nf = getenv("FGS_FUSED");
        e->no_fused = !(nf && nf[0] == '1');
        const char* nd_ = getenv("FGS_NO_DMA");
        e->no_dma = nd_ && nd_[0] == '1';
        const char* tr = getenv("FGS_PIPE_TRACE");
        if (tr && tr[0]) {
            e->pipe_trace = true;
            e->pipe_trace_path = std::string(tr) + "." + std::to_string(c->device);
        }
        const char* sm_ = getenv("FGS_SMALL");
        e->small_off = sm_ && sm_[0] == '0';
        const char* ov_ = getenv("FGS_OVERLAP"); // 0: pass A and the emit pass of consecutive submitted batches share one stream
        e->overlap_filter = !(ov_ && ov_[0] == '0');
        const char* sc_ = getenv("FGS_SMALL_CLUSTER"); // 0: one CTA per pod whatever the node count
        e->small_cluster = !(sc_ && sc_[0] == '0');
        const char* po_ = getenv("FGS_PATCH"); // 0: fgs_update_nodes always re-derives and re-analyses the whole table
        e->patch_off = po_ && po_[0] == '0';
        const char* sd_ = getenv("FGS_SMALL_DIRECT"); // 0: never compare in the small kernel, 1: always, unset: when the level tables are stale
        e->small_direct = sd_ && (sd_[0] == '0' || sd_[0] == '1') ? sd_[0] - '0' : -1;
        const char* fm = getenv("FGS_FILTER");
        e->filter_mode = fm ? (fm[0] == 'c' ? 1 : (fm[0] == 'l' ? 2 : 0)) : 0;
        const char* ht = getenv("FGS_HOST_THREADS");
        int nth = ht ? atoi(ht) : 4;
        const int hw = (int)std::thread::hardware_concurrency();
        if (hw > 0 && nth > hw) nth = hw;
        e->host_threads = nth < 1 ? 1 : nth;
    }
    if (c->stream) {
        e->stream = (cudaStream_t)c->stream;
        e->own_stream = false;
    } else {
        ce = cudaStreamCreateWithFlags(&e->stream, cudaStreamNonBlocking);
        if (ce != cudaSuccess) {
            g_create_error = std::string("cudaStreamCreate: ") + cudaGetErrorString(ce);
            delete e;
            return FGS_E_CUDA;
        }
        e->own_stream = true;
    }
    for (auto& ev : e->events) {
        ce = cudaEventCreate(&ev);
        if (ce != cudaSuccess) {
            g_create_error = std::string("cudaEventCreate: ") + cudaGetErrorString(ce);
            e->free_all();
            delete e;
            return FGS_E_CUDA;
        }
    }
    ce = eval_setup();
    if (ce != cudaSuccess) {
        g_create_error = std::string("kernel attribute setup: ") + cudaGetErrorString(ce) + " (libfgs is built for sm_100a only)";
        e->free_all();
        delete e;
        return FGS_E_CUDA;
    }
    // default plugin set = the flexgpu profile: FlexGPU filter+score (manifests/flexgpu/templates/configmap.yaml)
    fgs_plugin_cfg d;
    std::memset(&d, 0, sizeof(d));
    d.filter_mask = FGS_PL_FLEXGPU | FGS_PL_FIT;
    d.filter_order[0] = FGS_PL_FIT;
    d.filter_order[1] = FGS_PL_FLEXGPU;
    d.score_mask = FGS_PL_FLEXGPU | FGS_PL_FIT;
    d.weight_flex = d.weight_fit = d.weight_alloc = 1;
    d.fit_strategy = 0;
    d.fit_n_res = 2;
    d.fit_res[0] = FGS_RES_CPU;
    d.fit_res[1] = FGS_RES_MEM;
    d.fit_res_w[0] = d.fit_res_w[1] = 1; // vendor/.../apis/config/v1beta2/defaults.go:241-257
    d.alloc_mode = 0;
    d.alloc_n_res = 2;
    d.alloc_res[0] = FGS_RES_CPU;
    d.alloc_res_w[0] = 1 << 20; // pkg/noderesources/resource_allocation.go:38
    d.alloc_res[1] = FGS_RES_MEM;
    d.alloc_res_w[1] = 1;
    e->set_plugins(&d);
    *out = (fgs_engine*)e;
    return FGS_OK;
}

extern "C" void fgs_destroy(fgs_engine* eh) {
    if (!eh) return;
    Engine* e = (Engine*)eh;
    cudaSetDevice(e->device);
    for (cudaStream_t st : {e->stream, e->front_stream, e->copy_stream, e->d2h_stream, e->merge_stream})
        if (st || st == e->stream) cudaStreamSynchronize(st); // batches submitted and never collected: let their work and copies end
    e->trace_dump();
    if (e->trace_epoch) cudaEventDestroy(e->trace_epoch);
    e->free_all();
    delete e;
}

extern "C" int fgs_set_plugins(fgs_engine* eh, const fgs_plugin_cfg* c) {
    if (!eh) return FGS_E_INVALID;
    Engine* e = (Engine*)eh;
    FGS_ENTER(e);
    if (e->in_flight) return e->fail(FGS_E_STATE, "fgs_set_plugins while submitted batches are in flight: fgs_collect them first");
    return e->set_plugins(c);
}

// ---- int32 specialisation: can this (node table, pod batch) pair run on the narrow kernels? --------------
//
// Per resource dimension the shift is the number of trailing zero bits common to every value of that
// dimension on both sides (OR of all values), so dividing by 2^shift is exact; the batch qualifies when
// every shifted value stays below the limits the int32 arithmetic needs (2^24 where a value is multiplied
// by 100, 2^30 where it is only compared). Anything else runs on the int64 kernels.
static int tz64(uint64_t v) { return v ? __builtin_ctzll(v) : 62; }

static uint32_t narrow_spec(const Engine* e);

int Engine::decide_narrow(const fgs_pod_table* t, bool launch_pods) {
    const int rc0 = decide_narrow_once(t, launch_pods);
    if (rc0 || narrow_batch || !an_loose || !narrow_refused_by_nodes) return rc0;
    // Not eligible, under an analysis that row patches only ever widened (a row that left the table still counts in its OR
    // and MAX): take it again over the table as it is now, once
    const int rc1 = resolve_patch();
    if (rc1) return rc1;
    if (an_loose) {
        const int rc2 = analyze_nodes_now(false);
        if (rc2) return rc2;
    }
    return decide_narrow_once(t, launch_pods);
}

int Engine::decide_narrow_once(const fgs_pod_table* t, bool launch_pods) {
    narrow_batch = false;
    narrow_refused_by_nodes = false;
    if (force_wide || !node_an_valid || S < 2) return FGS_OK;
    const bool sc_flex = cfg.score_mask & FGS_PL_FLEXGPU, sc_fit = cfg.score_mask & FGS_PL_FIT, sc_alloc = cfg.score_mask & FGS_PL_ALLOCATABLE;
    EvalParams ep;
    fill_eval_params(&ep, nullptr);
    if (sc_fit && !ep.fit_fast) return FGS_OK;                 // scoring resources beyond cpu/memory
    if (sc_fit && (node_an.flags & 1)) {                       // a node with zero cpu/memory capacity drops out of the average
        narrow_refused_by_nodes = true;
        return FGS_OK;
    }
    const int64_t wsum = (sc_flex ? ep.w_flex : 0) + (sc_fit ? ep.w_fit : 0) + (sc_alloc ? ep.w_alloc : 0);
    if (ep.w_flex > 163 || ep.w_fit > 163 || ep.w_alloc > 163 || wsum * 100 + 1 >= (1 << 14)) return FGS_OK; // 14-bit score field of the warp key
    (void)t;
    uint64_t o[ND_DIMS] = {0};
    int64_t mx[ND_DIMS] = {0};
    auto take = [&](int d, int c) { // statistics collected while the column was staged (fgs_stage_pods)
        o[d] |= stage_or[c];
        mx[d] = stage_max[c] > mx[d] ? stage_max[c] : mx[d];
    };
    take(DIM_CPU, PD_REQ_CPU);
    take(DIM_CPU, PD_NZ_CPU);
    take(DIM_MEM, PD_REQ_MEM);
    take(DIM_MEM, PD_NZ_MEM);
    take(DIM_EPH, PD_REQ_EPH);
    take(DIM_S0, PD_FIXED_COLS + 0);
    take(DIM_S1, PD_FIXED_COLS + 1);
    take(DIM_FG, PD_GPU_LIM);
    take(DIM_FM, PD_MEM_LIM);
    const uint32_t extra = stage_present_or;
    if (extra & ~3u & ~cfg.fit_ignored_scalar_mask) return FGS_OK; // a pod asks for a scalar beyond the two flex slots
    NarrowShifts sh;
    for (int d = 0; d < ND_DIMS; ++d) {
        const uint64_t both = o[d] | node_an.or_bits[d];
        int s = d == DIM_FG ? 0 : tz64(both);
        if (s > 40) s = 40;
        const int64_t m = std::max(mx[d], (int64_t)node_an.max_val[d]) >> s;
        const int64_t limit = (d == DIM_EPH || d == DIM_S0 || d == DIM_S1) ? ((int64_t)1 << 30) : ((int64_t)1 << 24);
        if (m >= limit) {
            narrow_refused_by_nodes = true;
            return FGS_OK;
        }
        sh.s[d] = s;
    }
    // buffers
    if (p_pad > pn_cap) {
        if (d_pn) cudaFree(d_pn);
        d_pn = nullptr;
        FGS_CUDA(this, cudaMalloc(&d_pn, sizeof(int) * PN_COLS * p_pad));
        pn_cap = p_pad;
    }
    {
        const int rc = use_nn(sh, cfg.fit_strategy);
        if (rc) return rc;
    }
    if (launch_pods) {
        launch_narrow_pods((const int64_t*)d_pods, p_pad, S, sh, cfg.fit_ignored_scalar_mask, (int*)d_pn, stream);
        launches++;
    }
    FGS_CUDA(this, cudaGetLastError());
    narrow_batch = true;
    return FGS_OK;
}

void Engine::invalidate_nn() {
    for (auto& en : nn_ring) en.sig = -1;
    narrow_nodes_sig = -1;
}

int Engine::use_nn(const NarrowShifts& sh, int sig) {
    if (n_pad > nn_cap) { // the node table grew: every entry is too small
        FGS_CUDA(this, cudaDeviceSynchronize());
        for (auto& en : nn_ring) {
            if (en.d) cudaFree(en.d);
            en.d = nullptr;
            en.sig = -1;
        }
        d_nn = nullptr;
        nn_cur = -1;
        nn_cap = n_pad;
    }
    int hit = -1;
    for (int i = 0; i < kNnRing; ++i)
        if (nn_ring[i].sig == sig && nn_ring[i].gen == node_gen && std::memcmp(&nn_ring[i].sh, &sh, sizeof(sh)) == 0) hit = i;
    if (hit < 0) {
        // an entry no batch in flight reads: an empty one, else the least recently used
        bool in_use[kNnRing] = {false};
        for (const Slot& sl : slots)
            if (sl.busy && sl.nn_idx >= 0) in_use[sl.nn_idx] = true;
        int victim = -1;
        for (int i = 0; i < kNnRing; ++i) {
            if (in_use[i]) continue;
            if (nn_ring[i].sig < 0 || nn_ring[i].gen != node_gen) {
                victim = i;
                break;
            }
            if (victim < 0 || nn_ring[i].last_use < nn_ring[victim].last_use) victim = i;
        }
        if (victim < 0) return fail(FGS_E_STATE, "int32 node tables: every entry is in use by a batch in flight");
        NnEntry& en = nn_ring[victim];
        if (!en.d) FGS_CUDA(this, cudaMalloc(&en.d, sizeof(int) * NN_COLS * nn_cap));
        launch_narrow_nodes((const int64_t*)d_nd, n_pad, sh, sig, (int*)en.d, stream);
        launches++;
        FGS_CUDA(this, cudaGetLastError());
        en.sh = sh, en.sig = sig, en.gen = node_gen;
        hit = victim;
    }
    nn_ring[hit].last_use = ++nn_clock;
    if (hit != nn_cur || narrow_nodes_sig != sig) orders_valid = false; // the fused pass's node orders follow the narrow node table
    nn_cur = hit;
    d_nn = nn_ring[hit].d;
    narrow_sh = sh;
    narrow_nodes_sig = sig;
    slots[cur_slot].nn_idx = hit;
    return FGS_OK;
}

void Engine::fill_narrow_params(NarrowParams* p) const {
    EvalParams ep;
    fill_eval_params(&ep, nullptr);
    std::memset(p, 0, sizeof(*p));
    p->nn = (const int*)d_nn;
    p->pn = (const int*)d_pn;
    p->nd = (const int64_t*)d_nd;
    p->reduce_max = ep.reduce_max;
    p->reduce_sum = ep.reduce_sum;
    p->select_key = ep.select_key;
    p->bitmap = ep.bitmap;
    p->scores = ep.scores;
    p->n_nodes = n_nodes;
    p->n_pad = n_pad;
    p->node_base = node_base;
    p->n_pods = n_pods;
    p->p_pad = p_pad;
    p->words_per_row = words_per_row;
    p->score_stride = score_stride;
    p->pods_per_cta = ep.pods_per_cta;
    p->flex_filter_first = ep.flex_filter_first;
    p->sh = narrow_sh;
    p->w_flex = (int)ep.w_flex;
    p->w_fit = (int)ep.w_fit;
    p->w_alloc = (int)ep.w_alloc;
    p->fit_w_cpu = ep.fit_w_cpu;
    p->fit_w_mem = ep.fit_w_mem;
    p->fit_inv_wsum = ep.fit_inv_wsum;
    p->seed = seed;
}

static uint32_t narrow_spec(const Engine* e) {
    EvalParams ep;
    e->fill_eval_params(&ep, nullptr);
    return narrow_spec_bits(e->cfg.filter_mask & FGS_PL_FLEXGPU, e->cfg.filter_mask & FGS_PL_FIT, e->cfg.score_mask & FGS_PL_FLEXGPU,
                            e->cfg.score_mask & FGS_PL_FIT, e->cfg.fit_strategy == 1, e->cfg.score_mask & FGS_PL_ALLOCATABLE,
                            ep.fit_w_cpu == ep.fit_w_mem);
}

// ---- the fused pass -----------------------------------------------------------------------------------------------
// One pass over pod × node instead of two when nothing has to happen between Filter and Score: a single engine holding
// the whole cluster (no cross-shard merge), no node sampling, no nominated pods, default outputs, and no
// NodeResourcesAllocatable among the score plugins (its normaliser needs the feasible set first).
bool Engine::fused_eligible(const fgs_eval_opts* opts) const {
    if (no_fused || !narrow_batch || peer || nom_q > 0 || n_pods <= 0) return false;
    if (node_base != 0 || (total_nodes != 0 && total_nodes != n_nodes)) return false;
    if (opts && (opts->sample || opts->skip_pass_b || opts->emit_raw || (opts->emit_mask != 0 && opts->emit_mask != cfg.score_mask))) return false;
    if (cfg.score_mask & FGS_PL_ALLOCATABLE) return false;
    EvalParams ep;
    fill_eval_params(&ep, opts);
    if (ep.key_shift != 40) return false;
    return narrow_fused_available(narrow_spec(this));
}

int Engine::run_fused(const fgs_eval_opts* opts, int part) { // part: 1 = prepass, 2 = the pass itself, 3 = both
    cudaSetDevice(device);
    sampled = false;
    if (part == 2 && !fused_prepass_done) return fail(FGS_E_STATE, "fgs_pass_fused_main before fgs_pass_fused_prepass");
    if (!orders_valid) {
        if (n_pad > order_cap) {
            if (d_order) cudaFree(d_order);
            if (d_order_scratch) cudaFree(d_order_scratch);
            d_order = d_order_scratch = nullptr;
            order_scratch_bytes = narrow_orders_scratch_bytes(n_pad);
            FGS_CUDA(this, cudaMalloc(&d_order, sizeof(int) * 2 * (size_t)n_pad));
            FGS_CUDA(this, cudaMalloc(&d_order_scratch, order_scratch_bytes));
            order_cap = n_pad;
        }
        launch_narrow_orders((const int*)d_nn, n_nodes, n_pad, (int*)d_order, d_order_scratch, order_scratch_bytes, stream);
        launches += 3;
        FGS_CUDA(this, cudaGetLastError());
        orders_valid = true;
    }
    EvalParams ep;
    fill_eval_params(&ep, opts);
    last_key_shift = ep.key_shift;
    if (part & 1) {
        launch_init_perpod((int64_t*)d_perpod, p_pad, stream);
        launches++;
    }
    NarrowParams np;
    fill_narrow_params(&np);
    if (!launch_fused_narrow(np, narrow_spec(this), (const int*)d_order, (const int*)d_order + n_pad, stream, events[2], part))
        return fail(FGS_E_STATE, "fused pass: no instantiation for this plugin set");
    if ((part & 1) && (cfg.score_mask & FGS_PL_FLEXGPU)) launches++;
    if (part & 2) launches++;
    fused_prepass_done = part == 1;
    if (part == 1) {
        FGS_CUDA(this, cudaGetLastError());
        return FGS_OK;
    }
    if ((cfg.filter_mask & FGS_PL_FLEXGPU) && node_an_valid && node_an.n_special > 0) {
        launch_filter_panic_pass((const int*)d_nn, (const int*)d_pn, (const int64_t*)d_special, n_pad, node_an.n_special, n_pods, p_pad,
                                 (cfg.filter_mask & FGS_PL_FIT) ? 1 : 0, ep.flex_filter_first, (int64_t*)d_perpod, stream);
        launches++;
    }
    if ((cfg.score_mask & FGS_PL_FLEXGPU) && node_an_valid && node_an.n_special > 0) { // reads the bitmap the pass has just written
        launch_special_pass((const uint32_t*)d_bitmap, words_per_row, (const int64_t*)d_special, n_pad, node_an.n_special,
                            (const int64_t*)d_pods, (int64_t*)d_perpod, n_pods, p_pad, stream);
        launches++;
    }
    FGS_CUDA(this, cudaGetLastError());
    FGS_CUDA(this, cudaEventRecord(events[3], stream));
    reduce_was_narrow = emit_was_narrow = true;
    reduce_was_lookup = false;
    emitted = true;
    was_fused = true;
    return FGS_OK;
}

// ---- pass A as table look-ups ----------------------------------------------------------------------------------------
void Engine::ft_free() {
    if (ft) {
        ft->free_all();
        delete ft;
        ft = nullptr;
    }
}

// *use ← the level tables are current for this node table and profile (built here when they are not) and the batch's
// rows fit the look-up kernel; false → the compare kernels run
int Engine::ensure_filter_tables(bool* use) {
    *use = false;
    if (filter_mode == 1 || n_nodes <= 0) return FGS_OK;
    if (!ft) {
        ft = new (std::nothrow) FilterTables();
        if (!ft) return fail(FGS_E_INVALID, "out of host memory");
        const char* lim = getenv("FGS_FILTER_TABLE_MB");
        if (lim && atoll(lim) > 0) ft->limit_bytes = (size_t)atoll(lim) << 20;
    }
    const uint32_t fmask = cfg.filter_mask;
    const bool current = ft->node_gen == node_gen && ft->filter_mask == fmask && ft->n_pad == n_pad && ft->v.n_dims == 5 + S;
    if (!current || (!ft->valid && !ft->refused)) {
        FtBuildParams bp;
        bp.nd = (const int64_t*)d_nd;
        bp.n_nodes = n_nodes;
        bp.n_pad = n_pad;
        bp.S = S;
        bp.filter_mask = fmask;
        ft->v.n_dims = 5 + S;
        ft->n_pad = n_pad;
        const cudaError_t ce = ft_build(ft, bp, stream, &launches);
        if (ce != cudaSuccess) {
            cudaGetLastError();
            ft->valid = false;
            if (ce != cudaErrorMemoryAllocation) return fail(FGS_E_CUDA, std::string("filter tables: ") + cudaGetErrorString(ce));
            ft->refused = true; // no room for the tables: the compare kernels need none
        }
        ft->node_gen = node_gen;
        ft->filter_mask = fmask;
    }
    *use = ft->valid && ft_rows_fit(ft->v.W);
    if (!*use && filter_mode == 2) return fail(FGS_E_STATE, "FGS_FILTER=lookup: the level tables exceed their memory limit for this node table");
    return FGS_OK;
}

void Engine::trace_collect(int k) {
    Slot& sl = slots[k];
    if (!trace_epoch) return;
    for (cudaEvent_t ev : {sl.ev_t0, sl.ev_reduce, sl.ev_merge1, sl.ev_emit0, sl.ev_emit, sl.ev_merge2, sl.ev_done}) {
        float ms = -1.f;
        if (cudaEventElapsedTime(&ms, trace_epoch, ev) != cudaSuccess) {
            cudaGetLastError();
            ms = -1.f;
        }
        trace_rows.push_back(ms);
    }
}

void Engine::trace_dump() {
    if (!pipe_trace || trace_rows.empty()) return;
    if (FILE* f = fopen(pipe_trace_path.c_str(), "w")) {
        fprintf(f, "# ms since the first batch: filter_start filter_end merge1_end emit_start emit_end merge2_end d2h_end\n");
        for (size_t i = 0; i + 7 <= trace_rows.size(); i += 7)
            fprintf(f, "%.4f %.4f %.4f %.4f %.4f %.4f %.4f\n", trace_rows[i], trace_rows[i + 1], trace_rows[i + 2], trace_rows[i + 3],
                    trace_rows[i + 4], trace_rows[i + 5], trace_rows[i + 6]);
        fclose(f);
    }
}

void Engine::fill_derive_params(DeriveParams* dp) const {
    std::memset(dp, 0, sizeof(*dp));
    dp->raw = (const int64_t*)d_raw;
    dp->nd = (int64_t*)d_nd;
    dp->n_nodes = n_nodes;
    dp->n_pad = n_pad;
    dp->S = S;
    dp->D = D;
    dp->alloc_mode = cfg.alloc_mode;
    dp->alloc_n_res = (cfg.score_mask & FGS_PL_ALLOCATABLE) ? cfg.alloc_n_res : 0;
    for (int k = 0; k < dp->alloc_n_res; ++k) {
        dp->alloc_res[k] = cfg.alloc_res[k];
        dp->alloc_w[k] = cfg.alloc_res_w[k];
    }
}

// node-side analysis for the int32 specialisation (one small round trip)
int Engine::analyze_nodes_now(bool table_changed) {
    Engine* e = this;
    e->an_loose = false;
    e->node_an_valid = false;
    e->invalidate_nn();
    if (e->patch_spec) { // a full analysis supersedes what the row patches assumed
        FGS_CUDA(e, cudaStreamSynchronize(e->stream));
        e->patch_spec = false;
        *e->h_patch_flag = 0;
        FGS_CUDA(e, cudaMemsetAsync(e->d_patch_flag, 0, sizeof(int), e->stream));
    }
    if (table_changed) e->node_gen++; // upload, full re-derive, sequential commits (row patches count their own)
    if (!e->d_an) FGS_CUDA(e, cudaMalloc(&e->d_an, sizeof(NodeAnalysis)));
    NodeAnalysis init;
    std::memset(&init, 0, sizeof(init));
    init.araw_min = INT64_MAX;
    init.araw_max = INT64_MIN;
    FGS_CUDA(e, cudaMemcpyAsync(e->d_an, &init, sizeof(init), cudaMemcpyHostToDevice, e->stream));
    FGS_CUDA(e, cudaStreamSynchronize(e->stream)); // `init` lives on this stack frame
    if (e->special_cap < e->n_pad) {
        if (e->d_special) cudaFree(e->d_special);
        e->d_special = nullptr;
        FGS_CUDA(e, cudaMalloc(&e->d_special, sizeof(int64_t) * 2 * e->n_pad));
        e->special_cap = e->n_pad;
    }
    launch_analyze_nodes((const int64_t*)e->d_nd, e->n_nodes, e->n_pad, e->S, (NodeAnalysis*)e->d_an, (int64_t*)e->d_special, e->stream);
    e->launches++;
    FGS_CUDA(e, cudaGetLastError());
    FGS_CUDA(e, cudaMemcpyAsync(&e->node_an, e->d_an, sizeof(NodeAnalysis), cudaMemcpyDeviceToHost, e->stream));
    FGS_CUDA(e, cudaStreamSynchronize(e->stream));
    e->node_an_valid = true;
    return FGS_OK;
}

static int run_derive(Engine* e) {
    DeriveParams dp;
    e->fill_derive_params(&dp);
    launch_derive(dp, e->stream);
    e->launches++;
    FGS_CUDA(e, cudaGetLastError());
    e->nodes_derived = true;
    return e->analyze_nodes_now();
}

int Engine::derive_now() { return run_derive(this); }

// Rows patched by patch_nodes were assumed to stay inside the node analysis the host holds. Calls that read the analysis,
// the narrow node table or the special list wait for the patch here and redo the analysis when a row did not.
int Engine::resolve_patch() {
    if (!patch_spec) return FGS_OK;
    FGS_CUDA(this, cudaStreamSynchronize(stream));
    if (*h_patch_flag) {
        patch_redos++;
        return analyze_nodes_now(false);
    }
    patch_spec = false;
    return FGS_OK;
}

int Engine::nodes_ready() {
    if (!nodes_derived) return derive_now();
    return resolve_patch();
}

// one row of a node table in the raw column order (stage_node_columns for a single row, row-major)
static void node_row_values(const Engine* e, const fgs_node_table* t, int64_t src, int64_t* out) {
    const int S = e->S, D = e->D;
    out[NR_ALLOC_MCPU] = t->alloc_mcpu[src], out[NR_ALLOC_MEM] = t->alloc_mem[src], out[NR_ALLOC_EPH] = t->alloc_eph[src];
    out[NR_ALLOC_PODS] = t->alloc_pods[src], out[NR_REQ_MCPU] = t->req_mcpu[src], out[NR_REQ_MEM] = t->req_mem[src];
    out[NR_REQ_EPH] = t->req_eph[src], out[NR_NUM_PODS] = t->num_pods[src], out[NR_NZ_MCPU] = t->nz_mcpu[src], out[NR_NZ_MEM] = t->nz_mem[src];
    out[NR_FLEX_GPU_SUM] = t->flex_gpu_sum[src], out[NR_FLEX_MEM_SUM] = t->flex_mem_sum[src];
    out[NR_META] = (int64_t)(((uint64_t)t->monopoly_mask[src] << 32) | (uint64_t)t->flags[src]);
    out[NR_PRESENT] = (int64_t)t->alloc_scalar_present[src];
    for (int s = 0; s < S; ++s) {
        out[nr_alloc_scalar(s)] = t->alloc_scalar[(int64_t)s * t->n_nodes + src];
        out[nr_req_scalar(S, s)] = t->req_scalar[(int64_t)s * t->n_nodes + src];
    }
    for (int d = 0; d < D; ++d) out[nr_used(S, d)] = d < t->n_devices ? t->used_mem[(int64_t)d * t->n_nodes + src] : 0;
}

// fgs_update_nodes, asynchronous form: see patch_nodes (kernels_narrow.cu)
int Engine::patch_rows(const fgs_node_table* rows, const int64_t* dirty_idx) {
    Engine* e = this;
    const int64_t R_in = rows->n_nodes;
    const int ncols = nr_cols(S, D);
    // a node listed twice: the later row wins, as with the copies of the synchronous form (a thread per row would race)
    std::vector<int64_t> keep; // table rows to apply, when some were dropped
    {
        bool dup = false;
        if (R_in <= 8) {
            for (int64_t a = 0; a < R_in && !dup; ++a)
                for (int64_t b = a + 1; b < R_in; ++b)
                    if (dirty_idx[a] == dirty_idx[b]) dup = true;
        } else {
            std::vector<int64_t> srt(dirty_idx, dirty_idx + R_in);
            std::sort(srt.begin(), srt.end());
            dup = std::adjacent_find(srt.begin(), srt.end()) != srt.end();
        }
        if (dup) {
            std::vector<int64_t> order((size_t)R_in);
            for (int64_t r = 0; r < R_in; ++r) order[(size_t)r] = r;
            std::stable_sort(order.begin(), order.end(), [&](int64_t a, int64_t b) { return dirty_idx[a] < dirty_idx[b]; });
            for (int64_t k = 0; k < R_in; ++k)
                if (k + 1 == R_in || dirty_idx[order[(size_t)k]] != dirty_idx[order[(size_t)k + 1]]) keep.push_back(order[(size_t)k]);
        }
    }
    const int64_t R = keep.empty() ? R_in : (int64_t)keep.size();
    auto src_row = [&](int64_t r) { return keep.empty() ? r : keep[(size_t)r]; };
    if (!d_patch_flag) {
        FGS_CUDA(e, cudaMalloc(&d_patch_flag, sizeof(int)));
        FGS_CUDA(e, cudaMemset(d_patch_flag, 0, sizeof(int)));
        FGS_CUDA(e, cudaHostAlloc((void**)&h_patch_flag, sizeof(int), cudaHostAllocMapped));
        *h_patch_flag = 0;
    }
    PatchParams pp;
    fill_derive_params(&pp.dp);
    pp.rows = nullptr, pp.idx = nullptr;
    pp.n_dirty = (int)R, pp.ncols = ncols;
    pp.ref = node_an;
    pp.special = (long long*)d_special;
    pp.nn = (int*)d_nn;
    pp.sh = narrow_sh;
    pp.fit_most = cfg.fit_strategy == 1;
    pp.narrow_valid = d_nn && nn_cur >= 0 && nn_ring[nn_cur].gen == node_gen && narrow_nodes_sig == cfg.fit_strategy && nn_cap >= n_pad;
    pp.d_flag = d_patch_flag;
    pp.h_flag = h_patch_flag;
    if (R <= kPatchInline) {
        for (int64_t r = 0; r < R; ++r) {
            pp.inline_idx[r] = dirty_idx[src_row(r)];
            node_row_values(this, rows, src_row(r), (int64_t*)pp.inline_rows[r]);
        }
    } else {
        const size_t need = sizeof(int64_t) * (size_t)R * (size_t)(ncols + 1);
        if (need > patch_bytes) {
            FGS_CUDA(e, cudaStreamSynchronize(stream));
            if (h_patch) cudaFreeHost(h_patch);
            if (d_patch) cudaFree(d_patch);
            h_patch = d_patch = nullptr;
            patch_bytes = 0;
            FGS_CUDA(e, cudaMallocHost(&h_patch, need * 2));
            FGS_CUDA(e, cudaMalloc(&d_patch, need * 2));
            patch_bytes = need * 2;
        } else if (patch_spec) {
            FGS_CUDA(e, cudaStreamSynchronize(stream)); // the copy of the previous patch may still read the staging buffer
        }
        int64_t* h = (int64_t*)h_patch;
        for (int64_t r = 0; r < R; ++r) {
            node_row_values(this, rows, src_row(r), h + r * ncols);
            h[R * ncols + r] = dirty_idx[src_row(r)];
        }
        FGS_CUDA(e, cudaMemcpyAsync(d_patch, h_patch, need, cudaMemcpyHostToDevice, stream));
        pp.rows = (const long long*)d_patch;
        pp.idx = (const long long*)d_patch + R * ncols;
    }
    FGS_CUDA(e, launch_patch_nodes(pp, stream));
    launches++;
    node_gen++;
    if (pp.narrow_valid && nn_cur >= 0) nn_ring[nn_cur].gen = node_gen; // the table in use was patched along: it describes the new version
    patch_spec = true;
    an_loose = true;
    patches++;
    return FGS_OK;
}

// copy one table row-set into the pinned column-major staging buffer
static void stage_node_columns(Engine* e, const fgs_node_table* t, int64_t dst0, int64_t count, int64_t src0) {
    const int64_t np = e->n_pad;
    const int S = e->S, D = e->D;
    int64_t* h = (int64_t*)e->h_raw;
    auto col = [&](int c, const int64_t* src) { std::memcpy(h + (int64_t)c * np + dst0, src + src0, sizeof(int64_t) * (size_t)count); };
    col(NR_ALLOC_MCPU, t->alloc_mcpu);
    col(NR_ALLOC_MEM, t->alloc_mem);
    col(NR_ALLOC_EPH, t->alloc_eph);
    col(NR_ALLOC_PODS, t->alloc_pods);
    col(NR_REQ_MCPU, t->req_mcpu);
    col(NR_REQ_MEM, t->req_mem);
    col(NR_REQ_EPH, t->req_eph);
    col(NR_NUM_PODS, t->num_pods);
    col(NR_NZ_MCPU, t->nz_mcpu);
    col(NR_NZ_MEM, t->nz_mem);
    col(NR_FLEX_GPU_SUM, t->flex_gpu_sum);
    col(NR_FLEX_MEM_SUM, t->flex_mem_sum);
    for (int64_t i = 0; i < count; ++i) {
        h[(int64_t)NR_META * np + dst0 + i] = (int64_t)(((uint64_t)t->monopoly_mask[src0 + i] << 32) | (uint64_t)t->flags[src0 + i]);
        h[(int64_t)NR_PRESENT * np + dst0 + i] = (int64_t)t->alloc_scalar_present[src0 + i];
    }
    for (int s = 0; s < S; ++s) {
        col(nr_alloc_scalar(s), t->alloc_scalar + (int64_t)s * t->n_nodes);
        col(nr_req_scalar(S, s), t->req_scalar + (int64_t)s * t->n_nodes);
    }
    for (int d = 0; d < D; ++d) {
        if (d < t->n_devices) col(nr_used(S, d), t->used_mem + (int64_t)d * t->n_nodes);
        else std::memset(h + (int64_t)nr_used(S, d) * np + dst0, 0, sizeof(int64_t) * (size_t)count);
    }
}

static int validate_node_table(Engine* e, const fgs_node_table* t) {
    if (!t || t->n_nodes < 0) return e->fail(FGS_E_INVALID, "node table: null or negative size");
    if (t->n_scalars != e->S) return e->fail(FGS_E_INVALID, "node table: n_scalars differs from the engine's");
    if (t->n_devices < 1 || t->n_devices > FGS_MAX_DEVICES) return e->fail(FGS_E_TOO_LARGE, "node table: n_devices must be in [1, 16]");
    if (t->n_nodes >= ((int64_t)1 << 24) - 1) return e->fail(FGS_E_TOO_LARGE, "node table: at most 2^24-2 nodes per engine");
    // the selectHost key carries the GLOBAL node index in 24 bits
    if (e->node_base + t->n_nodes > ((int64_t)1 << 24) - 2) return e->fail(FGS_E_TOO_LARGE, "node table: node_index_base + n_nodes exceeds 2^24-2");
    return FGS_OK;
}

extern "C" int fgs_upload_nodes(fgs_engine* eh, const fgs_node_table* t) {
    if (!eh) return FGS_E_INVALID;
    Engine* e = (Engine*)eh;
    FGS_ENTER(e);
    if (e->in_flight) return e->fail(FGS_E_STATE, "fgs_upload_nodes while submitted batches are in flight: fgs_collect them first");
    int rc = validate_node_table(e, t);
    if (rc) return rc;
    cudaSetDevice(e->device);
    FGS_CUDA(e, cudaStreamSynchronize(e->stream)); // staging buffer reuse
    rc = e->ensure_nodes(t->n_nodes, t->n_devices);
    if (rc) return rc;
    e->n_nodes = t->n_nodes;
    e->D = std::max(e->D_cap, 1);
    std::memset(e->h_raw, 0, sizeof(int64_t) * nr_cols(e->S, e->D) * e->n_pad);
    stage_node_columns(e, t, 0, t->n_nodes, 0);
    FGS_CUDA(e, cudaMemcpyAsync(e->d_raw, e->h_raw, sizeof(int64_t) * nr_cols(e->S, e->D) * e->n_pad, cudaMemcpyHostToDevice, e->stream));
    e->have_nodes = true;
    e->synced_gen.clear(); // fgs_sync_nodes re-fills it when the upload came from there
    return run_derive(e);
}

extern "C" int fgs_update_nodes(fgs_engine* eh, const fgs_node_table* rows, const int64_t* dirty_idx) {
    if (!eh) return FGS_E_INVALID;
    Engine* e = (Engine*)eh;
    FGS_ENTER(e);
    if (!e->have_nodes) return e->fail(FGS_E_STATE, "fgs_update_nodes before fgs_upload_nodes");
    if (e->in_flight) return e->fail(FGS_E_STATE, "fgs_update_nodes while submitted batches are in flight: fgs_collect them first");
    int rc = validate_node_table(e, rows);
    if (rc) return rc;
    if (rows->n_devices > e->D) return e->fail(FGS_E_INVALID, "fgs_update_nodes: more device slots than the uploaded table; re-upload");
    if (rows->n_nodes && !dirty_idx) return e->fail(FGS_E_INVALID, "fgs_update_nodes: null dirty_idx");
    cudaSetDevice(e->device);
    for (int64_t r = 0; r < rows->n_nodes; ++r)
        if (dirty_idx[r] < 0 || dirty_idx[r] >= e->n_nodes) return e->fail(FGS_E_INVALID, "fgs_update_nodes: dirty index out of range");
    if (rows->n_nodes == 0) return FGS_OK;
    // a few rows of a table whose analysis the host holds: patched, derived and checked on the device, without waiting
    if (!e->patch_off && e->nodes_derived && e->node_an_valid && rows->n_nodes <= Engine::kPatchMaxRows) return e->patch_rows(rows, dirty_idx);
    FGS_CUDA(e, cudaStreamSynchronize(e->stream));
    const int ncols = nr_cols(e->S, e->D);
    for (int64_t r = 0; r < rows->n_nodes; ++r) {
        const int64_t n = dirty_idx[r];
        stage_node_columns(e, rows, n, 1, r);
        // one small strided copy per dirty row: ncols values, pitch = n_pad
        FGS_CUDA(e, cudaMemcpy2DAsync((int64_t*)e->d_raw + n, sizeof(int64_t) * e->n_pad, (int64_t*)e->h_raw + n, sizeof(int64_t) * e->n_pad,
                                      sizeof(int64_t), (size_t)ncols, cudaMemcpyHostToDevice, e->stream));
    }
    return run_derive(e);
}

// Cache.UpdateSnapshot's incremental walk for the device tables: only the rows whose NodeInfo.Generation moved
extern "C" int fgs_sync_nodes(fgs_engine* eh, fgs_packer* pk, const fgs_snapshot* snap, int64_t n0, int64_t n1, int64_t* n_repacked) {
    if (!eh || !pk || !snap) return FGS_E_INVALID;
    Engine* e = (Engine*)eh;
    FGS_ENTER(e);
    if (!snap->generation) return e->fail(FGS_E_INVALID, "fgs_sync_nodes: the snapshot carries no generation column");
    if (n0 < 0 || n1 < n0 || n1 > snap->n_nodes) return e->fail(FGS_E_INVALID, "fgs_sync_nodes: node range outside the snapshot");
    const int64_t N = n1 - n0;
    std::vector<int64_t> stale;
    bool full = !e->have_nodes || e->n_nodes != N || (int64_t)e->synced_gen.size() != N;
    if (!full) {
        // block-wise: a loop that may push_back does not vectorise, memcmp over 256 generations does, and almost every block is clean
        const int64_t* gnow = snap->generation + n0;
        const int64_t* gold = e->synced_gen.data();
        for (int64_t b0 = 0; b0 < N; b0 += 256) {
            const int64_t b1 = std::min<int64_t>(N, b0 + 256);
            if (std::memcmp(gnow + b0, gold + b0, sizeof(int64_t) * (size_t)(b1 - b0)) == 0) continue;
            for (int64_t i = b0; i < b1; ++i)
                if (gnow[i] != gold[i]) stale.push_back(n0 + i);
        }
        if ((int64_t)stale.size() * 2 > N) full = true;
    }
    const fgs_node_table* t = nullptr;
    int rc;
    if (!full && !stale.empty()) {
        rc = fgs_packer_node_rows(pk, snap, stale.data(), (int64_t)stale.size(), &t);
        if (rc) return e->fail(rc, std::string("fgs_sync_nodes: ") + fgs_packer_error(pk));
        if (t->n_devices > e->D) full = true; // a stale row grew more devices than the uploaded table has slots for
    }
    if (full) {
        rc = fgs_packer_nodes(pk, snap, n0, n1, &t);
        if (rc) return e->fail(rc, std::string("fgs_sync_nodes: ") + fgs_packer_error(pk));
        rc = fgs_upload_nodes(eh, t);
        if (rc) return rc;
        e->synced_gen.assign(snap->generation + n0, snap->generation + n1);
        if (n_repacked) *n_repacked = N;
        return FGS_OK;
    }
    if (!stale.empty()) {
        std::vector<int64_t> local(stale.size());
        for (size_t i = 0; i < stale.size(); ++i) local[i] = stale[i] - n0;
        rc = fgs_update_nodes(eh, t, local.data());
        if (rc) return rc;
        for (size_t i = 0; i < stale.size(); ++i) e->synced_gen[(size_t)local[i]] = snap->generation[stale[i]];
    }
    if (n_repacked) *n_repacked = (int64_t)stale.size();
    return FGS_OK;
}

// async: called by fgs_submit — the staging buffers belong to a pipeline slot that is known to be free, so the
// compute stream is not drained, and the H2D copy runs on the copy stream (it overlaps the previous batch's kernels)
// The pinned staging buffer is written once and read by the DMA engine only: a streaming store keeps it out of the
// caches and spares the read-for-ownership (with one engine per GPU and several GPUs per host the staging passes share
// the host's memory bandwidth).
static inline void stream_store(int64_t* p, int64_t v) {
#if defined(__x86_64__)
    _mm_stream_si64((long long*)p, (long long)v);
#else
    *p = v;
#endif
}

static int stage_pods_impl(Engine* e, const fgs_pod_table* t, bool async) {
    if (!e->have_nodes) return e->fail(FGS_E_STATE, "fgs_stage_pods before fgs_upload_nodes");
    if (!t || t->n_pods < 0) return e->fail(FGS_E_INVALID, "pod table: null or negative size");
    if (t->n_scalars != e->S) return e->fail(FGS_E_INVALID, "pod table: n_scalars differs from the engine's");
    cudaSetDevice(e->device);
    if (!async) FGS_CUDA(e, cudaStreamSynchronize(e->stream));
    int rc = e->ensure_pods(t->n_pods);
    if (rc) return rc;
    rc = e->nodes_ready(); // derived columns, and a settled node analysis (row patches are checked on the device)
    if (rc) return rc;
    e->n_pods = t->n_pods;
    const int64_t P = t->n_pods, pp = e->p_pad;
    const int S = e->S;
    int64_t* h = (int64_t*)e->h_pods;
    FGS_CUDA(e, cudaEventRecord(e->events[0], e->stream));
    // DMA path: the table brings its column statistics and lies in page-locked memory (fgs_pod_stats, fgs_packer_pin) —
    // every column goes to the GPU straight from the caller's arrays: no host pass, no staging copy
    if (t->stats && (t->stats->flags & FGS_POD_STATS_PINNED) && P > 0 && !e->no_dma) {
        cudaStream_t st = async ? e->copy_stream : e->stream;
        int64_t* d = (int64_t*)e->d_pods;
        const int64_t* cols[PD_FIXED_COLS] = {nullptr};
        cols[PD_GPU_LIM] = t->flex_gpu_limit, cols[PD_MEM_LIM] = t->flex_mem_limit, cols[PD_REQ_CPU] = t->req_mcpu, cols[PD_REQ_MEM] = t->req_mem;
        cols[PD_REQ_EPH] = t->req_eph, cols[PD_NZ_CPU] = t->nz_mcpu, cols[PD_NZ_MEM] = t->nz_mem;
        for (int c = 0; c < PD_FIXED_COLS; ++c)
            if (cols[c]) FGS_CUDA(e, cudaMemcpyAsync(d + (int64_t)c * pp, cols[c], sizeof(int64_t) * (size_t)P, cudaMemcpyHostToDevice, st));
        // the S scalar columns are one [S][P] block in the table and [S][pp] on the device
        FGS_CUDA(e, cudaMemcpy2DAsync(d + (int64_t)PD_FIXED_COLS * pp, sizeof(int64_t) * (size_t)pp, t->req_scalar, sizeof(int64_t) * (size_t)P,
                                      sizeof(int64_t) * (size_t)P, (size_t)S, cudaMemcpyHostToDevice, st));
        uint32_t* scratch = (uint32_t*)(d + (int64_t)pd_cols(S) * pp);
        FGS_CUDA(e, cudaMemcpyAsync(scratch, t->flags, sizeof(uint32_t) * (size_t)P, cudaMemcpyHostToDevice, st));
        FGS_CUDA(e, cudaMemcpyAsync(scratch + pp, t->req_scalar_present, sizeof(uint32_t) * (size_t)P, cudaMemcpyHostToDevice, st));
        if (pp > P) // the padding pods of every column
            FGS_CUDA(e, cudaMemset2DAsync(d + P, sizeof(int64_t) * (size_t)pp, 0, sizeof(int64_t) * (size_t)(pp - P), (size_t)pd_cols(S), st));
        launch_make_pod_meta(scratch, scratch + pp, d + (int64_t)PD_META * pp, P, pp, st);
        e->launches++;
        FGS_CUDA(e, cudaGetLastError());
        for (int c = 0; c < pd_cols(S); ++c) {
            e->stage_or[c] = t->stats->or_bits[c];
            e->stage_max[c] = t->stats->max_val[c];
        }
        e->stage_present_or = t->stats->present_or;
        if (async) {
            FGS_CUDA(e, cudaEventRecord(e->ev_h2d, e->copy_stream));
            FGS_CUDA(e, cudaStreamWaitEvent(e->stream, e->ev_h2d, 0));
        }
        int rc2 = e->decide_narrow(t);
        if (rc2) return rc2;
        rc2 = e->stage_nominated_aux(t);
        if (rc2) return rc2;
        FGS_CUDA(e, cudaEventRecord(e->events[1], e->stream));
        e->have_pods = true;
        e->staged_gen = e->node_gen;
        e->reduced = false;
        e->staged_by_dma = true;
        return FGS_OK;
    }
    e->staged_by_dma = false;
    // one pass over every column: copy into the pinned staging buffer and collect the OR / MAX that the int32
    // specialisation needs (decide_narrow); a few host threads share the work
    // ONE parallel region: every thread takes a contiguous range of pods through all the columns (one fork / join per
    // batch instead of one per column — with one engine per GPU and several GPUs per host the joins were the cost)
    const int nth = e->host_threads;
    constexpr int kMaxCols = PD_FIXED_COLS + FGS_MAX_SCALARS;
    const int64_t* src[kMaxCols] = {nullptr};
    src[PD_GPU_LIM] = t->flex_gpu_limit;
    src[PD_MEM_LIM] = t->flex_mem_limit;
    src[PD_REQ_CPU] = t->req_mcpu;
    src[PD_REQ_MEM] = t->req_mem;
    src[PD_REQ_EPH] = t->req_eph;
    src[PD_NZ_CPU] = t->nz_mcpu;
    src[PD_NZ_MEM] = t->nz_mem;
    for (int s = 0; s < S; ++s) src[PD_FIXED_COLS + s] = t->req_scalar + (int64_t)s * P;
    const int ncols = pd_cols(S);
    uint64_t acc_or[kMaxCols] = {0};
    int64_t acc_max[kMaxCols] = {0};
    uint32_t acc_pr = 0;
#pragma omp parallel num_threads(nth)
    {
        const int tn = omp_get_num_threads(), ti = omp_get_thread_num();
        const int64_t i0 = P * ti / tn, i1 = P * (ti + 1) / tn;
        uint64_t lo[kMaxCols] = {0};
        int64_t lm[kMaxCols] = {0};
        uint32_t lpr = 0;
        for (int c = 0; c < ncols; ++c) {
            int64_t* dst = h + (int64_t)c * pp;
            if (c == PD_META) {
                for (int64_t i = i0; i < i1; ++i) {
                    stream_store(dst + i, (int64_t)(((uint64_t)t->req_scalar_present[i] << 32) | (uint64_t)t->flags[i]));
                    lpr |= t->req_scalar_present[i];
                }
                continue;
            }
            const int64_t* sc = src[c];
            uint64_t o = 0;
            int64_t m = 0;
            for (int64_t i = i0; i < i1; ++i) {
                const int64_t v = sc[i];
                stream_store(dst + i, v);
                o |= (uint64_t)v;
                m = v > m ? v : m;
            }
            lo[c] = o;
            lm[c] = m;
        }
#if defined(__x86_64__)
        _mm_sfence(); // this thread's streaming stores are globally visible before the region ends
#endif
#pragma omp critical
        {
            for (int c = 0; c < ncols; ++c) {
                acc_or[c] |= lo[c];
                acc_max[c] = lm[c] > acc_max[c] ? lm[c] : acc_max[c];
            }
            acc_pr |= lpr;
        }
    }
    for (int c = 0; c < ncols; ++c) {
        std::memset(h + (int64_t)c * pp + P, 0, sizeof(int64_t) * (size_t)(pp - P));
        e->stage_or[c] = acc_or[c];
        e->stage_max[c] = acc_max[c];
    }
    e->stage_present_or = acc_pr;
    if (async) {
        FGS_CUDA(e, cudaMemcpyAsync(e->d_pods, e->h_pods, sizeof(int64_t) * pd_cols(S) * pp, cudaMemcpyHostToDevice, e->copy_stream));
        FGS_CUDA(e, cudaEventRecord(e->ev_h2d, e->copy_stream));
        FGS_CUDA(e, cudaStreamWaitEvent(e->stream, e->ev_h2d, 0));
    } else {
        FGS_CUDA(e, cudaMemcpyAsync(e->d_pods, e->h_pods, sizeof(int64_t) * pd_cols(S) * pp, cudaMemcpyHostToDevice, e->stream));
    }
    rc = e->decide_narrow(t); // int32 tables for this batch when every value qualifies
    if (rc) return rc;
    rc = e->stage_nominated_aux(t); // priority / uid columns for the nominated-pod filter pass, if any are set
    if (rc) return rc;
    FGS_CUDA(e, cudaEventRecord(e->events[1], e->stream));
    e->have_pods = true;
    e->staged_gen = e->node_gen;
    e->reduced = false;
    return FGS_OK;
}

extern "C" int fgs_stage_pods(fgs_engine* eh, const fgs_pod_table* t) {
    if (!eh) return FGS_E_INVALID;
    Engine* e = (Engine*)eh;
    FGS_ENTER(e);
    if (e->in_flight) return e->fail(FGS_E_STATE, "fgs_stage_pods while submitted batches are in flight: fgs_collect them first");
    return stage_pods_impl(e, t, false);
}

// The reference's node sampling on top of the reduce pass's full bitmap (sample.cu): walk, truncate, re-reduce.
int Engine::run_sampling(const fgs_eval_opts* opts) {
    Engine* e = this;
    if (e->sampled) return FGS_OK;
    if (e->node_base != 0 || (e->total_nodes != 0 && e->total_nodes != e->n_nodes))
        return e->fail(FGS_E_STATE, "node sampling needs the whole cluster on one engine (node_index_base 0)");
    if (e->n_pods == 0 || e->n_nodes == 0) return FGS_OK;
    if (e->p_pad > e->walk_cap) {
        if (e->d_walk) cudaFree(e->d_walk);
        if (e->h_walk) cudaFreeHost(e->h_walk);
        e->d_walk = e->h_walk = nullptr;
        e->walk_cap = 0;
        FGS_CUDA(e, cudaMalloc(&e->d_walk, sizeof(int64_t) * (size_t)(e->p_pad + 1)));
        FGS_CUDA(e, cudaMallocHost(&e->h_walk, sizeof(int64_t) * (size_t)(e->p_pad + 1)));
        e->walk_cap = e->p_pad;
    }
    SampleParams sp;
    std::memset(&sp, 0, sizeof(sp));
    sp.bitmap = (uint32_t*)e->d_bitmap;
    sp.nd = (const int64_t*)e->d_nd;
    sp.pods = (const int64_t*)e->d_pods;
    sp.reduce_max = (int64_t*)e->d_perpod;
    sp.reduce_sum = (int64_t*)e->d_perpod + (int64_t)FGS_REDUCE_ROWS * e->p_pad;
    sp.walk = e->d_walk;
    sp.next_start = (int64_t*)e->d_walk + e->p_pad; // int2 [p_pad] = int64 [p_pad], then the scalar
    sp.words_per_row = e->words_per_row;
    sp.n_nodes = e->n_nodes;
    sp.n_pad = e->n_pad;
    sp.n_pods = e->n_pods;
    sp.p_pad = e->p_pad;
    sp.k = num_feasible_nodes_to_find(e->n_nodes, opts->percentage_of_nodes_to_score);
    sp.start = opts->start_node_index;
    {   // the block maps of the composed walk (long batches on node tables of up to 16 384 rows)
        const int64_t need = fgs::sample_scratch_bytes(e->n_nodes, e->n_pods);
        if (need > e->sample_scratch_cap) {
            FGS_CUDA(e, cudaStreamSynchronize(e->stream));
            if (e->d_sample_scratch) cudaFree(e->d_sample_scratch);
            e->d_sample_scratch = nullptr;
            e->sample_scratch_cap = 0;
            FGS_CUDA(e, cudaMalloc(&e->d_sample_scratch, (size_t)need));
            e->sample_scratch_cap = need;
        }
        sp.scratch = need > 0 ? e->d_sample_scratch : nullptr;
        sp.scratch_bytes = e->sample_scratch_cap;
    }
    FGS_CUDA(e, launch_sampling(sp, e->stream));
    e->launches += sample_launch_count(e->n_nodes, e->n_pods, sp.scratch != nullptr);
    if ((e->cfg.score_mask & FGS_PL_FLEXGPU) && e->node_an_valid && e->node_an.n_special > 0) {
        launch_special_pass((const uint32_t*)e->d_bitmap, e->words_per_row, (const int64_t*)e->d_special, e->n_pad, e->node_an.n_special,
                            (const int64_t*)e->d_pods, (int64_t*)e->d_perpod, e->n_pods, e->p_pad, e->stream);
        e->launches++;
        FGS_CUDA(e, cudaGetLastError());
    }
    e->sampled = true;
    return FGS_OK;
}

extern "C" int fgs_pass_reduce(fgs_engine* eh) {
    if (!eh) return FGS_E_INVALID;
    Engine* e = (Engine*)eh;
    FGS_ENTER(e);
    if (!e->have_pods) return e->fail(FGS_E_STATE, "fgs_pass_reduce before fgs_stage_pods");
    if (e->staged_gen != e->node_gen)
        return e->fail(FGS_E_STATE, "fgs_pass_reduce: the node table changed since the batch was staged (its int32 columns follow the node table): stage it again");
    cudaSetDevice(e->device);
    {
        const int rc = e->ensure_slot_nn();
        if (rc) return rc;
    }
    e->sampled = false;
    e->masked = false;
    e->was_fused = false;
    EvalParams p;
    e->fill_eval_params(&p, nullptr);
    bool lookup = false;
    {
        const int rc = e->ensure_filter_tables(&lookup);
        if (rc) return rc;
    }
    bool done = false;
    e->reduce_was_lookup = lookup;
    if (lookup) { // Filter by table look-up: one warp per pod ANDs the level rows of its values (filter_tables.cu)
        FtRowsParams rp;
        std::memset(&rp, 0, sizeof(rp));
        rp.v = e->ft->v;
        rp.pods = (const int64_t*)e->d_pods;
        rp.nd = (const int64_t*)e->d_nd;
        rp.n_pods = e->n_pods;
        rp.p_pad = e->p_pad;
        rp.n_nodes = e->n_nodes;
        rp.n_pad = e->n_pad;
        rp.S = e->S;
        rp.fit_ignored = e->cfg.fit_ignored_scalar_mask;
        rp.bitmap = (uint32_t*)e->d_bitmap;
        rp.reduce_max = p.reduce_max;
        rp.reduce_sum = p.reduce_sum;
        rp.select_key = p.select_key;
        rp.sc_flex = (e->cfg.score_mask & FGS_PL_FLEXGPU) ? 1 : 0;
        rp.sc_alloc = (e->cfg.score_mask & FGS_PL_ALLOCATABLE) ? 1 : 0;
        // the per-pod rows the special nodes feed are written by the same kernel: no initialisation, no side passes
        rp.special = (const int64_t*)e->d_special;
        rp.n_special = e->node_an_valid ? e->node_an.n_special : 0;
        rp.filter_mask = e->cfg.filter_mask;
        rp.flex_first = p.flex_filter_first;
        const int64_t need_idx = (int64_t)rp.v.n_dims * e->p_pad;
        if (need_idx > e->ft->rowidx_cap) {
            if (e->ft->d_rowidx) cudaFree(e->ft->d_rowidx);
            e->ft->d_rowidx = nullptr;
            e->ft->rowidx_cap = 0;
            FGS_CUDA(e, cudaMalloc(&e->ft->d_rowidx, sizeof(int) * (size_t)need_idx));
            e->ft->rowidx_cap = need_idx;
        }
        FGS_CUDA(e, ft_launch_rows(rp, (int*)e->ft->d_rowidx, e->stream));
        e->launches += 2;
    } else {
        launch_init_perpod((int64_t*)e->d_perpod, e->p_pad, e->stream);
        e->launches++;
        if (e->narrow_batch) {
            NarrowParams np;
            e->fill_narrow_params(&np);
            done = launch_reduce_narrow(np, narrow_spec(e), e->stream);
        }
        if (!done) launch_eval_reduce(p, e->stream);
        e->launches++;
    }
    e->reduce_was_narrow = done;
    if (done && (e->cfg.filter_mask & FGS_PL_FLEXGPU) && e->node_an_valid && e->node_an.n_special > 0) { // the narrow loop leaves this flag to a side pass
        launch_filter_panic_pass((const int*)e->d_nn, (const int*)e->d_pn, (const int64_t*)e->d_special, e->n_pad, e->node_an.n_special, e->n_pods, e->p_pad,
                                 (e->cfg.filter_mask & FGS_PL_FIT) ? 1 : 0, p.flex_filter_first, (int64_t*)e->d_perpod, e->stream);
        e->launches++;
    }
    if (!lookup && (e->cfg.score_mask & FGS_PL_FLEXGPU) && e->node_an_valid && e->node_an.n_special > 0) {
        launch_special_pass((const uint32_t*)e->d_bitmap, e->words_per_row, (const int64_t*)e->d_special, e->n_pad, e->node_an.n_special,
                            (const int64_t*)e->d_pods, (int64_t*)e->d_perpod, e->n_pods, e->p_pad, e->stream);
        e->launches++;
    }
    FGS_CUDA(e, cudaGetLastError());
    {
        const int rc = e->run_nominated_pass(); // RunFilterPluginsWithNominatedPods' first pass, where pods are nominated
        if (rc) return rc;
    }
    FGS_CUDA(e, cudaEventRecord(e->events[2], e->stream));
    e->reduced = true;
    return FGS_OK;
}

extern "C" int fgs_fused_eligible(fgs_engine* eh, const fgs_eval_opts* opts) {
    if (!eh) return 0;
    Engine* e = (Engine*)eh;
    FGS_ENTER_RET(e, 0);
    return e->have_pods && e->fused_eligible(opts) ? 1 : 0;
}

extern "C" int fgs_pass_fused(fgs_engine* eh, const fgs_eval_opts* opts) {
    if (!eh) return FGS_E_INVALID;
    Engine* e = (Engine*)eh;
    FGS_ENTER(e);
    if (!e->have_pods) return e->fail(FGS_E_STATE, "fgs_pass_fused before fgs_stage_pods");
    if (!e->fused_eligible(opts)) return e->fail(FGS_E_STATE, "fgs_pass_fused: this batch / configuration needs the two passes (see fgs_fused_eligible)");
    return e->run_fused(opts, 3);
}

extern "C" int fgs_pass_fused_prepass(fgs_engine* eh, const fgs_eval_opts* opts) {
    if (!eh) return FGS_E_INVALID;
    Engine* e = (Engine*)eh;
    FGS_ENTER(e);
    if (!e->have_pods) return e->fail(FGS_E_STATE, "fgs_pass_fused_prepass before fgs_stage_pods");
    if (!e->fused_eligible(opts)) return e->fail(FGS_E_STATE, "fgs_pass_fused_prepass: this batch / configuration needs the two passes (see fgs_fused_eligible)");
    return e->run_fused(opts, 1);
}

extern "C" int fgs_pass_fused_main(fgs_engine* eh, const fgs_eval_opts* opts) {
    if (!eh) return FGS_E_INVALID;
    Engine* e = (Engine*)eh;
    FGS_ENTER(e);
    if (!e->have_pods) return e->fail(FGS_E_STATE, "fgs_pass_fused_main before fgs_stage_pods");
    if (!e->fused_eligible(opts)) return e->fail(FGS_E_STATE, "fgs_pass_fused_main: this batch / configuration needs the two passes (see fgs_fused_eligible)");
    return e->run_fused(opts, 2);
}

extern "C" int fgs_pass_emit(fgs_engine* eh, const fgs_eval_opts* opts) {
    if (!eh) return FGS_E_INVALID;
    Engine* e = (Engine*)eh;
    FGS_ENTER(e);
    if (!e->have_pods) return e->fail(FGS_E_STATE, "fgs_pass_emit before fgs_stage_pods");
    if (e->staged_gen != e->node_gen)
        return e->fail(FGS_E_STATE, "fgs_pass_emit: the node table changed since the batch was staged (its int32 columns follow the node table): stage it again");
    cudaSetDevice(e->device);
    {
        const int rc = e->ensure_slot_nn();
        if (rc) return rc;
    }
    if (opts && opts->sample) { // keep only the nodes the reference's walk would have found
        const int rc = e->run_sampling(opts);
        if (rc) return rc;
    }
    EvalParams p;
    e->fill_eval_params(&p, opts);
    e->last_key_shift = p.key_shift;
    bool done = false;
    const bool default_opts = !opts || ((opts->emit_mask == 0 || opts->emit_mask == e->cfg.score_mask) && !opts->emit_raw);
    if (e->narrow_batch && default_opts && p.key_shift == 40) {
        NarrowParams np;
        e->fill_narrow_params(&np);
        done = launch_emit_narrow(np, narrow_spec(e), e->stream);
    }
    if (!done) launch_eval_emit(p, e->stream);
    e->emit_was_narrow = done;
    e->launches += done ? narrow_last_emit_launches() : 1;
    FGS_CUDA(e, cudaGetLastError());
    FGS_CUDA(e, cudaEventRecord(e->events[3], e->stream));
    e->emitted = true;
    return FGS_OK;
}

// the compact rows finalize_pods wrote (16 bytes per pod) → the caller's arrays
static void unpack_results(const int64_t* h, int64_t P, int64_t pp, fgs_results* out) {
    // one plain loop per output array (no test inside: they vectorise); 16 bytes per pod in, 20 out
    int32_t* cnt = out->feasible_count;
    uint32_t* err = out->pod_err;
    int32_t* sel = out->selected_node;
    if (cnt)
        for (int64_t p = 0; p < P; ++p) cnt[p] = (int32_t)((uint32_t)(uint64_t)h[p] & 0x0fffffffu);
    if (err)
        for (int64_t p = 0; p < P; ++p) err[p] = (uint32_t)(uint64_t)h[p] >> 28;
    if (sel)
        for (int64_t p = 0; p < P; ++p) sel[p] = (int32_t)(uint32_t)((uint64_t)h[p] >> 32);
    if (out->selected_score) std::memcpy(out->selected_score, h + pp, sizeof(int64_t) * (size_t)P);
}

extern "C" int fgs_fetch_results(fgs_engine* eh, fgs_results* out) {
    if (!eh || !out) return FGS_E_INVALID;
    Engine* e = (Engine*)eh;
    FGS_ENTER(e);
    if (!e->have_pods) return e->fail(FGS_E_STATE, "fgs_fetch_results before an evaluation");
    cudaSetDevice(e->device);
    const int64_t P = e->n_pods, pp = e->p_pad;
    // per-pod epilogue on the GPU (key decode + error bits), then one compact D2H: 4 × int64 per pod
    int64_t* d_compact = (int64_t*)e->d_perpod + (int64_t)kCompactRow0 * pp;
    launch_finalize_pods((const int64_t*)e->d_perpod, d_compact, P, pp, e->last_key_shift, e->emitted ? 1 : 0,
                         (e->cfg.score_mask & FGS_PL_FLEXGPU) ? 1 : 0, e->stream);
    e->launches++;
    FGS_CUDA(e, cudaGetLastError());
    FGS_CUDA(e, cudaMemcpyAsync(e->h_perpod, d_compact, sizeof(int64_t) * 2 * pp, cudaMemcpyDeviceToHost, e->stream));
    if (out->feasible_bitmap)
        FGS_CUDA(e, cudaMemcpyAsync(out->feasible_bitmap, e->d_bitmap, sizeof(uint32_t) * P * e->words_per_row, cudaMemcpyDeviceToHost, e->stream));
    if (out->scores && e->emitted)
        FGS_CUDA(e, cudaMemcpyAsync(out->scores, e->d_scores, sizeof(int64_t) * P * e->score_stride, cudaMemcpyDeviceToHost, e->stream));
    if (e->sampled) // the walks: int2 {start, processed} per pod, then nextStartNodeIndex
        FGS_CUDA(e, cudaMemcpyAsync(e->h_walk, e->d_walk, sizeof(int64_t) * (size_t)(pp + 1), cudaMemcpyDeviceToHost, e->stream));
    if (e->peer) {
        const int rc = e->peer_fetch_error(Engine::kSlots, e->stream);
        if (rc) return rc;
    }
    FGS_CUDA(e, cudaEventRecord(e->events[4], e->stream));
    FGS_CUDA(e, cudaStreamSynchronize(e->stream));
    {
        const int rc = e->peer_check_error(Engine::kSlots); // a merge of this batch timed out: nothing to hand back
        if (rc) return rc;
    }
    out->next_start_node_index = -1;
    if (e->sampled) {
        out->next_start_node_index = ((const int64_t*)e->h_walk)[pp];
        if (out->processed_nodes)
            for (int64_t p = 0; p < P; ++p) out->processed_nodes[p] = ((const int32_t*)e->h_walk)[2 * p + 1];
    } else if (out->processed_nodes) {
        for (int64_t p = 0; p < P; ++p) out->processed_nodes[p] = (int32_t)e->n_nodes;
    }
    const int64_t* h = (const int64_t*)e->h_perpod;
    unpack_results(h, P, pp, out);
    out->words_per_row = e->words_per_row;
    out->score_row_stride = e->score_stride;
    // an event that was never recorded (the batch came through the one-launch path) makes cudaEventElapsedTime fail and
    // leaves the error behind for the next cudaGetLastError: read 0 and clear it
    auto elapsed = [](cudaEvent_t a, cudaEvent_t b) {
        float ms = 0.f;
        if (cudaEventElapsedTime(&ms, a, b) != cudaSuccess) {
            cudaGetLastError();
            ms = 0.f;
        }
        return ms;
    };
    e->tm.h2d_ms = elapsed(e->events[0], e->events[1]);
    e->tm.reduce_ms = elapsed(e->events[1], e->events[2]);
    if (e->emitted) {
        e->tm.emit_ms = elapsed(e->events[2], e->events[3]);
        e->tm.d2h_ms = elapsed(e->events[3], e->events[4]);
    } else {
        e->tm.emit_ms = 0.f;
        e->tm.d2h_ms = elapsed(e->events[2], e->events[4]);
    }
    return FGS_OK;
}

// The reference's call shape — one pod (or a handful) per cycle, sched/scheduler.go:425-455 — in ONE launch: see small_cycle.
// *done = false: the batch does not qualify (too many pods, options, a sharded engine, values outside the int32 range, level
// tables over their memory limit …) and nothing was changed; the caller takes the general path.
int Engine::evaluate_small(const fgs_pod_table* t, const fgs_eval_opts* opts, fgs_results* out, bool* done) {
    Engine* e = this;
    *done = false;
    if (small_off || !t || t->n_pods < 1 || t->n_pods > kSmallMaxPods || peer || nom_q > 0 || in_flight || !have_nodes || filter_mode == 1 || force_wide) return FGS_OK;
    if (t->n_scalars != S) return FGS_OK; // the general path reports it
    if (t->n_pods > 64 && t->n_pods * n_nodes > ((int64_t)4 << 20)) return FGS_OK; // a CTA per pod stops paying: the tiled kernels take over
    if (opts && (opts->sample || opts->skip_pass_b || opts->emit_raw || (opts->emit_mask != 0 && opts->emit_mask != cfg.score_mask))) return FGS_OK;
    cudaSetDevice(device);
    int rc;
    {   // every call that leaves work on the stream waits for it before it returns (or keeps in_flight > 0), so the stream is
        // idle here; only a reallocation below wants the device drained for certain
        const int64_t pp2 = round_up(std::max<int64_t>(t->n_pods, 1), kTileP);
        if (pp2 > p_pad_cap || pp2 > p_plane_cap || n_pad / 32 != words_per_row || n_pad != score_stride) FGS_CUDA(e, cudaStreamSynchronize(stream));
        rc = ensure_pods(t->n_pods);
        if (rc) return rc;
    }
    if (!nodes_derived) {
        rc = derive_now();
        if (rc) return rc;
    }
    // Level tables that describe THIS node table are used; stale ones (the caller's loop changed a node since: bind, delete,
    // NodeInfo update) are not rebuilt for a handful of pods — the kernel compares instead, which costs it one pass over
    // the node columns per pod (from L2) and nothing up front.
    const int64_t P = t->n_pods;
    const bool tables_current = ft && ft->valid && ft->node_gen == node_gen && ft->filter_mask == cfg.filter_mask && ft->n_pad == n_pad &&
                                ft->v.n_dims == 5 + S && ft_rows_fit(ft->v.W);
    // … unless the same node table keeps being asked: from the ninth call on the tables are worth their build.
    // A handful of pods against a long node list go to a cluster of CTAs per pod, which always compares. (More pods than
    // that gain nothing: a GPC places about two clusters of eight at a time, 16 pods took as long as with one CTA each.)
    const bool cluster = small_cluster && small_direct != 0 && filter_mode != 2 && P <= kSmallInlinePods && n_pad >= 2048;
    if (small_gen != node_gen) small_gen = node_gen, small_same = 0;
    else small_same++;
    bool direct = cluster || small_direct == 1 || (small_direct != 0 && filter_mode != 2 && !tables_current && P * n_nodes <= ((int64_t)1 << 19) && small_same < 8);
    if (!direct) {
        rc = resolve_patch(); // the look-up form reads the special list
        if (rc) return rc;
        bool lookup = false;
        rc = ensure_filter_tables(&lookup);
        if (rc) return rc;
        if (!lookup) {
            if (small_direct == 0 || filter_mode == 2 || P * n_nodes > ((int64_t)1 << 21)) return FGS_OK;
            direct = true; // no room for the tables
        }
    }
    const int ncols = pd_cols(S);
    if (!h_small) {
        FGS_CUDA(e, cudaHostAlloc(&h_small, sizeof(int64_t) * ((size_t)(PD_FIXED_COLS + FGS_MAX_SCALARS + 2) * kSmallMaxPods + 8), cudaHostAllocMapped));
        std::memset(h_small, 0, sizeof(int64_t) * ((size_t)(PD_FIXED_COLS + FGS_MAX_SCALARS + 2) * kSmallMaxPods + 8));
        FGS_CUDA(e, cudaMalloc(&d_small_done, 256));
        FGS_CUDA(e, cudaMemset(d_small_done, 0, 256));
    }
    int64_t* hp = (int64_t*)h_small;
    int64_t* hc = hp + (int64_t)(PD_FIXED_COLS + FGS_MAX_SCALARS) * kSmallMaxPods;
    volatile uint64_t* hflag = (volatile uint64_t*)(hc + 2 * kSmallMaxPods);
    const int64_t* src[PD_FIXED_COLS + FGS_MAX_SCALARS] = {nullptr};
    src[PD_GPU_LIM] = t->flex_gpu_limit, src[PD_MEM_LIM] = t->flex_mem_limit, src[PD_REQ_CPU] = t->req_mcpu, src[PD_REQ_MEM] = t->req_mem;
    src[PD_REQ_EPH] = t->req_eph, src[PD_NZ_CPU] = t->nz_mcpu, src[PD_NZ_MEM] = t->nz_mem;
    for (int s2 = 0; s2 < S; ++s2) src[PD_FIXED_COLS + s2] = t->req_scalar + (int64_t)s2 * P;
    stage_present_or = 0;
    for (int c = 0; c < ncols; ++c) {
        uint64_t o = 0;
        int64_t m = 0;
        for (int64_t i = 0; i < P; ++i) {
            int64_t v;
            if (c == PD_META) {
                v = (int64_t)(((uint64_t)t->req_scalar_present[i] << 32) | (uint64_t)t->flags[i]);
                stage_present_or |= t->req_scalar_present[i];
            } else {
                v = src[c][i];
                o |= (uint64_t)v;
                m = v > m ? v : m;
            }
            hp[(int64_t)c * kSmallMaxPods + i] = v;
        }
        stage_or[c] = c == PD_META ? 0 : o;
        stage_max[c] = c == PD_META ? 0 : m;
    }
    n_pods = P;
    rc = decide_narrow(t, false); // shifts for this batch (and the int32 node table when they moved); the pods' int32 values are the kernel's
    if (rc) return rc;
    if (!narrow_batch) return FGS_OK; // the general path stages again and takes the int64 kernels
    EvalParams ep;
    fill_eval_params(&ep, nullptr);
    if (ep.key_shift != 40) return FGS_OK;
    SmallParams sp;
    std::memset(&sp, 0, sizeof(sp));
    if (!direct) sp.v = ft->v;
    sp.direct = direct ? 1 : 0;
    sp.cluster = cluster ? kSmallCluster : 1;
    sp.n_inline = P <= kSmallInlinePods ? (int)P : 0;
    for (int64_t i = 0; i < sp.n_inline; ++i)
        for (int c = 0; c < ncols; ++c) sp.pods_inline[i][c] = hp[(int64_t)c * kSmallMaxPods + i];
    sp.h_pods = hp;
    sp.h_compact = hc;
    sp.d_pods = (int64_t*)d_pods;
    sp.d_pn = (int*)d_pn;
    sp.nn = (const int*)d_nn;
    sp.nd = (const int64_t*)d_nd;
    sp.perpod = (int64_t*)d_perpod;
    sp.bitmap = (uint32_t*)d_bitmap;
    sp.scores = (int64_t*)d_scores;
    sp.n_pods = P;
    sp.p_pad = p_pad;
    sp.n_nodes = n_nodes;
    sp.n_pad = n_pad;
    sp.node_base = node_base;
    sp.W = words_per_row;
    sp.score_stride = score_stride;
    sp.S = S;
    sp.fit_ignored = cfg.fit_ignored_scalar_mask;
    sp.filter_mask = cfg.filter_mask;
    sp.score_mask = cfg.score_mask;
    sp.flex_first = ep.flex_filter_first;
    sp.fit_most = cfg.fit_strategy == 1;
    sp.w_eq = ep.fit_w_cpu == ep.fit_w_mem;
    sp.w_flex = (int)ep.w_flex, sp.w_fit = (int)ep.w_fit, sp.w_alloc = (int)ep.w_alloc;
    sp.fit_w_cpu = ep.fit_w_cpu, sp.fit_w_mem = ep.fit_w_mem;
    sp.fit_inv_wsum = ep.fit_inv_wsum;
    for (int d2 = 0; d2 < ND_DIMS; ++d2) sp.sh[d2] = narrow_sh.s[d2];
    sp.seed = seed;
    sp.special = (const int64_t*)d_special;
    sp.n_special = node_an_valid ? node_an.n_special : 0;
    sp.d_done = (unsigned int*)d_small_done;
    sp.h_flag = hflag;
    sp.seq = ++small_seq;
    sp.spec_flag = patch_spec ? d_patch_flag : nullptr;
    FGS_CUDA(e, ft_launch_small(sp, stream));
    launches++;
    if (out->feasible_bitmap || out->scores) { // planes to the host: copies behind the kernel, then the stream
        if (out->feasible_bitmap)
            FGS_CUDA(e, cudaMemcpyAsync(out->feasible_bitmap, d_bitmap, sizeof(uint32_t) * P * words_per_row, cudaMemcpyDeviceToHost, stream));
        if (out->scores) FGS_CUDA(e, cudaMemcpyAsync(out->scores, d_scores, sizeof(int64_t) * P * score_stride, cudaMemcpyDeviceToHost, stream));
        FGS_CUDA(e, cudaStreamSynchronize(stream));
    } else { // per-pod results only: watch the completion word the last CTA stores (a few µs sooner than the stream reports)
        const auto t0 = std::chrono::steady_clock::now();
        for (unsigned spin = 1; (*hflag & ~kSmallDeclined) != sp.seq; ++spin) {
            if ((spin & 0xfff) == 0 && std::chrono::steady_clock::now() - t0 > std::chrono::milliseconds(20)) {
                FGS_CUDA(e, cudaStreamSynchronize(stream)); // something is wrong (or very slow): let the runtime say what
                if ((*hflag & ~kSmallDeclined) != sp.seq) return fail(FGS_E_CUDA, "small-batch kernel finished without reporting");
                break;
            }
        }
    }
    if (*hflag & kSmallDeclined) { // a patched row changed the node analysis under this launch: settle it and go again
        rc = resolve_patch();
        if (rc) return rc;
        if (patch_spec) return fail(FGS_E_STATE, "small-batch kernel declined without a raised patch flag");
        return evaluate_small(t, opts, out, done);
    }
    patch_spec = false; // the launch ran behind the patches and found the flag down
    unpack_results(hc, P, kSmallMaxPods, out);
    if (out->processed_nodes)
        for (int64_t i = 0; i < P; ++i) out->processed_nodes[i] = (int32_t)n_nodes;
    out->words_per_row = words_per_row;
    out->score_row_stride = score_stride;
    out->next_start_node_index = -1;
    staged_gen = node_gen;
    have_pods = true, reduced = true, emitted = true, sampled = false, masked = false, was_fused = false, staged_by_dma = false;
    reduce_was_lookup = !direct, reduce_was_narrow = false, emit_was_narrow = true, was_small = true;
    last_key_shift = 40;
    tm.h2d_ms = tm.reduce_ms = tm.emit_ms = tm.d2h_ms = 0.f;
    *done = true;
    return FGS_OK;
}

extern "C" int fgs_evaluate(fgs_engine* eh, const fgs_pod_table* pods, const fgs_eval_opts* opts, fgs_results* out) {
    if (!eh) return FGS_E_INVALID;
    Engine* e = (Engine*)eh;
    FGS_ENTER(e);
    e->was_small = false;
    if (out) {
        bool done = false;
        const int rc0 = e->evaluate_small(pods, opts, out, &done);
        if (rc0) return rc0;
        if (done) return FGS_OK;
    }
    int rc = fgs_stage_pods(eh, pods);
    if (rc) return rc;
    e->emitted = false;
    if (e->peer && opts && opts->sample) return e->fail(FGS_E_INVALID, "fgs_evaluate: node sampling walks ONE node list (single engine only)");
    e->was_fused = false;
    if (e->fused_eligible(opts)) { // Filter and Score in one pass over pod × node
        rc = e->run_fused(opts, 3);
        if (rc) return rc;
        return fgs_fetch_results(eh, out);
    }
    rc = fgs_pass_reduce(eh);
    if (rc) return rc;
    if (e->peer) { // node-sharded engine: the other shards' maxima and counts
        rc = e->peer_merge(0, e->stream);
        if (rc) return rc;
    }
    if (!(opts && opts->skip_pass_b)) {
        rc = fgs_pass_emit(eh, opts);
        if (rc) return rc;
        if (e->peer) { // … and their selectHost keys
            rc = e->peer_merge(1, e->stream);
            if (rc) return rc;
        }
    } else if (opts->sample) { // Filter only, truncated as the reference's walk would
        rc = e->run_sampling(opts);
        if (rc) return rc;
    }
    return fgs_fetch_results(eh, out);
}

// RunScorePlugins under the framework's feasible list (generic_scheduler.go:420-426, runtime/framework.go:861-933): the
// bitmap pass A left is ANDed with the caller's mask, NormalizeScore's reductions are redone over what is left (feasible
// counts, max FlexGPU score, Allocatable hi / lo, the special-node rows), the shards are merged, and pass B scores and
// selects over that list.
extern "C" int fgs_score_masked(fgs_engine* eh, const uint32_t* mask, int32_t mask_on_device, const fgs_eval_opts* opts, fgs_results* out) {
    if (!eh || !out) return FGS_E_INVALID;
    Engine* e = (Engine*)eh;
    FGS_ENTER(e);
    if (!mask) return e->fail(FGS_E_INVALID, "fgs_score_masked: null mask");
    if (e->in_flight) return e->fail(FGS_E_STATE, "fgs_score_masked while submitted batches are in flight: fgs_collect them first");
    if (!e->have_pods || !e->reduced)
        return e->fail(FGS_E_STATE, "fgs_score_masked: no filtered batch — run fgs_evaluate with skip_pass_b (or fgs_stage_pods + fgs_pass_reduce) first");
    if (e->staged_gen != e->node_gen)
        return e->fail(FGS_E_STATE, "fgs_score_masked: the node table changed since the batch was staged (its int32 columns follow the node table): stage it again");
    if (e->was_fused) return e->fail(FGS_E_STATE, "fgs_score_masked after the fused pass: its bitmap has been scored already; stage the batch again");
    if (opts && (opts->sample || opts->skip_pass_b)) return e->fail(FGS_E_INVALID, "fgs_score_masked: sample / skip_pass_b make no sense here");
    cudaSetDevice(e->device);
    const int64_t P = e->n_pods, W = e->words_per_row;
    if (P > 0) {
        const uint32_t* d_m = mask;
        if (!mask_on_device) {
            const size_t need = sizeof(uint32_t) * (size_t)P * (size_t)W;
            if (need > e->mask_bytes) {
                FGS_CUDA(e, cudaStreamSynchronize(e->stream));
                if (e->d_mask) cudaFree(e->d_mask);
                e->d_mask = nullptr;
                e->mask_bytes = 0;
                FGS_CUDA(e, cudaMalloc(&e->d_mask, need));
                e->mask_bytes = need;
            }
            FGS_CUDA(e, cudaMemcpyAsync(e->d_mask, mask, need, cudaMemcpyHostToDevice, e->stream));
            d_m = (const uint32_t*)e->d_mask;
        }
        if (!ft_rows_fit(W)) return e->fail(FGS_E_TOO_LARGE, "fgs_score_masked: bitmap rows beyond the kernel's shared memory (1.6 M nodes per engine)");
        FtMaskParams mp;
        std::memset(&mp, 0, sizeof(mp));
        const bool orders = e->ft && e->ft->valid && e->ft->node_gen == e->node_gen && e->ft->n_pad == e->n_pad;
        if (orders) mp.v = e->ft->v;
        mp.have_orders = orders ? 1 : 0;
        mp.pods = (const int64_t*)e->d_pods;
        mp.nd = (const int64_t*)e->d_nd;
        mp.n_pods = P;
        mp.p_pad = e->p_pad;
        mp.n_nodes = e->n_nodes;
        mp.n_pad = e->n_pad;
        mp.W = W;
        mp.bitmap = (uint32_t*)e->d_bitmap;
        mp.mask = d_m;
        mp.reduce_max = (int64_t*)e->d_perpod;
        mp.reduce_sum = (int64_t*)e->d_perpod + (int64_t)FGS_REDUCE_ROWS * e->p_pad;
        mp.select_key = (int64_t*)e->d_perpod + (int64_t)(FGS_REDUCE_ROWS + 1) * e->p_pad;
        mp.sc_flex = (e->cfg.score_mask & FGS_PL_FLEXGPU) ? 1 : 0;
        mp.sc_alloc = (e->cfg.score_mask & FGS_PL_ALLOCATABLE) ? 1 : 0;
        FGS_CUDA(e, ft_launch_masked(mp, e->stream));
        e->launches++;
        if ((e->cfg.score_mask & FGS_PL_FLEXGPU) && e->node_an_valid && e->node_an.n_special > 0) {
            launch_special_pass((const uint32_t*)e->d_bitmap, e->words_per_row, (const int64_t*)e->d_special, e->n_pad, e->node_an.n_special,
                                (const int64_t*)e->d_pods, (int64_t*)e->d_perpod, e->n_pods, e->p_pad, e->stream);
            e->launches++;
            FGS_CUDA(e, cudaGetLastError());
        }
    }
    e->emitted = false;
    e->masked = true; // the int64 emit kernel recomputes Filter: it must AND the bitmap in, as after node sampling
    int rc;
    if (e->peer) { // node-sharded: the other shards' maxima and counts over THEIR part of the list
        rc = e->peer_merge(0, e->stream);
        if (rc) return rc;
    }
    {
        fgs_eval_opts o;
        std::memset(&o, 0, sizeof(o));
        if (opts) o = *opts;
        o.sample = 0;
        rc = fgs_pass_emit(eh, &o);
        if (rc) return rc;
    }
    if (e->peer) {
        rc = e->peer_merge(1, e->stream);
        if (rc) return rc;
    }
    return fgs_fetch_results(eh, out);
}

extern "C" int fgs_get_device_view(fgs_engine* eh, fgs_device_view* v) {
    if (!eh || !v) return FGS_E_INVALID;
    Engine* e = (Engine*)eh;
    FGS_ENTER(e);
    if (!e->have_pods) return e->fail(FGS_E_STATE, "fgs_get_device_view before fgs_stage_pods");
    v->reduce_max = e->d_perpod;
    v->reduce_sum = (int64_t*)e->d_perpod + (int64_t)FGS_REDUCE_ROWS * e->p_pad;
    v->select_key = (int64_t*)e->d_perpod + (int64_t)(FGS_REDUCE_ROWS + 1) * e->p_pad;
    v->feasible_bitmap = e->d_bitmap;
    v->scores = e->d_scores;
    v->p_pad = e->p_pad;
    v->words_per_row = e->words_per_row;
    v->score_row_stride = e->score_stride;
    v->stream = (void*)e->stream;
    return FGS_OK;
}

extern "C" int fgs_synchronize(fgs_engine* eh) {
    if (!eh) return FGS_E_INVALID;
    Engine* e = (Engine*)eh;
    FGS_ENTER(e);
    cudaSetDevice(e->device);
    FGS_CUDA(e, cudaStreamSynchronize(e->stream));
    return FGS_OK;
}

extern "C" int fgs_get_timings(fgs_engine* eh, fgs_timings* t) {
    if (!eh || !t) return FGS_E_INVALID;
    Engine* e = (Engine*)eh;
    *t = e->tm;
    t->kernel_launches = e->launches;
    t->reduce_int32 = e->reduce_was_narrow ? 1 : 0;
    t->emit_int32 = e->emit_was_narrow ? 1 : 0;
    t->fused = e->was_fused ? 1 : 0;
    t->reduce_lookup = e->reduce_was_lookup ? 1 : 0;
    t->small = e->was_small ? 1 : 0;
    t->node_patches = (int64_t)e->patches;
    t->node_patch_redos = (int64_t)e->patch_redos;
    return FGS_OK;
}

// =================================================================================================
// fgs_submit / fgs_collect — the same evaluation with two batches in flight
// =================================================================================================
// A stream of batches through fgs_evaluate serialises, per batch, the host-side staging, the H2D copy, the two
// passes, the D2H of the results and their unpacking. Here the per-batch buffers exist twice (slots): while the
// kernels of batch k run, the host already stages batch k+1 into the other slot, its H2D copy runs on a copy
// stream, and batch k-1's results are unpacked. The bitmap / score planes are shared (the kernels of consecutive
// batches are ordered on the compute stream), so a submitted batch returns per-pod results only.

void Engine::bind_slot(int k) {
    Slot& o = slots[cur_slot];
    o.d_pods = d_pods, o.h_pods = h_pods, o.d_perpod = d_perpod, o.h_perpod = h_perpod, o.d_pn = d_pn;
    o.p_pad_cap = p_pad_cap, o.pn_cap = pn_cap, o.n_pods = n_pods, o.p_pad = p_pad;
    o.narrow_batch = narrow_batch, o.emitted = emitted, o.sampled = sampled, o.have_pods = have_pods, o.key_shift = last_key_shift;
    o.sh = narrow_sh, o.nn_sig = narrow_nodes_sig;
    const Slot& n = slots[k];
    d_pods = n.d_pods, h_pods = n.h_pods, d_perpod = n.d_perpod, h_perpod = n.h_perpod, d_pn = n.d_pn;
    p_pad_cap = n.p_pad_cap, pn_cap = n.pn_cap, n_pods = n.n_pods, p_pad = n.p_pad;
    narrow_batch = n.narrow_batch, emitted = n.emitted, sampled = n.sampled, have_pods = n.have_pods, last_key_shift = n.key_shift;
    cur_slot = k;
    // the int32 node table that goes with the slot's batch: its ring entry if that still holds it (always, while the batch is in
    // flight); a resident batch whose entry was recycled gets a new one from use_nn before its passes run (ensure_slot_nn)
    narrow_sh = n.sh, narrow_nodes_sig = n.nn_sig;
    nn_cur = -1, d_nn = nullptr;
    if (n.have_pods && n.narrow_batch && n.nn_idx >= 0) {
        const NnEntry& en = nn_ring[n.nn_idx];
        if (en.sig == n.nn_sig && en.sig >= 0 && en.gen == node_gen && std::memcmp(&en.sh, &n.sh, sizeof(n.sh)) == 0) nn_cur = n.nn_idx, d_nn = en.d;
    }
}

int Engine::ensure_slot_nn() {
    if (!have_pods || !narrow_batch || (nn_cur >= 0 && d_nn)) return FGS_OK;
    if (narrow_nodes_sig < 0) return fail(FGS_E_STATE, "the staged batch has no int32 node table: stage it again");
    const NarrowShifts sh = narrow_sh;
    return use_nn(sh, narrow_nodes_sig);
}

static int ensure_pipeline(Engine* e) {
    if (!e->copy_stream) {
        FGS_CUDA(e, cudaStreamCreateWithFlags(&e->copy_stream, cudaStreamNonBlocking));
        FGS_CUDA(e, cudaStreamCreateWithFlags(&e->d2h_stream, cudaStreamNonBlocking));
        FGS_CUDA(e, cudaEventCreateWithFlags(&e->ev_h2d, cudaEventDisableTiming));
        FGS_CUDA(e, cudaEventCreateWithFlags(&e->ev_kernels, cudaEventDisableTiming));
        for (auto& sl : e->slots) FGS_CUDA(e, cudaEventCreateWithFlags(&sl.ev_done, e->pipe_trace ? cudaEventDefault : cudaEventDisableTiming));
    }
    if (e->peer && !e->merge_stream) {
        int lo = 0, hi = 0;
        FGS_CUDA(e, cudaDeviceGetStreamPriorityRange(&lo, &hi));
        FGS_CUDA(e, cudaStreamCreateWithPriority(&e->merge_stream, cudaStreamNonBlocking, hi)); // its few CTAs go first
        for (auto& sl : e->slots)
            for (cudaEvent_t* ev : {&sl.ev_reduce, &sl.ev_merge1, &sl.ev_emit, &sl.ev_merge2, &sl.ev_t0, &sl.ev_emit0})
                FGS_CUDA(e, cudaEventCreateWithFlags(ev, e->pipe_trace ? cudaEventDefault : cudaEventDisableTiming));
    }
    return FGS_OK;
}

namespace {
// every other batch of the sharded pipeline evaluates into the second bitmap
struct BitmapFor {
    Engine* e;
    void* saved;
    BitmapFor(Engine* e_, int sel) : e(e_), saved(e_->d_bitmap) {
        if (sel == 1 && e->d_bitmap_alt) e->d_bitmap = e->d_bitmap_alt;
        if (sel == 2 && e->d_bitmap_alt2) e->d_bitmap = e->d_bitmap_alt2;
    }
    ~BitmapFor() { e->d_bitmap = saved; }
};
} // namespace

int Engine::ensure_front_stream() {
    if (front_stream) return FGS_OK;
    int lo = 0, hi = 0;
    FGS_CUDA(this, cudaDeviceGetStreamPriorityRange(&lo, &hi));
    FGS_CUDA(this, cudaStreamCreateWithPriority(&front_stream, cudaStreamNonBlocking, hi));
    for (cudaEvent_t* ev : {&ev_front, &ev_back, &ev_plane[0], &ev_plane[1], &ev_plane[2]}) FGS_CUDA(this, cudaEventCreateWithFlags(ev, cudaEventDisableTiming));
    return FGS_OK;
}

namespace {
// `stream` names the front stream while a submitted batch is staged and filtered (every launch in between follows it)
struct FrontWindow {
    Engine* e;
    cudaStream_t back;
    bool open;
    FrontWindow(Engine* e_, bool on) : e(e_), back(e_->stream), open(on) {
        if (open) {
            e->window_back = back;
            e->stream = e->front_stream;
        }
    }
    void close() {
        if (open) {
            e->stream = back;
            e->window_back = nullptr;
            open = false;
        }
    }
    ~FrontWindow() { close(); }
};
} // namespace

// The sharded pipeline (a peer group is attached): per batch k
//     compute stream   reduce(k)              emit(k-1)               reduce(k+1)   …
//     merge stream            merge1(k)              merge2(k-1), epilogue(k-1)   merge1(k+1) …
// so every cross-shard merge overlaps a pass of the other batch. The emit pass of batch k is therefore issued by the
// NEXT submit (after that batch's reduce pass) or, when nothing follows, by fgs_collect. Up to THREE batches are in
// flight here (k-1 being read back and unpacked, k waiting for its emit pass, k+1 just submitted): with two, the
// compute stream would run dry while the host collects. The bitmap exists twice (the emit pass of batch k-1 reads its
// rows while the reduce pass of batch k writes). Every rank must make the same sequence of submit / collect calls:
// the merge kernels meet their peers in issue order.
int Engine::issue_emit_sharded(int j) {
    const int cur = cur_slot;
    if (j != cur) bind_slot(j);
    Slot& sl = slots[j];
    auto body = [&]() -> int {
        Engine* e = this;
        FGS_CUDA(e, cudaStreamWaitEvent(stream, sl.ev_merge1, 0));
        if (pipe_trace) FGS_CUDA(e, cudaEventRecord(sl.ev_emit0, stream));
        {
            BitmapFor bm(this, sl.bitmap_sel);
            const int rc = fgs_pass_emit((fgs_engine*)this, sl.have_opts ? &sl.opts : nullptr);
            if (rc) return rc;
        }
        FGS_CUDA(e, cudaEventRecord(sl.ev_emit, stream));
        if (ev_plane[sl.bitmap_sel]) {
            FGS_CUDA(e, cudaEventRecord(ev_plane[sl.bitmap_sel], stream));
            plane_used[sl.bitmap_sel] = true;
        }
        FGS_CUDA(e, cudaStreamWaitEvent(merge_stream, sl.ev_emit, 0));
        const int rc = peer_merge(1, merge_stream, true);
        if (rc) return rc;
        int64_t* d_compact = (int64_t*)d_perpod + (int64_t)kCompactRow0 * p_pad;
        launch_finalize_pods((const int64_t*)d_perpod, d_compact, n_pods, p_pad, last_key_shift, 1, (cfg.score_mask & FGS_PL_FLEXGPU) ? 1 : 0, merge_stream);
        launches++;
        FGS_CUDA(e, cudaGetLastError());
        FGS_CUDA(e, cudaEventRecord(sl.ev_merge2, merge_stream));
        FGS_CUDA(e, cudaStreamWaitEvent(d2h_stream, sl.ev_merge2, 0));
        FGS_CUDA(e, cudaMemcpyAsync(h_perpod, d_compact, sizeof(int64_t) * 2 * p_pad, cudaMemcpyDeviceToHost, d2h_stream));
        {
            const int rc2 = peer_fetch_error(j, d2h_stream);
            if (rc2) return rc2;
        }
        FGS_CUDA(e, cudaEventRecord(sl.ev_done, d2h_stream));
        sl.emit_pending = false;
        return FGS_OK;
    };
    const int rc = body();
    if (j != cur) bind_slot(cur);
    return rc;
}

int Engine::submit_sharded(const fgs_pod_table* pods, const fgs_eval_opts* opts, bool resident) {
    Engine* e = this;
    const int k = in_flight == 0 ? (next_slot >= 0 && next_slot < ring() ? next_slot : cur_slot) : (oldest_slot + in_flight) % ring();
    const int newest = in_flight == 0 ? -1 : (oldest_slot + in_flight - 1) % ring();
    if (k != cur_slot) bind_slot(k);
    if (in_flight == 0) oldest_slot = k;
    next_slot = (k + 1) % ring();
    // Pass A of this batch goes to the front stream, beside the emit passes of the two batches before it (one running or
    // queued, one issued at the end of this call): three bitmaps are live then.
    const bool overlap = overlap_filter && nom_q == 0;
    if (overlap) {
        int rc = ensure_front_stream();
        if (rc) return rc;
        rc = nodes_ready();
        if (rc) return rc;
        if (in_flight == 0) {
            FGS_CUDA(e, cudaEventRecord(ev_back, stream));
            FGS_CUDA(e, cudaStreamWaitEvent(front_stream, ev_back, 0));
        }
    }
    FrontWindow win(this, overlap);
    if (resident) {
        if (!have_pods) return fail(FGS_E_STATE, "fgs_resubmit: no batch is resident in this pipeline slot (fgs_submit fills the slots in turn)");
        if (slots[k].node_gen != node_gen) return fail(FGS_E_STATE, "fgs_resubmit: the node table changed since this batch was staged: fgs_submit it again");
    } else {
        // A batch larger than any before makes ensure_pods replace the bitmap / score planes, which the emit pass of the
        // previous batch has still to read: issue that pass first and let everything in flight drain.
        if (pods && in_flight > 0 && round_up(std::max<int64_t>(pods->n_pods, 1), kTileP) > p_plane_cap) {
            if (newest >= 0 && slots[newest].emit_pending) {
                win.close();
                const int rc = issue_emit_sharded(newest);
                if (rc) return rc;
            }
            FGS_CUDA(e, cudaDeviceSynchronize());
            if (overlap && win.open == false) {
                window_back = stream;
                stream = front_stream;
                win.open = true;
            }
        }
        const int rc = stage_pods_impl(e, pods, true);
        if (rc) return rc;
    }
    emitted = false;
    slots[k].node_gen = node_gen;
    const int n_bitmaps = overlap ? 3 : 2;
    slots[k].bitmap_sel = (int)(submit_seq++ % (uint64_t)n_bitmaps);
    for (int which = 1; which <= 2; ++which) { // the second / third bitmap, as large as the first
        if (slots[k].bitmap_sel != which) continue;
        void*& buf = which == 1 ? d_bitmap_alt : d_bitmap_alt2;
        size_t& have = which == 1 ? bitmap_alt_bytes : bitmap_alt2_bytes;
        const size_t need = sizeof(uint32_t) * (size_t)p_plane_cap * (size_t)words_per_row;
        if (need > have) {
            FGS_CUDA(e, cudaDeviceSynchronize());
            if (buf) cudaFree(buf);
            buf = nullptr;
            have = 0;
            FGS_CUDA(e, cudaMalloc(&buf, need));
            have = need;
        }
    }
    Slot& sl = slots[k];
    if (pipe_trace) {
        if (!trace_epoch) {
            FGS_CUDA(e, cudaEventCreate(&trace_epoch));
            FGS_CUDA(e, cudaEventRecord(trace_epoch, stream));
        }
        FGS_CUDA(e, cudaEventRecord(sl.ev_t0, stream));
    }
    if (overlap && plane_used[sl.bitmap_sel]) FGS_CUDA(e, cudaStreamWaitEvent(stream, ev_plane[sl.bitmap_sel], 0)); // the emit pass that read this bitmap last
    {
        BitmapFor bm(this, sl.bitmap_sel);
        const int rc = fgs_pass_reduce((fgs_engine*)this);
        if (rc) return rc;
    }
    FGS_CUDA(e, cudaEventRecord(sl.ev_reduce, stream));
    win.close();
    FGS_CUDA(e, cudaStreamWaitEvent(merge_stream, sl.ev_reduce, 0));
    {
        const int rc = peer_merge(0, merge_stream, true);
        if (rc) return rc;
    }
    FGS_CUDA(e, cudaEventRecord(sl.ev_merge1, merge_stream));
    sl.emit_pending = true;
    sl.have_opts = opts != nullptr;
    if (opts) sl.opts = *opts;
    sl.busy = true;
    in_flight++;
    if (newest >= 0 && slots[newest].emit_pending) return issue_emit_sharded(newest);
    return FGS_OK;
}

extern "C" int fgs_resubmit(fgs_engine* eh, const fgs_eval_opts* opts) {
    if (!eh) return FGS_E_INVALID;
    Engine* e = (Engine*)eh;
    FGS_ENTER(e);
    if (e->in_flight >= e->max_in_flight())
        return e->fail(FGS_E_STATE, "fgs_resubmit: " + std::to_string(e->max_in_flight()) + " batches are in flight already; fgs_collect one first");
    if (e->nom_q > 0) return e->fail(FGS_E_STATE, "fgs_resubmit: nominated pods are set: use fgs_evaluate");
    if (opts && (opts->sample || opts->skip_pass_b)) return e->fail(FGS_E_INVALID, "fgs_resubmit: sample / skip_pass_b need fgs_evaluate");
    cudaSetDevice(e->device);
    const int rc = ensure_pipeline(e);
    if (rc) return rc;
    if (e->peer) return e->submit_sharded(nullptr, opts, true);
    return e->submit_single(nullptr, opts, true);
}

// More pipeline slots than batches in flight: every slot keeps the batch it was last handed resident on the device
// (pod tables, int32 tables, per-pod rows), so that fgs_resubmit can walk a whole job of n batches again and again
// without staging. 0 restores the default (as many slots as batches in flight).
extern "C" int fgs_set_resident_batches(fgs_engine* eh, int32_t n) {
    if (!eh) return FGS_E_INVALID;
    Engine* e = (Engine*)eh;
    FGS_ENTER(e);
    if (e->in_flight) return e->fail(FGS_E_STATE, "fgs_set_resident_batches while submitted batches are in flight: fgs_collect them first");
    if (n != 0 && (n < 3 || n > Engine::kSlots)) return e->fail(FGS_E_TOO_LARGE, "fgs_set_resident_batches: 0 (default) or 3.." + std::to_string(Engine::kSlots));
    e->resident_cfg = n;
    e->next_slot = -1;
    if (e->cur_slot >= e->ring()) e->bind_slot(0);
    return FGS_OK;
}

extern "C" int fgs_max_in_flight(fgs_engine* eh) {
    if (!eh) return FGS_E_INVALID;
    return ((Engine*)eh)->max_in_flight();
}

extern "C" int fgs_submit(fgs_engine* eh, const fgs_pod_table* pods, const fgs_eval_opts* opts) {
    if (!eh) return FGS_E_INVALID;
    Engine* e = (Engine*)eh;
    FGS_ENTER(e);
    if (e->in_flight >= e->max_in_flight())
        return e->fail(FGS_E_STATE, "fgs_submit: " + std::to_string(e->max_in_flight()) + " batches are in flight already; fgs_collect one first");
    if (opts && (opts->sample || opts->skip_pass_b)) return e->fail(FGS_E_INVALID, "fgs_submit: sample / skip_pass_b need fgs_evaluate");
    if (e->nom_q > 0) return e->fail(FGS_E_STATE, "fgs_submit: nominated pods are set (their per-batch columns are not double-buffered): use fgs_evaluate");
    cudaSetDevice(e->device);
    {
        const int rc = ensure_pipeline(e);
        if (rc) return rc;
    }
    if (e->peer) return e->submit_sharded(pods, opts, false);
    return e->submit_single(pods, opts, false);
}

// Single engine, per batch k:
//     front stream (high priority)   stage kernels(k), pass A(k)              stage kernels(k+1), pass A(k+1) …
//     compute stream                          wait A(k) → emit(k), finalize(k)          wait A(k+1) → emit(k+1) …
// Pass A is bound by L2 look-ups and instruction issue, the emit pass by HBM writes: side by side the first mostly hides
// behind the second. Consecutive batches filter into two bitmaps in turn (A(k+2) waits for emit(k) to have read its one);
// everything else a batch touches belongs to its pipeline slot or is written by one stream only.
int Engine::submit_single(const fgs_pod_table* pods, const fgs_eval_opts* opts, bool resident) {
    Engine* e = this;
    fgs_engine* eh = (fgs_engine*)this;
    // the next slot of the ring (the ring is walked round-robin, so a job of n batches keeps its slots)
    const int k = e->in_flight == 0 ? (e->next_slot >= 0 && e->next_slot < e->ring() ? e->next_slot : e->cur_slot) : (e->oldest_slot + e->in_flight) % e->ring();
    if (k != e->cur_slot) e->bind_slot(k);
    if (e->in_flight == 0) e->oldest_slot = k;
    e->next_slot = (k + 1) % e->ring();
    int rc;
    const bool overlap = e->overlap_filter && e->no_fused && e->nom_q == 0 && !(opts && (opts->sample || opts->skip_pass_b));
    if (overlap) {
        rc = e->ensure_front_stream();
        if (rc) return rc;
        rc = e->nodes_ready(); // a pending row patch is settled on the compute stream, before the window
        if (rc) return rc;
        if (e->in_flight == 0) { // whatever ran on the compute stream so far (uploads, patches, other calls) comes first
            FGS_CUDA(e, cudaEventRecord(e->ev_back, e->stream));
            FGS_CUDA(e, cudaStreamWaitEvent(e->front_stream, e->ev_back, 0));
        }
    }
    FrontWindow win(e, overlap);
    if (resident) {
        if (!e->have_pods) return fail(FGS_E_STATE, "fgs_resubmit: no batch is resident in this pipeline slot (fgs_submit fills the slots in turn)");
        if (e->slots[k].node_gen != e->node_gen) return fail(FGS_E_STATE, "fgs_resubmit: the node table changed since this batch was staged: fgs_submit it again");
        FGS_CUDA(e, cudaEventRecord(e->events[0], e->stream));
        FGS_CUDA(e, cudaEventRecord(e->events[1], e->stream));
    } else {
        rc = stage_pods_impl(e, pods, true);
        if (rc) return rc;
        e->slots[k].node_gen = e->node_gen;
    }
    e->emitted = false;
    e->was_fused = false;
    int sel = 0;
    if (overlap) {
        sel = (int)(e->submit_seq++ & 1);
        if (sel == 1) { // the second bitmap, as large as the first
            const size_t need = sizeof(uint32_t) * (size_t)e->p_plane_cap * (size_t)e->words_per_row;
            if (need > e->bitmap_alt_bytes) {
                FGS_CUDA(e, cudaDeviceSynchronize());
                if (e->d_bitmap_alt) cudaFree(e->d_bitmap_alt);
                e->d_bitmap_alt = nullptr;
                e->bitmap_alt_bytes = 0;
                FGS_CUDA(e, cudaMalloc(&e->d_bitmap_alt, need));
                e->bitmap_alt_bytes = need;
            }
        }
        if (e->plane_used[sel]) FGS_CUDA(e, cudaStreamWaitEvent(e->stream, e->ev_plane[sel], 0)); // the emit pass that read this bitmap last
    }
    if (!overlap && e->fused_eligible(opts)) {
        rc = e->run_fused(opts, 3);
        if (rc) return rc;
    } else {
        {
            BitmapFor bf(e, sel);
            rc = fgs_pass_reduce(eh);
        }
        if (rc) return rc;
        if (overlap) {
            FGS_CUDA(e, cudaEventRecord(e->ev_front, e->stream));
            win.close();
            FGS_CUDA(e, cudaStreamWaitEvent(e->stream, e->ev_front, 0));
        }
        {
            BitmapFor bf(e, sel);
            rc = fgs_pass_emit(eh, opts);
        }
        if (rc) return rc;
    }
    const int64_t P = e->n_pods, pp = e->p_pad;
    int64_t* d_compact = (int64_t*)e->d_perpod + (int64_t)kCompactRow0 * pp;
    launch_finalize_pods((const int64_t*)e->d_perpod, d_compact, P, pp, e->last_key_shift, 1, (e->cfg.score_mask & FGS_PL_FLEXGPU) ? 1 : 0, e->stream);
    e->launches++;
    FGS_CUDA(e, cudaGetLastError());
    if (overlap) {
        FGS_CUDA(e, cudaEventRecord(e->ev_plane[sel], e->stream));
        e->plane_used[sel] = true;
    }
    // the D2H of the per-pod results runs on its own stream: the next batch's kernels do not wait for it
    FGS_CUDA(e, cudaEventRecord(e->ev_kernels, e->stream));
    FGS_CUDA(e, cudaStreamWaitEvent(e->d2h_stream, e->ev_kernels, 0));
    FGS_CUDA(e, cudaMemcpyAsync(e->h_perpod, d_compact, sizeof(int64_t) * 2 * pp, cudaMemcpyDeviceToHost, e->d2h_stream));
    FGS_CUDA(e, cudaEventRecord(e->slots[k].ev_done, e->d2h_stream));
    e->slots[k].busy = true;
    e->in_flight++;
    return FGS_OK;
}

extern "C" int fgs_collect(fgs_engine* eh, fgs_results* out) {
    if (!eh || !out) return FGS_E_INVALID;
    Engine* e = (Engine*)eh;
    FGS_ENTER(e);
    if (e->in_flight == 0) return e->fail(FGS_E_STATE, "fgs_collect: nothing was submitted");
    cudaSetDevice(e->device);
    const int k = e->oldest_slot;
    if (e->slots[k].emit_pending) { // sharded pipeline, nothing was submitted after this batch: finish it now
        const int rc = e->issue_emit_sharded(k);
        if (rc) return rc;
    }
    FGS_CUDA(e, cudaEventSynchronize(e->slots[k].ev_done));
    if (e->pipe_trace && e->peer) e->trace_collect(k);
    if (e->peer) { // a merge of this batch timed out: drop the batch, report
        const int rc = e->peer_check_error(k);
        if (rc) {
            e->slots[k].busy = false;
            e->in_flight--;
            e->oldest_slot = (k + 1) % e->ring();
            return rc;
        }
    }
    // the oldest batch's host buffer: the members alias slot cur_slot, the other slot keeps its own copy
    const bool is_cur = k == e->cur_slot;
    const int64_t P = is_cur ? e->n_pods : e->slots[k].n_pods, pp = is_cur ? e->p_pad : e->slots[k].p_pad;
    const int64_t* h = (const int64_t*)(is_cur ? e->h_perpod : e->slots[k].h_perpod);
    unpack_results(h, P, pp, out);
    if (out->processed_nodes)
        for (int64_t p = 0; p < P; ++p) out->processed_nodes[p] = (int32_t)e->n_nodes;
    out->words_per_row = e->words_per_row;
    out->score_row_stride = e->score_stride;
    out->next_start_node_index = -1;
    e->slots[k].busy = false;
    e->in_flight--;
    e->oldest_slot = (k + 1) % e->ring();
    return FGS_OK;
}
