// engine.h — the object behind the opaque fgs_engine handle.
#pragma once
#include <cuda_runtime.h>

#include <atomic>
#include <string>
#include <vector>

#include "../../include/fgs.h"
#include "kernels.cuh"
#include "kernels_narrow.cuh"
#include "pack.h"

namespace fgs {

// per-pod device block: reduce_max[FGS_REDUCE_ROWS] | reduce_sum | select_key[3] | compact results[4]
constexpr int kCompactRow0 = FGS_REDUCE_ROWS + 1 + 3;
constexpr int kPerPodRows = kCompactRow0 + 4;

struct GangState; // gang_quota.cu
struct FilterTables; // filter_tables.cuh
struct SeqState;  // sequential.cu

struct Engine {
    int device = 0;
    int S = 2, D = 1, D_cap = 0;
    int64_t node_base = 0, total_nodes = 0;
    uint64_t seed = 0;
    cudaStream_t stream = nullptr;
    bool own_stream = false;
    cudaEvent_t events[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};
    std::string err;

    fgs_plugin_cfg cfg;
    bool have_cfg = false, have_nodes = false, nodes_derived = false, have_pods = false, emitted = false;
    int last_key_shift = 40;

    // nodes
    int64_t n_nodes = 0, n_pad = 0, n_pad_cap = 0;
    void* d_raw = nullptr;
    void* d_nd = nullptr;
    void* h_raw = nullptr;
    // pods / results
    int64_t n_pods = 0, p_pad = 0, p_pad_cap = 0, p_plane_cap = 0;
    int64_t words_per_row = 0, score_stride = 0;
    void* d_pods = nullptr;
    void* d_perpod = nullptr;
    void* d_bitmap = nullptr;
    void* d_scores = nullptr;
    void* h_pods = nullptr;
    void* h_perpod = nullptr;

    // int32 specialisation (kernels_narrow.cu): analysis of the node side, current shifts, tables
    NodeAnalysis node_an;
    bool node_an_valid = false;
    NarrowShifts narrow_sh;
    int narrow_nodes_sig = -1; // fit strategy the narrow node table was built for (-1 = stale)
    bool narrow_batch = false; // the staged pod batch qualifies
    int host_threads = 4;      // threads that stage the pod columns (FGS_HOST_THREADS)
    uint64_t stage_or[PD_FIXED_COLS + FGS_MAX_SCALARS] = {0};
    int64_t stage_max[PD_FIXED_COLS + FGS_MAX_SCALARS] = {0};
    uint32_t stage_present_or = 0;
    bool force_wide = false;   // FGS_FORCE_INT64=1: always take the int64 kernels (tests)
    bool no_dma = false;       // FGS_NO_DMA=1: always stage through the engine's own pinned buffer (tests)
    bool staged_by_dma = false; // the last staged batch came by DMA straight from the caller's table
    bool reduce_was_narrow = false, emit_was_narrow = false;
    void* d_nn = nullptr; // the int32 node table in use: one entry of nn_ring
    // The int32 node table depends on the batch's shifts, and several batches are in flight (or resident in pipeline slots):
    // a small ring of tables keyed by (node table version, fit strategy, shifts). A batch keeps its entry until it is
    // collected; an entry no batch in flight uses is recycled, least recently used first.
    static constexpr int kNnRing = 8;
    struct NnEntry {
        void* d = nullptr;
        NarrowShifts sh;
        int sig = -1; // fit strategy the table was built for, -1 = holds nothing
        uint64_t gen = 0, last_use = 0;
    };
    NnEntry nn_ring[kNnRing];
    int nn_cur = -1;
    uint64_t nn_clock = 0;
    int use_nn(const NarrowShifts& sh, int sig); // d_nn / narrow_sh / narrow_nodes_sig ← a table narrowed under sh (built if no entry holds one)
    void invalidate_nn();
    int ensure_slot_nn(); // after bind_slot: the bound batch's int32 node table is there (rebuilt if its ring entry was recycled)
    void* d_pn = nullptr;
    void* d_an = nullptr;
    void* d_special = nullptr; // int64 [2][n_pad]: nodes with a negative MemScore or on which NewGPUNode panics
    int64_t special_cap = 0;
    int64_t nn_cap = 0, pn_cap = 0;
    int decide_narrow(const fgs_pod_table* t, bool launch_pods = true);
    int decide_narrow_once(const fgs_pod_table* t, bool launch_pods);
    bool narrow_refused_by_nodes = false; // the last decide_narrow_once said no because of the node analysis
    // small batches (<= 64 pods): the whole cycle of every pod in one launch (filter_tables.cu: small_cycle); pod columns and
    // result records travel through one block of mapped host memory. FGS_SMALL=0 turns the path off (tests).
    void* h_small = nullptr; // mapped: int64 [pd_cols][64] pod columns, then int64 [2][64] result records, then the completion word
    void* d_small_done = nullptr; // device counter of finished pods
    uint64_t small_seq = 0;
    bool small_off = false;
    void* d_sample_scratch = nullptr; // node sampling: block maps of the composed walk
    int64_t sample_scratch_cap = 0;
    int small_direct = -1;
    bool small_cluster = true;
    uint64_t small_gen = ~0ull; // node_gen of the last small-batch call, and how many calls in a row it has served
    int small_same = 0;
    // row patches (patch_nodes): the flag a row outside the host's node analysis raises, the staging pair for many rows
    bool patch_off = false, patch_spec = false;
    bool an_loose = false; // rows were patched since the last full analysis: it may be wider than the table needs
    int* d_patch_flag = nullptr;
    volatile int* h_patch_flag = nullptr;
    void *h_patch = nullptr, *d_patch = nullptr;
    size_t patch_bytes = 0;
    uint64_t patches = 0, patch_redos = 0;
    uint64_t staged_gen = 0; // node_gen when the current batch was staged
    int resolve_patch();
    int nodes_ready();
    int patch_rows(const fgs_node_table* rows, const int64_t* dirty_idx);
    bool was_small = false;  // the last fgs_evaluate took it
    int evaluate_small(const fgs_pod_table* t, const fgs_eval_opts* opts, fgs_results* out, bool* done);
    void fill_narrow_params(NarrowParams* p) const;
    // the fused pass (eval_narrow.cu: filters + scores in ONE pass over pod × node, after a per-pod prepass for the
    // FlexGPU maximum): node orders by GPUScore / MemScore, rebuilt whenever the narrow node table is
    void* d_order = nullptr;          // int [2][n_pad]
    void* d_order_scratch = nullptr;
    size_t order_scratch_bytes = 0;
    int64_t order_cap = 0;
    bool orders_valid = false;
    bool no_fused = true;             // FGS_FUSED=1 turns the fused pass on (opt-in: slower than the two passes on C2)
    bool was_fused = false;           // the last evaluation took the fused pass
    bool fused_eligible(const fgs_eval_opts* opts) const;
    bool fused_prepass_done = false;
    int run_fused(const fgs_eval_opts* opts, int part); // init + prepass (1), fused kernel + special pass (2): what pass_reduce + pass_emit do

    // fgs_submit / fgs_collect: two batches in flight. A slot owns the per-batch buffers; the engine's members
    // above (d_pods, h_pods, d_perpod, h_perpod, d_pn, …) always alias the slot being filled.
    struct Slot {
        void *d_pods = nullptr, *h_pods = nullptr, *d_perpod = nullptr, *h_perpod = nullptr, *d_pn = nullptr;
        int64_t p_pad_cap = 0, pn_cap = 0, n_pods = 0, p_pad = 0;
        bool narrow_batch = false, emitted = false, sampled = false, have_pods = false, busy = false;
        uint64_t node_gen = 0; // Engine::node_gen when the batch was staged (fgs_resubmit refuses a stale batch)
        int key_shift = 40;
        NarrowShifts sh;     // the shifts its int32 pod table was narrowed under, and …
        int nn_sig = -1;     // … the fit strategy of the node table that goes with them
        int nn_idx = -1;     // the ring entry it uses while it is in flight
        cudaEvent_t ev_done = nullptr;
        // sharded pipeline (peer attached): the emit pass of a batch is issued with the NEXT submit (or by collect)
        bool emit_pending = false;
        int bitmap_sel = 0;
        fgs_eval_opts opts;
        bool have_opts = false;
        cudaEvent_t ev_reduce = nullptr, ev_merge1 = nullptr, ev_emit = nullptr, ev_merge2 = nullptr;
        cudaEvent_t ev_t0 = nullptr, ev_emit0 = nullptr; // FGS_PIPE_TRACE: start of the filter pass / of the emit pass
    };
    static constexpr int64_t kPatchMaxRows = 4096; // fgs_update_nodes beyond this re-derives the whole table
    static constexpr int kSlots = 32; // upper bound of fgs_set_resident_batches
    Slot slots[kSlots];
    // batches in flight: two on one GPU (a third buys nothing there while the tables are resident, tools/pipe_breakdown.py),
    // three in the sharded pipeline (see submit_sharded). The ring of slots is that long too unless
    // fgs_set_resident_batches asked for more: extra slots keep more batches RESIDENT for fgs_resubmit, and the first
    // of them also lets ONE more batch be in flight — its H2D copy is then issued a whole pass earlier, which is what
    // hides it when the results of the oldest batch are slow to come back (8 GPUs on one host).
    int base_in_flight() const { return peer ? 3 : 2; }
    int resident_cfg = 0;
    int ring() const { return resident_cfg > base_in_flight() ? resident_cfg : base_in_flight(); }
    int max_in_flight() const { return ring() > base_in_flight() ? base_in_flight() + 1 : base_in_flight(); }
    // FGS_PIPE_TRACE=<file>: the sharded pipeline's events are created with timing and every collected batch appends one
    // line (ms since the first batch's start): filter start, filter end, merge 1 end, emit start, emit end, merge 2 end, D2H end
    bool pipe_trace = false;
    std::string pipe_trace_path;
    cudaEvent_t trace_epoch = nullptr;
    std::vector<float> trace_rows;
    void trace_collect(int slot);
    void trace_dump();
    int next_slot = -1;      // slot the next submit takes (-1: the current one) — the ring is walked round-robin
    int submit_single(const fgs_pod_table* pods, const fgs_eval_opts* opts, bool resident);
    uint64_t submit_seq = 0; // batches submitted so far: the sharded pipeline alternates two bitmaps by its parity
    int cur_slot = 0;       // slot whose buffers the members alias
    int in_flight = 0;      // submitted, not yet collected
    int oldest_slot = 0;
    cudaStream_t copy_stream = nullptr, d2h_stream = nullptr;
    cudaEvent_t ev_h2d = nullptr, ev_kernels = nullptr;
    void bind_slot(int k);  // save the members into slots[cur_slot], load slots[k]
    // sharded pipeline: merges on their own (high-priority) stream so that they overlap the other batch's pass;
    // every other batch uses a second bitmap (the emit pass of batch k-1 reads its rows while the reduce pass of batch k writes)
    cudaStream_t merge_stream = nullptr;
    void* d_bitmap_alt = nullptr;
    void* d_bitmap_alt2 = nullptr; // sharded pipeline with pass A on the front stream: three batches' bitmaps are live at a time
    size_t bitmap_alt2_bytes = 0;
    size_t bitmap_alt_bytes = 0;
    // single-engine pipeline: pass A of batch k+1 on a second (higher-priority) stream next to the emit pass of batch k
    bool overlap_filter = true;
    cudaStream_t front_stream = nullptr, window_back = nullptr; // window_back: the compute stream while `stream` is swapped for front_stream
    cudaEvent_t ev_front = nullptr, ev_back = nullptr, ev_plane[3] = {nullptr, nullptr, nullptr};
    bool plane_used[3] = {false, false, false};
    int ensure_front_stream();
    int peer_merge(int which, cudaStream_t on, bool overlapped = false); // peer_merge.cu
    int peer_fetch_error(int slot, cudaStream_t on); // copy of the merge error word, behind the merges on `on`
    int peer_check_error(int slot);                  // FGS_E_STATE when that copy says a merge timed out
    int submit_sharded(const fgs_pod_table* pods, const fgs_eval_opts* opts, bool resident);
    int issue_emit_sharded(int slot);

    // nominated pods (nominated.cu): the PodNominator's content, grouped by node, and the per-batch columns
    // (priority, uid, dirty flag) the second filter pass needs
    void* d_nom = nullptr;      // int64 [NQ_COLS][nom_q] | int64 nodes[nom_m] | int64 off[nom_m + 1]
    int64_t nom_q = 0, nom_m = 0, nom_cap_bytes = 0;
    void* d_pod_aux = nullptr;  // int64 [3][aux_cap]: priority, uid, dirty
    void* h_pod_aux = nullptr;
    int64_t aux_cap = 0;
    bool aux_has_uid = false;
    int stage_nominated_aux(const fgs_pod_table* t); // called by the pod staging when nominated pods are set
    int run_nominated_pass();                        // after the reduce pass, before the merges

    // node sampling (sample.cu): per-pod walks of the current batch
    void* d_walk = nullptr;    // int2 [p_pad] + one int64 (nextStartNodeIndex)
    void* h_walk = nullptr;
    int64_t walk_cap = 0;
    bool sampled = false;      // the current batch's bitmap was truncated by the walk
    int run_sampling(const fgs_eval_opts* opts);

    // pass A as table look-ups (filter_tables.cu): level tables of the node side of the filters, rebuilt lazily when the
    // node table or the filter plugins change. FGS_FILTER=compare forces the compare kernels, =lookup refuses to fall back.
    FilterTables* ft = nullptr;
    int filter_mode = 0; // 0 auto (look-up when the tables fit), 1 compare kernels only, 2 look-up only (tests)
    bool reduce_was_lookup = false;
    int ensure_filter_tables(bool* use);
    void ft_free();

    std::vector<int64_t> synced_gen; // fgs_sync_nodes: NodeInfo.Generation of every row of the shard as last uploaded
    void* d_mask = nullptr;  // fgs_score_masked: the caller's feasible mask when it comes from the host
    size_t mask_bytes = 0;
    bool reduced = false;    // the staged batch has been through pass A (fgs_score_masked needs its bitmap)
    bool masked = false;     // … and its bitmap has been cut down to the caller's feasible list since

    GangState* gang = nullptr;
    SeqState* seq = nullptr;
    struct PeerState* peer = nullptr; // peer_merge.cu: the other shards' exchange buffers
    struct PreemptCache* pcache = nullptr; // preempt.cu: the snapshot's pods as device columns, kept between calls
    uint64_t node_gen = 0;                 // bumped whenever the node table changes (upload, patch, sequential commits)

    fgs_timings tm = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
    int64_t launches = 0;

    // one caller at a time: the thread inside an entry point owns the engine (entry points call each other, so the
    // ownership nests); a second thread gets FGS_E_STATE instead of corrupting the staging slots
    std::atomic<uint64_t> owner{0};
    int owner_depth = 0;

    int fail(int code, const std::string& msg);
    int ensure_nodes(int64_t n, int D);
    int ensure_pods(int64_t p);
    int set_plugins(const fgs_plugin_cfg* c);
    void fill_eval_params(EvalParams* p, const fgs_eval_opts* opts) const;
    void free_all();
    void gang_free();
    void seq_free();
    void peer_free();
    void preempt_free();
    void fill_derive_params(DeriveParams* dp) const;
    int derive_now();        // derive_node_columns over the whole table + analysis
    int analyze_nodes_now(bool table_changed = true); // node-side analysis for the int32 specialisation (after any change of d_nd)
};

// first statement of every entry point that takes an engine
struct EngineEntry {
    Engine* e;
    bool ok;
    explicit EngineEntry(Engine* e_);
    ~EngineEntry();
};
#define FGS_ENTER_RET(e, ret)      \
    fgs::EngineEntry _entry(e);    \
    if (!_entry.ok) return (ret)
#define FGS_ENTER(e) FGS_ENTER_RET(e, FGS_E_STATE)

} // namespace fgs
