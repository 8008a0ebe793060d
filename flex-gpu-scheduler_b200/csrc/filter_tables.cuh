// filter_tables.cuh — pass A as table look-ups: per-dimension level tables of the node side of the two filters.
//
// Every test FlexGPU.Filter (pkg/flexgpu/flex_gpu.go:54-116, folded into one threshold per pod kind by
// derive_node_columns) and Fit.Filter (vendor/…/noderesources/fit.go:256-329) make for a (pod,node) pair has the form
// "pod value <= node threshold" in one of 5+S dimensions (whole-GPU limit, fractional limit, cpu, memory,
// ephemeral-storage, scalar s). A comparison only depends on the RANK of the pod value among the distinct node
// thresholds of that dimension, so for every dimension d and level r the engine keeps one bit row
//     T_d[r] = { n : threshold_d(n) >= level_d[r] }      (uint32 [n_pad/32], bit n%32 of word n/32)
// and a pod's feasibility row is the AND of 5+S such rows — a few coalesced loads per 32 (pod,node) pairs instead of
// 5+S compares per pair. The tables change only with the node table (upload / patch / sequential commits) and stay in
// L2 for clusters of C2's size (7 × 5 000 rows × 640 B = 22 MB of the B200's 126 MB).
#pragma once
#include <cuda_runtime.h>

#include <cstdint>

#include "../../include/fgs.h"
#include "kernels.cuh"

namespace fgs {

constexpr int kFtMaxDims = 5 + FGS_MAX_SCALARS;
enum FtDim : int { FT_GPU = 0, FT_FMEM, FT_CPU, FT_MEM, FT_EPH, FT_SCALAR0 };
enum FtOrder : int { FT_ORD_GSCORE = 0, FT_ORD_MSCORE, FT_ORD_ARAW, FT_ORDERS };

// device-side description handed to the kernels (by value)
struct FtView {
    const uint32_t* table;   // rows of all dimensions, one after the other: [rows_total][W]
    const int64_t* levels;   // [n_dims][lev_stride]: the distinct thresholds of each dimension, ascending
    const int32_t* ord;      // [FT_ORDERS][n_pad]: node indices by GPUScore / MemScore / Allocatable raw score, highest first
    const int64_t* ord_val;  // [FT_ORDERS][n_pad]: the sorted values
    int64_t lev_stride;
    int64_t W;               // words per row = n_pad / 32
    int64_t ones_row;         // index of the all-ones row behind the last dimension
    int64_t row0[kFtMaxDims]; // first row of dimension d
    int32_t R[kFtMaxDims];    // levels of dimension d (its rows are 0..R[d]: row R[d] is empty)
    int32_t n_dims;
};

struct FilterTables {
    FtView v;
    uint64_t node_gen = ~0ull;      // Engine::node_gen the tables were built from
    uint32_t filter_mask = ~0u;     // … and the filter plugins in the profile
    bool valid = false;
    bool refused = false;           // over the memory limit for this node table: the compare kernels run instead
    int64_t n_pad = 0, n_nodes = 0;
    // device buffers (grow-only)
    void* d_table = nullptr;
    size_t table_bytes = 0;
    void* d_levels = nullptr;       // int64 [n_dims][n_cap + 1]
    void* d_node_rank = nullptr;    // int32 [n_dims][n_cap]
    void* d_ord = nullptr;          // int32 [FT_ORDERS][n_cap]
    void* d_ord_val = nullptr;      // int64 [FT_ORDERS][n_cap]
    void* d_thr = nullptr;          // int64 [n_dims + FT_ORDERS][n_cap]: sort keys
    void* d_sort = nullptr;         // sort scratch: keys out, values in / out, scan, cub temp
    size_t sort_bytes = 0;
    void* d_counts = nullptr;       // int32 [kFtMaxDims]
    void* d_rowidx = nullptr;       // int32 [n_dims][p_cap]: per batch, the table row of every (pod, dimension)
    int64_t rowidx_cap = 0;
    int64_t n_cap = 0;
    int dims_cap = 0;
    int64_t rows_total = 0;
    size_t limit_bytes = (size_t)24 << 30;
    double build_ms = 0;
    void free_all();
};

struct FtBuildParams {
    const int64_t* nd; // derived node table
    int64_t n_nodes, n_pad;
    int S;
    uint32_t filter_mask;
};

// (Re)builds the tables on `stream`; synchronises it once (the level counts size the table). Returns cudaSuccess and
// ft->valid, or ft->refused when the tables would exceed ft->limit_bytes.
cudaError_t ft_build(FilterTables* ft, const FtBuildParams& bp, cudaStream_t stream, int64_t* launches);

struct FtRowsParams {
    FtView v;
    const int64_t* pods;     // device pod table int64 [pd_cols(S)][p_pad]
    const int64_t* nd;       // derived node table (fallback of the order walk)
    int64_t n_pods, p_pad, n_nodes, n_pad;
    int S;
    uint32_t fit_ignored;
    uint32_t* bitmap;        // [P][W]
    int64_t* reduce_max;     // [FGS_REDUCE_ROWS][p_pad]
    int64_t* reduce_sum;     // [p_pad]
    int64_t* select_key;     // [3][p_pad]: cleared here (the emit pass max-accumulates into it)
    int sc_flex, sc_alloc;
    // the few nodes that feed a pod's error bits (negative MemScore, NewGPUNode panics): special[0][k] = node | panic << 40,
    // special[1][k] = |MemScore| when negative. With them the kernel also writes RR_FILTER_PANIC / RR_NEG_MAG /
    // RR_SCORE_PANIC, so that no separate per-pod pass (or initialisation) is needed.
    const int64_t* special;
    int n_special;
    uint32_t filter_mask;
    int flex_first;
};
// pass A: bitmap rows, feasible counts, max FlexGPU score, Allocatable hi / −lo of every pod — two launches: the row of
// every (pod, dimension) into rowidx (int32 [n_dims][p_pad], caller's scratch), then one warp per pod
cudaError_t ft_launch_rows(const FtRowsParams& p, int* rowidx, cudaStream_t stream);
bool ft_rows_fit(int64_t W); // the row of one pod fits the kernel's shared memory

// bitmap &= mask per pod, then count and maxima over what is left (fgs_score_masked). v.ord / v.ord_val are read only
// with have_orders; W = words per bitmap row.
struct FtMaskParams {
    FtView v;
    int have_orders;
    const int64_t* pods;
    const int64_t* nd;
    int64_t n_pods, p_pad, n_nodes, n_pad, W;
    uint32_t* bitmap;
    const uint32_t* mask;
    int64_t* reduce_max;
    int64_t* reduce_sum;
    int64_t* select_key; // [3][p_pad]: cleared, the emit pass that follows accumulates into it
    int sc_flex, sc_alloc;
};
cudaError_t ft_launch_masked(const FtMaskParams& p, cudaStream_t stream);

// ---- small batches: the whole cycle of a pod in ONE CTA, the batch in ONE launch ---------------------------------------------
// What the reference's scheduleOne does per pod (sched/scheduler.go:425-455 → generic_scheduler.go:93-167): up to 256 pods,
// one CTA each — the pod's columns are read straight from mapped host memory, Filter is the AND of the level rows, the
// normaliser reductions are block reductions, every node is scored with the int32 arithmetic of emit_narrow (same magic
// multipliers, same selectHost key) and the 16-byte result record goes straight back to mapped host memory.
constexpr int kSmallMaxPods = 256;
constexpr uint64_t kSmallDeclined = 1ull << 63;
constexpr int kSmallCluster = 8;    // portable cluster size
constexpr int kSmallInlinePods = 8; // pods that travel as kernel arguments
struct SmallParams {
    FtView v;
    const int64_t* h_pods;   // mapped host block int64 [pd_cols(S)][kSmallMaxPods]
    int64_t* h_compact;      // mapped host block int64 [2][kSmallMaxPods]: the result records
    int64_t* d_pods;         // device pod table [pd_cols(S)][p_pad], filled for the plane getters
    int* d_pn;               // int32 pod table [PN_COLS][p_pad], filled for later emit passes
    const int* nn;           // int32 node table
    const int64_t* nd;       // derived node table
    int64_t* perpod;         // per-pod block: reduce rows | count | select keys | compact rows
    uint32_t* bitmap;
    int64_t* scores;
    int64_t n_pods, p_pad, n_nodes, n_pad, node_base, W, score_stride;
    int S;
    uint32_t fit_ignored, filter_mask, score_mask;
    int flex_first, fit_most, w_eq;
    int w_flex, w_fit, w_alloc, fit_w_cpu, fit_w_mem;
    float fit_inv_wsum;
    int sh[8];               // NarrowShifts
    uint64_t seed;
    const int64_t* special;
    int n_special;
    // completion without a stream synchronisation: the CTA that finishes last (device counter) stores `seq` into the mapped
    // word behind the records once every record is visible to the host
    unsigned int* d_done;
    volatile uint64_t* h_flag;
    uint64_t seq;
    const int* spec_flag; // row patches not yet confirmed by the host (patch_nodes): non-zero → the node analysis this launch was
                          // parameterised by no longer holds; every CTA leaves and the completion word carries kSmallDeclined
    int direct; // 1: no level tables (stale or over their limit): Filter by comparing every node's thresholds, maxima by reduction
    // A handful of pods against a long node list: a thread-block CLUSTER per pod instead of one CTA — each CTA takes a slice of
    // the nodes (always the compare form), the maxima, the count and the selectHost keys are merged through distributed
    // shared memory. The pods then travel as kernel arguments (no read of host memory on the way in).
    int cluster;  // CTAs per pod: 1, or kSmallCluster
    int n_inline; // > 0: pods_inline holds the pod columns, h_pods is not read
    int64_t pods_inline[kSmallInlinePods][PD_FIXED_COLS + FGS_MAX_SCALARS];
};
cudaError_t ft_launch_small(const SmallParams& p, cudaStream_t stream);

// the pods that reach NewGPUNode's panic inside FlexGPU.Filter → RR_FILTER_PANIC, from the int64 tables (any S)
void ft_launch_filter_panic(const int64_t* nd, const int64_t* pods, const int64_t* special, int64_t n_pad, int n_special, int64_t n_pods,
                            int64_t p_pad, int S, uint32_t filter_mask, uint32_t fit_ignored, int flex_first, int64_t* reduce_max,
                            cudaStream_t stream);

} // namespace fgs
