// kernels.cuh — launch parameters and device-table layouts shared by kernels.cu and engine.cu.
#pragma once
#include <cuda_runtime.h>

#include <cstdint>

#include "../../include/fgs.h"

namespace fgs {

// ---- device node table ("derived" columns, int64 [ND_COLS][n_pad]) ---------------------------
// Written by derive_node_columns from the uploaded raw columns; read by the evaluation kernels.
enum NodeCol : int {
    ND_THR_GPU = 0,   // largest whole-GPU limit FlexGPU.Filter accepts on the node, INT64_MIN if none
    ND_THR_MEM,       // largest fractional limit accepted (min of node headroom and best device)
    ND_FREE_CPU,      // Allocatable − Requested (fit.go:277-303)
    ND_FREE_MEM,
    ND_FREE_EPH,
    ND_FLAGS,         // low 32 bits: FGS_NF_* | derived bits below
    ND_CAP_CPU,       // Allocatable.MilliCPU
    ND_NZR_CPU,       // NonZeroRequested.MilliCPU
    ND_CAP_MEM,
    ND_NZR_MEM,
    ND_GPU_SCORE,     // gpuNode.GPUScore (gpu_node.go:179-187)
    ND_MEM_SCORE,     // gpuNode.MemScore (gpu_node.go:189-199)
    ND_ALLOC_RAW,     // Allocatable.Score raw value (allocatable.go:120-143)
    ND_ALLOC_EPH,     // Allocatable.EphemeralStorage (generic Fit scoring only)
    ND_LEFT_PODS,     // AllowedPodNumber − len(Pods) (core.go:367)
    ND_MEM_EACH,      // memory of one device, memCnt / gpuCnt (gpu_node.go:55); 0 where NewGPUNode panics
    ND_FIXED_COLS,
    // then S columns free_scalar[s] = Allocatable[s] − Requested[s], then S columns alloc_scalar[s]
};
constexpr uint32_t NDF_PODS_OK = 1u << 8;      // len(Pods)+1 <= AllowedPodNumber (fit.go:259)
constexpr uint32_t NDF_MEMSCORE_NEG = 1u << 9; // MemScore < 0 (an over-committed device)

inline __host__ __device__ int nd_cols(int S) { return ND_FIXED_COLS + 2 * S; }
inline __host__ __device__ int nd_free_scalar(int s) { return ND_FIXED_COLS + s; }
inline __host__ __device__ int nd_alloc_scalar(int S, int s) { return ND_FIXED_COLS + S + s; }

// ---- device raw node table (uploaded as is, int64 [NR_COLS][n_pad]) -----------------------------
enum RawNodeCol : int {
    NR_ALLOC_MCPU = 0, NR_ALLOC_MEM, NR_ALLOC_EPH, NR_ALLOC_PODS, NR_REQ_MCPU, NR_REQ_MEM, NR_REQ_EPH, NR_NUM_PODS,
    NR_NZ_MCPU, NR_NZ_MEM, NR_FLEX_GPU_SUM, NR_FLEX_MEM_SUM,
    NR_META,          // low 32: flags, high 32: monopoly mask
    NR_PRESENT,       // alloc_scalar_present
    NR_FIXED_COLS,
    // then S alloc_scalar, S req_scalar, D used_mem columns
};
inline __host__ __device__ int nr_cols(int S, int D) { return NR_FIXED_COLS + 2 * S + D; }
inline __host__ __device__ int nr_alloc_scalar(int s) { return NR_FIXED_COLS + s; }
inline __host__ __device__ int nr_req_scalar(int S, int s) { return NR_FIXED_COLS + S + s; }
inline __host__ __device__ int nr_used(int S, int d) { return NR_FIXED_COLS + 2 * S + d; }

// ---- device pod table (int64 [PD_COLS][p_pad]) -----------------------------------------------------
enum PodCol : int {
    PD_GPU_LIM = 0, PD_MEM_LIM, PD_REQ_CPU, PD_REQ_MEM, PD_REQ_EPH, PD_NZ_CPU, PD_NZ_MEM,
    PD_META,          // low 32: FGS_PF_* flags, high 32: req_scalar_present
    PD_FIXED_COLS,
    // then S req_scalar columns
};
inline __host__ __device__ int pd_cols(int S) { return PD_FIXED_COLS + S; }

// rows of reduce_max
// RR_NEG_MAG: largest |FlexGPU score| among feasible nodes whose score is negative (an over-committed device);
// RR_SCORE_PANIC: a feasible node on which FlexGPU.Score's NewGPUNode would panic. Both are turned into the
// pod's error bits by finalize_pods once the feasible count and the flex maximum are known.
enum ReduceRow : int { RR_FLEX_MAX = 0, RR_ALLOC_HI, RR_ALLOC_NEGLO, RR_FILTER_PANIC, RR_NEG_MAG, RR_SCORE_PANIC };

constexpr int kThreads = 256;       // threads per CTA of the evaluation kernels
constexpr int kNodesPerThread = 2;
constexpr int kTileN = kThreads * kNodesPerThread; // 512 nodes per CTA
constexpr int kTileP = 64;          // pods per shared-memory tile
constexpr int kWarps = kThreads / 32;
constexpr int kNodePad = 1024;       // node tables are padded to this (a multiple of every kernel's node tile)

struct DeriveParams {
    const int64_t* raw;
    int64_t* nd;
    int64_t n_nodes, n_pad;
    int S, D;
    int alloc_mode, alloc_n_res;
    int alloc_res[FGS_MAX_SCORE_RES];
    int64_t alloc_w[FGS_MAX_SCORE_RES];
};

struct EvalParams {
    const int64_t* nd;
    const int64_t* pods;
    int64_t* reduce_max;   // [FGS_REDUCE_ROWS][p_pad]
    int64_t* reduce_sum;   // [p_pad]
    int64_t* select_key;   // [2][p_pad]
    uint32_t* bitmap;      // [P][words_per_row]
    int64_t* scores;       // [P][score_stride]
    int64_t n_nodes, n_pad, node_base;
    int64_t n_pods, p_pad;
    int64_t words_per_row, score_stride;
    int pods_per_cta;      // multiple of kTileP
    int S;
    uint32_t filter_mask, score_mask, emit_mask, emit_raw;
    int flex_filter_first; // FlexGPU.Filter runs before Fit.Filter in the profile
    int64_t w_flex, w_fit, w_alloc;
    int fit_strategy, fit_n_res;
    int fit_res[FGS_MAX_SCORE_RES];
    int64_t fit_w[FGS_MAX_SCORE_RES];
    uint32_t fit_ignored;
    int fit_fast;          // 1: score resources ⊆ {cpu, memory}: register-only path
    int fit_w_cpu, fit_w_mem, fit_wsum;
    float fit_inv_wsum;
    uint64_t seed;
    int key_shift;         // bits below the score field of the select key (40, or 24 without hash)
    int feas_from_bitmap;  // EMIT: AND the recomputed Filter verdict with the bitmap (node sampling truncated it)
};

// sets the dynamic shared memory limits of the evaluation kernels on the current device
cudaError_t eval_setup();
// reduce_max rows := INT64_MIN (panic row := 0), reduce_sum := 0, select_key rows := 0
void launch_init_perpod(int64_t* perpod, int64_t p_pad, void* stream);
// per-pod results in the form the host reads back — int64 [2][p_pad], 16 bytes per pod:
// [0] feasible count (low 28 bits) | FGS_PODERR_* bits << 28 | selected node (global index, -1 none) << 32, [1] selected score
void launch_finalize_pods(const int64_t* perpod, int64_t* compact, int64_t n_pods, int64_t p_pad, int key_shift, int emitted, int sc_flex,
                          void* stream);
void launch_make_pod_meta(const uint32_t* flags, const uint32_t* present, int64_t* meta, int64_t n_pods, int64_t p_pad, void* stream);
void launch_derive(const DeriveParams& p, void* stream);
void launch_eval_reduce(const EvalParams& p, void* stream);
void launch_eval_emit(const EvalParams& p, void* stream);
size_t eval_smem_bytes(int S);

} // namespace fgs
