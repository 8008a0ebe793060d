// derive.cuh — pod-independent columns of ONE node row, shared by derive_node_columns (whole table, at upload)
// and the sequential-commit kernel (the single row a placement changes).
//   NewGPUNode / constructGPUs      pkg/flexgpu/gpu_node.go:30-120   (memEach, per-device free memory)
//   GPUScore / MemScore             pkg/flexgpu/gpu_node.go:179-199
//   FlexGPU.Filter steps 3-5 and 7  pkg/flexgpu/flex_gpu.go:64-116   → one threshold per pod kind
//   Allocatable raw score           pkg/noderesources/allocatable.go:120-143
#pragma once
#include "kernels.cuh"

namespace fgs {

// raw row: raw[c * rs], derived row written to nd[c * ns] (global tables: pointer to the node's element, stride
// n_pad; the sequential kernel's shared-memory copy: stride = nodes per CTA)
// DYNAMIC_ONLY: the row was derived before and only what a placement can change is redone — Allocatable and with
// it memEach (read back from ND_MEM_EACH, saving the 64-bit division) and the Allocatable raw score stay.
template <bool DYNAMIC_ONLY = false>
__device__ __forceinline__ void derive_row(const DeriveParams& prm, const long long* __restrict__ raw, long long rs, long long* __restrict__ nd, long long ns) {
    typedef long long i64;
    typedef unsigned long long u64;
    constexpr i64 kI64Min = (i64)0x8000000000000000ull;
    constexpr i64 kI64Max = (i64)0x7fffffffffffffffull;
    auto wmul = [](i64 a, i64 b) { return (i64)((u64)a * (u64)b); };
    auto wadd = [](i64 a, i64 b) { return (i64)((u64)a + (u64)b); };
    const int S = prm.S, D = prm.D;
    auto R = [&](int c) { return raw[(i64)c * rs]; };
    const u64 meta = (u64)R(NR_META);
    uint32_t flags = (uint32_t)meta;
    const uint32_t mono = (uint32_t)(meta >> 32);

    const i64 gpuAlloc = R(nr_alloc_scalar(FGS_SCALAR_FLEX_GPU));
    const i64 memAlloc = R(nr_alloc_scalar(FGS_SCALAR_FLEX_MEM));
    const i64 gpu_headroom = gpuAlloc - R(NR_FLEX_GPU_SUM); // flex_gpu.go:85-93
    const i64 mem_headroom = memAlloc - R(NR_FLEX_MEM_SUM); // :95-98
    const bool has_flex = flags & FGS_NF_HAS_FLEX;
    const bool panic = flags & FGS_NF_REF_PANIC;

    i64 max_free = kI64Max, gpu_score = 0, mem_score = 0, mem_each = 0;
    if (!panic) {
        const i64 gpuCnt = gpuAlloc; // > 0 here (REF_PANIC covers <= 0 and a missing key)
        const i64 memCnt = (flags & FGS_NF_HAS_MEM_RES) ? memAlloc : 0;
        const i64 memEach = DYNAMIC_ONLY ? nd[(i64)ND_MEM_EACH * ns] : memCnt / gpuCnt; // gpu_node.go:55
        mem_each = memEach;
        max_free = kI64Min;
        i64 used_sum = 0;
        for (int d = 0; d < D; ++d) {
            if ((i64)d >= gpuCnt) break;
            const i64 used = R(nr_used(S, d));
            const bool m = (mono >> d) & 1u;
            used_sum += used;
            if (!m) {
                const i64 fr = memEach - used; // MemAssumeFitIndexes' test, gpu_node.go:136
                if (fr > max_free) max_free = fr;
                if (used == 0) gpu_score++; // gpu_node.go:181-186
            }
        }
        mem_score = gpuCnt * memEach - used_sum; // gpu_node.go:189-199
        if (mem_score < 0) flags |= NDF_MEMSCORE_NEG;
    }
    i64 thr_gpu = kI64Min, thr_mem = kI64Min;
    if (has_flex) {
        if (mem_headroom >= 0) thr_gpu = gpu_headroom;
        if (gpu_headroom >= 0) thr_mem = mem_headroom < max_free ? mem_headroom : max_free;
    }
    if (R(NR_NUM_PODS) + 1 <= R(NR_ALLOC_PODS)) flags |= NDF_PODS_OK; // fit.go:259

    // Allocatable raw score: Σ mode(alloc_r)·w_r / Σ w_r with Go's wrapping int64 (allocatable.go:120-143)
    i64 alloc_raw = 0;
    if (!DYNAMIC_ONLY) {
        i64 node_score = 0, weight_sum = 0;
        for (int k = 0; k < prm.alloc_n_res; ++k) {
            const int res = prm.alloc_res[k];
            i64 a = 0;
            if (res == FGS_RES_CPU) a = R(NR_ALLOC_MCPU);
            else if (res == FGS_RES_MEM) a = R(NR_ALLOC_MEM);
            else if (res == FGS_RES_EPH) a = R(NR_ALLOC_EPH);
            else if (res >= FGS_RES_SCALAR0 && res < FGS_RES_SCALAR0 + S) a = R(nr_alloc_scalar(res - FGS_RES_SCALAR0));
            const i64 rs_ = prm.alloc_mode == 0 ? wmul(-1, a) : a;
            node_score = wadd(node_score, wmul(rs_, prm.alloc_w[k]));
            weight_sum = wadd(weight_sum, prm.alloc_w[k]);
        }
        alloc_raw = weight_sum != 0 ? node_score / weight_sum : 0;
    }

    auto W = [&](int c, i64 v) { nd[(i64)c * ns] = v; };
    W(ND_THR_GPU, thr_gpu);
    W(ND_THR_MEM, thr_mem);
    W(ND_FREE_CPU, R(NR_ALLOC_MCPU) - R(NR_REQ_MCPU));
    W(ND_FREE_MEM, R(NR_ALLOC_MEM) - R(NR_REQ_MEM));
    W(ND_FREE_EPH, R(NR_ALLOC_EPH) - R(NR_REQ_EPH));
    W(ND_FLAGS, (i64)flags);
    W(ND_NZR_CPU, R(NR_NZ_MCPU));
    W(ND_NZR_MEM, R(NR_NZ_MEM));
    W(ND_GPU_SCORE, gpu_score);
    W(ND_MEM_SCORE, mem_score);
    W(ND_LEFT_PODS, R(NR_ALLOC_PODS) - R(NR_NUM_PODS));
    for (int s = 0; s < S; ++s) W(nd_free_scalar(s), R(nr_alloc_scalar(s)) - R(nr_req_scalar(S, s)));
    if (!DYNAMIC_ONLY) {
        W(ND_CAP_CPU, R(NR_ALLOC_MCPU));
        W(ND_CAP_MEM, R(NR_ALLOC_MEM));
        W(ND_ALLOC_RAW, alloc_raw);
        W(ND_ALLOC_EPH, R(NR_ALLOC_EPH));
        W(ND_MEM_EACH, mem_each);
        for (int s = 0; s < S; ++s) W(nd_alloc_scalar(S, s), R(nr_alloc_scalar(s)));
    }
}

__device__ __forceinline__ void derive_one_node(const DeriveParams& prm, long long n) {
    typedef long long i64;
    const i64 np = prm.n_pad;
    i64* nd = (i64*)prm.nd;
    if (n >= prm.n_nodes) { // padding rows: never feasible, all columns zero
        const int ncols = nd_cols(prm.S);
        for (int c = 0; c < ncols; ++c) nd[(i64)c * np + n] = 0;
        nd[(i64)ND_THR_GPU * np + n] = (i64)0x8000000000000000ull;
        nd[(i64)ND_THR_MEM * np + n] = (i64)0x8000000000000000ull;
        return;
    }
    derive_row(prm, (const i64*)prm.raw + n, np, nd + n, np);
}

} // namespace fgs
