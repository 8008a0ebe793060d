// kernels_narrow.cuh — int32 table layouts and launchers of the narrow evaluation kernels.
#pragma once
#include <cstdint>

namespace fgs {

// resource dimensions that get one exact power-of-two shift each
enum NarrowDim : int {
    DIM_CPU = 0, // milli-cores: free cpu, capacity, non-zero requested, pod requests
    DIM_MEM,     // bytes
    DIM_EPH,
    DIM_S0,      // scalar slot 0 (nvidia.flex.com/gpu) as a Fit.Filter request
    DIM_S1,      // scalar slot 1 (nvidia.flex.com/memory) as a Fit.Filter request
    DIM_FG,      // whole-GPU limits / GPUScore (never shifted; tracked for the range check)
    DIM_FM,      // fractional memory limits / per-device free memory / MemScore
    ND_DIMS
};

struct NarrowShifts {
    int s[ND_DIMS];
};

struct NodeAnalysis {
    unsigned long long or_bits[ND_DIMS];
    long long max_val[ND_DIMS];
    long long araw_min, araw_max;
    int flags; // bit0: a valid node has cpu or memory capacity <= 0; bit1: a MemScore is negative
    int n_special; // valid nodes that feed a pod's error bits: negative MemScore, or NewGPUNode panics on them
};

// Division by an invariant (Granlund–Montgomery, round-up variant) for dividends x < 2^31 and divisors 1 <= d < 2^31:
// with l = ceil(log2 d) and m = floor(2^(31+l) / d) + 1 (m < 2^32), floor(x / d) = floor(x·m / 2^(31+l)) because
// m·d = 2^(31+l) + e with 0 < e <= d <= 2^l and x·e < 2^(31+l). The kernels keep 2·x (< 2^32) and take
// umulhi(2x, m) >> l: two instructions per exact quotient. d == 0 → m = 0: the quotient reads 0, as the reference's
// "capacity == 0 → score 0" (least_allocated.go:48-50, most_allocated.go:52-54).
#if defined(__CUDACC__)
__host__ __device__
#endif
inline uint32_t narrow_magic(uint32_t d, int* l) {
    if (d == 0) {
        *l = 0;
        return 0;
    }
    int lg = 0;
    while ((1ull << lg) < (unsigned long long)d) ++lg;
    *l = lg;
    return (uint32_t)((1ull << (31 + lg)) / d + 1ull);
}

// The same pair without the search loop and the 64-bit division — it sits on every emit CTA's start-up path, where 64 threads
// derive the pods' multipliers while the other warps wait: ceil(log2 d) from the leading zeros, and floor(2^(31+l) / d) from a
// double quotient (within one of it, the quotient being below 2^32) settled by the exact remainder. Equal to narrow_magic
// for every d (tests/cpp/magic_test.cpp: all d up to 2^24 and random ones up to 2^31).
#if defined(__CUDACC__)
__host__ __device__
#endif
inline uint32_t narrow_magic_fast(uint32_t d, int* l) {
    if (d == 0) {
        *l = 0;
        return 0;
    }
#if defined(__CUDA_ARCH__)
    const int lg = d == 1 ? 0 : 32 - __clz((int)(d - 1));
#else
    const int lg = d == 1 ? 0 : 32 - __builtin_clz(d - 1);
#endif
    *l = lg;
    const double q = (double)(1ull << (31 + lg)) / (double)d; // in [2^31, 2^32)
    uint32_t m = (uint32_t)q;
    const long long r = (long long)((1ull << (31 + lg)) - (unsigned long long)m * (unsigned long long)d);
    if (r < 0) --m;
    else if (r >= (long long)d) ++m;
    return m + 1u;
}

// Allocatable.NormalizeScore outside the int32 range (allocatable.go:160: score = (raw − lowest) · 100 / (highest − lowest)):
// (x · 100) / range with Go's int64 semantics — the product wraps, the quotient truncates toward zero. A 64-bit division is a
// ~100-instruction subroutine and the divisor is one per POD, so the pod carries m = floor((2^64 − 1) / range) (range > 0;
// m = 0 says "no reciprocal") and a node multiplies: for a = |x| · 100 < 2^63, q = hi64(a · m) is Q, Q − 1 or Q − 2
// (a · m / 2^64 > a / range − (a / 2^64)(1 + 1 / range) > Q − 2), and the exact remainder settles which. Every other input
// takes the division itself: the result is exact for all of them (tests/cpp/magic_test.cpp).
#if defined(__CUDACC__)
__host__ __device__
#endif
inline unsigned long long alloc_recip(long long range) { return range > 0 ? ~0ull / (unsigned long long)range : 0ull; }

#if defined(__CUDACC__)
__host__ __device__
#endif
inline int alloc_div100(long long x, long long range, unsigned long long m) {
    if (m != 0 && x > -(1ll << 56) && x < (1ll << 56)) {
        const unsigned long long a = (unsigned long long)(x < 0 ? -x : x) * 100ull;
#if defined(__CUDA_ARCH__)
        unsigned long long q = __umul64hi(a, m);
#else
        unsigned long long q = (unsigned long long)(((unsigned __int128)a * m) >> 64);
#endif
        unsigned long long r = a - q * (unsigned long long)range;
        if (r >= (unsigned long long)range) ++q, r -= (unsigned long long)range;
        if (r >= (unsigned long long)range) ++q;
        return (int)(x < 0 ? -(long long)q : (long long)q);
    }
    return (int)((long long)((unsigned long long)x * 100ull) / range);
}

enum NarrowNodeCol : int {
    NN_THR_GPU = 0, NN_THR_MEM, NN_FREE_CPU, NN_FREE_MEM, NN_FREE_EPH, NN_FS0, NN_FS1, NN_FLAGS,
    NN_CAP_C, NN_X_C, NN_CAP_M, NN_X_M, NN_GSCORE, NN_MSCORE,
    NN_MAG_C, NN_MAG_M, NN_MAG_SH, // division by the node's cpu / memory capacity: magic multipliers, shifts (cpu | mem << 8)
    NN_COLS
};

enum NarrowPodCol : int { PN_LIM_G = 0, PN_LIM_M, PN_RC, PN_RM, PN_RE, PN_RS0, PN_RS1, PN_NZC, PN_NZM, PN_KIND, PN_COLS };

struct NarrowParams {
    const int* nn;          // [NN_COLS][n_pad]
    const int* pn;          // [PN_COLS][p_pad]
    const int64_t* nd;      // derived int64 node table (Allocatable raw score column)
    int64_t* reduce_max;
    int64_t* reduce_sum;
    int64_t* select_key;
    uint32_t* bitmap;
    int64_t* scores;
    int64_t n_nodes, n_pad, node_base;
    int64_t n_pods, p_pad;
    int64_t words_per_row, score_stride;
    int pods_per_cta;
    int tile_first; // emit pass: the node tile of blockIdx.x == 0 (the tail tile is launched on its own when it is mostly empty)
    int tail_groups; // emit pass, tail launch: warps per node group A in {1, 2, 4}: the 8 warps split the pods 8/A ways
    int flex_filter_first;
    NarrowShifts sh;
    int w_flex, w_fit, w_alloc;
    int fit_w_cpu, fit_w_mem;
    float fit_inv_wsum;
    uint64_t seed;
    uint32_t fused_spec; // fused pass only: which filters run (SP_FLEX_F / SP_FIT_F), a run-time choice there
};

// special[0][k] = node index | (panics ? 1<<40 : 0), special[1][k] = |MemScore| when it is negative, else 0
void launch_analyze_nodes(const int64_t* nd, int64_t n_nodes, int64_t n_pad, int S, NodeAnalysis* out, int64_t* special, void* stream);
// per pod: RR_NEG_MAG / RR_SCORE_PANIC rows of reduce_max from the feasibility bits of the special nodes
void launch_special_pass(const uint32_t* bitmap, int64_t words_per_row, const int64_t* special, int64_t n_pad, int n_special,
                         const int64_t* pods, int64_t* reduce_max, int64_t n_pods, int64_t p_pad, void* stream);
// the pods that reach NewGPUNode's panic inside FlexGPU.Filter → RR_FILTER_PANIC (narrow tables; after either filter pass)
void launch_filter_panic_pass(const int* nn, const int* pn, const int64_t* special, int64_t n_pad, int n_special, int64_t n_pods, int64_t p_pad,
                              int fit_f, int flex_first, int64_t* reduce_max, void* stream);
void launch_narrow_nodes(const int64_t* nd, int64_t n_pad, const NarrowShifts& sh, int fit_most, int* nn, void* stream);
void launch_narrow_pods(const int64_t* pods, int64_t p_pad, int S, const NarrowShifts& sh, uint32_t fit_ignored, int* pn, void* stream);
uint32_t narrow_spec_bits(bool flex_f, bool fit_f, bool sc_flex, bool sc_fit, bool fit_most, bool sc_alloc, bool w_eq);
// false = no instantiation for that plugin combination (the caller falls back to the int64 kernels)
bool launch_reduce_narrow(const NarrowParams& p, uint32_t spec, void* stream);
bool launch_emit_narrow(const NarrowParams& p, uint32_t spec, void* stream);
int narrow_last_emit_launches(); // 1, or 2 when the tail node tile was launched on its own
// The fused pass (one pass over pod × node; no Allocatable among the score plugins): node orders, then prepass + kernel.
// ev_between (cudaEvent_t, optional) is recorded between the prepass and the main kernel.
size_t narrow_orders_scratch_bytes(int64_t n_nodes);
void launch_narrow_orders(const int* nn, int64_t n_nodes, int64_t n_pad, int* ord, void* scratch, size_t scratch_bytes, void* stream);
bool narrow_fused_available(uint32_t spec);
// part: 1 = the prepass only, 2 = the main kernel only, 3 = both
bool launch_fused_narrow(const NarrowParams& p, uint32_t spec, const int* ord_g, const int* ord_m, void* stream, void* ev_between, int part);

} // namespace fgs
