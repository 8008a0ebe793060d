// seq_kernel.cuh — device side of the sequential scheduling cycles (see sequential.cu for the overview):
// parameters, PTX wrappers, FlexGPU.Reserve on a raw row, the persistent seq_kernel and reserve_kernel.
// Included by sequential.cu only, inside namespace fgs.
#pragma once
namespace {

constexpr i64 kMin64 = (i64)0x8000000000000000ull;
constexpr int kSeqThreads = 256;              // node-owning threads per CTA
constexpr int kSeqWarps = kSeqThreads / 32;  // node-owning warps
constexpr int kSeqBlock = kSeqThreads + 32;   // + the communication warp (warp 0)
constexpr int kSeqTileP = 64;
constexpr int kMaxK = 32;        // nodes per thread (bit mask of feasible nodes)
constexpr int kWordsA = 8;       // mailbox words per CTA, phase A (7 used)
constexpr int kWordsB = 4;       // phase B (3 used)
constexpr int kMaxCluster = 16;  // CTAs of the single-cluster variant

enum SeqExtra : int { SX_ACCT_CPU = 0, SX_ACCT_MEM, SX_ANNOT, SX_NS, SX_COLS };
enum SeqOut : int { SO_COUNT = 0, SO_NODE, SO_SCORE, SO_ERR, SO_DEVICE, SO_COMMITTED, SO_QUOTA, SO_COLS };

struct SeqParams {
    DeriveParams dv;       // raw + derived node tables, plugin args of the derivation
    const i64* pods;       // [pd_cols(S)][p_pad]
    const i64* extra;      // [SX_COLS][p_pad]
    i64* out;              // [SO_COLS][p_pad]
    u64* mail_a;           // [2][G][kWordsA] self-tagged words
    u64* mail_b;           // [2][G][kWordsB]
    u64* prof;             // [G][16] clock sums per section (FGS_SEQ_PROFILE=1), else NULL
    // ElasticQuota tables (CapacityScheduling in the profile), device memory; q_has_eq == NULL: not in the profile
    const uint8_t* q_has_eq; // [NS]
    const i64* q_min;        // [NS][2+S]
    const i64* q_max;
    i64* q_used;             // rewritten by CapacityScheduling.Reserve
    uint32_t* q_used_present;
    i64 n_ns;
    int* status;           // [0]: index of the first pod whose commit left the supported range, INT_MAX none
    i64 n_pods, p_pad;
    int K, state_smem;
    uint32_t filter_mask, score_mask;
    int flex_filter_first;
    i64 w_flex, w_fit, w_alloc;
    int fit_strategy, fit_n_res;
    int fit_res[FGS_MAX_SCORE_RES];
    i64 fit_w[FGS_MAX_SCORE_RES];
    uint32_t fit_ignored;
    u64 seed;
    int key_shift, reserve_flex;
};

// ---- small PTX wrappers -------------------------------------------------------------------------
__device__ __forceinline__ uint32_t s_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(s_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(s_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    uint32_t ok = 0;
    while (!ok) {
        asm volatile(
            "{\n"
            ".reg .pred p;\n"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
            "selp.u32 %0, 1, 0, p;\n"
            "}\n"
            : "=r"(ok)
            : "r"(s_u32(bar)), "r"(parity)
            : "memory");
    }
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(s_u32(dst)),
                 "l"(__cvta_generic_to_global(src)), "r"(bytes), "r"(s_u32(bar))
                 : "memory");
}
// mailbox words: relaxed, gpu scope, never cached in L1
__device__ __forceinline__ void st_word(u64* p, uint32_t payload, uint32_t tag) {
    const u64 w = ((u64)tag << 32) | payload;
    asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(w) : "memory");
}
__device__ __forceinline__ u64 ld_word(const u64* p) {
    u64 w;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(w) : "l"(p) : "memory");
    return w;
}

// the same words through distributed shared memory (one thread-block cluster): remote store into a peer CTA's
// mailbox, local polling
__device__ __forceinline__ uint32_t mapa_u32(uint32_t saddr, uint32_t rank) {
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(saddr), "r"(rank));
    return r;
}
__device__ __forceinline__ void st_word_peer(uint32_t raddr, uint32_t payload, uint32_t tag) {
    const u64 w = ((u64)tag << 32) | payload;
    asm volatile("st.relaxed.cluster.shared::cluster.u64 [%0], %1;" ::"r"(raddr), "l"(w) : "memory");
}
__device__ __forceinline__ u64 ld_word_local(const u64* p) {
    u64 w;
    asm volatile("ld.relaxed.cluster.shared::cta.u64 %0, [%1];" : "=l"(w) : "r"(s_u32(p)) : "memory");
    return w;
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// a real consumer of `v` that cannot be removed: the clock read after it cannot issue before v has arrived
__device__ __forceinline__ void consume(i64 v) {
    asm volatile("{\n.reg .pred q;\nsetp.eq.s64 q, %0, 0x7ff0123456789abc;\n@q trap;\n}\n" ::"l"(v));
}

__device__ __forceinline__ i64 wmul(i64 a, i64 b) { return (i64)((u64)a * (u64)b); }
__device__ __forceinline__ i64 wadd(i64 a, i64 b) { return (i64)((u64)a + (u64)b); }
__device__ __forceinline__ i64 wsub(i64 a, i64 b) { return (i64)((u64)a - (u64)b); }

__device__ __forceinline__ i64 warp_max_i64(i64 v) {
    const int hi = (int)(v >> 32);
    const uint32_t lo = (uint32_t)v;
    const int mhi = __reduce_max_sync(0xffffffffu, hi);
    const uint32_t mlo = __reduce_max_sync(0xffffffffu, hi == mhi ? lo : 0u);
    return (i64)(((u64)(uint32_t)mhi << 32) | mlo);
}

__device__ __forceinline__ float u64_to_float(u64 a) {
    return fmaf(__uint2float_rn((uint32_t)(a >> 32)), 4294967296.0f, __uint2float_rn((uint32_t)a));
}

// Go's int64 x / d (truncation toward zero, wrapping), d != 0. The common case — both positive, quotient below
// 2^18 — avoids the 64-bit division routine: a float estimate is within 1/8 of the true quotient there, so its
// rounding k is the quotient or one above, and the sign of the exact remainder x − k·d decides.
__device__ __forceinline__ i64 go_div(i64 x, i64 d) {
    if (x >= 0 && d > 0 && x < ((i64)1 << 62) && d < ((i64)1 << 62)) {
        const float est = __fdividef(u64_to_float((u64)x), u64_to_float((u64)d));
        if (est < 262144.0f) {
            const int k = __float2int_rn(est);
            const i64 r = x - (i64)((u64)(uint32_t)k * (u64)d);
            return (i64)k + (r >> 63);
        }
    }
    if (d == -1) return (i64)(0ull - (u64)x);
    return x / d;
}

} // namespace

// =================================================================================================
// seq_kernel
// =================================================================================================
namespace {

struct SeqSmem {
    int bars, ring, part_a, fin_a, part_b, fin_b, mail_a, mail_b, fit, state, total;
    int npc; // pod columns per tile
};

__host__ __device__ inline SeqSmem seq_smem(int S, int D, int K, int state_smem) {
    SeqSmem L;
    L.npc = pd_cols(S) + SX_COLS;
    int off = 0;
    L.bars = off;
    off += 64;
    L.ring = off;
    off += 2 * L.npc * kSeqTileP * 8;
    L.part_a = off;
    off += kSeqWarps * 4 * 8;
    L.fin_a = off;
    off += 4 * 8;
    L.part_b = off;
    off += kSeqWarps * 2 * 8;
    L.fin_b = off;
    off += 2 * 8;
    L.mail_a = off; // cluster variant: u64 [2][kMaxCluster][kWordsA]
    off += 2 * kMaxCluster * kWordsA * 8;
    L.mail_b = off;
    off += 2 * kMaxCluster * kWordsB * 8;
    L.fit = off;
    off += K * kSeqThreads * 8;
    off = (off + 127) & ~127;
    L.state = off;
    off += state_smem * (nd_cols(S) + nr_cols(S, D)) * kSeqThreads * K * 8; // state_smem = versions of the state held (0, 1, 2)
    L.total = off;
    return L;
}

// FlexGPU.Reserve on one raw node row (flex_gpu.go:178-223): the device index, or FGS_RESERVE_*.
// mem_each = memCnt / gpuCnt as NewGPUNode computes it (gpu_node.go:55), taken from the derived row.
__device__ __forceinline__ int reserve_on_row(const i64* raw, i64 rs, i64 mem_each, int S, int D, uint32_t kind, i64 mem_limit) {
    if (kind == 0) return FGS_RESERVE_NONE;     // :187-189
    if (kind == 3) return FGS_RESERVE_CONFLICT; // :191-193
    auto R = [&](int c) { return raw[(i64)c * rs]; };
    const u64 meta = (u64)R(NR_META);
    const uint32_t flags = (uint32_t)meta, mono = (uint32_t)(meta >> 32);
    if (flags & FGS_NF_REF_PANIC) return FGS_RESERVE_PANIC; // NewGPUNode, :195
    const i64 gpuCnt = R(nr_alloc_scalar(FGS_SCALAR_FLEX_GPU)); // > 0: REF_PANIC covers the rest
    int dev = FGS_RESERVE_FAIL;
    for (int d = D - 1; d >= 0; --d) { // lowest fitting index wins; no early exit so that the loads overlap
        const i64 used = R(nr_used(S, d));
        const bool m = (mono >> d) & 1u;
        // GPUAssumeFitIndexes, gpu_node.go:163-177 / MemAssumeFitIndexes, :122-161 (indexes[0] is the first fit)
        const bool fit = kind == 1 ? (!m && used == 0) : (!m && mem_each >= used + mem_limit);
        if (fit && (i64)d < gpuCnt) dev = d;
    }
    return dev;
}

// assume + Reserve of one pod on one node row: FlexGPU.Reserve's device index from the CURRENT row (src), then
// NodeInfo.AddPod (framework/types.go:583-609) and what constructGPUs will read next cycle (gpu_node.go:66-120),
// written as a complete row into dst, and dst's derived columns redone. dst may be src (in place) or the node's
// other version. All loads happen before the first store.
struct CommitOut {
    int dev;       // device index or FGS_RESERVE_*
    bool ok;       // Reserve did not fail: the pod can be committed here
    bool in_range; // every aggregate stayed inside |v| < 2^53
};
__device__ __forceinline__ CommitOut commit_row(const SeqParams& prm, const i64* src_raw, const i64* src_nd, i64* dst_raw, i64* dst_nd,
                                                i64 stride, uint32_t kind, i64 gpu_lim, i64 mem_lim, i64 rc, i64 rm, i64 re,
                                                const i64* pod_col0 /* &ring[0][jp], column stride kSeqTileP */) {
    const int S = prm.dv.S, D = prm.dv.D;
    CommitOut co;
    co.dev = FGS_RESERVE_SKIPPED;
    co.ok = true;
    co.in_range = true;
    if (prm.reserve_flex) {
        co.dev = reserve_on_row(src_raw, stride, src_nd[(i64)ND_MEM_EACH * stride], S, D, kind, mem_lim);
        co.ok = co.dev >= 0 || co.dev == FGS_RESERVE_NONE;
    }
    if (!co.ok) return co;
    auto A = [&](int c) { return src_raw[(i64)c * stride]; };
    auto PCc = [&](int c) { return pod_col0[c * kSeqTileP]; };
    const i64 n_rc = A(NR_REQ_MCPU) + rc, n_rm = A(NR_REQ_MEM) + rm, n_re = A(NR_REQ_EPH) + re;
    const i64 n_nzc = A(NR_NZ_MCPU) + PCc(pd_cols(S) + SX_ACCT_CPU), n_nzm = A(NR_NZ_MEM) + PCc(pd_cols(S) + SX_ACCT_MEM);
    const i64 n_np = A(NR_NUM_PODS) + 1;
    // what nodeResourceLimitSum (flex_gpu.go:133-140) will see
    const i64 n_gs = A(NR_FLEX_GPU_SUM) + ((kind & 1u) ? gpu_lim : 0), n_ms = A(NR_FLEX_MEM_SUM) + ((kind & 2u) ? mem_lim : 0);
    u64 meta = (u64)A(NR_META);
    const i64 gpuCnt = A(nr_alloc_scalar(FGS_SCALAR_FLEX_GPU)); // 0 when the key is missing
    // the annotation constructGPUs reads: the index Reserve just set, else the pod's own
    const i64 annot = kind == 0 ? (i64)FGS_ANNOT_NONE : (co.dev >= 0 ? (i64)co.dev : PCc(pd_cols(S) + SX_ANNOT));
    const bool annot_ok = annot != (i64)FGS_ANNOT_NONE && annot >= 0 && annot < gpuCnt;
    if (annot != (i64)FGS_ANNOT_NONE) {
        if (!annot_ok) meta |= FGS_NF_REF_PANIC; // gpus[index] out of range (gpu_node.go:110,115)
        else if (kind & 1u) meta |= 1ull << (32 + (int)annot);
    }
    i64 n_sc[FGS_MAX_SCALARS], n_used[FGS_MAX_DEVICES];
#pragma unroll
    for (int s = 0; s < FGS_MAX_SCALARS; ++s) n_sc[s] = s < S ? A(nr_req_scalar(S, s)) + PCc(PD_FIXED_COLS + s) : 0;
#pragma unroll
    for (int d = 0; d < FGS_MAX_DEVICES; ++d)
        n_used[d] = d < D ? A(nr_used(S, d)) + ((annot_ok && (kind & 2u) && (i64)d == annot) ? mem_lim : 0) : 0;
    // ---- stores: every column a placement can change, so that dst is complete whichever version it is ----
    auto put = [&](int c, i64 v) {
        dst_raw[(i64)c * stride] = v;
        co.in_range = co.in_range && v > -kMaxQuantity && v < kMaxQuantity;
    };
    put(NR_REQ_MCPU, n_rc);
    put(NR_REQ_MEM, n_rm);
    put(NR_REQ_EPH, n_re);
    put(NR_NZ_MCPU, n_nzc);
    put(NR_NZ_MEM, n_nzm);
    put(NR_NUM_PODS, n_np);
    put(NR_FLEX_GPU_SUM, n_gs);
    put(NR_FLEX_MEM_SUM, n_ms);
    dst_raw[(i64)NR_META * stride] = (i64)meta;
#pragma unroll
    for (int s = 0; s < FGS_MAX_SCALARS; ++s)
        if (s < S) put(nr_req_scalar(S, s), n_sc[s]);
#pragma unroll
    for (int d = 0; d < FGS_MAX_DEVICES; ++d)
        if (d < D) put(nr_used(S, d), n_used[d]);
    derive_row<true>(prm.dv, dst_raw, stride, dst_nd, stride);
    return co;
}

// The same commit built by a whole warp for ONE node (the warp's selectHost candidate): lane l owns one of the
// columns a placement can change — 0-7 the fixed aggregates, 8.. the scalar requests, 16.. the devices — so the
// sums, FlexGPU.Reserve's first-fit search (a ballot) and the re-derivation (GPUScore = popcount of a ballot,
// MemScore and the best device by warp reductions) run side by side instead of as one thread's ~450
// dependent instructions. Every lane passes the same arguments. src / dst: column 0 of the node in the two versions.
__device__ __forceinline__ CommitOut commit_row_warp(const SeqParams& prm, const i64* src_raw, const i64* src_nd, i64* dst_raw, i64* dst_nd,
                                                     i64 stride, uint32_t kind, i64 gpu_lim, i64 mem_lim, i64 rc, i64 rm, i64 re,
                                                     const i64* pod_col0, int lane) {
    const int S = prm.dv.S, D = prm.dv.D;
    auto A = [&](int c) { return src_raw[(i64)c * stride]; };
    auto PCc = [&](int c) { return pod_col0[c * kSeqTileP]; };
    // ---- my column, its increment, and the Allocatable value its derived column needs -------------------------
    const bool dev_lane = lane >= 16 && lane - 16 < D;
    const bool sc_lane = lane >= 8 && lane - 8 < S;
    int col = -1, acol = -1;
    i64 delta = 0;
    switch (lane) {
        case 0: col = NR_REQ_MCPU, acol = NR_ALLOC_MCPU, delta = rc; break;
        case 1: col = NR_REQ_MEM, acol = NR_ALLOC_MEM, delta = rm; break;
        case 2: col = NR_REQ_EPH, acol = NR_ALLOC_EPH, delta = re; break;
        case 3: col = NR_NZ_MCPU, delta = PCc(pd_cols(S) + SX_ACCT_CPU); break;
        case 4: col = NR_NZ_MEM, delta = PCc(pd_cols(S) + SX_ACCT_MEM); break;
        case 5: col = NR_NUM_PODS, acol = NR_ALLOC_PODS, delta = 1; break;
        case 6: col = NR_FLEX_GPU_SUM, acol = nr_alloc_scalar(FGS_SCALAR_FLEX_GPU), delta = (kind & 1u) ? gpu_lim : 0; break;
        case 7: col = NR_FLEX_MEM_SUM, acol = nr_alloc_scalar(FGS_SCALAR_FLEX_MEM), delta = (kind & 2u) ? mem_lim : 0; break;
        default: break;
    }
    if (sc_lane) col = nr_req_scalar(S, lane - 8), acol = nr_alloc_scalar(lane - 8), delta = PCc(PD_FIXED_COLS + lane - 8);
    if (dev_lane) col = nr_used(S, lane - 16);
    const i64 v = col >= 0 ? A(col) : 0;
    const i64 al = acol >= 0 ? A(acol) : 0;
    u64 meta = (u64)A(NR_META);
    const i64 gpuCnt = A(nr_alloc_scalar(FGS_SCALAR_FLEX_GPU)); // 0 when the key is missing
    const i64 mem_each = src_nd[(i64)ND_MEM_EACH * stride];
    const int d = lane - 16;
    const bool dev_valid = dev_lane && (i64)d < gpuCnt;

    // ---- FlexGPU.Reserve (flex_gpu.go:178-223): first device in index order that fits ---------------------------
    CommitOut co;
    co.dev = FGS_RESERVE_SKIPPED;
    co.ok = true;
    co.in_range = true;
    if (prm.reserve_flex) {
        const bool m = ((uint32_t)(meta >> 32) >> (d & 31)) & 1u;
        const bool fit = dev_valid && !m && (kind == 1 ? v == 0 : mem_each >= v + mem_lim);
        const uint32_t fits = __ballot_sync(0xffffffffu, fit) >> 16;
        if (kind == 0) co.dev = FGS_RESERVE_NONE;
        else if (kind == 3) co.dev = FGS_RESERVE_CONFLICT;
        else if ((uint32_t)meta & FGS_NF_REF_PANIC) co.dev = FGS_RESERVE_PANIC;
        else co.dev = fits ? __ffs(fits) - 1 : FGS_RESERVE_FAIL;
        co.ok = co.dev >= 0 || co.dev == FGS_RESERVE_NONE;
    }
    if (!co.ok) return co; // uniform

    // ---- assume: NodeInfo.AddPod + the annotation constructGPUs will read ---------------------------------------------
    const i64 annot = kind == 0 ? (i64)FGS_ANNOT_NONE : (co.dev >= 0 ? (i64)co.dev : PCc(pd_cols(S) + SX_ANNOT));
    const bool annot_ok = annot != (i64)FGS_ANNOT_NONE && annot >= 0 && annot < gpuCnt;
    if (annot != (i64)FGS_ANNOT_NONE) {
        if (!annot_ok) meta |= FGS_NF_REF_PANIC;
        else if (kind & 1u) meta |= 1ull << (32 + (int)annot);
    }
    if (dev_lane) delta = (annot_ok && (kind & 2u) && (i64)d == annot) ? mem_lim : 0;
    const i64 nv = v + delta;
    if (col >= 0) dst_raw[(i64)col * stride] = nv;
    if (lane == 15) dst_raw[(i64)NR_META * stride] = (i64)meta;
    co.in_range = !__ballot_sync(0xffffffffu, col >= 0 && !(nv > -kMaxQuantity && nv < kMaxQuantity));

    // ---- re-derive (derive.cuh, DYNAMIC_ONLY) ---------------------------------------------------------------------------
    const uint32_t flags0 = (uint32_t)meta;
    const bool panic = flags0 & FGS_NF_REF_PANIC;
    const bool m2 = ((uint32_t)(meta >> 32) >> (d & 31)) & 1u;
    const bool live = dev_valid && !panic;
    // 64-bit sum of the devices' used memory: lanes 16..31 only
    i64 used_sum = live ? nv : 0;
#pragma unroll
    for (int off = 1; off < 16; off <<= 1) used_sum += __shfl_xor_sync(0xffffffffu, used_sum, off);
    used_sum = __shfl_sync(0xffffffffu, used_sum, 16);
    constexpr i64 kMinI = (i64)0x8000000000000000ull, kMaxI = (i64)0x7fffffffffffffffull;
    i64 max_free = warp_max_i64((live && !m2) ? mem_each - nv : kMinI); // MemAssumeFitIndexes' test, gpu_node.go:136
    if (panic) max_free = kMaxI;
    const i64 gpu_score = __popc(__ballot_sync(0xffffffffu, live && !m2 && nv == 0)); // gpu_node.go:181-186
    const i64 mem_score = panic ? 0 : gpuCnt * mem_each - used_sum;                       // :189-199
    const i64 mine = al - nv; // free cpu / mem / eph / scalar, left pods, gpu / memory headroom
    const i64 gpu_head = __shfl_sync(0xffffffffu, mine, 6), mem_head = __shfl_sync(0xffffffffu, mine, 7);
    const bool pods_ok = __shfl_sync(0xffffffffu, (int)(nv + 1 <= al), 5);
    auto W = [&](int c, i64 x) { dst_nd[(i64)c * stride] = x; };
    const bool has_flex = flags0 & FGS_NF_HAS_FLEX;
    switch (lane) {
        case 0: W(ND_FREE_CPU, mine); break;
        case 1: W(ND_FREE_MEM, mine); break;
        case 2: W(ND_FREE_EPH, mine); break;
        case 3: W(ND_NZR_CPU, nv); break;
        case 4: W(ND_NZR_MEM, nv); break;
        case 5:
            W(ND_LEFT_PODS, mine);
            W(ND_FLAGS, (i64)(flags0 | (mem_score < 0 ? NDF_MEMSCORE_NEG : 0u) | (pods_ok ? NDF_PODS_OK : 0u)));
            break;
        case 6:
            W(ND_THR_GPU, (has_flex && mem_head >= 0) ? gpu_head : kMinI);
            W(ND_GPU_SCORE, gpu_score);
            break;
        case 7:
            W(ND_THR_MEM, (has_flex && gpu_head >= 0) ? (mem_head < max_free ? mem_head : max_free) : kMinI);
            W(ND_MEM_SCORE, mem_score);
            break;
        default: break;
    }
    if (sc_lane) W(nd_free_scalar(lane - 8), mine);
    return co;
}

template <bool SMEM, bool CLUSTER, bool SPEC>
__global__ void __launch_bounds__(kSeqBlock, 1) seq_kernel(const __grid_constant__ SeqParams prm) {
    extern __shared__ __align__(128) unsigned char smem[];
    const int S = prm.dv.S, D = prm.dv.D;
    const int K = prm.K;
    const SeqSmem L = seq_smem(S, D, K, SMEM ? (SPEC ? 2 : 1) : 0);
    uint64_t* bars = (uint64_t*)(smem + L.bars);
    i64* s_ring = (i64*)(smem + L.ring);
    i64* s_part_a = (i64*)(smem + L.part_a);
    i64* s_fin_a = (i64*)(smem + L.fin_a);
    u64* s_part_b = (u64*)(smem + L.part_b);
    u64* s_fin_b = (u64*)(smem + L.fin_b);
    i64* s_fit = (i64*)(smem + L.fit);
    u64* s_mail_a = (u64*)(smem + L.mail_a);
    u64* s_mail_b = (u64*)(smem + L.mail_b);
    int* s_dirty = (int*)(smem + L.bars + 32); // local index of the node whose row the last commit rewrote, -1 none

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const bool comm = warp == 0;        // warp 0 talks to the other CTAs; warps 1..8 own the nodes
    const int ct = tid - 32, cw = warp - 1;
    const int G = (int)gridDim.x, cta = (int)blockIdx.x;
    const i64 np = prm.dv.n_pad, n_nodes = prm.dv.n_nodes;
    const i64 cta_nodes = (i64)kSeqThreads * K;
    const i64 node0 = (i64)cta * cta_nodes;
    const int ncols = nd_cols(S);
    const int nrc = nr_cols(S, D);
    const int npc = L.npc;
    const i64 n_tiles = (prm.n_pods + kSeqTileP - 1) / kSeqTileP;

    auto issue_pod_tile = [&](i64 it) { // one elected thread
        const int stage = (int)(it & 1);
        const i64 p0 = it * kSeqTileP;
        i64* dst = s_ring + (size_t)stage * npc * kSeqTileP;
        mbar_expect_tx(&bars[stage], (uint32_t)(npc * kSeqTileP * 8));
        for (int c = 0; c < pd_cols(S); ++c) bulk_g2s(dst + c * kSeqTileP, prm.pods + (i64)c * prm.p_pad + p0, kSeqTileP * 8, &bars[stage]);
        for (int c = 0; c < SX_COLS; ++c)
            bulk_g2s(dst + (pd_cols(S) + c) * kSeqTileP, prm.extra + (i64)c * prm.p_pad + p0, kSeqTileP * 8, &bars[stage]);
    };

    if (tid == 0) {
        mbar_init(&bars[0], 1);
        mbar_init(&bars[1], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        *s_dirty = -1;
        s_dirty[1] = 0;
    }
    __syncthreads();
    if (tid == 0 && n_tiles > 0) issue_pod_tile(0);

    // ---- node state of the nodes this CTA owns: st = derived columns, rw = raw columns; element (c, i) at
    // [c * stride + i], i = node − node0 ----------------------------------------------------------------------
    // SPEC (one node per thread, both versions in shared memory): a node has two row versions; the commit of a
    // cycle is BUILT into the inactive one while exchange (B) is in flight, and the winner only flips its `cur`.
    i64 *st, *rw;
    i64 st_stride;
    const i64 vstride = (i64)(ncols + nrc) * cta_nodes; // elements per version
    int cur = 0;
    if (SMEM) {
        st = (i64*)(smem + L.state);
        rw = st + (i64)ncols * cta_nodes;
        st_stride = cta_nodes;
        for (int c = 0; c < ncols + nrc; ++c)
            for (i64 i = tid; i < cta_nodes; i += kSeqBlock) {
                const i64 n = node0 + i;
                const i64* src = c < ncols ? (const i64*)prm.dv.nd + (i64)c * np : (const i64*)prm.dv.raw + (i64)(c - ncols) * np;
                const i64 v = n < n_nodes ? src[n] : 0;
                st[(i64)c * st_stride + i] = v;
                if (SPEC) st[vstride + (i64)c * st_stride + i] = v; // the second version of every row
            }
    } else {
        st = (i64*)prm.dv.nd + node0;
        rw = (i64*)prm.dv.raw + node0;
        st_stride = np;
    }
    if (CLUSTER) {
        for (int k = tid; k < 2 * kMaxCluster * (kWordsA + kWordsB); k += kSeqBlock) s_mail_a[k] = 0; // mail_b follows mail_a
        cluster_sync_all(); // no peer writes into a mailbox that is not cleared yet
    }
    __syncthreads();

    const bool use_flex_filter = prm.filter_mask & FGS_PL_FLEXGPU;
    const bool use_fit_filter = prm.filter_mask & FGS_PL_FIT;
    const bool sc_flex = prm.score_mask & FGS_PL_FLEXGPU;
    const bool sc_fit = prm.score_mask & FGS_PL_FIT;
    const bool sc_alloc = prm.score_mask & FGS_PL_ALLOCATABLE;

    long long pc[6] = {0, 0, 0, 0, 0, 0}, pt = 0;
    const bool prof = prm.prof != nullptr && tid == 32;
    auto tick = [&](int k) {
        if (prof) {
            const long long now = clock64();
            pc[k] += now - pt;
            pt = now;
        }
    };
    const bool cprof = prm.prof != nullptr && tid == 0;
    long long cc[4] = {0, 0, 0, 0}, cpt = 0;
    auto ctick = [&](int k) {
        if (cprof) {
            const long long now = clock64();
            cc[k] += now - cpt;
            cpt = now;
        }
    };
    if (prof) pt = clock64();
    if (cprof) cpt = clock64();
    // communication warp, shared-memory state only: copy the row the last commit rewrote to the global tables
    // (off the critical path: the owner thread only flags it)
    auto flush_dirty = [&]() {
        const int dv = *(volatile int*)s_dirty;
        if (dv >= 0) {
            const int di = dv & 0xffff;
            const i64* vs = (const i64*)(smem + L.state) + (i64)(dv >> 16) * vstride; // the row's current version
            const i64 n = node0 + di;
            for (int c = lane; c < ncols + nrc; c += 32) {
                if (c < ncols) ((i64*)prm.dv.nd)[(i64)c * np + n] = vs[(i64)c * st_stride + di];
                else ((i64*)prm.dv.raw)[(i64)(c - ncols) * np + n] = vs[(i64)c * st_stride + di];
            }
            __syncwarp();
            if (lane == 0) *s_dirty = -1;
        }
    };

    // ---- ElasticQuota state: owned by the communication warp of CTA 0 (lane c holds resource column c) ---------
    const bool quota_on = prm.q_has_eq != nullptr;
    const bool quota_warp = quota_on && comm && cta == 0;
    const int QW = 2 + S;
    i64 agg_used = 0, agg_min = 0; // Σ over namespaces with a quota (aggregatedUsedOverMinWith, elasticquota.go:40-51)
    uint32_t agg_present = 0;
    bool pend_has = false, pend_key = false; // the previous pod's reservation, applied once its commit is known
    int pend_ns = -1;
    i64 pend_v = 0;
    uint32_t pend_keys = 0;
    if (quota_warp) {
        for (i64 k = 0; k < prm.n_ns; ++k) {
            if (!prm.q_has_eq[k]) continue;
            const uint32_t up = prm.q_used_present[k];
            agg_present |= up;
            if (lane < QW) {
                if (lane < 2 || ((up >> (lane - 2)) & 1u)) agg_used += prm.q_used[k * QW + lane];
                agg_min += prm.q_min[k * QW + lane];
            }
        }
    }
    int* s_commit = s_dirty + 1; // 1: the previous cycle's pod was committed by a thread of this CTA

    // with quotas one more (drain) iteration carries the last pod's commit flag to the quota warp
    const i64 n_iter = prm.n_pods + (quota_on ? 1 : 0);
    for (i64 p = 0; p < n_iter; ++p) {
        const bool drain = p >= prm.n_pods;
        const i64 it = p / kSeqTileP;
        const int jp = (int)(p % kSeqTileP), stage = (int)(it & 1);
        const uint32_t tag = (uint32_t)(p + 1);
        const int par = (int)(p & 1);
        const i64* praw = s_ring + (size_t)stage * npc * kSeqTileP;
        auto PC = [&](int c) { return praw[c * kSeqTileP + jp]; };

        if (comm) {
            // ================= the communication warp =====================================================
            // CapacityScheduling.PreFilter of this pod (capacity_scheduling.go:201-275, no nominated pods), while
            // the compute warps filter: once against the quota table as it is and once with the previous pod's
            // reservation added — which of the two holds is known when that pod's commit flag arrives with (A)
            uint32_t qbits = 0;
            bool q_has = false, q_key = false;
            int q_ns = -1;
            i64 q_v = 0;
            uint32_t q_keys = 0;
            if (quota_warp && !drain) {
                if (jp == 0) mbar_wait(&bars[stage], (uint32_t)((it >> 1) & 1));
                q_ns = (int)PC(pd_cols(S) + SX_NS);
                q_keys = (uint32_t)((u64)PC(PD_META) >> 32);
                q_v = lane == 0 ? PC(PD_REQ_CPU) : (lane == 1 ? PC(PD_REQ_MEM) : (lane < QW ? PC(PD_FIXED_COLS + lane - 2) : 0));
                q_has = q_ns >= 0 && q_ns < prm.n_ns && prm.q_has_eq[q_ns];
                q_key = lane < 2 || (lane < QW && ((q_keys >> (lane - 2)) & 1u));
                if (q_has) { // a namespace without an ElasticQuota passes (:210-217)
                    const uint32_t up = *(volatile uint32_t*)(prm.q_used_present + q_ns);
                    const bool u_key = lane < 2 || (lane < QW && ((up >> (lane - 2)) & 1u));
                    const i64 u = u_key ? *(volatile i64*)(prm.q_used + (i64)q_ns * QW + lane) : 0;
                    const i64 mx = lane < QW ? prm.q_max[(i64)q_ns * QW + lane] : 0;
                    const i64 d_same = (pend_has && pend_ns == q_ns && pend_key) ? pend_v : 0;
                    const i64 d_all = (pend_has && pend_key) ? pend_v : 0;
                    // usedOverMaxWith = cmp2(request, Used, Max): the request's scalar keys (elasticquota.go:94-96,165-181)
                    const uint32_t m0 = __ballot_sync(0xffffffffu, q_key && q_v + u > mx);
                    const uint32_t m1 = __ballot_sync(0xffffffffu, q_key && q_v + u + d_same > mx);
                    // aggregatedUsedOverMinWith: keys of Σused ∪ the request's (:40-51)
                    const uint32_t ak0 = agg_present | q_keys, ak1 = ak0 | (pend_has ? pend_keys : 0u);
                    const bool a0 = lane < 2 || (lane < QW && ((ak0 >> (lane - 2)) & 1u));
                    const bool a1 = lane < 2 || (lane < QW && ((ak1 >> (lane - 2)) & 1u));
                    const uint32_t n0 = __ballot_sync(0xffffffffu, a0 && agg_used + q_v > agg_min);
                    const uint32_t n1 = __ballot_sync(0xffffffffu, a1 && agg_used + d_all + q_v > agg_min);
                    const uint32_t code0 = m0 ? 1u : (n0 ? 2u : 0u), code1 = m1 ? 1u : (n1 ? 2u : 0u);
                    qbits = (code0 << 1) | (code1 << 3);
                }
            }
            __syncthreads(); // (1) the compute warps' phase-A partials are in shared memory
            if (lane == 0 && !drain && jp == 0 && it + 1 < n_tiles) { // every thread is past its last read of the previous tile
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                issue_pod_tile(it + 1);
            }
            const bool w = lane < kSeqWarps;
            const i64 cf = w ? s_part_a[lane * 4 + 0] : 0;
            if (cprof) consume(cf);
            ctick(0); // commit + Filter of the slowest thread of this CTA
            uint32_t c2 = __reduce_add_sync(0xffffffffu, (uint32_t)cf & 0xffffffu);
            uint32_t f2 = __reduce_or_sync(0xffffffffu, (uint32_t)(cf >> 24));
            if (quota_on) { // bits 1-4: the two PreFilter verdicts (CTA 0); bit 5: the previous pod was committed here
                const int cm = *(volatile int*)s_commit;
                f2 |= qbits | (cm ? 32u : 0u);
                __syncwarp();
                if (lane == 0 && cm) *s_commit = 0;
            }
            i64 a1 = warp_max_i64(w ? s_part_a[lane * 4 + 1] : kMin64);
            i64 a2 = warp_max_i64(w ? s_part_a[lane * 4 + 2] : kMin64);
            i64 a3 = warp_max_i64(w ? s_part_a[lane * 4 + 3] : kMin64);
            if (G > 1) {
                uint32_t pay = 0;
                if (lane == 0) pay = c2 | (f2 << 24);
                else if (lane < 7) {
                    const i64 v = lane <= 2 ? a1 : (lane <= 4 ? a2 : a3);
                    pay = (lane & 1) ? (uint32_t)v : (uint32_t)((u64)v >> 32);
                }
                if (CLUSTER) {
                    const uint32_t slot = s_u32(s_mail_a + ((size_t)par * kMaxCluster + cta) * kWordsA + lane);
                    if (lane < 7)
                        for (int dst = 0; dst < G; ++dst) st_word_peer(mapa_u32(slot, (uint32_t)dst), pay, tag);
                } else if (lane < 7) {
                    st_word(prm.mail_a + ((size_t)par * G + cta) * kWordsA + lane, pay, tag);
                }
                uint32_t tc = 0, tf = 0;
                i64 t1 = kMin64, t2 = kMin64, t3 = kMin64;
                for (int c = lane; c < G; c += 32) {
                    const u64* m = CLUSTER ? s_mail_a + ((size_t)par * kMaxCluster + c) * kWordsA : prm.mail_a + ((size_t)par * G + c) * kWordsA;
                    u64 wv[7];
                    bool ok;
                    do {
                        ok = true;
#pragma unroll
                        for (int k = 0; k < 7; ++k) {
                            wv[k] = CLUSTER ? ld_word_local(m + k) : ld_word(m + k);
                            ok = ok && (uint32_t)(wv[k] >> 32) == tag;
                        }
                    } while (!ok);
                    tc += (uint32_t)wv[0] & 0xffffffu;
                    tf |= ((uint32_t)wv[0] >> 24) & 0xffu;
                    const i64 x1 = (i64)((wv[2] << 32) | (wv[1] & 0xffffffffull));
                    const i64 x2 = (i64)((wv[4] << 32) | (wv[3] & 0xffffffffull));
                    const i64 x3 = (i64)((wv[6] << 32) | (wv[5] & 0xffffffffull));
                    t1 = x1 > t1 ? x1 : t1;
                    t2 = x2 > t2 ? x2 : t2;
                    t3 = x3 > t3 ? x3 : t3;
                }
                c2 = __reduce_add_sync(0xffffffffu, tc);
                f2 = __reduce_or_sync(0xffffffffu, tf);
                a1 = warp_max_i64(t1);
                a2 = warp_max_i64(t2);
                a3 = warp_max_i64(t3);
            }
            if (cprof) consume((i64)c2);
            ctick(1); // exchange A, including the wait for the slowest CTA
            uint32_t qcode = 0;
            if (quota_on) {
                const bool committed_prev = (f2 >> 5) & 1u;
                qcode = committed_prev ? (f2 >> 3) & 3u : (f2 >> 1) & 3u;
                if (quota_warp) {
                    if (pend_has && committed_prev) { // CapacityScheduling.Reserve → reserveResource (elasticquota.go:74-80)
                        if (pend_key) {
                            const uint32_t up = *(volatile uint32_t*)(prm.q_used_present + pend_ns);
                            const bool u_key = lane < 2 || ((up >> (lane - 2)) & 1u);
                            i64* up_ = prm.q_used + (i64)pend_ns * QW + lane;
                            *(volatile i64*)up_ = (u_key ? *(volatile i64*)up_ : 0) + pend_v;
                            agg_used += pend_v;
                        }
                        __syncwarp();
                        if (lane == 0) *(volatile uint32_t*)(prm.q_used_present + pend_ns) |= pend_keys;
                        __syncwarp();
                        agg_present |= pend_keys;
                    }
                    pend_has = q_has && qcode == 0 && !drain;
                    pend_ns = q_ns;
                    pend_v = q_v;
                    pend_keys = q_keys;
                    pend_key = q_key;
                }
                if (qcode) c2 = 0; // rejected in PreFilter: no Filter/Score results, nothing to select
                f2 = (f2 & 1u) | (qcode << 8);
            }
            if (lane == 0) {
                s_fin_a[0] = (i64)(((u64)f2 << 32) | c2);
                s_fin_a[1] = a1;
                s_fin_a[2] = a2;
                s_fin_a[3] = a3;
            }
            __syncthreads(); // (2)
            if (SMEM) flush_dirty();
            if (c2 > 0) {    // uniform over the grid: phase B runs
                __syncthreads(); // (3) the compute warps' keys are in shared memory
                u64 k2 = (u64)warp_max_i64(w ? (i64)s_part_b[lane * 2 + 0] : 0);
                if (cprof) consume((i64)k2);
                ctick(2); // Fit score + NormalizeScore of the slowest thread of this CTA
                uint32_t e2 = __reduce_or_sync(0xffffffffu, w ? (uint32_t)s_part_b[lane * 2 + 1] : 0u);
                if (G > 1) {
                    const uint32_t pay = lane == 0 ? (uint32_t)k2 : (lane == 1 ? (uint32_t)(k2 >> 32) : e2);
                    if (CLUSTER) {
                        const uint32_t slot = s_u32(s_mail_b + ((size_t)par * kMaxCluster + cta) * kWordsB + lane);
                        if (lane < 3)
                            for (int dst = 0; dst < G; ++dst) st_word_peer(mapa_u32(slot, (uint32_t)dst), pay, tag);
                    } else if (lane < 3) {
                        st_word(prm.mail_b + ((size_t)par * G + cta) * kWordsB + lane, pay, tag);
                    }
                    u64 tk = 0;
                    uint32_t te = 0;
                    for (int c = lane; c < G; c += 32) {
                        const u64* m = CLUSTER ? s_mail_b + ((size_t)par * kMaxCluster + c) * kWordsB : prm.mail_b + ((size_t)par * G + c) * kWordsB;
                        u64 w0, w1, w2;
                        do {
                            w0 = CLUSTER ? ld_word_local(m + 0) : ld_word(m + 0);
                            w1 = CLUSTER ? ld_word_local(m + 1) : ld_word(m + 1);
                            w2 = CLUSTER ? ld_word_local(m + 2) : ld_word(m + 2);
                        } while ((uint32_t)(w0 >> 32) != tag || (uint32_t)(w1 >> 32) != tag || (uint32_t)(w2 >> 32) != tag);
                        const u64 x = (w1 << 32) | (w0 & 0xffffffffull);
                        tk = x > tk ? x : tk;
                        te |= (uint32_t)w2;
                    }
                    k2 = (u64)warp_max_i64((i64)tk);
                    e2 = __reduce_or_sync(0xffffffffu, te);
                }
                if (cprof) consume((i64)k2);
                ctick(3); // exchange B
                if (lane == 0) {
                    s_fin_b[0] = k2;
                    s_fin_b[1] = e2;
                }
                __syncthreads(); // (4)
            }
            continue;
        }

        // ================= the compute warps ===============================================================
        if (!drain && jp == 0) mbar_wait(&bars[stage], (uint32_t)((it >> 1) & 1));
        // per-pod values (broadcast reads)
        const u64 pmeta = (u64)PC(PD_META);
        const uint32_t pf = (uint32_t)pmeta;
        const uint32_t kind = pf & 3u;
        const bool req_none = pf & FGS_PF_REQ_NONE;
        const uint32_t smask = (uint32_t)(pmeta >> 32) & ~prm.fit_ignored;
        const i64 gpu_lim = PC(PD_GPU_LIM), mem_lim = PC(PD_MEM_LIM);
        const i64 lim = kind == 1 ? gpu_lim : mem_lim;
        const i64 rc = PC(PD_REQ_CPU), rm = PC(PD_REQ_MEM), re = PC(PD_REQ_EPH);
        const i64 nzc = PC(PD_NZ_CPU), nzm = PC(PD_NZ_MEM);

        // ---- phase A: Filter on my nodes ---------------------------------------------------------------------
        uint32_t feas_mask = 0, cnt = 0, fl = 0;
        i64 vflex = kMin64, vhi = kMin64, vnlo = kMin64;
        const bool flex_pod = use_flex_filter && kind != 0;
        const bool fit_req = use_fit_filter && !req_none;
        const int Kc = drain ? 0 : K; // the drain iteration only carries flags
        for (int j = 0; j < Kc; ++j) {
            const i64 i = (i64)j * kSeqThreads + ct;
            if (node0 + i >= n_nodes) break;
            auto C = [&](int c) { return st[(i64)c * st_stride + i]; };
            // every column the verdict can need, loaded before any of them is used (one shared-memory latency)
            const uint32_t nf = (uint32_t)C(ND_FLAGS);
            const i64 thr = C(kind == 1 ? ND_THR_GPU : ND_THR_MEM);
            const i64 fc = C(ND_FREE_CPU), fm_ = C(ND_FREE_MEM), fe = C(ND_FREE_EPH);
            const i64 sflex = C((kind & 1u) ? ND_GPU_SCORE : ND_MEM_SCORE); // a gpu limit wins over a memory limit (flex_gpu.go:157-159)
            const i64 ar = C(ND_ALLOC_RAW);
            // FlexGPU.Filter, flex_gpu.go:41-119, folded into one threshold per pod kind (derive.cuh)
            const bool flex_ok = !flex_pod || (kind != 3 && lim <= thr);
            // Fit.Filter, fit.go:256-329
            bool fit_ok = !use_fit_filter || (nf & NDF_PODS_OK);
            if (fit_req) {
                fit_ok = fit_ok & (rc <= fc) & (rm <= fm_) & (re <= fe);
                uint32_t m = smask;
                while (m) {
                    const int s = __ffs(m) - 1;
                    m &= m - 1;
                    if (PC(PD_FIXED_COLS + s) > C(nd_free_scalar(s))) fit_ok = false;
                }
            }
            const bool valid = nf & FGS_NF_VALID;
            // the reference panics inside NewGPUNode (flex_gpu.go:101) when FlexGPU.Filter gets that far
            const bool reach = flex_pod && kind != 3 && flex_ok && (nf & FGS_NF_REF_PANIC) && (prm.flex_filter_first || fit_ok) && valid;
            if (reach) fl |= 1u;
            if (flex_ok && fit_ok && valid && !reach) {
                feas_mask |= 1u << j;
                cnt++;
                const i64 s = kind != 0 ? sflex : 0;
                vflex = s > vflex ? s : vflex;
                vhi = ar > vhi ? ar : vhi;
                vnlo = -ar > vnlo ? -ar : vnlo;
            }
        }
        cnt = __reduce_add_sync(0xffffffffu, cnt);
        fl = __reduce_or_sync(0xffffffffu, fl);
        if (sc_flex) vflex = warp_max_i64(vflex);
        if (sc_alloc) {
            vhi = warp_max_i64(vhi);
            vnlo = warp_max_i64(vnlo);
        }
        if (lane == 0) {
            s_part_a[cw * 4 + 0] = (i64)(cnt | (fl << 24));
            s_part_a[cw * 4 + 1] = vflex;
            s_part_a[cw * 4 + 2] = vhi;
            s_part_a[cw * 4 + 3] = vnlo;
        }
        tick(0); // pod fetch + Filter on my nodes
        __syncthreads(); // (1)

        // ---- while the communication warp exchanges the phase-A results: the part of Score that needs none of
        // them — Fit's Least/MostAllocated (least_allocated.go:29-57, most_allocated.go:29-61,
        // resource_allocation.go:45-117) on my feasible nodes -----------------------------------------------
        if (sc_fit) {
            uint32_t fm = feas_mask;
            while (fm) {
                const int j = __ffs(fm) - 1;
                fm &= fm - 1;
                const i64 i = (i64)j * kSeqThreads + ct;
                auto C = [&](int c) { return st[(i64)c * st_stride + i]; };
                i64 node_score = 0, weight_sum = 0;
                for (int k = 0; k < prm.fit_n_res; ++k) {
                    const int res = prm.fit_res[k];
                    i64 alloc = 0, req = 0;
                    if (res == FGS_RES_CPU) {
                        alloc = C(ND_CAP_CPU);
                        req = C(ND_NZR_CPU) + nzc;
                    } else if (res == FGS_RES_MEM) {
                        alloc = C(ND_CAP_MEM);
                        req = C(ND_NZR_MEM) + nzm;
                    } else if (res == FGS_RES_EPH) {
                        alloc = C(ND_ALLOC_EPH);
                        req = (alloc - C(ND_FREE_EPH)) + re;
                    } else if (res >= FGS_RES_SCALAR0 && res < FGS_RES_SCALAR0 + S) {
                        const int s = res - FGS_RES_SCALAR0;
                        const i64 pr = PC(PD_FIXED_COLS + s);
                        if (pr != 0) { // resource_allocation.go:93-97
                            alloc = C(nd_alloc_scalar(S, s));
                            req = (alloc - C(nd_free_scalar(s))) + pr;
                        }
                    }
                    if (alloc == 0) continue; // resource_allocation.go:64-67
                    i64 rs;
                    if (prm.fit_strategy == 0) rs = req > alloc ? 0 : go_div(wmul(alloc - req, 100), alloc);
                    else rs = go_div(wmul(req > alloc ? alloc : req, 100), alloc);
                    node_score = wadd(node_score, wmul(rs, prm.fit_w[k]));
                    weight_sum = wadd(weight_sum, prm.fit_w[k]);
                }
                s_fit[j * kSeqThreads + ct] = weight_sum == 0 ? 0 : go_div(node_score, weight_sum);
            }
        }
        tick(1); // Fit score, overlapped with exchange A
        __syncthreads(); // (2)
        const i64 count = (i64)(uint32_t)s_fin_a[0];
        if (prof) consume(count);
        tick(2); // waiting for exchange A
        const uint32_t qcode = (uint32_t)(s_fin_a[0] >> 40) & 3u; // CapacityScheduling.PreFilter: 0 pass, 1 over Max, 2 over Σmin
        const bool fpanic = ((s_fin_a[0] >> 32) & 1) && qcode == 0;
        const i64 g_flex = s_fin_a[1], g_hi = s_fin_a[2], g_nlo = s_fin_a[3];

        // ---- phase B: NormalizeScore / weights / selectHost key on my feasible nodes -------------------------
        u64 key = 0;     // 0 = no candidate
        uint32_t eb = 0; // FGS_PODERR_* raised while scoring
        CommitOut spec;  // SPEC: the commit of this pod on my node, built ahead of the verdict
        spec.dev = FGS_RESERVE_SKIPPED;
        spec.ok = false;
        spec.in_range = true;
        if (count > 0) { // uniform over the grid
            const i64 fmax = g_flex > 0 ? g_flex : 0; // maxCount starts at 0 (normalize_score.go:27)
            const i64 alo = -g_nlo;
            const i64 arange = wsub(g_hi, alo);
            uint32_t ph = (uint32_t)(prm.seed ^ (prm.seed >> 32)) ^ ((uint32_t)p * 0x9E3779B1u);
            ph ^= ph >> 16;
            ph *= 0x85EBCA6Bu;
            ph ^= ph >> 13;
            uint32_t fm = feas_mask;
            while (fm) {
                const int j = __ffs(fm) - 1;
                fm &= fm - 1;
                const i64 i = (i64)j * kSeqThreads + ct;
                auto C = [&](int c) { return st[(i64)c * st_stride + i]; };
                i64 t = 0;
                if (sc_flex) {
                    // FlexGPU.Score (flex_gpu.go:142-166) calls NewGPUNode first, for every pod kind (:148)
                    if ((uint32_t)C(ND_FLAGS) & FGS_NF_REF_PANIC) eb |= FGS_PODERR_REF_PANIC;
                    const i64 s = (kind & 1u) ? C(ND_GPU_SCORE) : (kind == 2 ? C(ND_MEM_SCORE) : 0);
                    // DefaultNormalizeScore, reverse (normalize_score.go:26-54)
                    const i64 v = fmax == 0 ? 100 : 100 - go_div(wmul(100, s), fmax);
                    if (v > 100 || v < 0) eb |= FGS_PODERR_SCORE_RANGE; // runtime/framework.go:920-923
                    t = wadd(t, wmul(v, prm.w_flex));
                }
                if (sc_fit) {
                    const i64 v = s_fit[j * kSeqThreads + ct];
                    if (v > 100 || v < 0) eb |= FGS_PODERR_SCORE_RANGE;
                    t = wadd(t, wmul(v, prm.w_fit));
                }
                if (sc_alloc) { // allocatable.go:61-72,146-171
                    const i64 v = arange == 0 ? 0 : go_div(wmul(wsub(C(ND_ALLOC_RAW), alo), 100), arange);
                    if (v > 100 || v < 0) eb |= FGS_PODERR_SCORE_RANGE;
                    t = wadd(t, wmul(v, prm.w_alloc));
                }
                // selectHost candidate (generic_scheduler.go:146-167): same key as the batched kernels
                const uint32_t ng = (uint32_t)(node0 + i);
                u64 k;
                if (prm.key_shift == 40) {
                    uint32_t h = (ph ^ ng) * 0x9E3779B1u;
                    h ^= h >> 15;
                    k = ((u64)t << 40) | ((u64)(h & 0xffffu) << 24) | ng;
                } else {
                    k = ((u64)t << 24) | ng;
                }
                k += 1;
                key = k > key ? k : key;
            }
            const u64 my_key = key;
            key = (u64)warp_max_i64((i64)key);
            eb = __reduce_or_sync(0xffffffffu, eb);
            if (lane == 0) {
                s_part_b[cw * 2 + 0] = key;
                s_part_b[cw * 2 + 1] = eb;
            }
            tick(3); // NormalizeScore + key on my feasible nodes
            __syncthreads(); // (3)
            if (SPEC) {
                // this warp's candidate (its maximal key) is the only one of its nodes that can win: build that node's
                // row after this pod into the node's inactive version while the communication warp exchanges the keys;
                // it becomes current only if the node wins
                if (key != 0) { // warp-uniform after the reduction
                    const int ol = __ffs(__ballot_sync(0xffffffffu, my_key == key)) - 1;
                    const int ocur = __shfl_sync(0xffffffffu, cur, ol);
                    const i64 oi = (i64)cw * 32 + ol; // one node per thread: local index = compute-thread index
                    const i64* sb = (const i64*)(smem + L.state) + (i64)ocur * vstride;
                    i64* nb = (i64*)(smem + L.state) + (i64)(ocur ^ 1) * vstride;
                    const CommitOut co = commit_row_warp(prm, sb + (i64)ncols * cta_nodes + oi, sb + oi, nb + (i64)ncols * cta_nodes + oi, nb + oi,
                                                         st_stride, kind, gpu_lim, mem_lim, rc, rm, re, praw + jp, lane);
                    if (lane == ol) spec = co;
                    __syncwarp(); // the owner reads what the other lanes stored once it has flipped
                }
            }
            __syncthreads(); // (4)
            key = s_fin_b[0];
            if (prof) consume((i64)key);
            tick(4); // waiting for exchange B
            eb = (uint32_t)s_fin_b[1];
        }

        // ---- verdict of the cycle (every compute thread computes the same) ------------------------------------
        uint32_t err = fpanic ? FGS_PODERR_REF_PANIC : 0u;
        if (count >= 2) err |= eb; // scoring only runs with >= 2 feasible nodes (generic_scheduler.go:112-127)
        i64 sel = -1, score = 0;
        if (key > 0) {
            const u64 k = key - 1;
            sel = (i64)(k & 0xffffffu);
            score = (i64)(k >> prm.key_shift);
        }
        const bool mine = sel >= node0 && sel < node0 + cta_nodes && (int)((sel - node0) % kSeqThreads) == ct;
        if (!drain && (sel < 0 ? (cta == 0 && ct == 0) : mine)) {
            int dev = FGS_RESERVE_SKIPPED;
            bool commit = sel >= 0 && err == 0;
            if (commit) {
                const i64 i = sel - node0;
                CommitOut co;
                if (SPEC) co = spec; // built by my warp while the keys were exchanged (my only node: i == ct)
                else co = commit_row(prm, rw + i, st + i, rw + i, st + i, st_stride, kind, gpu_lim, mem_lim, rc, rm, re, praw + jp);
                dev = co.dev;
                if (dev == FGS_RESERVE_PANIC) err |= FGS_PODERR_REF_PANIC;
                commit = co.ok;
                if (commit) {
                    if (SPEC) { // the version built ahead becomes the node's row
                        cur ^= 1;
                        st = (i64*)(smem + L.state) + (i64)cur * vstride;
                        rw = st + (i64)ncols * cta_nodes;
                    }
                    if (!co.in_range) atomicMin(prm.status, (int)(p < 0x7fffffff ? p : 0x7ffffffe));
                    if (SMEM) *s_dirty = (int)i | (cur << 16); // the communication warp copies the row to the global tables
                    if (quota_on) *s_commit = 1; // → CapacityScheduling.Reserve of this pod, applied by the quota warp
                }
            }
            prm.out[(i64)SO_COUNT * prm.p_pad + p] = count;
            prm.out[(i64)SO_NODE * prm.p_pad + p] = sel;
            prm.out[(i64)SO_SCORE * prm.p_pad + p] = score;
            prm.out[(i64)SO_ERR * prm.p_pad + p] = (i64)err;
            prm.out[(i64)SO_DEVICE * prm.p_pad + p] = (i64)dev;
            prm.out[(i64)SO_COMMITTED * prm.p_pad + p] = commit ? 1 : 0;
            prm.out[(i64)SO_QUOTA * prm.p_pad + p] = (i64)qcode;
        }
        tick(5); // verdict + commit (this thread's share)
    }
    __syncthreads();
    if (SMEM && comm) flush_dirty();
    if (prof)
        for (int k = 0; k < 6; ++k) prm.prof[(size_t)cta * 16 + k] = (u64)pc[k];
    if (cprof)
        for (int k = 0; k < 4; ++k) prm.prof[(size_t)cta * 16 + 8 + k] = (u64)cc[k];
    if (CLUSTER) cluster_sync_all(); // no CTA leaves while a peer could still address its shared memory
}

// FlexGPU.Reserve for a batch of (pod, node) pairs, no commit
__global__ void reserve_kernel(const i64* __restrict__ raw, const i64* __restrict__ nd, i64 np, i64 n_nodes, int S, int D, const i64* __restrict__ pods, i64 p_pad,
                               i64 n_pods, const int* __restrict__ node, int* __restrict__ dev) {
    const i64 p = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n_pods) return;
    const int n = node[p];
    if (n < 0 || n >= n_nodes) {
        dev[p] = FGS_RESERVE_SKIPPED;
        return;
    }
    const uint32_t kind = (uint32_t)pods[(i64)PD_META * p_pad + p] & 3u;
    dev[p] = reserve_on_row(raw + n, np, nd[(i64)ND_MEM_EACH * np + n], S, D, kind, pods[(i64)PD_MEM_LIM * p_pad + p]);
}

} // namespace
