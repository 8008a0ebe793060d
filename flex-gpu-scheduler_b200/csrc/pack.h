// pack.h — host packer: object snapshot → SoA int64 tables (the new PreFilter/PreScore stage's
// "pack" step, north_star). Product code.
#pragma once
#include <string>
#include <vector>

#include "../../include/fgs.h"

namespace fgs {

struct PackedNodes {
    int64_t n = 0;
    int S = 2, D = 1;
    std::vector<int64_t> alloc_mcpu, alloc_mem, alloc_eph, alloc_pods, req_mcpu, req_mem, req_eph, num_pods, nz_mcpu,
        nz_mem, alloc_scalar, req_scalar, flex_gpu_sum, flex_mem_sum, used_mem;
    std::vector<uint32_t> alloc_scalar_present, monopoly_mask, flags;
    fgs_node_table view;
    void make_view();
};

struct PackedPods {
    int64_t n = 0;
    int S = 2;
    std::vector<int64_t> flex_gpu_limit, flex_mem_limit, req_mcpu, req_mem, req_eph, req_scalar, nz_mcpu, nz_mem, acct_nz_mcpu,
        acct_nz_mem;
    std::vector<uint32_t> flags, req_scalar_present;
    std::vector<int32_t> ns, pg, priority, index_annot;
    fgs_pod_table view;
    fgs_pod_stats stats; // OR / MAX of the int64 columns, collected while packing (view.stats points here)
    void make_view();
};

// Both return FGS_OK or a negative FGS_E_* with `err` set.
int pack_nodes(const fgs_snapshot* snap, int64_t n0, int64_t n1, PackedNodes* out, std::string* err);
// table row i = snapshot row rows[i]: the rows a generation diff found stale (fgs_sync_nodes)
int pack_node_rows(const fgs_snapshot* snap, const int64_t* rows, int64_t k, PackedNodes* out, std::string* err);
int pack_pods(const fgs_podset* pods, int64_t p0, int64_t p1, PackedPods* out, std::string* err);

// Range rule shared by the packer and the table-level entry points: every quantity must satisfy
// |v| < 2^53 so that the sums the reference forms in int64 (and in arbitrary-precision Quantity,
// pkg/flexgpu/flex_gpu.go:85-98) cannot wrap where Go would not.
constexpr int64_t kMaxQuantity = (int64_t)1 << 53;
inline bool in_range(int64_t v) { return v > -kMaxQuantity && v < kMaxQuantity; }

} // namespace fgs
