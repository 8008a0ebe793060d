// pack.cpp — object snapshot → SoA tables. One walk over each node's pod list replaces the walks
// the reference repeats for every (pod,node) pair:
//   nodeResourceLimitSum        pkg/flexgpu/flex_gpu.go:133-140   → flex_gpu_sum / flex_mem_sum
//   NewGPUNode / constructGPUs  pkg/flexgpu/gpu_node.go:30-120    → used_mem[D][N], monopoly_mask, REF_PANIC flag
//   len(nodeInfo.Pods)          vendor/.../noderesources/fit.go:259 → num_pods
// and each pending pod is reduced once to the numbers the plugins derive from it on every call.
#include "pack.h"

#include <cuda_runtime_api.h>

#include <utility>

#include <cstring>
#include <new>

#include "podmath.h"

namespace fgs {

void PackedNodes::make_view() {
    std::memset(&view, 0, sizeof(view));
    view.n_nodes = n;
    view.n_scalars = S;
    view.n_devices = D;
    view.alloc_mcpu = alloc_mcpu.data();
    view.alloc_mem = alloc_mem.data();
    view.alloc_eph = alloc_eph.data();
    view.alloc_pods = alloc_pods.data();
    view.req_mcpu = req_mcpu.data();
    view.req_mem = req_mem.data();
    view.req_eph = req_eph.data();
    view.num_pods = num_pods.data();
    view.nz_mcpu = nz_mcpu.data();
    view.nz_mem = nz_mem.data();
    view.alloc_scalar = alloc_scalar.data();
    view.req_scalar = req_scalar.data();
    view.alloc_scalar_present = alloc_scalar_present.data();
    view.flex_gpu_sum = flex_gpu_sum.data();
    view.flex_mem_sum = flex_mem_sum.data();
    view.used_mem = used_mem.data();
    view.monopoly_mask = monopoly_mask.data();
    view.flags = flags.data();
}

void PackedPods::make_view() {
    std::memset(&view, 0, sizeof(view));
    view.n_pods = n;
    view.n_scalars = S;
    view.flex_gpu_limit = flex_gpu_limit.data();
    view.flex_mem_limit = flex_mem_limit.data();
    view.flags = flags.data();
    view.req_mcpu = req_mcpu.data();
    view.req_mem = req_mem.data();
    view.req_eph = req_eph.data();
    view.req_scalar = req_scalar.data();
    view.req_scalar_present = req_scalar_present.data();
    view.nz_mcpu = nz_mcpu.data();
    view.nz_mem = nz_mem.data();
    view.ns = ns.data();
    view.pg = pg.data();
    view.priority = priority.data();
    view.acct_nz_mcpu = acct_nz_mcpu.data();
    view.acct_nz_mem = acct_nz_mem.data();
    view.index_annot = index_annot.data();
    view.stats = &stats;
}

// rows == NULL: every row of `snap` in order; else table row i = snapshot row rows[i] (k of them)
static int pack_nodes_all(const fgs_snapshot* snap, PackedNodes* out, std::string* err, const int64_t* rows = nullptr, int64_t k = 0);

int pack_node_rows(const fgs_snapshot* snap, const int64_t* rows, int64_t k, PackedNodes* out, std::string* err) {
    if (!snap || snap->n_nodes < 0 || snap->n_scalars < 2 || snap->n_scalars > FGS_MAX_SCALARS) {
        *err = "pack_node_rows: bad snapshot header";
        return FGS_E_INVALID;
    }
    if (k < 0 || (k > 0 && !rows)) {
        *err = "pack_node_rows: bad row list";
        return FGS_E_INVALID;
    }
    for (int64_t i = 0; i < k; ++i)
        if (rows[i] < 0 || rows[i] >= snap->n_nodes) {
            *err = "pack_node_rows: row outside the snapshot";
            return FGS_E_INVALID;
        }
    static const int64_t none = 0;
    return pack_nodes_all(snap, out, err, k ? rows : &none, k);
}

int pack_nodes(const fgs_snapshot* snap, int64_t n0, int64_t n1, PackedNodes* out, std::string* err) {
    if (!snap || snap->n_nodes < 0 || snap->n_scalars < 2 || snap->n_scalars > FGS_MAX_SCALARS) {
        *err = "pack_nodes: bad snapshot header";
        return FGS_E_INVALID;
    }
    if (n0 < 0 || n1 < n0 || n1 > snap->n_nodes) {
        *err = "pack_nodes: node range outside the snapshot";
        return FGS_E_INVALID;
    }
    // a shard is the same snapshot with the per-node arrays advanced; pod_off stays absolute
    fgs_snapshot sub = *snap;
    const int S = snap->n_scalars;
    sub.n_nodes = n1 - n0;
    sub.node_valid = snap->node_valid + n0;
    sub.alloc = snap->alloc + n0 * (4 + S);
    sub.alloc_scalar_present = snap->alloc_scalar_present + n0;
    sub.requested = snap->requested + n0 * (3 + S);
    sub.nonzero = snap->nonzero + n0 * 2;
    sub.pod_off = snap->pod_off + n0;
    return pack_nodes_all(&sub, out, err);
}

static int pack_nodes_all(const fgs_snapshot* snap, PackedNodes* out, std::string* err, const int64_t* rows, int64_t k) {
    const int64_t N = rows ? k : snap->n_nodes;
    const int S = snap->n_scalars, AW = 4 + S, RW = 3 + S;
    // device slots = largest gpu count on any node
    int D = 1;
    for (int64_t i = 0; i < N; ++i) {
        const int64_t n = rows ? rows[i] : i;
        if (!((snap->alloc_scalar_present[n] >> FGS_SCALAR_FLEX_GPU) & 1u)) continue;
        const int64_t g = snap->alloc[n * AW + 4 + FGS_SCALAR_FLEX_GPU];
        if (g > FGS_MAX_DEVICES) {
            *err = "pack_nodes: node " + std::to_string(n) + " has " + std::to_string(g) + " flex gpus; at most " +
                   std::to_string(FGS_MAX_DEVICES) + " device slots are supported";
            return FGS_E_TOO_LARGE;
        }
        if (g > D) D = (int)g;
    }
    PackedNodes& t = *out;
    t.n = N;
    t.S = S;
    t.D = D;
    auto z = [&](std::vector<int64_t>& v, int64_t k) { v.assign((size_t)k, 0); };
    z(t.alloc_mcpu, N), z(t.alloc_mem, N), z(t.alloc_eph, N), z(t.alloc_pods, N);
    z(t.req_mcpu, N), z(t.req_mem, N), z(t.req_eph, N), z(t.num_pods, N), z(t.nz_mcpu, N), z(t.nz_mem, N);
    z(t.alloc_scalar, N * S), z(t.req_scalar, N * S), z(t.flex_gpu_sum, N), z(t.flex_mem_sum, N), z(t.used_mem, N * D);
    t.alloc_scalar_present.assign((size_t)N, 0);
    t.monopoly_mask.assign((size_t)N, 0);
    t.flags.assign((size_t)N, 0);

    const fgs_podset* ps = &snap->pods;
    for (int64_t i = 0; i < N; ++i) {
        const int64_t n = rows ? rows[i] : i; // snapshot row → table row i
        const int64_t* al = snap->alloc + n * AW;
        const int64_t* rq = snap->requested + n * RW;
        for (int k = 0; k < AW; ++k)
            if (!in_range(al[k])) {
                *err = "pack_nodes: allocatable of node " + std::to_string(n) + " out of range";
                return FGS_E_RANGE;
            }
        for (int k = 0; k < RW; ++k)
            if (!in_range(rq[k])) {
                *err = "pack_nodes: requested of node " + std::to_string(n) + " out of range";
                return FGS_E_RANGE;
            }
        t.alloc_mcpu[(size_t)i] = al[0];
        t.alloc_mem[(size_t)i] = al[1];
        t.alloc_eph[(size_t)i] = al[2];
        t.alloc_pods[(size_t)i] = al[3];
        t.req_mcpu[(size_t)i] = rq[0];
        t.req_mem[(size_t)i] = rq[1];
        t.req_eph[(size_t)i] = rq[2];
        t.nz_mcpu[(size_t)i] = snap->nonzero[n * 2 + 0];
        t.nz_mem[(size_t)i] = snap->nonzero[n * 2 + 1];
        if (!in_range(t.nz_mcpu[(size_t)i]) || !in_range(t.nz_mem[(size_t)i])) {
            *err = "pack_nodes: non-zero requested of node " + std::to_string(n) + " out of range";
            return FGS_E_RANGE;
        }
        const uint32_t present = snap->alloc_scalar_present[n];
        t.alloc_scalar_present[(size_t)i] = present;
        for (int s = 0; s < S; ++s) {
            t.alloc_scalar[(size_t)(s * N + i)] = ((present >> s) & 1u) ? al[4 + s] : 0;
            t.req_scalar[(size_t)(s * N + i)] = rq[3 + s];
        }
        const int64_t p0 = snap->pod_off[n], p1 = snap->pod_off[n + 1];
        t.num_pods[(size_t)i] = p1 - p0;

        uint32_t fl = 0;
        if (snap->node_valid[n]) fl |= FGS_NF_VALID;
        const bool has_gpu = (present >> FGS_SCALAR_FLEX_GPU) & 1u, has_mem = (present >> FGS_SCALAR_FLEX_MEM) & 1u;
        if (has_gpu) fl |= FGS_NF_HAS_GPU_RES;
        if (has_mem) fl |= FGS_NF_HAS_MEM_RES;
        if (has_gpu && has_mem) fl |= FGS_NF_HAS_FLEX;
        // gpu count as NewGPUNode reads it: a missing key is the zero Quantity (gpu_node.go:37-42)
        const int64_t gpu_cnt = has_gpu ? al[4 + FGS_SCALAR_FLEX_GPU] : 0;
        if (gpu_cnt <= 0) fl |= FGS_NF_REF_PANIC; // divide by zero (:55) / negative make (:68)

        int64_t gsum = 0, msum = 0;
        uint32_t mono = 0;
        for (int64_t q = p0; q < p1; ++q) {
            const PodLimits l = pod_limits(ps, q);
            if (!in_range(l.gpu) || !in_range(l.mem)) {
                *err = "pack_nodes: flex limit of a pod on node " + std::to_string(n) + " out of range";
                return FGS_E_RANGE;
            }
            gsum = wrap_add(gsum, l.gpu);
            msum = wrap_add(msum, l.mem);
            if (!l.gpu_exist && !l.mem_exist) continue;       // gpu_node.go:85-88
            const int32_t idx = ps->index_annot[q];
            if (idx == FGS_ANNOT_NONE) continue;              // :91-96
            if (idx < 0 || idx >= gpu_cnt) {                  // :110 / :115 index out of range
                fl |= FGS_NF_REF_PANIC;
                continue;
            }
            if (l.gpu_exist) mono |= 1u << idx;
            if (l.mem_exist) {
                int64_t& u = t.used_mem[(size_t)((int64_t)idx * N + i)];
                u = wrap_add(u, l.mem);
            }
        }
        if (!in_range(gsum) || !in_range(msum)) {
            *err = "pack_nodes: flex limit sum of node " + std::to_string(n) + " out of range";
            return FGS_E_RANGE;
        }
        for (int d = 0; d < D; ++d)
            if (!in_range(t.used_mem[(size_t)((int64_t)d * N + i)])) {
                *err = "pack_nodes: per-device used memory of node " + std::to_string(n) + " out of range";
                return FGS_E_RANGE;
            }
        t.flex_gpu_sum[(size_t)i] = gsum;
        t.flex_mem_sum[(size_t)i] = msum;
        t.monopoly_mask[(size_t)i] = mono;
        t.flags[(size_t)i] = fl;
    }
    t.make_view();
    return FGS_OK;
}

int pack_pods(const fgs_podset* pods, int64_t p0, int64_t p1, PackedPods* out, std::string* err) {
    if (!pods || pods->n_pods < 0 || pods->n_scalars < 2 || pods->n_scalars > FGS_MAX_SCALARS) {
        *err = "pack_pods: bad podset header";
        return FGS_E_INVALID;
    }
    if (p0 < 0 || p1 < p0 || p1 > pods->n_pods) {
        *err = "pack_pods: pod range outside the podset";
        return FGS_E_INVALID;
    }
    const int64_t P = p1 - p0;
    const int S = pods->n_scalars;
    PackedPods& t = *out;
    t.n = P;
    t.S = S;
    auto z = [&](std::vector<int64_t>& v, int64_t k) { v.assign((size_t)k, 0); };
    z(t.flex_gpu_limit, P), z(t.flex_mem_limit, P), z(t.req_mcpu, P), z(t.req_mem, P), z(t.req_eph, P);
    z(t.req_scalar, P * S), z(t.nz_mcpu, P), z(t.nz_mem, P), z(t.acct_nz_mcpu, P), z(t.acct_nz_mem, P);
    t.index_annot.assign((size_t)P, FGS_ANNOT_NONE);
    t.flags.assign((size_t)P, 0);
    t.req_scalar_present.assign((size_t)P, 0);
    t.ns.assign((size_t)P, 0);
    t.pg.assign((size_t)P, -1);
    t.priority.assign((size_t)P, 0);
    std::memset(&t.stats, 0, sizeof(t.stats));
    auto acc = [&](int c, int64_t v) { // every quantity is >= 0 here (checked below)
        t.stats.or_bits[c] |= (uint64_t)v;
        if (v > t.stats.max_val[c]) t.stats.max_val[c] = v;
    };
    for (int64_t i = 0; i < P; ++i) {
        const int64_t src = p0 + i;
        const PodLimits l = pod_limits(pods, src);
        const PodRequest r = pod_request(pods, src);
        // API validation forbids negative quantities; the kernels rely on requests and limits being >= 0
        bool ok = in_range(l.gpu) && in_range(l.mem) && in_range(r.score_mcpu) && in_range(r.score_mem) && l.gpu >= 0 && l.mem >= 0 &&
                  r.score_mcpu >= 0 && r.score_mem >= 0 && in_range(r.nz_mcpu) && in_range(r.nz_mem) && r.nz_mcpu >= 0 && r.nz_mem >= 0;
        for (int k = 0; k < 3 + S; ++k) ok = ok && in_range(r.v[k]) && r.v[k] >= 0;
        if (!ok) {
            *err = "pack_pods: a quantity of pod " + std::to_string(src) + " is negative or out of range";
            return FGS_E_RANGE;
        }
        uint32_t fl = 0;
        if (l.gpu_exist) fl |= FGS_PF_GPU_LIMIT;
        if (l.mem_exist) fl |= FGS_PF_MEM_LIMIT;
        // fit.go:270-275: MilliCPU == 0 && Memory == 0 && EphemeralStorage == 0 && len(ScalarResources) == 0
        if (r.v[0] == 0 && r.v[1] == 0 && r.v[2] == 0 && r.scalar_keys == 0) fl |= FGS_PF_REQ_NONE;
        t.flags[(size_t)i] = fl;
        t.flex_gpu_limit[(size_t)i] = l.gpu;
        t.flex_mem_limit[(size_t)i] = l.mem;
        t.req_mcpu[(size_t)i] = r.v[0];
        t.req_mem[(size_t)i] = r.v[1];
        t.req_eph[(size_t)i] = r.v[2];
        for (int s = 0; s < S; ++s) t.req_scalar[(size_t)(s * P + i)] = r.v[3 + s];
        t.req_scalar_present[(size_t)i] = r.scalar_keys;
        t.nz_mcpu[(size_t)i] = r.score_mcpu;
        t.nz_mem[(size_t)i] = r.score_mem;
        t.acct_nz_mcpu[(size_t)i] = r.nz_mcpu;
        t.acct_nz_mem[(size_t)i] = r.nz_mem;
        t.index_annot[(size_t)i] = pods->index_annot ? pods->index_annot[src] : FGS_ANNOT_NONE;
        t.ns[(size_t)i] = pods->ns ? pods->ns[src] : 0;
        t.pg[(size_t)i] = pods->pg ? pods->pg[src] : -1;
        t.priority[(size_t)i] = pods->priority ? pods->priority[src] : 0;
        acc(0, l.gpu), acc(1, l.mem), acc(2, r.v[0]), acc(3, r.v[1]), acc(4, r.v[2]), acc(5, r.score_mcpu), acc(6, r.score_mem);
        for (int s = 0; s < S; ++s) acc(8 + s, r.v[3 + s]);
        t.stats.present_or |= r.scalar_keys;
    }
    t.make_view();
    return FGS_OK;
}

} // namespace fgs

// ---- C ABI: the packer object (host only) ---------------------------------------------------------

struct fgs_packer {
    int S = 2;
    fgs::PackedNodes nodes;
    fgs::PackedPods pods;
    std::string err;
    // fgs_packer_pin: the pod-table columns the engine copies, as currently page-locked (pointer, bytes)
    bool pin = false;
    std::vector<std::pair<void*, size_t>> locked;
    void unlock_all() {
        for (auto& r : locked) cudaHostUnregister(r.first);
        locked.clear();
    }
    int lock_columns();
};

int fgs_packer::lock_columns() {
    fgs::PackedPods& t = pods;
    std::vector<std::pair<void*, size_t>> want;
    auto add = [&](const void* p, size_t bytes) {
        if (p && bytes) want.push_back({(void*)p, bytes});
    };
    const size_t P = (size_t)t.n;
    add(t.flex_gpu_limit.data(), P * 8), add(t.flex_mem_limit.data(), P * 8), add(t.req_mcpu.data(), P * 8), add(t.req_mem.data(), P * 8);
    add(t.req_eph.data(), P * 8), add(t.nz_mcpu.data(), P * 8), add(t.nz_mem.data(), P * 8), add(t.req_scalar.data(), P * 8 * (size_t)t.S);
    add(t.flags.data(), P * 4), add(t.req_scalar_present.data(), P * 4);
    t.stats.flags &= ~FGS_POD_STATS_PINNED;
    if (want != locked) { // the vectors moved or changed size: lock them anew
        unlock_all();
        for (auto& r : want) {
            const cudaError_t ce = cudaHostRegister(r.first, r.second, cudaHostRegisterDefault);
            if (ce != cudaSuccess) {
                cudaGetLastError();
                err = std::string("fgs_packer_pin: cudaHostRegister: ") + cudaGetErrorString(ce);
                unlock_all();
                return FGS_E_CUDA;
            }
            locked.push_back(r);
        }
    }
    if (!want.empty()) t.stats.flags |= FGS_POD_STATS_PINNED;
    return FGS_OK;
}

extern "C" int fgs_packer_pin(fgs_packer* p) {
    if (!p) return FGS_E_INVALID;
    p->pin = true;
    return p->lock_columns();
}

extern "C" int fgs_packer_create(int32_t n_scalars, fgs_packer** out) {
    if (!out || n_scalars < 2 || n_scalars > FGS_MAX_SCALARS) return FGS_E_INVALID;
    fgs_packer* p = new (std::nothrow) fgs_packer();
    if (!p) return FGS_E_INVALID;
    p->S = n_scalars;
    *out = p;
    return FGS_OK;
}

extern "C" void fgs_packer_destroy(fgs_packer* p) {
    if (p) p->unlock_all();
    delete p;
}

extern "C" const char* fgs_packer_error(const fgs_packer* p) { return p ? p->err.c_str() : "null packer"; }

extern "C" int fgs_packer_nodes(fgs_packer* p, const fgs_snapshot* snap, int64_t n0, int64_t n1, const fgs_node_table** out) {
    if (!p || !out) return FGS_E_INVALID;
    if (snap && snap->n_scalars != p->S) {
        p->err = "fgs_packer_nodes: snapshot n_scalars differs from the packer's";
        return FGS_E_INVALID;
    }
    const int rc = fgs::pack_nodes(snap, n0, n1, &p->nodes, &p->err);
    if (rc) return rc;
    *out = &p->nodes.view;
    return FGS_OK;
}

extern "C" int fgs_packer_node_rows(fgs_packer* p, const fgs_snapshot* snap, const int64_t* rows, int64_t n_rows, const fgs_node_table** out) {
    if (!p || !out) return FGS_E_INVALID;
    if (snap && snap->n_scalars != p->S) {
        p->err = "fgs_packer_node_rows: snapshot n_scalars differs from the packer's";
        return FGS_E_INVALID;
    }
    const int rc = fgs::pack_node_rows(snap, rows, n_rows, &p->nodes, &p->err);
    if (rc) return rc;
    *out = &p->nodes.view;
    return FGS_OK;
}

extern "C" int fgs_packer_pods(fgs_packer* p, const fgs_podset* pods, int64_t p0, int64_t p1, const fgs_pod_table** out) {
    if (!p || !out) return FGS_E_INVALID;
    if (pods && pods->n_scalars != p->S) {
        p->err = "fgs_packer_pods: podset n_scalars differs from the packer's";
        return FGS_E_INVALID;
    }
    const int rc = fgs::pack_pods(pods, p0, p1, &p->pods, &p->err);
    if (rc) return rc;
    if (p->pin) {
        const int rc2 = p->lock_columns();
        if (rc2) return rc2;
    }
    *out = &p->pods.view;
    return FGS_OK;
}
