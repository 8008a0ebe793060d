// podmath.h — host-side pod/request arithmetic shared by the packer and the synthetic-cluster
// generator (product code; independent of oracle/).
//
// Reference semantics followed (paths relative to the reference repo):
//   computePodResourceRequest   vendor/k8s.io/kubernetes/pkg/scheduler/framework/plugins/noderesources/fit.go:162-180
//   calculateResource           vendor/k8s.io/kubernetes/pkg/scheduler/framework/types.go:715-745
//   calculatePodResourceRequest vendor/.../plugins/noderesources/resource_allocation.go:119-143
//   GetRequestForResource       vendor/k8s.io/kubernetes/pkg/scheduler/util/pod_resources.go:50-85
//   podResourceLimit            pkg/flexgpu/flex_gpu.go:121-131
#pragma once
#include <algorithm>
#include <cstdint>

#include "../../include/fgs_types.h"

namespace fgs {

constexpr int64_t kDefaultMilliCPURequest = 100;
constexpr int64_t kDefaultMemoryRequest = 200ll * 1024 * 1024;

inline int64_t wrap_add(int64_t a, int64_t b) { return (int64_t)((uint64_t)a + (uint64_t)b); }
inline int64_t wrap_sub(int64_t a, int64_t b) { return (int64_t)((uint64_t)a - (uint64_t)b); }

// One pod's aggregated request in slot order {mcpu, mem, eph, scalars…}
struct PodRequest {
    int64_t v[3 + FGS_MAX_SCALARS] = {0};
    uint32_t scalar_keys = 0;   // ScalarResources keys that exist after Add/SetMaxResource
    int64_t nz_mcpu = 0;        // non-zero cpu/memory (types.go:715-745 non0CPU/non0Mem)
    int64_t nz_mem = 0;
    int64_t score_mcpu = 0;     // calculatePodResourceRequest(cpu) incl. the overhead Value() quirk
    int64_t score_mem = 0;
};

struct PodLimits {
    int64_t gpu = 0, mem = 0;
    bool gpu_exist = false, mem_exist = false;
};

inline void regular_and_init(const fgs_podset* ps, int64_t i, int64_t& r0, int64_t& r1, int64_t& i0,
                             int64_t& i1) {
    r0 = ps->ctr_off[i];
    i1 = ps->ctr_off[i + 1];
    i0 = i1 - (ps->n_init ? ps->n_init[i] : 0);
    r1 = i0;
}

inline PodRequest pod_request(const fgs_podset* ps, int64_t i) {
    const int S = ps->n_scalars, W = 3 + S;
    PodRequest out;
    int64_t r0, r1, i0, i1;
    regular_and_init(ps, i, r0, r1, i0, i1);
    for (int64_t k = r0; k < r1; ++k) {
        const int64_t* row = ps->c_req + k * W;
        const uint32_t pr = ps->c_req_present[k];
        for (int c = 0; c < 3; ++c)
            if ((pr >> c) & 1u) out.v[c] = wrap_add(out.v[c], row[c]);
        for (int s = 0; s < S; ++s)
            if ((pr >> (3 + s)) & 1u) {
                out.v[3 + s] = wrap_add(out.v[3 + s], row[3 + s]);
                out.scalar_keys |= 1u << s;
            }
        const int64_t c_cpu = (pr & 1u) ? row[0] : kDefaultMilliCPURequest;
        const int64_t c_mem = (pr & 2u) ? row[1] : kDefaultMemoryRequest;
        out.nz_mcpu = wrap_add(out.nz_mcpu, c_cpu);
        out.nz_mem = wrap_add(out.nz_mem, c_mem);
    }
    out.score_mcpu = out.nz_mcpu;
    out.score_mem = out.nz_mem;
    for (int64_t k = i0; k < i1; ++k) {
        const int64_t* row = ps->c_req + k * W;
        const uint32_t pr = ps->c_req_present[k];
        for (int c = 0; c < 3; ++c)
            if ((pr >> c) & 1u) out.v[c] = std::max(out.v[c], row[c]);
        for (int s = 0; s < S; ++s)
            if ((pr >> (3 + s)) & 1u) {
                out.v[3 + s] = std::max(out.v[3 + s], row[3 + s]);
                out.scalar_keys |= 1u << s;
            }
        const int64_t c_cpu = (pr & 1u) ? row[0] : kDefaultMilliCPURequest;
        const int64_t c_mem = (pr & 2u) ? row[1] : kDefaultMemoryRequest;
        out.nz_mcpu = std::max(out.nz_mcpu, c_cpu);
        out.nz_mem = std::max(out.nz_mem, c_mem);
        out.score_mcpu = std::max(out.score_mcpu, c_cpu);
        out.score_mem = std::max(out.score_mem, c_mem);
    }
    if (ps->overhead_present && (ps->overhead_present[i] >> 31)) {
        const uint32_t pr = ps->overhead_present[i] & 0x7fffffffu;
        const int64_t* row = ps->overhead + i * W;
        for (int c = 0; c < 3; ++c)
            if ((pr >> c) & 1u) out.v[c] = wrap_add(out.v[c], row[c]);
        for (int s = 0; s < S; ++s)
            if ((pr >> (3 + s)) & 1u) {
                out.v[3 + s] = wrap_add(out.v[3 + s], row[3 + s]);
                out.scalar_keys |= 1u << s;
            }
        if (pr & 1u) {
            out.nz_mcpu = wrap_add(out.nz_mcpu, row[0]);
            // the scorer adds Overhead[cpu].Value() — whole cores rounded up — to a milli-core
            // sum (resource_allocation.go:137-139); kept because results must match bit for bit
            const int64_t cores = row[0] >= 0 ? (row[0] + 999) / 1000 : -((-row[0]) / 1000);
            out.score_mcpu = wrap_add(out.score_mcpu, cores);
        }
        if (pr & 2u) {
            out.nz_mem = wrap_add(out.nz_mem, row[1]);
            out.score_mem = wrap_add(out.score_mem, row[1]);
        }
    }
    return out;
}

inline PodLimits pod_limits(const fgs_podset* ps, int64_t i) {
    PodLimits l;
    int64_t r0, r1, i0, i1;
    regular_and_init(ps, i, r0, r1, i0, i1);
    for (int64_t k = r0; k < r1; ++k) {
        const uint32_t pr = ps->c_lim_present[k];
        if (pr & 1u) {
            l.gpu_exist = true;
            l.gpu = wrap_add(l.gpu, ps->c_lim_flex[k * 2 + 0]);
        }
        if (pr & 2u) {
            l.mem_exist = true;
            l.mem = wrap_add(l.mem, ps->c_lim_flex[k * 2 + 1]);
        }
    }
    return l;
}

} // namespace fgs
