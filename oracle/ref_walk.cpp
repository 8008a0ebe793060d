// ref_walk.cpp — CPU oracle: restatement of the reference Go plugins on the flattened object
// snapshot. TEST INFRASTRUCTURE ONLY (see fgs_oracle.h). Every function cites the reference
// file:line it follows; paths are relative to /root/reference, "vendor/…/scheduler" abbreviates
// vendor/k8s.io/kubernetes/pkg/scheduler.
//
// The walk is deliberately object-shaped: per (pod,node) it re-sums the limits of every pod on
// the node and rebuilds the per-device table exactly like the Go code does, including the
// pointer-aliasing behaviour of MemAssumeFitIndexes. It is not optimised.
#include "fgs_oracle.h"

#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstring>
#include <functional>
#include <mutex>
#include <thread>
#include <vector>

namespace {

typedef int64_t i64;
typedef uint64_t u64;

// Go's int64 arithmetic wraps; C++ signed overflow is undefined, so wrap explicitly.
inline i64 wadd(i64 a, i64 b) { return (i64)((u64)a + (u64)b); }
inline i64 wsub(i64 a, i64 b) { return (i64)((u64)a - (u64)b); }
inline i64 wmul(i64 a, i64 b) { return (i64)((u64)a * (u64)b); }

struct RefPanic {
    const char* what;
};

const i64 kMaxNodeScore = 100;                       // vendor/…/framework/interface.go:97
const i64 kDefaultMilliCPU = 100;                    // vendor/…/scheduler/util/pod_resources.go:35
const i64 kDefaultMemory = 200 * 1024 * 1024;        // :37

// ---- pod helpers ---------------------------------------------------------------------------

struct CtrRange {
    i64 reg0, reg1;   // regular containers
    i64 init0, init1; // init containers
};

inline CtrRange ctrs(const fgs_podset* ps, i64 i) {
    CtrRange r;
    r.reg0 = ps->ctr_off[i];
    r.init1 = ps->ctr_off[i + 1];
    r.init0 = r.init1 - (ps->n_init ? ps->n_init[i] : 0);
    r.reg1 = r.init0;
    return r;
}

// framework.Resource (vendor/…/framework/types.go:410-419)
struct Resource {
    i64 mcpu = 0, mem = 0, eph = 0;
    i64 allowedPods = 0;
    i64 scalar[FGS_MAX_SCALARS] = {0};
    uint32_t scalarPresent = 0; // keys of ScalarResources
};

// Resource.Add(ResourceList) types.go:429-453, on a request row {mcpu,mem,eph,scalars}
inline void resAdd(Resource& r, const i64* row, uint32_t present, int S) {
    if (present & 1u) r.mcpu = wadd(r.mcpu, row[0]);
    if (present & 2u) r.mem = wadd(r.mem, row[1]);
    if (present & 4u) r.eph = wadd(r.eph, row[2]);
    for (int s = 0; s < S; ++s)
        if (present & (8u << s)) { // AddScalar → SetScalar: the key exists afterwards (:475-487)
            r.scalar[s] = wadd(r.scalar[s], row[3 + s]);
            r.scalarPresent |= 1u << s;
        }
}

// Resource.SetMaxResource types.go:489-508
inline void resSetMax(Resource& r, const i64* row, uint32_t present, int S) {
    if (present & 2u) r.mem = std::max(r.mem, row[1]);
    if (present & 1u) r.mcpu = std::max(r.mcpu, row[0]);
    if (present & 4u) r.eph = std::max(r.eph, row[2]);
    for (int s = 0; s < S; ++s)
        if (present & (8u << s)) {
            r.scalar[s] = std::max(r.scalar[s], row[3 + s]);
            r.scalarPresent |= 1u << s;
        }
}

// computePodResourceRequest — vendor/…/plugins/noderesources/fit.go:162-180 and
// pkg/capacityscheduling/capacity_scheduling.go:833-850 (same body).
Resource computePodResourceRequest(const fgs_podset* ps, i64 i) {
    const int S = ps->n_scalars, W = 3 + S;
    Resource r;
    CtrRange c = ctrs(ps, i);
    for (i64 k = c.reg0; k < c.reg1; ++k) resAdd(r, ps->c_req + k * W, ps->c_req_present[k], S);
    for (i64 k = c.init0; k < c.init1; ++k) resSetMax(r, ps->c_req + k * W, ps->c_req_present[k], S);
    if (ps->overhead_present && (ps->overhead_present[i] >> 31))
        resAdd(r, ps->overhead + i * W, ps->overhead_present[i] & 0x7fffffffu, S);
    return r;
}

// calculateResource — vendor/…/framework/types.go:715-745 (used by AddPod/RemovePod)
void calculateResource(const fgs_podset* ps, i64 i, Resource& res, i64& non0cpu, i64& non0mem) {
    const int S = ps->n_scalars, W = 3 + S;
    res = Resource();
    non0cpu = non0mem = 0;
    CtrRange c = ctrs(ps, i);
    auto nonzero = [&](i64 k, i64& cpu, i64& mem) { // GetNonzeroRequests pod_resources.go:42-45
        uint32_t pr = ps->c_req_present[k];
        cpu = (pr & 1u) ? ps->c_req[k * W + 0] : kDefaultMilliCPU;
        mem = (pr & 2u) ? ps->c_req[k * W + 1] : kDefaultMemory;
    };
    for (i64 k = c.reg0; k < c.reg1; ++k) {
        resAdd(res, ps->c_req + k * W, ps->c_req_present[k], S);
        i64 a, b;
        nonzero(k, a, b);
        non0cpu = wadd(non0cpu, a);
        non0mem = wadd(non0mem, b);
    }
    for (i64 k = c.init0; k < c.init1; ++k) {
        resSetMax(res, ps->c_req + k * W, ps->c_req_present[k], S);
        i64 a, b;
        nonzero(k, a, b);
        non0cpu = std::max(non0cpu, a);
        non0mem = std::max(non0mem, b);
    }
    if (ps->overhead_present && (ps->overhead_present[i] >> 31)) {
        uint32_t pr = ps->overhead_present[i] & 0x7fffffffu;
        resAdd(res, ps->overhead + i * W, pr, S);
        if (pr & 1u) non0cpu = wadd(non0cpu, ps->overhead[i * W + 0]);
        if (pr & 2u) non0mem = wadd(non0mem, ps->overhead[i * W + 1]);
    }
}

// calculatePodResourceRequest — vendor/…/plugins/noderesources/resource_allocation.go:119-143
// with useRequested == false (Least/MostAllocated never set it, :36-39; only BalancedAllocation
// does). `res` is a resource slot id.
i64 scorePodRequest(const fgs_podset* ps, i64 i, int res) {
    const int S = ps->n_scalars, W = 3 + S;
    CtrRange c = ctrs(ps, i);
    auto get = [&](i64 k) -> i64 { // GetRequestForResource pod_resources.go:50-85, nonZero=true
        uint32_t pr = ps->c_req_present[k];
        bool found = (pr >> res) & 1u;
        if (res == FGS_RES_CPU) return found ? ps->c_req[k * W + 0] : kDefaultMilliCPU;
        if (res == FGS_RES_MEM) return found ? ps->c_req[k * W + 1] : kDefaultMemory;
        return found ? ps->c_req[k * W + res] : 0;
    };
    i64 podRequest = 0;
    for (i64 k = c.reg0; k < c.reg1; ++k) podRequest = wadd(podRequest, get(k));
    for (i64 k = c.init0; k < c.init1; ++k) {
        i64 v = get(k);
        if (podRequest < v) podRequest = v;
    }
    if (ps->overhead_present && (ps->overhead_present[i] >> 31)) {
        uint32_t pr = ps->overhead_present[i] & 0x7fffffffu;
        if ((pr >> res) & 1u) {
            i64 q = ps->overhead[i * W + res];
            // :137-139 adds quantity.Value() for EVERY resource, cpu included: whole cores,
            // rounded up (quantity.go:728-733), added to a milli-core sum. Restated as is.
            if (res == FGS_RES_CPU) q = (q >= 0) ? (q + 999) / 1000 : -((-q) / 1000);
            podRequest = wadd(podRequest, q);
        }
    }
    return podRequest;
}

// podResourceLimit — pkg/flexgpu/flex_gpu.go:121-131 (regular containers only)
inline i64 podResourceLimit(const fgs_podset* ps, i64 i, int flexSlot, bool& exist) {
    exist = false;
    i64 sum = 0;
    CtrRange c = ctrs(ps, i);
    for (i64 k = c.reg0; k < c.reg1; ++k)
        if ((ps->c_lim_present[k] >> flexSlot) & 1u) {
            exist = true;
            sum = wadd(sum, ps->c_lim_flex[k * 2 + flexSlot]);
        }
    return sum;
}

// ---- node view -----------------------------------------------------------------------------

struct Snap {
    const fgs_snapshot* s;
    int S, AW, RW;
    explicit Snap(const fgs_snapshot* sn) : s(sn), S(sn->n_scalars), AW(4 + sn->n_scalars), RW(3 + sn->n_scalars) {}
    i64 podBegin(i64 n) const { return s->pod_off[n]; }
    i64 podEnd(i64 n) const { return s->pod_off[n + 1]; }
    const i64* alloc(i64 n) const { return s->alloc + n * AW; }
    const i64* req(i64 n) const { return s->requested + n * RW; }
    bool allocHasScalar(i64 n, int sc) const { return (s->alloc_scalar_present[n] >> sc) & 1u; }
    // addNominatedPods (vendor/…/framework/runtime/framework.go:795-823): nodeInfo.Clone() + AddPodInfo for the
    // nominated pods the evaluated pod must yield to — modelled as extra pods of ONE node, taken from another podset
    i64 ovNode = -1;
    const fgs_podset* ovPs = nullptr;
    std::vector<i64> ovPods;
    bool overlaid(i64 n) const { return n == ovNode && !ovPods.empty(); }
    // NodeInfo.RemovePod of the dry-run preemption (capacity_scheduling.go:486-495): pods of the snapshot marked as
    // removed (indexed by their row in s->pods); nullptr = none
    const std::vector<uint8_t>* removed = nullptr;
    bool gone(i64 q) const { return removed && (*removed)[(size_t)q]; }
};

// nodeResourceLimitSum — pkg/flexgpu/flex_gpu.go:133-140
i64 nodeResourceLimitSum(const Snap& sn, i64 n, int flexSlot) {
    i64 sum = 0;
    for (i64 q = sn.podBegin(n); q < sn.podEnd(n); ++q) {
        if (sn.gone(q)) continue;
        bool e;
        sum = wadd(sum, podResourceLimit(&sn.s->pods, q, flexSlot, e));
    }
    if (sn.overlaid(n))
        for (i64 q : sn.ovPods) {
            bool e;
            sum = wadd(sum, podResourceLimit(sn.ovPs, q, flexSlot, e));
        }
    return sum;
}

// gpuNode / gpu — pkg/flexgpu/gpu_node.go:13-24. `memory` is ONE shared cell: every gpus[i].memory
// points at the same *Quantity (gpu_node.go:55,70-75); modelled with memEachShared.
struct Gpu {
    int index;
    bool monopoly;
    i64 usedMemory;
};
struct GpuNode {
    i64 memEachShared;
    std::vector<Gpu> gpus;
};

// NewGPUNode + constructGPUs — pkg/flexgpu/gpu_node.go:30-65,67-120
GpuNode newGPUNode(const Snap& sn, i64 n) {
    // Allocatable[...] of a missing key is the zero Quantity → AsInt64() == 0, ok (:37-49)
    i64 gpuCnt = sn.allocHasScalar(n, FGS_SCALAR_FLEX_GPU) ? sn.alloc(n)[4 + FGS_SCALAR_FLEX_GPU] : 0;
    i64 memCnt = sn.allocHasScalar(n, FGS_SCALAR_FLEX_MEM) ? sn.alloc(n)[4 + FGS_SCALAR_FLEX_MEM] : 0;
    if (gpuCnt == 0) throw RefPanic{"integer divide by zero (gpu_node.go:55)"};
    i64 memEach = memCnt / gpuCnt; // Go int64 '/' truncates toward zero like C++
    if (gpuCnt < 0) throw RefPanic{"makeslice: len out of range (gpu_node.go:68)"};
    GpuNode g;
    g.memEachShared = memEach;
    g.gpus.resize((size_t)gpuCnt);
    for (i64 i = 0; i < gpuCnt; ++i) g.gpus[(size_t)i] = Gpu{(int)i, false, 0};
    auto visit = [&](const fgs_podset* ps, i64 q) {
        bool gpuExist, memExist;
        podResourceLimit(ps, q, FGS_SCALAR_FLEX_GPU, gpuExist);
        i64 memLimit = podResourceLimit(ps, q, FGS_SCALAR_FLEX_MEM, memExist);
        if (!gpuExist && !memExist) return; // :85-88
        // :90-96 Atoi of a missing ("") or malformed annotation errors → pod skipped
        int32_t idx = ps->index_annot[q];
        if (idx == FGS_ANNOT_NONE) return;
        if (gpuExist) {
            if (idx < 0 || idx >= gpuCnt) throw RefPanic{"index out of range (gpu_node.go:110)"};
            g.gpus[(size_t)idx].monopoly = true;
        }
        if (memExist) {
            if (idx < 0 || idx >= gpuCnt) throw RefPanic{"index out of range (gpu_node.go:115)"};
            g.gpus[(size_t)idx].usedMemory = wadd(g.gpus[(size_t)idx].usedMemory, memLimit);
        }
    };
    for (i64 q = sn.podBegin(n); q < sn.podEnd(n); ++q)
        if (!sn.gone(q)) visit(&sn.s->pods, q);
    if (sn.overlaid(n))
        for (i64 q : sn.ovPods) visit(sn.ovPs, q);
    return g;
}

// MemAssumeFitIndexes — pkg/flexgpu/gpu_node.go:122-161, aliasing included:
//  * `assumed := u.usedMemory; assumed.Add(*memLimit)` mutates the device's usedMemory (:134-135)
//  * `remain := u.memory; remain.Sub(*assumed)` mutates the SHARED memory cell (:139-140)
//  * every fit.remain is that same cell → the comparator compares a value with itself, never
//    "less"; sort.Slice on ≤ 12 elements is insertion sort → order unchanged (:152-154).
std::vector<int> memAssumeFitIndexes(GpuNode& g, i64 memLimit) {
    std::vector<int> fits;
    for (Gpu& u : g.gpus) {
        u.usedMemory = wadd(u.usedMemory, memLimit);
        if (!u.monopoly && g.memEachShared >= u.usedMemory) {
            g.memEachShared = wsub(g.memEachShared, u.usedMemory);
            fits.push_back(u.index);
        }
    }
    return fits;
}

// GPUAssumeFitIndexes — gpu_node.go:163-177
std::vector<int> gpuAssumeFitIndexes(const GpuNode& g) {
    std::vector<int> fits;
    for (const Gpu& u : g.gpus)
        if (!u.monopoly && u.usedMemory == 0) fits.push_back(u.index);
    return fits;
}

// GPUScore / MemScore — gpu_node.go:179-199
i64 gpuScore(const GpuNode& g) {
    i64 cnt = 0;
    for (const Gpu& u : g.gpus)
        if (!u.monopoly && u.usedMemory == 0) cnt++;
    return cnt;
}
i64 memScore(const GpuNode& g) {
    i64 remain = 0;
    for (const Gpu& u : g.gpus) {
        remain = wadd(remain, g.memEachShared);
        remain = wsub(remain, u.usedMemory);
    }
    return remain;
}

// ---- per-pod precomputed state (what PreFilter writes into CycleState) ----------------------

struct PodCtx {
    const fgs_podset* ps;
    i64 i;
    Resource fitReq;            // Fit.PreFilter fit.go:182-185
    i64 scoreReq[FGS_MAX_SCORE_RES]; // per configured Fit score resource
};

// FlexGPU.Filter — pkg/flexgpu/flex_gpu.go:41-119. Returns a framework.Code.
// reason (optional): which of FlexGPU.Filter's messages the Status carries (FGS_FLEX_REASON_*)
int flexFilter(const Snap& sn, const PodCtx& pc, i64 n, int* reason = nullptr) {
    int dummy;
    if (!reason) reason = &dummy;
    *reason = FGS_FLEX_REASON_NONE;
    bool gpuExist, memExist;
    i64 gpuLimit = podResourceLimit(pc.ps, pc.i, FGS_SCALAR_FLEX_GPU, gpuExist); // :52
    i64 memLimit = podResourceLimit(pc.ps, pc.i, FGS_SCALAR_FLEX_MEM, memExist); // :53
    if (!gpuExist && !memExist) return FGS_CODE_SUCCESS;                         // :54-56
    if (gpuExist && memExist) return *reason = FGS_FLEX_REASON_CONFLICT, FGS_CODE_UNRESOLVABLE; // :58-61
    if (!sn.allocHasScalar(n, FGS_SCALAR_FLEX_GPU)) return *reason = FGS_FLEX_REASON_UNKNOWN_RESOURCE, FGS_CODE_UNRESOLVABLE; // :64-68
    if (!sn.allocHasScalar(n, FGS_SCALAR_FLEX_MEM)) return *reason = FGS_FLEX_REASON_UNKNOWN_RESOURCE, FGS_CODE_UNRESOLVABLE; // :70-74
    i64 gpuAllocatable = sn.alloc(n)[4 + FGS_SCALAR_FLEX_GPU];
    i64 memAllocatable = sn.alloc(n)[4 + FGS_SCALAR_FLEX_MEM];
    i64 gpuAssume = wadd(nodeResourceLimitSum(sn, n, FGS_SCALAR_FLEX_GPU), gpuLimit); // :82-87
    i64 memAssume = wadd(nodeResourceLimitSum(sn, n, FGS_SCALAR_FLEX_MEM), memLimit);
    if (gpuAssume > gpuAllocatable) return *reason = FGS_FLEX_REASON_INSUFFICIENT_GPU, FGS_CODE_UNSCHEDULABLE; // :90-93
    if (memAssume > memAllocatable) return *reason = FGS_FLEX_REASON_INSUFFICIENT_MEM, FGS_CODE_UNSCHEDULABLE; // :95-98
    GpuNode no = newGPUNode(sn, n);                                // :101 (may panic)
    if (memExist) {                                                // :109-116
        if (memAssumeFitIndexes(no, memLimit).empty()) return *reason = FGS_FLEX_REASON_NO_FIT_INDEX, FGS_CODE_UNSCHEDULABLE;
    }
    return FGS_CODE_SUCCESS;
}

// FlexGPU.Score — flex_gpu.go:142-166 (NewGPUNode first, for every pod kind)
i64 flexScore(const Snap& sn, const PodCtx& pc, i64 n) {
    GpuNode no = newGPUNode(sn, n); // :148
    bool gpuExist, memExist;
    podResourceLimit(pc.ps, pc.i, FGS_SCALAR_FLEX_GPU, gpuExist);
    podResourceLimit(pc.ps, pc.i, FGS_SCALAR_FLEX_MEM, memExist);
    if (!gpuExist && !memExist) return 0;
    if (gpuExist) return gpuScore(no);
    return memScore(no);
}

// fitsRequest — vendor/…/plugins/noderesources/fit.go:256-329. Returns #insufficient resources.
// mask (optional): the InsufficientResource list as FGS_FIT_REASON_* bits
int fitsRequest(const Snap& sn, const PodCtx& pc, i64 n, uint32_t ignoredScalarMask, uint32_t* mask = nullptr) {
    int insufficient = 0;
    uint32_t dummy;
    if (!mask) mask = &dummy;
    *mask = 0;
    const i64* al = sn.alloc(n);
    i64 rqv[3 + FGS_MAX_SCALARS];
    for (int k = 0; k < sn.RW; ++k) rqv[k] = sn.req(n)[k];
    i64 numPods = sn.podEnd(n) - sn.podBegin(n);
    if (sn.removed) // NodeInfo.RemovePod (framework/types.go:640-690): Requested -= calculateResource, one pod fewer
        for (i64 q = sn.podBegin(n); q < sn.podEnd(n); ++q) {
            if (!sn.gone(q)) continue;
            Resource res;
            i64 a, b;
            calculateResource(&sn.s->pods, q, res, a, b);
            rqv[0] = wsub(rqv[0], res.mcpu);
            rqv[1] = wsub(rqv[1], res.mem);
            rqv[2] = wsub(rqv[2], res.eph);
            for (int sc = 0; sc < sn.S; ++sc)
                if ((res.scalarPresent >> sc) & 1u) rqv[3 + sc] = wsub(rqv[3 + sc], res.scalar[sc]);
            numPods--;
        }
    if (sn.overlaid(n)) // NodeInfo.AddPodInfo of the nominated pods: Requested += calculateResource, one more pod each
        for (i64 q : sn.ovPods) {
            Resource res;
            i64 a, b;
            calculateResource(sn.ovPs, q, res, a, b);
            rqv[0] = wadd(rqv[0], res.mcpu);
            rqv[1] = wadd(rqv[1], res.mem);
            rqv[2] = wadd(rqv[2], res.eph);
            for (int sc = 0; sc < sn.S; ++sc)
                if ((res.scalarPresent >> sc) & 1u) rqv[3 + sc] = wadd(rqv[3 + sc], res.scalar[sc]);
            numPods++;
        }
    const i64* rq = rqv;
    if (numPods + 1 > al[3]) insufficient++, *mask |= FGS_FIT_REASON_PODS; // :259-268
    const Resource& r = pc.fitReq;
    if (r.mcpu == 0 && r.mem == 0 && r.eph == 0 && r.scalarPresent == 0) return insufficient; // :270-275
    if (r.mcpu > wsub(al[0], rq[0])) insufficient++, *mask |= FGS_FIT_REASON_CPU; // :277-285
    if (r.mem > wsub(al[1], rq[1])) insufficient++, *mask |= FGS_FIT_REASON_MEM; // :286-294
    if (r.eph > wsub(al[2], rq[2])) insufficient++, *mask |= FGS_FIT_REASON_EPH; // :295-303
    for (int s = 0; s < sn.S; ++s) {                 // :305-326
        if (!((r.scalarPresent >> s) & 1u)) continue;
        if ((ignoredScalarMask >> s) & 1u) continue;
        if (r.scalar[s] > wsub(al[4 + s], rq[3 + s])) insufficient++, *mask |= FGS_FIT_REASON_SCALAR0 << s;
    }
    return insufficient;
}

// leastRequestedScore / mostRequestedScore — least_allocated.go:46-57, most_allocated.go:49-61
inline i64 leastRequestedScore(i64 requested, i64 capacity) {
    if (capacity == 0) return 0;
    if (requested > capacity) return 0;
    return wmul(wsub(capacity, requested), kMaxNodeScore) / capacity;
}
inline i64 mostRequestedScore(i64 requested, i64 capacity) {
    if (capacity == 0) return 0;
    if (requested > capacity) requested = capacity;
    return wmul(requested, kMaxNodeScore) / capacity;
}

// Fit.Score → resourceAllocationScorer.score + calculateResourceAllocatableRequest +
// least/mostResourceScorer — vendor/…/noderesources/fit.go:332-339,
// resource_allocation.go:49-114, least_allocated.go:29-44, most_allocated.go:29-45
i64 fitScore(const Snap& sn, const PodCtx& pc, i64 n, const fgs_plugin_cfg* cfg) {
    i64 nodeScore = 0, weightSum = 0;
    for (int k = 0; k < cfg->fit_n_res; ++k) {
        int res = cfg->fit_res[k];
        i64 podRequest = pc.scoreReq[k];
        bool isScalar = res >= FGS_RES_SCALAR0;
        i64 alloc = 0, req = 0;
        if (res == FGS_RES_UNKNOWN) {
            alloc = 0; // :103-113 falls through to (0,0)
        } else if (podRequest == 0 && isScalar) {
            alloc = 0; // :93-97
        } else if (res == FGS_RES_CPU) {
            alloc = sn.alloc(n)[0];
            req = wadd(sn.s->nonzero[n * 2 + 0], podRequest);
        } else if (res == FGS_RES_MEM) {
            alloc = sn.alloc(n)[1];
            req = wadd(sn.s->nonzero[n * 2 + 1], podRequest);
        } else if (res == FGS_RES_EPH) {
            alloc = sn.alloc(n)[2];
            req = wadd(sn.req(n)[2], podRequest);
        } else {
            int s = res - FGS_RES_SCALAR0;
            if (sn.allocHasScalar(n, s)) { // :105-107
                alloc = sn.alloc(n)[4 + s];
                req = wadd(sn.req(n)[3 + s], podRequest);
            }
        }
        if (alloc == 0) continue; // :64-67 only non-zero allocatable enters the maps
        i64 rs = cfg->fit_strategy == 0 ? leastRequestedScore(req, alloc) : mostRequestedScore(req, alloc);
        nodeScore = wadd(nodeScore, wmul(rs, cfg->fit_res_w[k]));
        weightSum = wadd(weightSum, cfg->fit_res_w[k]);
    }
    if (weightSum == 0) return 0;
    return nodeScore / weightSum;
}

// Allocatable.Score → score → resourceScorer — pkg/noderesources/allocatable.go:61-72,120-143,
// pkg/noderesources/resource_allocation.go:51-101. `requested` is computed by the reference but
// never read by the scorer, so it is not restated.
i64 allocatableScore(const Snap& sn, i64 n, const fgs_plugin_cfg* cfg) {
    i64 nodeScore = 0, weightSum = 0;
    for (int k = 0; k < cfg->alloc_n_res; ++k) {
        int res = cfg->alloc_res[k];
        i64 alloc = 0;
        if (res == FGS_RES_CPU) alloc = sn.alloc(n)[0];
        else if (res == FGS_RES_MEM) alloc = sn.alloc(n)[1];
        else if (res == FGS_RES_EPH) alloc = sn.alloc(n)[2];
        else if (res >= FGS_RES_SCALAR0) alloc = sn.alloc(n)[4 + (res - FGS_RES_SCALAR0)]; // missing key → 0
        i64 rs = cfg->alloc_mode == 0 ? wmul(-1, alloc) : alloc; // allocatable.go:133-139
        nodeScore = wadd(nodeScore, wmul(rs, cfg->alloc_res_w[k]));
        weightSum = wadd(weightSum, cfg->alloc_res_w[k]);
    }
    return nodeScore / weightSum; // weights validated > 0 (allocatable.go:51-58)
}

// helper.DefaultNormalizeScore(MaxNodeScore, reverse=true) —
// vendor/…/plugins/helper/normalize_score.go:26-54, as called by flex_gpu.go:172-176
void defaultNormalizeScoreReverse(std::vector<i64>& scores) {
    i64 maxCount = 0;
    for (i64 s : scores)
        if (s > maxCount) maxCount = s;
    if (maxCount == 0) {
        for (i64& s : scores) s = kMaxNodeScore;
        return;
    }
    for (i64& s : scores) {
        i64 v = wmul(kMaxNodeScore, s) / maxCount;
        s = kMaxNodeScore - v;
    }
}

// Allocatable.NormalizeScore — pkg/noderesources/allocatable.go:146-171
void allocatableNormalize(std::vector<i64>& scores) {
    i64 highest = -INT64_MAX, lowest = INT64_MAX;
    for (i64 s : scores) {
        if (s > highest) highest = s;
        if (s < lowest) lowest = s;
    }
    i64 oldRange = wsub(highest, lowest);
    for (i64& s : scores) {
        if (oldRange == 0) s = 0;
        else s = wmul(wsub(s, lowest), 100) / oldRange;
    }
}

// chunkSizeFor — vendor/…/framework/parallelize/parallelism.go:42-51
int chunkSizeFor(int n, int parallelism) {
    int s = (int)std::sqrt((double)n);
    int r = n / parallelism + 1;
    if (s > r) s = r;
    else if (s < 1) s = 1;
    return s;
}

// workqueue.ParallelizeUntil — vendor/k8s.io/client-go/util/workqueue/parallelizer.go:45-98.
// Worker threads persist across calls (goroutines are cheap; threads are not).
class Pool {
  public:
    explicit Pool(int workers) : workers_(workers) {
        for (int w = 1; w < workers_; ++w) threads_.emplace_back([this] { loop(); });
    }
    ~Pool() {
        stop_.store(true);
        gen_.fetch_add(1);
        for (auto& t : threads_) t.join();
    }
    void until(int pieces, const std::function<void(int)>& fn) {
        if (pieces == 0) return;
        chunk_ = chunkSizeFor(pieces, workers_);
        pieces_ = pieces;
        chunks_ = (pieces + chunk_ - 1) / chunk_;
        fn_ = &fn;
        next_.store(0);
        if (workers_ == 1 || chunks_ == 1) {
            work();
            return;
        }
        done_.store(0);
        gen_.fetch_add(1);
        work();
        while (done_.load(std::memory_order_acquire) < (int)threads_.size()) std::this_thread::yield();
    }

  private:
    void work() {
        for (;;) {
            int c = next_.fetch_add(1);
            if (c >= chunks_) break;
            int a = c * chunk_, b = std::min(pieces_, a + chunk_);
            for (int p = a; p < b; ++p) (*fn_)(p);
        }
    }
    void loop() {
        uint64_t seen = 0;
        for (;;) {
            int spins = 0;
            while (gen_.load(std::memory_order_acquire) == seen) {
                if (++spins > 2000) std::this_thread::yield();
            }
            seen = gen_.load();
            if (stop_.load()) return;
            work();
            done_.fetch_add(1, std::memory_order_release);
        }
    }
    int workers_;
    std::vector<std::thread> threads_;
    std::atomic<uint64_t> gen_{0};
    std::atomic<bool> stop_{false};
    std::atomic<int> next_{0}, done_{0};
    int chunk_ = 1, pieces_ = 0, chunks_ = 0;
    const std::function<void(int)>* fn_ = nullptr;
};

} // namespace

// =============================================================================================

// numFeasibleNodesToFind — vendor/k8s.io/kubernetes/pkg/scheduler/generic_scheduler.go:171-191
// (minFeasibleNodesToFind = 100, minFeasibleNodesPercentageToFind = 5, :43-53); int32 arithmetic as in Go.
extern "C" int32_t fgs_oracle_num_feasible_nodes_to_find(int32_t numAllNodes, int32_t percentageOfNodesToScore) {
    const int32_t minFeasibleNodesToFind = 100, minFeasibleNodesPercentageToFind = 5;
    if (numAllNodes < minFeasibleNodesToFind || percentageOfNodesToScore >= 100) return numAllNodes;
    int32_t adaptivePercentage = percentageOfNodesToScore;
    if (adaptivePercentage <= 0) {
        const int32_t basePercentageOfNodesToScore = 50;
        adaptivePercentage = basePercentageOfNodesToScore - numAllNodes / 125;
        if (adaptivePercentage < minFeasibleNodesPercentageToFind) adaptivePercentage = minFeasibleNodesPercentageToFind;
    }
    const int32_t numNodes = numAllNodes * adaptivePercentage / 100;
    if (numNodes < minFeasibleNodesToFind) return minFeasibleNodesToFind;
    return numNodes;
}

namespace {
// sampling state of one evaluation call (percentage < 0: every node is evaluated)
struct Sampling {
    int percentage = -1;
    int64_t* start = nullptr;   // g.nextStartNodeIndex, advanced per pod
    int32_t* processed = nullptr;
    // the PodNominator's content (nil = none): pods nominated to nodes by an earlier preemption
    const fgs_podset* nominated = nullptr;
    const int32_t* nomNode = nullptr;   // [Q] snapshot index of NominatedNodeName, < 0 none
    const int64_t* nomUid = nullptr;    // [Q] optional
    const int64_t* pendingUid = nullptr; // [P] optional
    // the verdicts of the filter plugins OUTSIDE this path (the framework's other Filter plugins): [p1-p0][N], 0 = some
    // other plugin rejected the node, so it is not in the list RunScorePlugins is handed (generic_scheduler.go:420-426)
    const uint8_t* extMask = nullptr;
};
int evaluateImpl(const fgs_snapshot* snap, const fgs_podset* pending, const fgs_plugin_cfg* cfg, int64_t p0,
                 int64_t p1, int workers, fgs_oracle_out* out, const Sampling& smp);
} // namespace

extern "C" int fgs_oracle_evaluate(const fgs_snapshot* snap, const fgs_podset* pending,
                                   const fgs_plugin_cfg* cfg, int64_t p0, int64_t p1, int workers,
                                   fgs_oracle_out* out) {
    return evaluateImpl(snap, pending, cfg, p0, p1, workers, out, Sampling());
}

// fgs_oracle_evaluate with a PodNominator: RunFilterPluginsWithNominatedPods (runtime/framework.go:752-823)
extern "C" int fgs_oracle_evaluate_nominated(const fgs_snapshot* snap, const fgs_podset* pending, const fgs_plugin_cfg* cfg,
                                             int64_t p0, int64_t p1, int workers, const fgs_podset* nominated,
                                             const int32_t* nom_node, const int64_t* nom_uid, const int64_t* pending_uid,
                                             fgs_oracle_out* out) {
    Sampling smp;
    if (nominated && nominated->n_pods > 0 && nom_node) {
        smp.nominated = nominated;
        smp.nomNode = nom_node;
        smp.nomUid = nom_uid;
        smp.pendingUid = pending_uid;
    }
    return evaluateImpl(snap, pending, cfg, p0, p1, workers, out, smp);
}

// findNodesThatPassFilters with the reference's node sampling, as it runs with ONE worker
// (generic_scheduler.go:266-338; with 16 workers which nodes are found is racy, :302-308)
extern "C" int fgs_oracle_evaluate_sampled(const fgs_snapshot* snap, const fgs_podset* pending,
                                           const fgs_plugin_cfg* cfg, int64_t p0, int64_t p1, int percentage,
                                           int64_t* start_index, int32_t* processed, fgs_oracle_out* out) {
    if (!start_index || percentage < 0) return -1;
    Sampling smp;
    smp.percentage = percentage;
    smp.start = start_index;
    smp.processed = processed;
    return evaluateImpl(snap, pending, cfg, p0, p1, 1, out, smp);
}

extern "C" int fgs_oracle_evaluate_masked(const fgs_snapshot* snap, const fgs_podset* pending, const fgs_plugin_cfg* cfg, int64_t p0,
                                          int64_t p1, int workers, const uint8_t* ext_mask, fgs_oracle_out* out) {
    Sampling smp;
    smp.extMask = ext_mask;
    return evaluateImpl(snap, pending, cfg, p0, p1, workers, out, smp);
}

namespace {
int evaluateImpl(const fgs_snapshot* snap, const fgs_podset* pending, const fgs_plugin_cfg* cfg, int64_t p0,
                 int64_t p1, int workers, fgs_oracle_out* out, const Sampling& smp) {
    Snap sn(snap);
    const i64 N = snap->n_nodes;
    if (workers < 1) workers = 1;
    Pool pool(workers);
    std::vector<uint8_t> status((size_t)N);
    std::vector<uint8_t> panicked((size_t)N);
    std::vector<int> feasible;
    std::vector<i64> sFlex, sFit, sAlloc, rFlex, rFit, rAlloc, tot;
    i64 wFlex = cfg->weight_flex ? cfg->weight_flex : 1; // runtime/framework.go:405-408
    i64 wFit = cfg->weight_fit ? cfg->weight_fit : 1;
    i64 wAlloc = cfg->weight_alloc ? cfg->weight_alloc : 1;

    for (i64 p = p0; p < p1; ++p) {
        const i64 row = p - p0;
        // PreFilter: Fit computes the pod request once (fit.go:182-185)
        PodCtx pc;
        pc.ps = pending;
        pc.i = p;
        pc.fitReq = computePodResourceRequest(pending, p);
        for (int k = 0; k < cfg->fit_n_res; ++k)
            pc.scoreReq[k] = cfg->fit_res[k] == FGS_RES_UNKNOWN ? 0 : scorePodRequest(pending, p, cfg->fit_res[k]);
        uint32_t err = 0;

        // findNodesThatPassFilters — vendor/…/scheduler/generic_scheduler.go:266-338 with
        // numFeasibleNodesToFind == N; RunFilterPlugins — runtime/framework.go:667-693: plugins in
        // profile order, early exit on the first non-success.
        std::function<void(int)> checkNode = [&](int n) {
            uint8_t st = FGS_CODE_SUCCESS;
            panicked[(size_t)n] = 0;
            // A NodeInfo whose Node() is nil is not an element of the list this loop walks: UpdateSnapshot builds
            // nodeInfoList from the node tree and drops entries without a Node
            // (vendor/…/scheduler/internal/cache/cache.go:254,290,314-325). The flat snapshot can still carry
            // such rows (the gang path skips them explicitly, core.go:324-326): report Error with no plugin.
            if (!sn.s->node_valid[n]) {
                status[(size_t)n] = FGS_CODE_ERROR;
                return;
            }
            // RunFilterPlugins — runtime/framework.go:667-693 — on a node view
            auto runFilters = [&](const Snap& view) {
                uint8_t r = FGS_CODE_SUCCESS;
                for (int k = 0; k < 2 && cfg->filter_order[k]; ++k) {
                    int pl = cfg->filter_order[k];
                    if (!(cfg->filter_mask & (uint32_t)pl)) continue;
                    int code = FGS_CODE_SUCCESS;
                    bool pan = false;
                    if (pl == FGS_PL_FLEXGPU) {
                        try {
                            code = flexFilter(view, pc, n);
                        } catch (const RefPanic&) {
                            pan = true;
                            panicked[(size_t)n] = 1;
                            code = FGS_CODE_ERROR;
                        }
                    } else if (pl == FGS_PL_FIT) {
                        code = fitsRequest(view, pc, n, cfg->fit_ignored_scalar_mask) ? FGS_CODE_UNSCHEDULABLE : FGS_CODE_SUCCESS;
                    }
                    if (code != FGS_CODE_SUCCESS) {
                        r = (uint8_t)(((pan ? 0xF : pl) << 4) | code);
                        break;
                    }
                }
                return r;
            };
            // RunFilterPluginsWithNominatedPods — runtime/framework.go:752-793: first with the nominated pods of equal
            // or higher priority added to the node (addNominatedPods :795-823), and only if that passes once more
            // without them; the status of the last pass that ran is the node's status
            bool podsAdded = false;
            if (smp.nominated) {
                Snap view = sn;
                view.ovNode = n;
                view.ovPs = smp.nominated;
                const int32_t myPrio = pending->priority ? pending->priority[p] : 0;
                for (i64 q = 0; q < smp.nominated->n_pods; ++q) {
                    if (smp.nomNode[q] != n) continue;
                    const int32_t pr = smp.nominated->priority ? smp.nominated->priority[q] : 0;
                    if (pr < myPrio) continue;
                    if (smp.nomUid && smp.pendingUid && smp.nomUid[q] == smp.pendingUid[p]) continue; // :812
                    view.ovPods.push_back(q);
                }
                if (!view.ovPods.empty()) {
                    podsAdded = true;
                    st = runFilters(view);
                }
            }
            if (!podsAdded || st == FGS_CODE_SUCCESS) st = runFilters(sn);
            status[(size_t)n] = st;
        };
        if (smp.percentage < 0) {
            pool.until((int)N, checkNode);
        } else if (N > 0) {
            // one worker: nodes[(nextStartNodeIndex+i) % len] for i = 0,1,…; the (K+1)-th success cancels the walk
            // and is itself dropped (:296-308); processedNodes = len(feasible) + len(NodeToStatusMap) (:329-330)
            const i64 K = fgs_oracle_num_feasible_nodes_to_find((int32_t)N, smp.percentage);
            const i64 start = ((*smp.start % N) + N) % N;
            std::fill(status.begin(), status.end(), (uint8_t)0xEE); // not visited
            std::fill(panicked.begin(), panicked.end(), (uint8_t)0);
            i64 found = 0, failed = 0;
            for (i64 i = 0; i < N; ++i) {
                const i64 n = (start + i) % N;
                checkNode((int)n);
                if (panicked[(size_t)n]) break; // status.Code() == Error: SendErrorWithCancel (:290-293)
                if (status[(size_t)n] == FGS_CODE_SUCCESS) {
                    if (found + 1 > K) {
                        status[(size_t)n] = 0xEE; // found but not kept: neither feasible nor in the status map
                        break;
                    }
                    found++;
                } else {
                    failed++;
                }
            }
            const i64 proc = found + failed;
            if (smp.processed) smp.processed[row] = (int32_t)proc;
            *smp.start = (start + proc) % N;
        }

        feasible.clear();
        for (i64 n = 0; n < N; ++n) {
            // a node another filter plugin of the profile rejected: Unschedulable with a plugin id outside this path
            if (smp.extMask && status[(size_t)n] == FGS_CODE_SUCCESS && !smp.extMask[row * N + n]) status[(size_t)n] = 0xE0 | FGS_CODE_UNSCHEDULABLE;
            if (status[(size_t)n] == FGS_CODE_SUCCESS) feasible.push_back((int)n);
            if (panicked[(size_t)n]) err |= FGS_PODERR_REF_PANIC;
        }
        const int F = (int)feasible.size();
        if (out->status) std::memcpy(out->status + row * N, status.data(), (size_t)N);
        if (out->feasible_count) out->feasible_count[row] = F;

        // prioritizeNodes → RunScorePlugins — generic_scheduler.go:398-447,
        // runtime/framework.go:861-933. Scoring errors only exist when it runs, i.e. F >= 2
        // (generic_scheduler.go:112-127).
        sFlex.assign((size_t)F, 0);
        sFit.assign((size_t)F, 0);
        sAlloc.assign((size_t)F, 0);
        std::atomic<int> scorePanic{0};
        std::function<void(int)> scoreNode = [&](int j) {
            int n = feasible[(size_t)j];
            if (cfg->score_mask & FGS_PL_FLEXGPU) {
                try {
                    sFlex[(size_t)j] = flexScore(sn, pc, n);
                } catch (const RefPanic&) {
                    scorePanic.store(1);
                }
            }
            if (cfg->score_mask & FGS_PL_FIT) sFit[(size_t)j] = fitScore(sn, pc, n, cfg);
            if (cfg->score_mask & FGS_PL_ALLOCATABLE) sAlloc[(size_t)j] = allocatableScore(sn, n, cfg);
        };
        pool.until(F, scoreNode);
        if (scorePanic.load() && F >= 2) err |= FGS_PODERR_REF_PANIC;
        rFlex = sFlex;
        rFit = sFit;
        rAlloc = sAlloc;
        if (F > 0) {
            if (cfg->score_mask & FGS_PL_FLEXGPU) defaultNormalizeScoreReverse(sFlex); // flex_gpu.go:172-176
            // Fit.ScoreExtensions() == nil (fit.go:87-89): no normalisation
            if (cfg->score_mask & FGS_PL_ALLOCATABLE) allocatableNormalize(sAlloc);
        }
        tot.assign((size_t)F, 0);
        bool range = false;
        for (int j = 0; j < F; ++j) {
            i64 t = 0;
            auto apply = [&](uint32_t pl, i64 s, i64 w) { // runtime/framework.go:912-927
                if (!(cfg->score_mask & pl)) return;
                if (s > kMaxNodeScore || s < 0) range = true;
                t = wadd(t, wmul(s, w));
            };
            apply(FGS_PL_FLEXGPU, sFlex[(size_t)j], wFlex);
            apply(FGS_PL_FIT, sFit[(size_t)j], wFit);
            apply(FGS_PL_ALLOCATABLE, sAlloc[(size_t)j], wAlloc);
            tot[(size_t)j] = t; // generic_scheduler.go:440-447
        }
        if (range && F >= 2) err |= FGS_PODERR_SCORE_RANGE;

        // selectHost — generic_scheduler.go:146-167: the tie set of the maximum
        i64 best = INT64_MIN;
        int ties = 0, first = -1;
        for (int j = 0; j < F; ++j) {
            if (tot[(size_t)j] > best) {
                best = tot[(size_t)j];
                ties = 1;
                first = feasible[(size_t)j];
            } else if (tot[(size_t)j] == best) {
                ties++;
            }
        }
        if (out->pod_err) out->pod_err[row] = err;
        if (out->max_total) out->max_total[row] = best;
        if (out->tie_count) out->tie_count[row] = ties;
        if (out->first_max) out->first_max[row] = first;
        auto scatter = [&](int64_t* dst, const std::vector<i64>& v) {
            if (!dst) return;
            int64_t* d = dst + row * N;
            for (i64 n = 0; n < N; ++n) d[n] = 0;
            for (int j = 0; j < F; ++j) d[feasible[(size_t)j]] = v[(size_t)j];
        };
        scatter(out->raw_flex, rFlex);
        scatter(out->raw_fit, rFit);
        scatter(out->raw_alloc, rAlloc);
        scatter(out->norm_flex, sFlex);
        scatter(out->norm_fit, sFit);
        scatter(out->norm_alloc, sAlloc);
        scatter(out->total, tot);
    }
    return 0;
}
} // namespace

// The reason inside each plugin's Filter status for pod p on every node: flex[n] = FGS_FLEX_REASON_* (0xF1 where
// NewGPUNode panics), fit[n] = FGS_FIT_REASON_* bits. Either may be NULL.
extern "C" int fgs_oracle_filter_reasons(const fgs_snapshot* snap, const fgs_podset* pending, const fgs_plugin_cfg* cfg,
                                         int64_t p, uint16_t* flex, uint16_t* fit) {
    Snap sn(snap);
    PodCtx pc;
    pc.ps = pending;
    pc.i = p;
    pc.fitReq = computePodResourceRequest(pending, p);
    for (i64 n = 0; n < snap->n_nodes; ++n) {
        if (flex) {
            int reason = 0;
            try {
                flexFilter(sn, pc, n, &reason);
            } catch (const RefPanic&) {
                reason = FGS_FLEX_REASON_PANIC;
            }
            flex[n] = (uint16_t)reason;
        }
        if (fit) {
            uint32_t mask = 0;
            fitsRequest(sn, pc, n, cfg->fit_ignored_scalar_mask, &mask);
            fit[n] = (uint16_t)mask;
        }
    }
    return 0;
}

// FlexGPU.Reserve — pkg/flexgpu/flex_gpu.go:178-223
extern "C" int fgs_oracle_reserve_index(const fgs_snapshot* snap, const fgs_podset* pending, int64_t p,
                                        int64_t n) {
    Snap sn(snap);
    bool gpuExist, memExist;
    podResourceLimit(pending, p, FGS_SCALAR_FLEX_GPU, gpuExist);
    i64 memLimit = podResourceLimit(pending, p, FGS_SCALAR_FLEX_MEM, memExist);
    if (!gpuExist && !memExist) return -2;
    if (gpuExist && memExist) return -3;
    try {
        GpuNode no = newGPUNode(sn, n);
        if (gpuExist) {
            std::vector<int> idx = gpuAssumeFitIndexes(no);
            return idx.empty() ? -1 : idx[0];
        }
        std::vector<int> idx = memAssumeFitIndexes(no, memLimit);
        return idx.empty() ? -1 : idx[0];
    } catch (const RefPanic&) {
        return -4;
    }
}

// ---- coscheduling ---------------------------------------------------------------------------

namespace {

// getNodeResource — pkg/coscheduling/core/core.go:349-382: clone the NodeInfo, RemovePod every pod
// whose full group name equals the desired one (vendor/…/framework/types.go:640-690), then
// left = allocatable − requested; pods left = AllowedPodNumber − len(Pods).
// Returned as an {mcpu, mem, eph, pods, scalars} row plus the set of scalar keys.
void getNodeResource(const Snap& sn, i64 n, int32_t desiredPG, i64* left, uint32_t& scalarKeys) {
    const int S = sn.S;
    i64 req[3 + FGS_MAX_SCALARS];
    for (int k = 0; k < 3 + S; ++k) req[k] = sn.req(n)[k];
    i64 numPods = sn.podEnd(n) - sn.podBegin(n);
    const fgs_podset* ps = &sn.s->pods;
    for (i64 q = sn.podBegin(n); q < sn.podEnd(n); ++q) {
        // util.GetPodGroupFullName (pkg/util/podgroup.go:55-62): "" when the pod has no label, and
        // the callers pass "" for "no group", so only labelled pods can match.
        if (ps->pg[q] < 0 || ps->pg[q] != desiredPG) continue;
        Resource res;
        i64 a, b;
        calculateResource(ps, q, res, a, b);
        req[0] = wsub(req[0], res.mcpu);
        req[1] = wsub(req[1], res.mem);
        req[2] = wsub(req[2], res.eph);
        for (int s = 0; s < S; ++s)
            if ((res.scalarPresent >> s) & 1u) req[3 + s] = wsub(req[3 + s], res.scalar[s]);
        numPods--;
    }
    const i64* al = sn.alloc(n);
    left[0] = wsub(al[0], req[0]);
    left[1] = wsub(al[1], req[1]);
    left[2] = wsub(al[2], req[2]);
    left[3] = wsub(al[3], numPods);
    scalarKeys = sn.s->alloc_scalar_present[n];
    for (int s = 0; s < S; ++s) left[4 + s] = ((scalarKeys >> s) & 1u) ? wsub(al[4 + s], req[3 + s]) : 0;
}

// CheckClusterResource — pkg/coscheduling/core/core.go:322-342. resourceRequest is a map: `rem`
// holds the remaining quantity per key, `live` the keys still in the map.
bool checkClusterResource(const Snap& sn, const i64* request, uint32_t keys, int32_t desiredPG) {
    const int K = 4 + sn.S;
    i64 rem[4 + FGS_MAX_SCALARS];
    for (int k = 0; k < K; ++k) rem[k] = request[k];
    uint32_t live = keys;
    for (i64 n = 0; n < sn.s->n_nodes; ++n) {
        if (!sn.s->node_valid[n]) continue; // :324-326
        i64 left[4 + FGS_MAX_SCALARS];
        uint32_t sk;
        getNodeResource(sn, n, desiredPG, left, sk);
        for (int k = 0; k < K; ++k) {
            if (!((live >> k) & 1u)) continue;
            // util.ResourceList always has cpu/memory/pods/ephemeral (pkg/util/resource.go:28-43);
            // a scalar the node lacks reads as the zero Quantity.
            i64 have = (k < 4 || ((sk >> (k - 4)) & 1u)) ? left[k] : 0;
            rem[k] = wsub(rem[k], have);
            if (rem[k] <= 0) live &= ~(1u << k); // :331-334
        }
        if (live == 0) return true; // :337-339
    }
    return false;
}

// the `pods` list of core.go:158-160: every pod the lister knows in the namespace with the label.
// Here: pods on nodes plus pending pods.
// `unlisted` (optional) marks pending pods the lister does not know: the reference's unit test
// evaluates a pod that was never added to the informer store (core_test.go:159-164).
i64 siblingCount(const fgs_snapshot* snap, const fgs_podset* pending, int32_t pg, const uint8_t* unlisted) {
    i64 c = 0;
    for (i64 q = 0; q < snap->pods.n_pods; ++q)
        if (snap->pods.pg[q] == pg) c++;
    for (i64 q = 0; q < pending->n_pods; ++q)
        if (pending->pg[q] == pg && !(unlisted && unlisted[q])) c++;
    return c;
}

} // namespace

extern "C" int fgs_oracle_check_cluster_resource(const fgs_snapshot* snap, const fgs_podgroups* pgs,
                                                 int64_t g) {
    Snap sn(snap);
    const int K = 4 + sn.S;
    i64 request[4 + FGS_MAX_SCALARS];
    for (int k = 0; k < K; ++k) request[k] = pgs->min_res[g * K + k];
    uint32_t keys = pgs->min_res_present[g];
    request[3] = pgs->min_member[g]; // core.go:185-187
    keys |= 1u << 3;
    return checkClusterResource(sn, request, keys, (int32_t)g) ? 1 : 0;
}

extern "C" int fgs_oracle_check_cluster_resource_raw(const fgs_snapshot* snap, const int64_t* request,
                                                     uint32_t keys, int32_t desired_pg) {
    Snap sn(snap);
    return checkClusterResource(sn, request, keys, desired_pg) ? 1 : 0;
}

extern "C" int fgs_oracle_gang_prefilter_listed(const fgs_snapshot* snap, const fgs_podset* pending,
                                                const fgs_podgroups* pgs, int64_t p0, int64_t p1, uint8_t* code,
                                                const uint8_t* unlisted);

extern "C" int fgs_oracle_gang_prefilter(const fgs_snapshot* snap, const fgs_podset* pending,
                                         const fgs_podgroups* pgs, int64_t p0, int64_t p1, uint8_t* code) {
    return fgs_oracle_gang_prefilter_listed(snap, pending, pgs, p0, p1, code, nullptr);
}

extern "C" int fgs_oracle_gang_prefilter_listed(const fgs_snapshot* snap, const fgs_podset* pending,
                                                const fgs_podgroups* pgs, int64_t p0, int64_t p1, uint8_t* code,
                                                const uint8_t* unlisted) {
    for (i64 p = p0; p < p1; ++p) {
        int32_t g = pending->pg[p];
        uint8_t c = 0;
        if (g < 0 || !pgs->exists[g]) { // core.go:151-154
            c = 0;
        } else if (pgs->denied[g]) { // :155-157
            c = 1;
        } else if (siblingCount(snap, pending, g, unlisted) < pgs->min_member[g]) { // :158-167
            c = 2;
        } else if (!pgs->has_min_resources[g]) { // :169-171
            c = 0;
        } else if (pgs->permitted[g]) { // :176-178
            c = 0;
        } else {
            c = fgs_oracle_check_cluster_resource(snap, pgs, g) ? 0 : 3; // :180-193
        }
        code[p - p0] = c;
    }
    return 0;
}

extern "C" int fgs_oracle_gang_permit(const fgs_snapshot* snap, const fgs_podset* pending,
                                      const fgs_podgroups* pgs, int64_t p0, int64_t p1, uint8_t* status,
                                      int32_t* assigned) {
    // CalculateAssignedPods — core.go:301-318: pods in ALL NodeInfos of the snapshot with the group
    // label, the group's namespace and a non-empty Spec.NodeName.
    auto calc = [&](int32_t g) {
        int cnt = 0;
        for (i64 q = 0; q < snap->pods.n_pods; ++q)
            if (snap->pods.pg[q] == g && snap->pods.ns[q] == pgs->ns[g] && snap->pods.has_node_name[q]) cnt++;
        return cnt;
    };
    if (assigned)
        for (i64 g = 0; g < pgs->n_groups; ++g) assigned[g] = calc((int32_t)g);
    for (i64 p = p0; p < p1; ++p) {
        int32_t g = pending->pg[p];
        uint8_t st;
        if (g < 0) st = 2;                  // PodGroupNotSpecified :201-203
        else if (!pgs->exists[g]) st = 3;   // PodGroupNotFound :204-207
        else st = (calc(g) + 1 >= pgs->min_member[g]) ? 0 : 1; // :209-215
        if (status) status[p - p0] = st;
    }
    return 0;
}

// ---- capacity scheduling ----------------------------------------------------------------------

namespace {

struct QRes { // framework.Resource restricted to what cmp2 reads
    i64 mcpu = 0, mem = 0;
    i64 scalar[FGS_MAX_SCALARS] = {0};
    uint32_t scalarPresent = 0;
};

// cmp2 — pkg/capacityscheduling/elasticquota.go:165-181: the scalar loop ranges over x1's keys
// only; a key missing from x2 or y reads 0.
bool cmp2(const QRes& x1, const QRes& x2, const QRes& y, int S) {
    if (wadd(x1.mcpu, x2.mcpu) > y.mcpu) return true;
    if (wadd(x1.mem, x2.mem) > y.mem) return true;
    for (int s = 0; s < S; ++s)
        if ((x1.scalarPresent >> s) & 1u)
            if (wadd(x1.scalar[s], x2.scalar[s]) > y.scalar[s]) return true;
    return false;
}

QRes rowToQRes(const i64* row, uint32_t present, int S) {
    QRes r;
    r.mcpu = row[0];
    r.mem = row[1];
    for (int s = 0; s < S; ++s) r.scalar[s] = ((present >> s) & 1u) ? row[2 + s] : 0; // missing key reads 0
    r.scalarPresent = present;
    return r;
}

// Resource.Add(util.ResourceList(r)) — util.ResourceList (pkg/util/resource.go:28-43) emits
// cpu, memory, pods, ephemeral and every scalar key; Add copies the scalar keys over.
void qAdd(QRes& acc, const Resource& r, int S) {
    acc.mcpu = wadd(acc.mcpu, r.mcpu);
    acc.mem = wadd(acc.mem, r.mem);
    for (int s = 0; s < S; ++s)
        if ((r.scalarPresent >> s) & 1u) {
            acc.scalar[s] = wadd(acc.scalar[s], r.scalar[s]);
            acc.scalarPresent |= 1u << s;
        }
}
void qAddQ(QRes& acc, const QRes& r, int S) {
    acc.mcpu = wadd(acc.mcpu, r.mcpu);
    acc.mem = wadd(acc.mem, r.mem);
    for (int s = 0; s < S; ++s)
        if ((r.scalarPresent >> s) & 1u) {
            acc.scalar[s] = wadd(acc.scalar[s], r.scalar[s]);
            acc.scalarPresent |= 1u << s;
        }
}

} // namespace

// ElasticQuotaInfo.reserveResource / unreserveResource — pkg/capacityscheduling/elasticquota.go:74-88, with the
// request addPodIfNotPresent / deletePodIfPresent hand them (computePodResourceRequest of the pod, :126-147):
// Used.Memory, Used.MilliCPU and every scalar KEY of the request (SetScalar creates the key in Used).
// used = one namespace's row {mcpu, mem, scalars}; *used_present = its scalar keys. op: +1 reserve, -1 unreserve.
extern "C" int fgs_oracle_quota_reserve(const fgs_podset* pods, int64_t p, int op, int64_t* used,
                                        uint32_t* used_present) {
    const int S = pods->n_scalars;
    Resource r = computePodResourceRequest(pods, p);
    auto app = [&](i64 a, i64 b) { return op >= 0 ? wadd(a, b) : wsub(a, b); };
    used[0] = app(used[0], r.mcpu);
    used[1] = app(used[1], r.mem);
    for (int s = 0; s < S; ++s)
        if ((r.scalarPresent >> s) & 1u) {
            const i64 cur = ((*used_present >> s) & 1u) ? used[2 + s] : 0; // a missing map key reads as 0
            used[2 + s] = app(cur, r.scalar[s]);
            *used_present |= 1u << s;
        }
    return 0;
}

extern "C" int fgs_oracle_quota_prefilter(const fgs_podset* pending, const fgs_quotas* q,
                                          const fgs_podset* nominated, const int64_t* pending_uid,
                                          const int64_t* nominated_uid, int64_t p0, int64_t p1,
                                          uint8_t* code) {
    const int S = q->n_scalars, W = 2 + S;
    const uint32_t allScalars = (S >= 32) ? 0xffffffffu : ((1u << S) - 1u);
    for (i64 p = p0; p < p1; ++p) {
        // capacity_scheduling.go:204-217
        Resource podReq = computePodResourceRequest(pending, p);
        int32_t ns = pending->ns[p];
        if (ns < 0 || ns >= q->n_ns || !q->has_eq[ns]) {
            code[p - p0] = 0;
            continue;
        }
        QRes inEQ, all; // nominatedPodsReqInEQWithPodReq, nominatedPodsReqWithPodReq (:219-231)
        if (nominated) {
            // :233-255 — the per-node NominatedPodsForNode lists, concatenated
            for (i64 k = 0; k < nominated->n_pods; ++k) {
                if (pending_uid && nominated_uid && nominated_uid[k] == pending_uid[p]) continue; // :237
                int32_t pns = nominated->ns[k];
                if (pns < 0 || pns >= q->n_ns || !q->has_eq[pns]) continue; // :241-242
                Resource pr = computePodResourceRequest(nominated, k);
                if (pns == ns && nominated->priority[k] >= pending->priority[p]) { // :248-250
                    qAdd(inEQ, pr, S);
                    qAdd(all, pr, S);
                } else if (pns != ns) {
                    // info.usedOverMin() = cmp(Used, Min) (elasticquota.go:98-100,161-163)
                    QRes used = rowToQRes(q->used + (i64)pns * W, q->used_scalar_present[pns], S);
                    QRes mn = rowToQRes(q->min + (i64)pns * W, allScalars, S);
                    if (!cmp2(used, QRes(), mn, S)) qAdd(all, pr, S); // :251-253
                }
            }
        }
        qAdd(inEQ, podReq, S); // :257-258
        qAdd(all, podReq, S);
        // eq.usedOverMaxWith — elasticquota.go:94-96: cmp2(podRequest, e.Used, e.Max)
        QRes used = rowToQRes(q->used + (i64)ns * W, q->used_scalar_present[ns], S);
        QRes mx = rowToQRes(q->max + (i64)ns * W, allScalars, S);
        if (cmp2(inEQ, used, mx, S)) { // :266-268
            code[p - p0] = 1;
            continue;
        }
        // aggregatedUsedOverMinWith — elasticquota.go:40-51
        QRes aggUsed, aggMin;
        for (i64 k = 0; k < q->n_ns; ++k) {
            if (!q->has_eq[k]) continue;
            qAddQ(aggUsed, rowToQRes(q->used + k * W, q->used_scalar_present[k], S), S);
            qAddQ(aggMin, rowToQRes(q->min + k * W, allScalars, S), S);
        }
        qAddQ(aggUsed, all, S);
        code[p - p0] = cmp2(aggUsed, QRes(), aggMin, S) ? 2 : 0; // :270-272
    }
    return 0;
}

// ---- dry-run preemption ----------------------------------------------------------------------------------------
//
// preemptor.SelectVictimsOnNode — pkg/capacityscheduling/capacity_scheduling.go:465-644, for every node of the
// snapshot (preemption.Evaluator.DryRunPreemption, vendor/…/framework/preemption/preemption.go:546-597, clones the
// NodeInfo and the CycleState per node, so nodes are independent). No PodDisruptionBudgets (every victim is
// "non-violating", :593) and an empty PodNominator (PreFilterState.nominatedPodsReq* == podReq, as in the
// reference's TestDryRunPreemption :311-318).
// tracks_pods: the CapacityScheduling PreFilterExtensions AddPod / RemovePod (:277-318) move a pod's request in and
// out of its quota's Used — only when the pod is in the quota's pod set, which is the case for every assigned pod
// in production and for none in the reference's unit test (its ElasticQuotaInfo literals carry no pod set and the
// test framework registers no CapacityScheduling PreFilter plugin).
// sort.Slice is restated as insertion sort (what Go runs below 12 elements); with more pods per node Go's
// quicksort may order EQUALLY important pods differently — pods of distinct (priority, start time) are unaffected.
extern "C" int fgs_oracle_select_victims(const fgs_snapshot* snap, const fgs_podset* pending, int64_t p,
                                         const fgs_plugin_cfg* cfg, const fgs_quotas* q, int tracks_pods,
                                         uint8_t* node_status, uint8_t* victim, int32_t* n_victims) {
    Snap sn(snap);
    const i64 N = snap->n_nodes;
    const int S = snap->n_scalars, W = 2 + S;
    const fgs_podset* ps = &snap->pods;
    const uint32_t allScalars = (S >= 32) ? 0xffffffffu : ((1u << S) - 1u);
    PodCtx pc;
    pc.ps = pending;
    pc.i = p;
    pc.fitReq = computePodResourceRequest(pending, p);
    QRes podReq;
    qAdd(podReq, pc.fitReq, S);
    const int32_t podNs = pending->ns[p], podPrio = pending->priority ? pending->priority[p] : 0;
    auto hasEQ = [&](int32_t ns) { return q && ns >= 0 && ns < q->n_ns && q->has_eq[ns]; };
    auto prio = [&](i64 a) { return ps->priority ? ps->priority[a] : 0; };
    auto start = [&](i64 a) { return ps->start_time ? ps->start_time[a] : (i64)0; };
    // util.MoreImportantPod (vendor/…/scheduler/util/utils.go): higher priority, then the earlier start time
    auto moreImportant = [&](i64 a, i64 b) { return prio(a) != prio(b) ? prio(a) > prio(b) : start(a) < start(b); };
    auto insertionSort = [](std::vector<i64>& d, const std::function<bool(i64, i64)>& less) {
        for (size_t i = 1; i < d.size(); ++i)
            for (size_t j = i; j > 0 && less(d[j], d[j - 1]); --j) std::swap(d[j], d[j - 1]);
    };
    std::vector<uint8_t> removed((size_t)std::max<i64>(ps->n_pods, 1), 0);
    if (victim) std::memset(victim, 0, (size_t)ps->n_pods);
    for (i64 n = 0; n < N; ++n) {
        if (n_victims) n_victims[n] = 0;
        if (!snap->node_valid[n]) {
            node_status[n] = FGS_PREEMPT_NOT_A_NODE;
            continue;
        }
        // the per-node clone of the ElasticQuotaSnapshotState
        std::vector<i64> used;
        std::vector<uint32_t> present;
        if (q) {
            used.assign(q->used, q->used + q->n_ns * W);
            present.assign(q->used_scalar_present, q->used_scalar_present + q->n_ns);
        }
        auto usedQ = [&](int32_t ns) { return rowToQRes(used.data() + (i64)ns * W, present[(size_t)ns], S); };
        auto quotaMove = [&](i64 a, int op) { // RemovePod / AddPod extensions → unreserveResource / reserveResource
            const int32_t ns = ps->ns[a];
            if (!tracks_pods || !hasEQ(ns)) return;
            fgs_oracle_quota_reserve(ps, a, op, used.data() + (i64)ns * W, &present[(size_t)ns]);
        };
        auto removePod = [&](i64 a) {
            removed[(size_t)a] = 1;
            quotaMove(a, -1);
        };
        auto addPod = [&](i64 a) {
            removed[(size_t)a] = 0;
            quotaMove(a, +1);
        };
        Snap view = sn;
        view.removed = &removed;
        bool panicked = false;
        auto fits = [&]() { // RunFilterPluginsWithNominatedPods with an empty nominator = RunFilterPlugins (:667-693)
            for (int k = 0; k < 2 && cfg->filter_order[k]; ++k) {
                const int pl = cfg->filter_order[k];
                if (!(cfg->filter_mask & (uint32_t)pl)) continue;
                int code = FGS_CODE_SUCCESS;
                if (pl == FGS_PL_FLEXGPU) {
                    try {
                        code = flexFilter(view, pc, n);
                    } catch (const RefPanic&) {
                        panicked = true;
                        return false;
                    }
                } else if (pl == FGS_PL_FIT) {
                    code = fitsRequest(view, pc, n, cfg->fit_ignored_scalar_mask) ? FGS_CODE_UNSCHEDULABLE : FGS_CODE_SUCCESS;
                }
                if (code != FGS_CODE_SUCCESS) return false;
            }
            return true;
        };
        auto overQuota = [&]() { // usedOverMaxWith(podReq) || aggregatedUsedOverMinWith(podReq) (:580-585, :606-607)
            if (cmp2(podReq, usedQ(podNs), rowToQRes(q->max + (i64)podNs * W, allScalars, S), S)) return true;
            QRes aggUsed, aggMin;
            for (i64 k = 0; k < q->n_ns; ++k) {
                if (!q->has_eq[k]) continue;
                qAddQ(aggUsed, usedQ((int32_t)k), S);
                qAddQ(aggMin, rowToQRes(q->min + k * W, allScalars, S), S);
            }
            qAddQ(aggUsed, podReq, S);
            return cmp2(aggUsed, QRes(), aggMin, S);
        };

        // :505 sort the pods in node by the priority class (least important first)
        std::vector<i64> pods;
        for (i64 a = sn.podBegin(n); a < sn.podEnd(n); ++a) pods.push_back(a);
        insertionSort(pods, [&](i64 a, i64 b) { return !moreImportant(a, b); });
        std::vector<i64> potential;
        const bool withEQ = hasEQ(podNs);
        if (withEQ) { // :508-543
            const bool moreThanMin = cmp2(podReq, usedQ(podNs), rowToQRes(q->min + (i64)podNs * W, allScalars, S), S);
            for (i64 a : pods) {
                const int32_t ns = ps->ns[a];
                if (!hasEQ(ns)) continue;
                if (moreThanMin) {
                    if (ns == podNs && prio(a) < podPrio) potential.push_back(a), removePod(a);
                } else {
                    const bool overMin = cmp2(usedQ(ns), QRes(), rowToQRes(q->min + (i64)ns * W, allScalars, S), S); // usedOverMin()
                    if (ns != podNs && overMin) potential.push_back(a), removePod(a);
                }
            }
        } else { // :544-557
            for (i64 a : pods) {
                if (hasEQ(ps->ns[a])) continue;
                if (prio(a) < podPrio) potential.push_back(a), removePod(a);
            }
        }
        auto finish = [&](uint8_t st, const std::vector<i64>& victims) {
            node_status[n] = st;
            if (st == FGS_PREEMPT_CANDIDATE) {
                if (victim)
                    for (i64 a : victims) victim[a] = 1;
                if (n_victims) n_victims[n] = (int32_t)victims.size();
            }
            for (i64 a : pods) removed[(size_t)a] = 0;
        };
        if (potential.empty()) {
            finish(FGS_PREEMPT_NO_VICTIMS, {});
            continue;
        }
        if (!fits()) { // :573-575
            finish(panicked ? FGS_PREEMPT_PANIC : FGS_PREEMPT_STILL_UNFIT, {});
            continue;
        }
        if (withEQ && overQuota()) { // :580-585
            finish(FGS_PREEMPT_QUOTA, {});
            continue;
        }
        insertionSort(potential, [&](i64 a, i64 b) { return moreImportant(a, b); }); // :589-591 most important first
        std::vector<i64> victims;
        uint8_t st = FGS_PREEMPT_CANDIDATE;
        for (i64 a : potential) { // reprievePod :595-615
            addPod(a);
            const bool ok = fits();
            if (panicked) {
                st = FGS_PREEMPT_PANIC;
                break;
            }
            if (!ok) {
                removePod(a);
                victims.push_back(a);
            }
            if (withEQ && overQuota()) {
                if (!ok) { // NodeInfo.RemovePod of a pod that is no longer on the node: "no corresponding pod …"
                    st = FGS_PREEMPT_ERROR;
                    break;
                }
                removePod(a);
                victims.push_back(a);
            }
        }
        if (st == FGS_PREEMPT_CANDIDATE && victims.empty()) st = FGS_PREEMPT_ERROR; // preemption.go:578-580 "expected at least one victim pod"
        finish(st, victims);
    }
    return 0;
}

// pickOneNodeForPreemption, preemption.go:397-505, stage by stage as the reference narrows minNodes1 / minNodes2.
// victims.Pods is in reprieve order (most important first), so Pods[0] is the victim with the highest priority — taken
// here as the maximum, which is the same pod's priority.
extern "C" int64_t fgs_oracle_pick_preemption_node(const fgs_snapshot* snap, const uint8_t* node_status, const uint8_t* victim,
                                                   uint8_t* tie_set) {
    const i64 N = snap->n_nodes;
    const fgs_podset* ps = &snap->pods;
    auto prio = [&](i64 a) { return (i64)(ps->priority ? ps->priority[a] : 0); };
    auto start = [&](i64 a) { return ps->start_time ? ps->start_time[a] : (i64)0; };
    std::memset(tie_set, 0, (size_t)N);
    std::vector<i64> nodes1, nodes2;
    for (i64 n = 0; n < N; ++n) // :404-415 — NumPDBViolations == 0 on every candidate
        if (node_status[n] == FGS_PREEMPT_CANDIDATE) nodes1.push_back(n);
    if (nodes1.empty()) return 0;
    auto victimsOf = [&](i64 n) {
        std::vector<i64> v;
        for (i64 a = snap->pod_off[n]; a < snap->pod_off[n + 1]; ++a)
            if (victim[a]) v.push_back(a);
        return v;
    };
    auto finish = [&](const std::vector<i64>& set) {
        for (i64 n : set) tie_set[n] = 1;
        return (i64)set.size();
    };
    if (nodes1.size() == 1) return finish(nodes1);
    // :422-440 minimum highest-priority victim
    i64 minHighest = INT64_MAX;
    for (i64 n : nodes1) {
        i64 hi = INT64_MIN;
        for (i64 a : victimsOf(n)) hi = std::max(hi, prio(a));
        if (hi < minHighest) {
            minHighest = hi;
            nodes2.clear();
        }
        if (hi == minHighest) nodes2.push_back(n);
    }
    if (nodes2.size() == 1) return finish(nodes2);
    // :444-466 smallest sum of (priority + MaxInt32 + 1)
    i64 minSum = INT64_MAX;
    nodes1.clear();
    for (i64 n : nodes2) {
        i64 sum = 0;
        for (i64 a : victimsOf(n)) sum += prio(a) + ((i64)INT32_MAX + 1);
        if (sum < minSum) {
            minSum = sum;
            nodes1.clear();
        }
        if (sum == minSum) nodes1.push_back(n);
    }
    if (nodes1.size() == 1) return finish(nodes1);
    // :470-486 fewest victims
    i64 minPods = INT64_MAX;
    nodes2.clear();
    for (i64 n : nodes1) {
        const i64 c = (i64)victimsOf(n).size();
        if (c < minPods) {
            minPods = c;
            nodes2.clear();
        }
        if (c == minPods) nodes2.push_back(n);
    }
    if (nodes2.size() == 1) return finish(nodes2);
    // :490-511 latest earliest-start-time of the highest-priority victims (util.GetEarliestPodStartTime, utils.go:56-78);
    // a node is replaced only by a strictly later one, so every node with the latest time may be the one returned
    auto earliestStart = [&](i64 n) {
        const std::vector<i64> v = victimsOf(n);
        i64 earliest = start(v[0]), maxPrio = prio(v[0]);
        for (i64 a : v) {
            if (prio(a) == maxPrio) {
                if (start(a) < earliest) earliest = start(a);
            } else if (prio(a) > maxPrio) {
                maxPrio = prio(a);
                earliest = start(a);
            }
        }
        return earliest;
    };
    i64 latest = INT64_MIN;
    for (i64 n : nodes2) latest = std::max(latest, earliestStart(n));
    nodes1.clear();
    for (i64 n : nodes2)
        if (earliestStart(n) == latest) nodes1.push_back(n);
    return finish(nodes1);
}
