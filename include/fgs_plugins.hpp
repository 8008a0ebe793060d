// fgs_plugins.hpp — the host side above the C ABI, in the shape of the reference's plugins.
//
// The reference is Go; this image has no Go toolchain, so the host layer a maintainer would write in Go over cgo
//  is written here in C++17 over the same C ABI (include/fgs.h), header-only, with the reference's
// names, argument meaning, status codes and messages:
//
//   framework::Status / Code / NodeScore / CycleState   vendor/k8s.io/kubernetes/pkg/scheduler/framework/interface.go:45-120, cycle_state.go
//   FlexGPU            Name, Filter, Score, NormalizeScore, Reserve            pkg/flexgpu/flex_gpu.go:17,41-223
//   Fit                Name, Filter, Score                                      vendor/…/plugins/noderesources/fit.go:42,221-339
//   Allocatable        Name, New (args validation), Score, NormalizeScore       pkg/noderesources/allocatable.go:43-171
//   Coscheduling       Name, PreFilter, Permit                                  pkg/coscheduling/coscheduling.go:58,124-215
//   CapacityScheduling Name, PreFilter, PostFilter (dry-run preemption + pick)    pkg/capacityscheduling/capacity_scheduling.go:103,201-275,320-338,465-644
//   Framework          RunPreFilterPlugins (the new pack + evaluate stage), RunFilterPlugins, RunScorePlugins, selectHost
//
// A "node" is its index in the fgs_snapshot (the reference keys by node name), a "pod" an index into an
// fgs_podset. Nothing here computes a verdict or a score: PreFilter runs the pod through libfgs and the other
// methods answer from the planes it returned. Without a CUDA device construction fails (no CPU fallback).
#pragma once

#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <map>
#include <memory>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

#include "fgs.h"

namespace fgs {
namespace framework {

// interface.go:60-80
enum class Code { Success = 0, Error = 1, Unschedulable = 2, UnschedulableAndUnresolvable = 3, Wait = 4, Skip = 5 };
constexpr int64_t MaxNodeScore = 100; // interface.go:88-95
constexpr int64_t MinNodeScore = 0;

class Status { // interface.go:98-190; a nil *Status is Success
  public:
    Status() = default;
    Status(Code c, std::vector<std::string> reasons = {}) : code_(c), reasons_(std::move(reasons)) {}
    Status(Code c, const std::string& reason) : code_(c), reasons_{reason} {}
    Code GetCode() const { return code_; }
    bool IsSuccess() const { return code_ == Code::Success; }
    bool IsUnschedulable() const { return code_ == Code::Unschedulable || code_ == Code::UnschedulableAndUnresolvable; }
    const std::vector<std::string>& Reasons() const { return reasons_; }
    std::string Message() const { // strings.Join(reasons, ", ")
        std::string m;
        for (size_t i = 0; i < reasons_.size(); ++i) m += (i ? ", " : "") + reasons_[i];
        return m;
    }
    const std::string& FailedPlugin() const { return plugin_; }
    Status& WithFailedPlugin(const std::string& p) {
        plugin_ = p;
        return *this;
    }

  private:
    Code code_ = Code::Success;
    std::vector<std::string> reasons_;
    std::string plugin_;
};

struct NodeScore { // interface.go:45-48 (Name → snapshot index)
    int64_t Node;
    int64_t Score;
};
using NodeScoreList = std::vector<NodeScore>;

} // namespace framework

// What PreFilter leaves in the CycleState for one pod: the planes libfgs returned, one entry per node.
struct PodEvaluation {
    int64_t n_nodes = 0;
    std::vector<uint8_t> flex_code, fit_code, merged_code; // framework.Code per node (fgs_status_plane)
    std::vector<uint16_t> flex_reason, fit_reason;         // fgs_reason_plane
    std::vector<int64_t> flex_raw, flex_norm, fit_score, alloc_raw, alloc_norm, total;
    std::vector<uint32_t> feasible;                        // bitmap
    int32_t feasible_count = 0, selected_node = -1;
    int64_t selected_score = 0;
    uint32_t pod_err = 0;
    bool Feasible(int64_t n) const { return (feasible[(size_t)(n >> 5)] >> (n & 31)) & 1u; }
};

class CycleState { // cycle_state.go: Write / Read keyed by plugin name
  public:
    void Write(const std::string& key, std::shared_ptr<PodEvaluation> v) { data_[key] = std::move(v); }
    std::shared_ptr<PodEvaluation> Read(const std::string& key) const {
        auto it = data_.find(key);
        if (it == data_.end()) throw std::runtime_error(key + " is not found"); // ErrNotFound
        return it->second;
    }

  private:
    std::map<std::string, std::shared_ptr<PodEvaluation>> data_;
};

// Scalar resource names by slot, for the messages ("Insufficient <name>").
struct ResourceNames {
    std::vector<std::string> scalar = {"nvidia.flex.com/gpu", "nvidia.flex.com/memory"}; // flex_gpu.go:18-19
    std::string Scalar(int s) const { return s < (int)scalar.size() ? scalar[(size_t)s] : "scalar-" + std::to_string(s); }
};

// framework.Handle + SnapshotSharedLister: owns the engine and the packer, holds the current snapshot.
class Handle {
  public:
    explicit Handle(int device = 0, int n_scalars = 2) {
        fgs_config c{};
        c.device = device;
        c.n_scalars = n_scalars;
        c.select_seed = 0x5EED;
        if (fgs_create(&c, &e_) != FGS_OK) throw std::runtime_error(std::string("fgs_create: ") + fgs_last_error(nullptr));
        if (fgs_packer_create(n_scalars, &pk_) != FGS_OK) throw std::runtime_error("fgs_packer_create");
    }
    ~Handle() {
        if (pk_) fgs_packer_destroy(pk_);
        if (e_) fgs_destroy(e_);
    }
    Handle(const Handle&) = delete;
    Handle& operator=(const Handle&) = delete;
    fgs_engine* engine() const { return e_; }
    fgs_packer* packer() const { return pk_; }
    const fgs_snapshot* snapshot() const { return snap_; }
    // Cache.UpdateSnapshot's role (internal/cache/cache.go:197-276): pack + upload — incrementally, by
    // NodeInfo.Generation, when the snapshot carries the column (fgs_sync_nodes); returns the rows that went up
    int64_t UpdateSnapshot(const fgs_snapshot* snap) {
        int64_t repacked = snap->n_nodes;
        if (snap->generation) {
            Check(fgs_sync_nodes(e_, pk_, snap, 0, snap->n_nodes, &repacked));
        } else {
            const fgs_node_table* t = nullptr;
            if (fgs_packer_nodes(pk_, snap, 0, snap->n_nodes, &t) != FGS_OK) throw std::runtime_error(std::string("pack: ") + fgs_packer_error(pk_));
            Check(fgs_upload_nodes(e_, t));
        }
        snap_ = snap;
        return repacked;
    }
    void Check(int rc) const {
        if (rc != FGS_OK) throw std::runtime_error(fgs_last_error(e_));
    }
    ResourceNames names;

  private:
    fgs_engine* e_ = nullptr;
    fgs_packer* pk_ = nullptr;
    const fgs_snapshot* snap_ = nullptr;
};

// ---- plugin args (KubeSchedulerConfiguration) -----------------------------------------------------------------------
struct ResourceSpec { // apis/config/types.go: ResourceSpec {Name, Weight}
    std::string Name;
    int64_t Weight;
};
struct NodeResourcesAllocatableArgs { // pkg/apis/config/types.go:50-62
    std::vector<ResourceSpec> Resources;
    std::string Mode; // "Least" | "Most" | "" (default Least)
};
struct NodeResourcesFitArgs { // vendor/…/apis/config/types.go: ScoringStrategy {Type, Resources}
    std::string ScoringStrategyType = "LeastAllocated";
    std::vector<ResourceSpec> Resources = {{"cpu", 1}, {"memory", 1}}; // v1beta2/defaults.go:241-257
};
struct ScorePluginRef {
    std::string Name;
    int64_t Weight = 1; // runtime/framework.go:405-408: 0 means 1
};
struct Profile { // the profile's plugin lists, in the order the configuration lists them
    std::vector<std::string> Filter = {"NodeResourcesFit", "FlexGPU"};
    std::vector<ScorePluginRef> Score = {{"FlexGPU", 1}, {"NodeResourcesFit", 1}};
    NodeResourcesFitArgs FitArgs;
    NodeResourcesAllocatableArgs AllocatableArgs;
};

namespace detail {
inline int32_t resource_slot(const ResourceNames& names, const std::string& r) {
    if (r == "cpu") return FGS_RES_CPU;
    if (r == "memory") return FGS_RES_MEM;
    if (r == "ephemeral-storage") return FGS_RES_EPH;
    for (size_t s = 0; s < names.scalar.size(); ++s)
        if (names.scalar[s] == r) return FGS_RES_SCALAR0 + (int32_t)s;
    return FGS_RES_UNKNOWN; // a name the scorer knows nothing about contributes (0, 0) (resource_allocation.go:93-101)
}
} // namespace detail

// NewAllocatable's validation (pkg/noderesources/allocatable.go:51-58,79-118): returns the error text, "" if fine
inline std::string ValidateAllocatableArgs(const NodeResourcesAllocatableArgs& a) {
    if (!a.Mode.empty() && a.Mode != "Least" && a.Mode != "Most") return "invalid mode, got " + a.Mode;
    for (const auto& r : a.Resources)
        if (r.Weight <= 0) return "resource Weight of " + r.Name + " should be a positive value, got " + std::to_string(r.Weight);
    return "";
}

// ---- the framework: runs the plugin set for one pod ------------------------------------------------------------------
class Framework {
  public:
    static constexpr const char* kStateKey = "PreFilterFGS"; // CycleState key of the PodEvaluation

    Framework(Handle& h, Profile p) : h_(h), profile_(std::move(p)) {
        const std::string err = ValidateAllocatableArgs(profile_.AllocatableArgs);
        if (!err.empty()) throw std::invalid_argument(err);
        cfg_ = fgs_plugin_cfg{};
        int k = 0;
        for (const auto& f : profile_.Filter) {
            const uint32_t id = f == "FlexGPU" ? FGS_PL_FLEXGPU : (f == "NodeResourcesFit" ? FGS_PL_FIT : 0u);
            if (!id) throw std::invalid_argument("unknown filter plugin " + f);
            cfg_.filter_mask |= id;
            if (k < 2) cfg_.filter_order[k++] = (int32_t)id;
        }
        cfg_.weight_flex = cfg_.weight_fit = cfg_.weight_alloc = 1;
        for (const auto& s : profile_.Score) {
            if (s.Name == "FlexGPU") cfg_.score_mask |= FGS_PL_FLEXGPU, cfg_.weight_flex = s.Weight;
            else if (s.Name == "NodeResourcesFit") cfg_.score_mask |= FGS_PL_FIT, cfg_.weight_fit = s.Weight;
            else if (s.Name == "NodeResourcesAllocatable") cfg_.score_mask |= FGS_PL_ALLOCATABLE, cfg_.weight_alloc = s.Weight;
            else throw std::invalid_argument("unknown score plugin " + s.Name);
        }
        const auto& fa = profile_.FitArgs;
        if (fa.ScoringStrategyType != "LeastAllocated" && fa.ScoringStrategyType != "MostAllocated")
            throw std::invalid_argument("scoring strategy " + fa.ScoringStrategyType + " is not supported");
        cfg_.fit_strategy = fa.ScoringStrategyType == "MostAllocated" ? 1 : 0;
        cfg_.fit_n_res = (int32_t)fa.Resources.size();
        for (size_t i = 0; i < fa.Resources.size() && i < FGS_MAX_SCORE_RES; ++i) {
            cfg_.fit_res[i] = detail::resource_slot(h_.names, fa.Resources[i].Name);
            cfg_.fit_res_w[i] = fa.Resources[i].Weight;
        }
        auto res = profile_.AllocatableArgs.Resources;
        if (res.empty()) res = {{"cpu", 1 << 20}, {"memory", 1}}; // defaultResourcesToWeightMap, resource_allocation.go:38
        cfg_.alloc_mode = profile_.AllocatableArgs.Mode == "Most" ? 1 : 0;
        cfg_.alloc_n_res = (int32_t)res.size();
        for (size_t i = 0; i < res.size() && i < FGS_MAX_SCORE_RES; ++i) {
            cfg_.alloc_res[i] = detail::resource_slot(h_.names, res[i].Name);
            cfg_.alloc_res_w[i] = res[i].Weight;
        }
        if (fgs_set_plugins(h_.engine(), &cfg_) != FGS_OK) throw std::invalid_argument(fgs_last_error(h_.engine()));
    }

    Handle& handle() const { return h_; }
    const fgs_plugin_cfg& config() const { return cfg_; }
    bool Scores(uint32_t plugin) const { return cfg_.score_mask & plugin; }
    bool Filters(uint32_t plugin) const { return cfg_.filter_mask & plugin; }
    int64_t Weight(uint32_t plugin) const {
        const int64_t w = plugin == FGS_PL_FLEXGPU ? cfg_.weight_flex : (plugin == FGS_PL_FIT ? cfg_.weight_fit : cfg_.weight_alloc);
        return w ? w : 1;
    }

    // The new PreFilter stage of the drop-in: flatten the pod, run it through libfgs against
    // the uploaded snapshot, keep every plane in the CycleState. All later Filter / Score calls of the cycle
    // answer from there.
    framework::Status RunPreFilterPlugins(CycleState& state, const fgs_podset* pods, int64_t i) {
        using framework::Code;
        using framework::Status;
        const fgs_pod_table* t = nullptr;
        if (fgs_packer_pods(h_.packer(), pods, i, i + 1, &t) != FGS_OK) return Status(Code::Error, fgs_packer_error(h_.packer()));
        auto ev = std::make_shared<PodEvaluation>();
        const int64_t N = h_.snapshot()->n_nodes, n_pad = (std::max<int64_t>(N, 1) + 1023) / 1024 * 1024;
        ev->n_nodes = N;
        ev->feasible.assign((size_t)(n_pad / 32), 0);
        std::vector<int64_t> plane((size_t)n_pad);
        fgs_results r{};
        r.feasible_count = &ev->feasible_count;
        r.selected_node = &ev->selected_node;
        r.selected_score = &ev->selected_score;
        r.pod_err = &ev->pod_err;
        r.feasible_bitmap = ev->feasible.data();
        r.scores = plane.data();
        fgs_engine* e = h_.engine();
        if (fgs_evaluate(e, t, nullptr, &r) != FGS_OK) return Status(Code::Error, fgs_last_error(e));
        ev->total.assign(plane.begin(), plane.begin() + N);
        auto codes = [&](uint32_t plugin, std::vector<uint8_t>& out) {
            out.assign((size_t)std::max<int64_t>(N, 1), 0);
            h_.Check(fgs_status_plane(e, plugin, out.data()));
        };
        auto reasons = [&](uint32_t plugin, std::vector<uint16_t>& out) {
            out.assign((size_t)std::max<int64_t>(N, 1), 0);
            h_.Check(fgs_reason_plane(e, plugin, out.data()));
        };
        codes(FGS_PL_FLEXGPU, ev->flex_code);
        codes(FGS_PL_FIT, ev->fit_code);
        codes(0, ev->merged_code);
        reasons(FGS_PL_FLEXGPU, ev->flex_reason);
        reasons(FGS_PL_FIT, ev->fit_reason);
        // one plugin's Score / NormalizeScore output at a time (fgs_eval_opts.emit_mask / emit_raw)
        auto emit = [&](uint32_t plugin, bool raw, std::vector<int64_t>& out) {
            fgs_eval_opts o{};
            o.emit_mask = plugin;
            o.emit_raw = raw ? 1 : 0;
            h_.Check(fgs_pass_emit(e, &o));
            fgs_results rr{};
            rr.scores = plane.data();
            h_.Check(fgs_fetch_results(e, &rr));
            out.assign(plane.begin(), plane.begin() + N);
            if (!raw) // the emitted value is normalised × weight: NormalizeScore's own output is the quotient
                for (auto& v : out) v /= Weight(plugin);
        };
        if (Scores(FGS_PL_FLEXGPU)) emit(FGS_PL_FLEXGPU, true, ev->flex_raw), emit(FGS_PL_FLEXGPU, false, ev->flex_norm);
        if (Scores(FGS_PL_FIT)) emit(FGS_PL_FIT, true, ev->fit_score);
        if (Scores(FGS_PL_ALLOCATABLE)) emit(FGS_PL_ALLOCATABLE, true, ev->alloc_raw), emit(FGS_PL_ALLOCATABLE, false, ev->alloc_norm);
        state.Write(kStateKey, ev);
        if (ev->pod_err & FGS_PODERR_REF_PANIC) return Status(Code::Error, "the reference panics in NewGPUNode on a node this pod reaches");
        return Status();
    }

    // RunScorePlugins (runtime/framework.go:861-933) as prioritizeNodes calls it (generic_scheduler.go:420-426): over the
    // nodes that passed EVERY filter plugin of the profile — the engine's own verdicts AND those of the plugins outside
    // this path. `nodes` = that list (snapshot indices). Every NormalizeScore then runs over it: the totals, the
    // selected node and its score in the CycleState are replaced by those of the masked list (fgs_score_masked).
    // Call after RunPreFilterPlugins of the same pod; returns Error when a normalised score leaves [0, 100].
    framework::Status RunScorePlugins(CycleState& state, const std::vector<int64_t>& nodes) {
        using framework::Code;
        using framework::Status;
        auto ev = state.Read(kStateKey);
        const int64_t N = ev->n_nodes, n_pad = (std::max<int64_t>(N, 1) + 1023) / 1024 * 1024;
        std::vector<uint32_t> mask((size_t)(n_pad / 32), 0u);
        for (int64_t n : nodes)
            if (n >= 0 && n < N) mask[(size_t)(n >> 5)] |= 1u << (n & 31);
        std::vector<int64_t> plane((size_t)n_pad);
        fgs_results r{};
        r.feasible_count = &ev->feasible_count;
        r.selected_node = &ev->selected_node;
        r.selected_score = &ev->selected_score;
        r.pod_err = &ev->pod_err;
        r.feasible_bitmap = ev->feasible.data();
        r.scores = plane.data();
        fgs_engine* e = h_.engine();
        if (fgs_score_masked(e, mask.data(), 0, nullptr, &r) != FGS_OK) return Status(Code::Error, fgs_last_error(e));
        ev->total.assign(plane.begin(), plane.begin() + N);
        if (ev->pod_err & FGS_PODERR_SCORE_RANGE) return Status(Code::Error, "a plugin returns an invalid score, it should in the range of [0, 100] after normalizing");
        if (ev->pod_err & FGS_PODERR_REF_PANIC) return Status(Code::Error, "the reference panics in NewGPUNode on a node this pod reaches");
        return Status();
    }

    // selectHost (generic_scheduler.go:146-167): the node libfgs picked out of the max-score tie set
    int64_t SelectHost(const CycleState& state) const { return state.Read(kStateKey)->selected_node; }

    // framework.FitError.Error() (framework/types.go:221-250) for the pod RunPreFilterPlugins evaluated last, when no node
    // took it: "0/<N> nodes are available: <count> <reason>, … ." — the histogram over the NodeToStatusMap comes from the
    // device (fgs_fit_error_histogram), the "<count> <reason>" strings are sorted as sort.Strings sorts them
    std::string FitError() const {
        int32_t c[FGS_DIAG_SLOTS] = {0};
        h_.Check(fgs_fit_error_histogram(h_.engine(), c));
        std::vector<std::string> parts;
        auto add = [&](int slot, const std::string& reason) {
            if (c[slot]) parts.push_back(std::to_string(c[slot]) + " " + reason);
        };
        add(FGS_DIAG_FLEX0 + FGS_FLEX_REASON_CONFLICT - 1, "pod conflict resources");                                   // flex_gpu.go:60
        add(FGS_DIAG_FLEX0 + FGS_FLEX_REASON_UNKNOWN_RESOURCE - 1, "unknown resource type");                            // :67,73
        add(FGS_DIAG_FLEX0 + FGS_FLEX_REASON_INSUFFICIENT_GPU - 1, "insufficient resource nvidia.flex.com/gpu");        // :92
        add(FGS_DIAG_FLEX0 + FGS_FLEX_REASON_INSUFFICIENT_MEM - 1, "insufficient resource nvidia.flex.com/memory");     // :97
        add(FGS_DIAG_FLEX0 + FGS_FLEX_REASON_NO_FIT_INDEX - 1, "no fit indexes resource nvidia.flex.com/memory");       // :114
        add(FGS_DIAG_FIT0 + 0, "Too many pods");                                                                        // fit.go:259-268
        add(FGS_DIAG_FIT0 + 1, "Insufficient cpu");
        add(FGS_DIAG_FIT0 + 2, "Insufficient memory");
        add(FGS_DIAG_FIT0 + 3, "Insufficient ephemeral-storage");
        for (int s = 0; s < FGS_MAX_SCALARS; ++s) add(FGS_DIAG_FIT0 + 4 + s, "Insufficient " + h_.names.Scalar(s));
        std::sort(parts.begin(), parts.end());
        std::string msg = "0/" + std::to_string(h_.snapshot() ? h_.snapshot()->n_nodes : 0) + " nodes are available: ";
        for (size_t i = 0; i < parts.size(); ++i) msg += (i ? ", " : "") + parts[i];
        return msg + ".";
    }

  private:
    Handle& h_;
    Profile profile_;
    fgs_plugin_cfg cfg_;
};

// ---- FlexGPU — pkg/flexgpu/flex_gpu.go ------------------------------------------------------------------------------
class FlexGPU {
  public:
    static constexpr const char* Name = "FlexGPU";                         // :17
    static constexpr const char* GPUResourceName = "nvidia.flex.com/gpu";  // :18
    static constexpr const char* MemResourceName = "nvidia.flex.com/memory";
    static constexpr const char* GPUIndexAnnotationKey = "nvidia.flex.com/index";
    explicit FlexGPU(Framework& f) : f_(f) {}

    // Filter(ctx, state, pod, nodeInfo) — :41-119
    framework::Status Filter(const CycleState& state, int64_t node) const {
        using framework::Code;
        using framework::Status;
        const auto ev = state.Read(Framework::kStateKey);
        switch (ev->flex_reason[(size_t)node]) {
            case FGS_FLEX_REASON_NONE: return Status();
            case FGS_FLEX_REASON_CONFLICT: return Status(Code::UnschedulableAndUnresolvable, "pod conflict resources");            // :60
            case FGS_FLEX_REASON_UNKNOWN_RESOURCE: return Status(Code::UnschedulableAndUnresolvable, "unknown resource type");      // :67,73
            case FGS_FLEX_REASON_INSUFFICIENT_GPU: return Status(Code::Unschedulable, std::string("insufficient resource ") + GPUResourceName); // :92
            case FGS_FLEX_REASON_INSUFFICIENT_MEM: return Status(Code::Unschedulable, std::string("insufficient resource ") + MemResourceName); // :97
            case FGS_FLEX_REASON_NO_FIT_INDEX: return Status(Code::Unschedulable, std::string("no fit indexes resource ") + MemResourceName);   // :114
            default: return Status(Code::Error, "NewGPUNode panics on this node (gpu_node.go:40-58,109,115)");
        }
    }
    // Score(ctx, state, pod, nodeName) — :142-166
    std::pair<int64_t, framework::Status> Score(const CycleState& state, int64_t node) const {
        return {state.Read(Framework::kStateKey)->flex_raw[(size_t)node], framework::Status()};
    }
    // NormalizeScore(ctx, state, pod, scores) — :172-176, helper.DefaultNormalizeScore(MaxNodeScore, reverse=true):
    // the normalisation runs over the feasible set inside libfgs; this writes its result into the list
    framework::Status NormalizeScore(const CycleState& state, framework::NodeScoreList& scores) const {
        const auto ev = state.Read(Framework::kStateKey);
        for (auto& s : scores) s.Score = ev->flex_norm[(size_t)s.Node];
        return framework::Status();
    }
    // Reserve(ctx, state, pod, nodeName) — :178-223: on success *index is what the reference writes into
    // pod.Annotations[GPUIndexAnnotationKey] (-1: the pod has no flex limit, nothing is written)
    framework::Status Reserve(const fgs_podset* pods, int64_t i, int64_t node, int* index) const {
        using framework::Code;
        using framework::Status;
        Handle& h = f_.handle();
        const fgs_pod_table* t = nullptr;
        if (fgs_packer_pods(h.packer(), pods, i, i + 1, &t) != FGS_OK) return Status(Code::Error, fgs_packer_error(h.packer()));
        const int32_t n = (int32_t)node;
        int32_t dev = FGS_RESERVE_SKIPPED;
        if (fgs_reserve(h.engine(), t, &n, &dev) != FGS_OK) return Status(Code::Error, fgs_last_error(h.engine()));
        *index = -1;
        if (dev >= 0) {
            *index = dev;
            return Status();
        }
        const bool gpu = t->flags[0] & FGS_PF_GPU_LIMIT;
        switch (dev) {
            case FGS_RESERVE_NONE: return Status();                                                                  // :187-189
            case FGS_RESERVE_CONFLICT: return Status(Code::UnschedulableAndUnresolvable, "pod conflict resources"); // :191-193
            case FGS_RESERVE_FAIL:
                return Status(Code::Unschedulable, std::string("allocate index fail ") + (gpu ? GPUResourceName : MemResourceName)); // :200,211
            default: return Status(Code::Error, "NewGPUNode panics on this node");
        }
    }

  private:
    Framework& f_;
};

// ---- NodeResourcesFit — vendor/…/plugins/noderesources/fit.go -----------------------------------------------------------
class Fit {
  public:
    static constexpr const char* Name = "NodeResourcesFit"; // :42
    explicit Fit(Framework& f) : f_(f) {}
    // Filter — :221-238: Unschedulable with one reason per insufficient resource (fitsRequest :256-329)
    framework::Status Filter(const CycleState& state, int64_t node) const {
        const auto ev = state.Read(Framework::kStateKey);
        const uint16_t m = ev->fit_reason[(size_t)node];
        if (!m) return framework::Status();
        std::vector<std::string> why;
        if (m & FGS_FIT_REASON_PODS) why.push_back("Too many pods");
        if (m & FGS_FIT_REASON_CPU) why.push_back("Insufficient cpu");
        if (m & FGS_FIT_REASON_MEM) why.push_back("Insufficient memory");
        if (m & FGS_FIT_REASON_EPH) why.push_back("Insufficient ephemeral-storage");
        for (int s = 0; s < FGS_MAX_SCALARS; ++s)
            if (m & (FGS_FIT_REASON_SCALAR0 << s)) why.push_back("Insufficient " + f_.handle().names.Scalar(s));
        return framework::Status(framework::Code::Unschedulable, why);
    }
    // Score — :332-339 (Least/MostAllocated; ScoreExtensions() == nil, :87-89)
    std::pair<int64_t, framework::Status> Score(const CycleState& state, int64_t node) const {
        return {state.Read(Framework::kStateKey)->fit_score[(size_t)node], framework::Status()};
    }

  private:
    Framework& f_;
};

// ---- NodeResourcesAllocatable — pkg/noderesources/allocatable.go ------------------------------------------------------------
class Allocatable {
  public:
    static constexpr const char* Name = "NodeResourcesAllocatable"; // :43
    explicit Allocatable(Framework& f) : f_(f) {}
    std::pair<int64_t, framework::Status> Score(const CycleState& state, int64_t node) const { // :61-72
        return {state.Read(Framework::kStateKey)->alloc_raw[(size_t)node], framework::Status()};
    }
    framework::Status NormalizeScore(const CycleState& state, framework::NodeScoreList& scores) const { // :146-171
        const auto ev = state.Read(Framework::kStateKey);
        for (auto& s : scores) s.Score = ev->alloc_norm[(size_t)s.Node];
        return framework::Status();
    }

  private:
    Framework& f_;
};

// RunFilterPlugins (runtime/framework.go:667-693): profile order, first failure wins
inline framework::Status RunFilterPlugins(Framework& f, const CycleState& state, int64_t node) {
    for (int k = 0; k < 2 && f.config().filter_order[k]; ++k) {
        framework::Status s = f.config().filter_order[k] == (int32_t)FGS_PL_FLEXGPU ? FlexGPU(f).Filter(state, node) : Fit(f).Filter(state, node);
        if (!s.IsSuccess()) return s.WithFailedPlugin(f.config().filter_order[k] == (int32_t)FGS_PL_FLEXGPU ? FlexGPU::Name : Fit::Name);
    }
    return framework::Status();
}

// ---- Coscheduling — pkg/coscheduling/coscheduling.go, core/core.go ------------------------------------------------------------
class Coscheduling {
  public:
    static constexpr const char* Name = "Coscheduling"; // coscheduling.go:58
    explicit Coscheduling(Handle& h) : h_(h) {}
    // PreFilter + Permit of every pending pod in one call (the PodGroupManager walks the same snapshot for each)
    void Evaluate(const fgs_podset* pending, const fgs_podgroups* pgs) {
        const size_t P = (size_t)std::max<int64_t>(pending->n_pods, 1);
        pre_.assign(P, 0);
        permit_.assign(P, 0);
        post_.assign(P, 0);
        h_.Check(fgs_gang_check(h_.engine(), h_.snapshot(), pending, pgs, pre_.data(), permit_.data(), nullptr));
        h_.Check(fgs_gang_postfilter(h_.engine(), h_.snapshot(), pending, pgs, post_.data(), nullptr));
    }
    // PreFilter — coscheduling.go:124-137 → PodGroupManager.PreFilter core.go:149-196; any failure is
    // UnschedulableAndUnresolvable "to avoid any preemption attempts"
    framework::Status PreFilter(int64_t pod) const {
        using framework::Code;
        using framework::Status;
        switch (pre_[(size_t)pod]) {
            case 0: return Status();
            case 1: return Status(Code::UnschedulableAndUnresolvable, "pod with pgName last failed, deny");                    // core.go:156
            case 2: return Status(Code::UnschedulableAndUnresolvable, "pre-filter pod cannot find enough sibling pods");        // :165
            default: return Status(Code::UnschedulableAndUnresolvable, "resource gap");                                          // :341
        }
    }
    // Permit — coscheduling.go:186-215
    framework::Status Permit(int64_t pod) const {
        using framework::Code;
        using framework::Status;
        switch (permit_[(size_t)pod]) {
            case 0: return Status(Code::Success);
            case 1: return Status(Code::Wait);
            case 2: return Status(Code::Success); // PodGroupNotSpecified
            default: return Status(Code::Unschedulable, "PodGroup not found");
        }
    }

    // PostFilter — coscheduling.go:140-176: every outcome is Unschedulable; *reject_group tells the caller to reject the
    // group's waiting pods, AddDeniedPodGroup and DeletePermittedPodGroup (:164-172)
    framework::Status PostFilter(int64_t pod, const std::string& pg_full_name, const std::string& pod_name, bool* reject_group) const {
        using framework::Code;
        using framework::Status;
        if (reject_group) *reject_group = false;
        switch (post_[(size_t)pod]) {
            case FGS_GANG_POSTFILTER_NO_GROUP: return Status(Code::Unschedulable, "can not find pod group");          // :143-147
            case FGS_GANG_POSTFILTER_ENOUGH_ASSIGNED: return Status(Code::Unschedulable);                              // :151-155
            case FGS_GANG_POSTFILTER_SMALL_GAP: return Status(Code::Unschedulable);                                    // :159-163
            default:
                if (reject_group) *reject_group = true;
                return Status(Code::Unschedulable,
                              "PodGroup " + pg_full_name + " gets rejected due to Pod " + pod_name + " is unschedulable even after PostFilter"); // :173-175
        }
    }

  private:
    Handle& h_;
    std::vector<uint8_t> pre_, permit_, post_;
};

// ---- CapacityScheduling — pkg/capacityscheduling/capacity_scheduling.go ---------------------------------------------------------
class CapacityScheduling {
  public:
    static constexpr const char* Name = "CapacityScheduling"; // :103
    explicit CapacityScheduling(Handle& h) : h_(h) {}
    // PreFilter — :201-275 for every pending pod against the ElasticQuotaInfos
    void Evaluate(const fgs_podset* pending, const fgs_quotas* q) {
        const fgs_pod_table* t = nullptr;
        if (fgs_packer_pods(h_.packer(), pending, 0, pending->n_pods, &t) != FGS_OK) throw std::runtime_error(fgs_packer_error(h_.packer()));
        code_.assign((size_t)std::max<int64_t>(pending->n_pods, 1), 0);
        h_.Check(fgs_quota_check(h_.engine(), t, q, nullptr, nullptr, nullptr, code_.data()));
    }
    framework::Status PreFilter(int64_t pod, const std::string& ns, const std::string& name) const {
        using framework::Code;
        using framework::Status;
        switch (code_[(size_t)pod]) {
            case 0: return Status();
            case 1: return Status(Code::Unschedulable, "Pod " + ns + "/" + name + " is rejected in PreFilter because ElasticQuota " + ns + " is more than Max"); // :267
            default: return Status(Code::Unschedulable, "Pod " + ns + "/" + name + " is rejected in PreFilter because total ElasticQuota used is more than min"); // :271
        }
    }


    // PostFilter — :320-338 → preemption.Evaluator.Preempt (vendor/…/framework/preemption/preemption.go:138-195) up to and
    // including SelectCandidate: the nodes where preemption might help (nodesWherePreemptionMightHelp, :335-350: not
    // UnschedulableAndUnresolvable in the filter pass), SelectVictimsOnNode on every one of them in ONE call, and
    // pickOneNodeForPreemption. prepareCandidate (evictions, clearing nominations: API-server work) stays with the caller,
    // who gets the victims. `eval` is the pod's PreFilter result (nullptr: every node might help). No PodDisruptionBudgets,
    // no extenders; every potential node is dry-run (the reference stops after GetOffsetAndNumCandidates' quota).
    struct PostFilterResult {
        int64_t NominatedNode = -1;       // framework.PostFilterResult.NominatedNodeName as a snapshot index; -1 = ""
        std::vector<int64_t> Victims;     // rows of snapshot->pods to evict on that node, most important first is NOT implied
        int64_t Candidates = 0;           // nodes on which preemption would work
    };
    std::pair<PostFilterResult, framework::Status> PostFilter(const fgs_podset* pending, int64_t pod, const fgs_quotas* q,
                                                             const PodEvaluation* eval = nullptr, bool tracks_pods = true) const {
        using framework::Code;
        using framework::Status;
        const fgs_snapshot* snap = h_.snapshot();
        if (!snap || snap->n_nodes == 0) return {PostFilterResult{}, Status(Code::Error, "no nodes available")}; // :199-201
        const int64_t N = snap->n_nodes, T = snap->pods.n_pods;
        std::vector<uint8_t> st((size_t)N), vic((size_t)std::max<int64_t>(T, 1));
        std::vector<int32_t> cnt((size_t)N);
        int32_t picked = -1;
        h_.Check(fgs_select_victims(h_.engine(), snap, pending, pod, q, tracks_pods ? 1 : 0, st.data(), vic.data(), cnt.data(), &picked));
        auto mightHelp = [&](int64_t n) { return !eval || eval->merged_code[(size_t)n] != (uint8_t)Code::UnschedulableAndUnresolvable; };
        PostFilterResult r;
        bool dropped_candidate = false;
        for (int64_t n = 0; n < N; ++n)
            if (st[(size_t)n] == FGS_PREEMPT_CANDIDATE) {
                if (mightHelp(n)) r.Candidates++;
                else dropped_candidate = true;
            }
        if (r.Candidates == 0) // :162-173 — the FitError text needs the per-node reasons; the code and the cleared nomination are the contract
            return {r, Status(Code::Unschedulable, "preemption: 0/" + std::to_string(N) + " nodes are available")};
        if (dropped_candidate && !mightHelp(picked)) picked = (int32_t)PickOnHost(snap, st, vic, mightHelp);
        r.NominatedNode = picked;
        for (int64_t a = snap->pod_off[picked]; a < snap->pod_off[picked + 1]; ++a)
            if (vic[(size_t)a]) r.Victims.push_back(a);
        return {r, Status()};
    }

  private:
    // pickOneNodeForPreemption (:386-505) among the candidates `keep` lets through — only when the device's pick over
    // all candidates landed on a node the filter pass had ruled out for good
    template <typename Keep>
    static int64_t PickOnHost(const fgs_snapshot* snap, const std::vector<uint8_t>& st, const std::vector<uint8_t>& vic, Keep keep) {
        const fgs_podset& ps = snap->pods;
        int64_t best = -1, bhi = 0, bsum = 0, bcnt = 0, bstart = 0;
        for (int64_t n = 0; n < snap->n_nodes; ++n) {
            if (st[(size_t)n] != FGS_PREEMPT_CANDIDATE || !keep(n)) continue;
            int64_t hi = INT64_MIN, sum = 0, c = 0, start = 0;
            for (int64_t a = snap->pod_off[n]; a < snap->pod_off[n + 1]; ++a) {
                if (!vic[(size_t)a]) continue;
                const int64_t pr = ps.priority ? ps.priority[a] : 0, t0 = ps.start_time ? ps.start_time[a] : 0;
                sum += pr + 2147483648ll;
                c++;
                if (pr > hi) hi = pr, start = t0;
                else if (pr == hi && t0 < start) start = t0;
            }
            const bool better = best < 0 || (hi != bhi ? hi < bhi : sum != bsum ? sum < bsum : c != bcnt ? c < bcnt : start > bstart);
            if (better) best = n, bhi = hi, bsum = sum, bcnt = c, bstart = start;
        }
        return best;
    }
    Handle& h_;
    std::vector<uint8_t> code_;
};

} // namespace fgs
