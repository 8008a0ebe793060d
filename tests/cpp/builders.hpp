// builders.hpp — test fixtures in C++: node / pod descriptions → the flat object snapshot of include/fgs_types.h.
// They play the role of the reference's st.MakeNode().Capacity(…) / st.MakePod().Req(…) / NewFakeSharedLister
// (test/util/fake.go:38-98); NodeInfo aggregates are accumulated the way NodeInfo.AddPod does
// (framework/types.go:583-609,715-745) with the product's podmath.h.
#pragma once
#include <cstdint>
#include <map>
#include <memory>
#include <string>
#include <vector>

#include "fgs_types.h"
#include "podmath.h"

namespace fgs {
namespace testing {

// resource slots: 0 cpu (milli), 1 memory, 2 ephemeral-storage, 3+s scalar s
struct Ctr {
    std::map<int, int64_t> requests;
    bool has_gpu_limit = false, has_mem_limit = false;
    int64_t gpu_limit = 0, mem_limit = 0;
};
inline Ctr MakeCtr(std::map<int, int64_t> req) {
    Ctr c;
    c.requests = std::move(req);
    return c;
}
inline Ctr GpuLimit(Ctr c, int64_t v) {
    c.has_gpu_limit = true;
    c.gpu_limit = v;
    return c;
}
inline Ctr MemLimit(Ctr c, int64_t v) {
    c.has_mem_limit = true;
    c.mem_limit = v;
    return c;
}
struct Pod {
    std::vector<Ctr> containers, init_containers;
    int32_t ns = 0, pg = -1, priority = 0, index = FGS_ANNOT_NONE;
    bool node_name = false;
    int64_t start = 0; // pod.Status.StartTime
};
inline Pod MakePod(std::vector<Ctr> ctrs, int32_t ns = 0, int32_t pg = -1, int32_t priority = 0) {
    Pod p;
    p.containers = std::move(ctrs);
    p.ns = ns;
    p.pg = pg;
    p.priority = priority;
    return p;
}
struct Node {
    int64_t cpu = 0, memory = 0, ephemeral = 0, pods = 110;
    std::map<int, int64_t> scalars; // scalar slot → allocatable
    std::vector<Pod> on;            // NodeInfo.Pods
    bool valid = true;
};

class Podset {
  public:
    Podset(const std::vector<Pod>& pods, int S) : S_(S) {
        const int W = 3 + S;
        ctr_off_.push_back(0);
        for (const Pod& p : pods) {
            auto add = [&](const Ctr& c) {
                std::vector<int64_t> row((size_t)W, 0);
                uint32_t pres = 0;
                for (const auto& kv : c.requests) row[(size_t)kv.first] = kv.second, pres |= 1u << kv.first;
                c_req_.insert(c_req_.end(), row.begin(), row.end());
                c_pres_.push_back(pres);
                c_lim_.push_back(c.gpu_limit);
                c_lim_.push_back(c.mem_limit);
                c_lpres_.push_back((c.has_gpu_limit ? 1u : 0u) | (c.has_mem_limit ? 2u : 0u));
            };
            for (const Ctr& c : p.containers) add(c);
            for (const Ctr& c : p.init_containers) add(c);
            ctr_off_.push_back((int64_t)c_pres_.size());
            n_init_.push_back((int32_t)p.init_containers.size());
            ns_.push_back(p.ns);
            pg_.push_back(p.pg);
            prio_.push_back(p.priority);
            named_.push_back(p.node_name ? 1 : 0);
            annot_.push_back(p.index);
            start_.push_back(p.start);
            ovp_.push_back(0);
        }
        ov_.assign(std::max<size_t>(pods.size(), 1) * (size_t)W, 0);
        auto pad = [](auto& v) {
            if (v.empty()) v.push_back({});
        };
        pad(ns_), pad(pg_), pad(prio_), pad(named_), pad(annot_), pad(n_init_), pad(ovp_), pad(c_pres_), pad(c_lpres_);
        if (c_req_.empty()) c_req_.assign((size_t)W, 0);
        if (c_lim_.empty()) c_lim_.assign(2, 0);
        if (start_.empty()) start_.push_back(0);
        ps_ = fgs_podset{};
        ps_.n_pods = (int64_t)pods.size();
        ps_.n_scalars = S;
        ps_.ns = ns_.data(), ps_.pg = pg_.data(), ps_.priority = prio_.data();
        ps_.has_node_name = named_.data(), ps_.index_annot = annot_.data();
        ps_.start_time = start_.data();
        ps_.ctr_off = ctr_off_.data(), ps_.n_init = n_init_.data();
        ps_.overhead_present = ovp_.data(), ps_.overhead = ov_.data();
        ps_.n_ctrs = pods.empty() ? 0 : ctr_off_.back();
        ps_.c_req = c_req_.data(), ps_.c_req_present = c_pres_.data();
        ps_.c_lim_flex = c_lim_.data(), ps_.c_lim_present = c_lpres_.data();
    }
    const fgs_podset* get() const { return &ps_; }
    const fgs_podset& ref() const { return ps_; }

  private:
    int S_;
    std::vector<int32_t> ns_, pg_, prio_, annot_, n_init_;
    std::vector<uint8_t> named_;
    std::vector<uint32_t> ovp_, c_pres_, c_lpres_;
    std::vector<int64_t> ctr_off_, ov_, c_req_, c_lim_, start_;
    fgs_podset ps_;
};

class Snapshot {
  public:
    Snapshot(const std::vector<Node>& nodes, int S = 2) {
        const int AW = 4 + S, RW = 3 + S;
        std::vector<Pod> all;
        pod_off_.push_back(0);
        for (const Node& n : nodes) {
            for (const Pod& p : n.on) all.push_back(p);
            pod_off_.push_back((int64_t)all.size());
        }
        pods_ = std::make_unique<Podset>(all, S);
        const size_t N = std::max<size_t>(nodes.size(), 1);
        alloc_.assign(N * (size_t)AW, 0);
        req_.assign(N * (size_t)RW, 0);
        nz_.assign(N * 2, 0);
        present_.assign(N, 0);
        valid_.assign(N, 1);
        for (size_t i = 0; i < nodes.size(); ++i) {
            const Node& n = nodes[i];
            valid_[i] = n.valid ? 1 : 0;
            int64_t* a = &alloc_[i * (size_t)AW];
            a[0] = n.cpu, a[1] = n.memory, a[2] = n.ephemeral, a[3] = n.pods;
            for (const auto& kv : n.scalars) a[4 + kv.first] = kv.second, present_[i] |= 1u << kv.first;
            for (int64_t q = pod_off_[i]; q < pod_off_[i + 1]; ++q) { // NodeInfo.AddPod
                const PodRequest r = pod_request(pods_->get(), q);
                for (int k = 0; k < RW; ++k) req_[i * (size_t)RW + (size_t)k] += r.v[k];
                nz_[i * 2 + 0] += r.nz_mcpu;
                nz_[i * 2 + 1] += r.nz_mem;
            }
        }
        sn_ = fgs_snapshot{};
        sn_.n_nodes = (int64_t)nodes.size();
        sn_.n_scalars = S;
        sn_.node_valid = valid_.data();
        sn_.alloc = alloc_.data();
        sn_.alloc_scalar_present = present_.data();
        sn_.requested = req_.data();
        sn_.nonzero = nz_.data();
        sn_.pod_off = pod_off_.data();
        sn_.pods = pods_->ref();
    }
    const fgs_snapshot* get() const { return &sn_; }

  private:
    std::unique_ptr<Podset> pods_;
    std::vector<int64_t> alloc_, req_, nz_, pod_off_;
    std::vector<uint32_t> present_;
    std::vector<uint8_t> valid_;
    fgs_snapshot sn_;
};

} // namespace testing
} // namespace fgs
