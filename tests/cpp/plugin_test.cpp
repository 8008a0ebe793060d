// plugin_test.cpp — the reference's own plugin tests, replayed against the C++ host layer (include/fgs_plugins.hpp)
// over libfgs. Each block names the Go test it follows; nodes and pods are built with tests/cpp/builders.hpp.
//   TestNodeResourcesAllocatable        pkg/noderesources/allocatable_test.go:112-236
//   TestPreFilter (capacity scheduling) pkg/capacityscheduling/capacity_scheduling_test.go:52-164
//   TestPermit / TestPreFilter (gang)   pkg/coscheduling/core/core_test.go:42-234 (representative cases)
//   FlexGPU has no tests in the reference: its Filter messages, Score / NormalizeScore and Reserve indexes are
//   checked against values worked out by hand from pkg/flexgpu/flex_gpu.go and gpu_node.go.
// Needs a CUDA device (run by tests/test_gpu_cpp_plugins.py).
#include <cstdio>
#include <cstdlib>
#include <string>
#include <vector>

#include "builders.hpp"
#include "fgs_plugins.hpp"

using namespace fgs;
using namespace fgs::testing;
using fgs::framework::Code;
using fgs::framework::NodeScoreList;
using fgs::framework::Status;

static int g_checks = 0, g_failed = 0;
#define CHECK(cond)                                                              \
    do {                                                                         \
        ++g_checks;                                                              \
        if (!(cond)) {                                                           \
            ++g_failed;                                                          \
            std::fprintf(stderr, "FAILED %s:%d: %s\n", __FILE__, __LINE__, #cond); \
        }                                                                        \
    } while (0)

static Node makeNode(int64_t milliCPU, int64_t memory) { // makeNodeInfo, allocatable_test.go:303-319
    Node n;
    n.cpu = milliCPU;
    n.memory = memory;
    n.pods = 110;
    return n;
}

// ---- TestNodeResourcesAllocatable ------------------------------------------------------------------------------------
static void TestNodeResourcesAllocatable(Handle& h) {
    struct Case {
        const char* name;
        std::vector<Node> nodes;
        NodeResourcesAllocatableArgs args;
        std::vector<int64_t> expected;
        const char* wantErr;
    };
    const std::vector<ResourceSpec> dflt = {{"cpu", 1 << 20}, {"memory", 1}};
    const int64_t Mi = 1048576;
    std::vector<Case> cases = {
        {"nothing scheduled, nothing requested", {makeNode(4000, 10000), makeNode(4000, 10000)}, {dflt, "Least"}, {0, 0}, nullptr},
        {"differently sized machines, least mode", {makeNode(4000, 10000), makeNode(6000, 10000)}, {dflt, "Least"}, {100, 0}, nullptr},
        {"differently sized machines, most mode", {makeNode(4000, 10000), makeNode(6000, 10000)}, {dflt, "Most"}, {0, 100}, nullptr},
        {"requested more than the node, least mode", {makeNode(4000, 1000), makeNode(5000, 1000)}, {dflt, "Least"}, {100, 0}, nullptr},
        {"cpu weight 1<<30, least", {makeNode(1000, 2000), makeNode(1005, 1000)}, {{{"cpu", 1 << 30}, {"memory", 1}}, "Least"}, {100, 0}, nullptr},
        {"3 differently sized machines, least mode", {makeNode(1000, 1000 * Mi), makeNode(2000, 2000 * Mi), makeNode(3000, 3000 * Mi)}, {dflt, "Least"}, {100, 50, 0}, nullptr},
        {"3 differently sized machines, most mode", {makeNode(1000, 1000 * Mi), makeNode(2000, 2000 * Mi), makeNode(3000, 3000 * Mi)}, {dflt, "Most"}, {0, 50, 100}, nullptr},
        {"resource with negative weight", {makeNode(4000, 10000)}, {{{"memory", -1}, {"cpu", 1}}, "Least"}, {}, "resource Weight of memory should be a positive value, got -1"},
        {"resource with zero weight", {makeNode(4000, 10000)}, {{{"memory", 1}, {"cpu", 0}}, "Least"}, {}, "resource Weight of cpu should be a positive value, got 0"},
        {"invalid mode", {makeNode(4000, 10000)}, {dflt, "Balanced"}, {}, "invalid mode, got Balanced"},
    };
    for (auto& c : cases) {
        Profile p;
        p.Filter = {};
        p.Score = {{Allocatable::Name, 1}};
        p.AllocatableArgs = c.args;
        std::string err;
        try {
            Framework fw(h, p); // NewAllocatable
            if (c.wantErr) {
                CHECK(!"expected an error");
                continue;
            }
            Snapshot snap(c.nodes);
            h.UpdateSnapshot(snap.get());
            Podset pods({MakePod({MakeCtr({})})}, 2);
            CycleState state;
            CHECK(fw.RunPreFilterPlugins(state, pods.get(), 0).IsSuccess());
            Allocatable pl(fw);
            NodeScoreList list;
            for (int64_t n = 0; n < (int64_t)c.nodes.size(); ++n) {
                auto sc = pl.Score(state, n); // :224-229 of the test
                CHECK(sc.second.IsSuccess());
                list.push_back({n, sc.first});
            }
            CHECK(pl.NormalizeScore(state, list).IsSuccess());
            for (size_t n = 0; n < c.nodes.size(); ++n) {
                CHECK(list[n].Score == c.expected[n]);
                if (list[n].Score != c.expected[n]) std::fprintf(stderr, "  %s: node %zu got %lld want %lld\n", c.name, n, (long long)list[n].Score, (long long)c.expected[n]);
            }
        } catch (const std::invalid_argument& e) {
            err = e.what();
            CHECK(c.wantErr && err == c.wantErr);
            if (!(c.wantErr && err == c.wantErr)) std::fprintf(stderr, "  %s: error '%s'\n", c.name, err.c_str());
        }
    }
}

// ---- FlexGPU: Filter / Score / NormalizeScore / Reserve ------------------------------------------------------------------
static void TestFlexGPU(Handle& h) {
    const int64_t GiB = 1ll << 30;
    auto gpuNode = [&](std::vector<Pod> on) {
        Node n = makeNode(32000, 128 * GiB);
        n.ephemeral = 100 * GiB;
        n.scalars = {{0, 2}, {1, 32 * GiB}}; // 2 gpus x 16 GiB
        n.on = std::move(on);
        return n;
    };
    auto placed = [&](Ctr c, int32_t index) {
        Pod p = MakePod({c});
        p.node_name = true;
        p.index = index;
        return p;
    };
    std::vector<Node> nodes = {
        gpuNode({}),                                                                                    // 0: empty
        gpuNode({placed(GpuLimit(MakeCtr({{0, 100}}), 1), 0)}),                                          // 1: device 0 held by a whole-GPU pod
        gpuNode({placed(MemLimit(MakeCtr({{0, 100}}), 10 * GiB), 0), placed(MemLimit(MakeCtr({{0, 100}}), 12 * GiB), 1)}), // 2: 6 and 4 GiB left
        makeNode(32000, 128 * GiB),                                                                       // 3: no flex resources
    };
    Snapshot snap(nodes);
    h.UpdateSnapshot(snap.get());
    Profile prof;
    prof.Filter = {Fit::Name, FlexGPU::Name};
    prof.Score = {{FlexGPU::Name, 1}, {Fit::Name, 1}};
    Framework fw(h, prof);
    FlexGPU flex(fw);
    Podset pods({
                    MakePod({MemLimit(MakeCtr({{0, 500}, {1, GiB}}), 5 * GiB)}),                      // 0 fractional 5 GiB
                    MakePod({GpuLimit(MakeCtr({{0, 500}, {1, GiB}}), 1)}),                            // 1 whole GPU
                    MakePod({MemLimit(GpuLimit(MakeCtr({{0, 500}}), 1), GiB)}),                       // 2 both limits
                    MakePod({MakeCtr({{0, 500}})}),                                                   // 3 no flex limit
                    MakePod({MemLimit(MakeCtr({{0, 500}}), 7 * GiB)}),                                // 4 fractional 7 GiB
                    MakePod({MakeCtr({{0, 64000}})}),                                                 // 5 too much cpu
                },
                2);
    { // fractional 5 GiB
        CycleState st;
        CHECK(fw.RunPreFilterPlugins(st, pods.get(), 0).IsSuccess());
        CHECK(flex.Filter(st, 0).IsSuccess() && flex.Filter(st, 1).IsSuccess() && flex.Filter(st, 2).IsSuccess());
        Status s3 = flex.Filter(st, 3);
        CHECK(s3.GetCode() == Code::UnschedulableAndUnresolvable && s3.Message() == "unknown resource type");
        // MemScore = Σ(memEach − used): 32, 32, 10 GiB (gpu_node.go:189-199); reverse-normalised over {0,1,2}
        CHECK(flex.Score(st, 0).first == 32 * GiB && flex.Score(st, 2).first == 10 * GiB);
        NodeScoreList l = {{0, flex.Score(st, 0).first}, {1, flex.Score(st, 1).first}, {2, flex.Score(st, 2).first}};
        flex.NormalizeScore(st, l);
        CHECK(l[0].Score == 0 && l[1].Score == 0 && l[2].Score == 100 - (100 * 10) / 32);
        int idx = -9;
        CHECK(flex.Reserve(pods.get(), 0, 2, &idx).IsSuccess() && idx == 0); // first fit: device 0 has 6 GiB
        CHECK(flex.Reserve(pods.get(), 0, 1, &idx).IsSuccess() && idx == 1); // device 0 is a monopoly
        CHECK(st.Read(Framework::kStateKey)->feasible_count == 3);
        CHECK(RunFilterPlugins(fw, st, 3).FailedPlugin() == FlexGPU::Name);
        // prioritizeNodes hands RunScorePlugins the nodes that passed EVERY filter plugin (generic_scheduler.go:420-426): say
        // another plugin of the profile rejected node 1. NormalizeScore now runs over {0, 2}: MemScore 32 and 10 GiB →
        // 100 − 100·32/32 = 0 and 100 − 100·10/32 = 69, so node 2 wins by far more than the LeastAllocated terms differ.
        auto ev = st.Read(Framework::kStateKey);
        CHECK(fw.RunScorePlugins(st, {0, 2}).IsSuccess());
        CHECK(ev->feasible_count == 2 && ev->Feasible(0) && !ev->Feasible(1) && ev->Feasible(2) && !ev->Feasible(3));
        CHECK(fw.SelectHost(st) == 2 && ev->selected_score == ev->total[2] && ev->total[2] - ev->total[0] > 60);
        CHECK(fw.RunScorePlugins(st, {3}).IsSuccess()); // a node the engine's own filters rejected stays out, whatever the list says
        CHECK(ev->feasible_count == 0 && fw.SelectHost(st) == -1);
    }
    { // whole GPU
        CycleState st;
        CHECK(fw.RunPreFilterPlugins(st, pods.get(), 1).IsSuccess());
        CHECK(flex.Filter(st, 0).IsSuccess() && flex.Filter(st, 1).IsSuccess());
        CHECK(flex.Filter(st, 2).IsSuccess()); // limit sums allow it (0 + 1 <= 2) although no device is idle …
        int idx = -9;
        Status r = flex.Reserve(pods.get(), 1, 2, &idx); // … and Reserve finds none (gpu_node.go:163-177)
        CHECK(r.GetCode() == Code::Unschedulable && r.Message() == "allocate index fail nvidia.flex.com/gpu");
        CHECK(flex.Reserve(pods.get(), 1, 1, &idx).IsSuccess() && idx == 1);
        CHECK(flex.Score(st, 0).first == 2 && flex.Score(st, 1).first == 1 && flex.Score(st, 2).first == 0); // GPUScore
    }
    { // both limits
        CycleState st;
        fw.RunPreFilterPlugins(st, pods.get(), 2);
        Status s = flex.Filter(st, 0);
        CHECK(s.GetCode() == Code::UnschedulableAndUnresolvable && s.Message() == "pod conflict resources");
        int idx = -9;
        CHECK(flex.Reserve(pods.get(), 2, 0, &idx).Message() == "pod conflict resources");
    }
    { // no flex limit: every node passes FlexGPU, Reserve has nothing to do
        CycleState st;
        fw.RunPreFilterPlugins(st, pods.get(), 3);
        CHECK(flex.Filter(st, 3).IsSuccess());
        int idx = -9;
        CHECK(flex.Reserve(pods.get(), 3, 3, &idx).IsSuccess() && idx == -1);
    }
    { // fractional 7 GiB: node 2's limit sum allows it (22 + 7 <= 32 GiB) but its devices have 6 and 4 GiB left
        CycleState st;
        fw.RunPreFilterPlugins(st, pods.get(), 4);
        Status s = flex.Filter(st, 2);
        CHECK(s.GetCode() == Code::Unschedulable && s.Message() == "no fit indexes resource nvidia.flex.com/memory");
        Podset big({MakePod({MemLimit(MakeCtr({{0, 500}}), 12 * GiB)})}, 2); // 22 + 12 > 32 GiB: the limit sum fails first
        CycleState st2;
        fw.RunPreFilterPlugins(st2, big.get(), 0);
        CHECK(flex.Filter(st2, 2).Message() == "insufficient resource nvidia.flex.com/memory");
    }
    { // Fit's reasons
        CycleState st;
        fw.RunPreFilterPlugins(st, pods.get(), 5);
        Status s = Fit(fw).Filter(st, 0);
        CHECK(s.GetCode() == Code::Unschedulable && s.Message() == "Insufficient cpu");
        CHECK(RunFilterPlugins(fw, st, 0).FailedPlugin() == Fit::Name);
        CHECK(st.Read(Framework::kStateKey)->feasible_count == 0 && fw.SelectHost(st) == -1);
    }
}

// ---- CapacityScheduling.PreFilter (capacity_scheduling_test.go:52-164) ---------------------------------------------------------
static void TestCapacityPreFilter(Handle& h) {
    CapacityScheduling cs(h);
    auto quotas = [](int64_t n_ns, std::vector<uint8_t>& has, std::vector<int64_t>& mn, std::vector<int64_t>& mx, std::vector<int64_t>& used,
                     std::vector<uint32_t>& pres) {
        fgs_quotas q{};
        q.n_ns = n_ns;
        q.n_scalars = 2;
        q.has_eq = has.data(), q.min = mn.data(), q.max = mx.data(), q.used = used.data(), q.used_scalar_present = pres.data();
        return q;
    };
    { // "pod subjects to ElasticQuota": min 1000 / max 2000 / used 300 (memory); requests 500 → Success, 1800 → over min sum
        std::vector<uint8_t> has = {0, 1};
        std::vector<int64_t> mn = {0, 0, 0, 0, 0, 1000, 0, 0}, mx = {0, 0, 0, 0, 0, 2000, 0, 0}, used = {0, 0, 0, 0, 0, 300, 0, 0};
        std::vector<uint32_t> pres = {0, 0};
        fgs_quotas q = quotas(2, has, mn, mx, used, pres);
        Podset pods({MakePod({MakeCtr({{1, 500}})}, 1), MakePod({MakeCtr({{1, 1800}})}, 1)}, 2);
        cs.Evaluate(pods.get(), &q);
        CHECK(cs.PreFilter(0, "ns1", "t1-p1").IsSuccess());
        Status s = cs.PreFilter(1, "ns1", "t1-p2");
        CHECK(s.GetCode() == Code::Unschedulable);
        CHECK(s.Message() == "Pod ns1/t1-p2 is rejected in PreFilter because ElasticQuota ns1 is more than Max" ||
              s.Message() == "Pod ns1/t1-p2 is rejected in PreFilter because total ElasticQuota used is more than min");
    }
    { // "the sum of used is bigger than the sum of min": 1800 + 200 used, Σmin 2000, request 500
        std::vector<uint8_t> has = {0, 1, 1};
        std::vector<int64_t> mn = {0, 0, 0, 0, 0, 1000, 0, 0, 0, 1000, 0, 0}, mx = {0, 0, 0, 0, 0, 2000, 0, 0, 0, 2000, 0, 0},
                             used = {0, 0, 0, 0, 0, 1800, 0, 0, 0, 200, 0, 0};
        std::vector<uint32_t> pres = {0, 0, 0};
        fgs_quotas q = quotas(3, has, mn, mx, used, pres);
        Podset pods({MakePod({MakeCtr({{1, 500}})}, 2)}, 2);
        cs.Evaluate(pods.get(), &q);
        Status s = cs.PreFilter(0, "ns2", "t1-p1");
        CHECK(s.GetCode() == Code::Unschedulable && s.Message() == "Pod ns2/t1-p1 is rejected in PreFilter because total ElasticQuota used is more than min");
    }
}

// ---- CapacityScheduling.PostFilter: dry-run preemption on every node + pickOneNodeForPreemption ----------------------------------
static void TestCapacityPostFilter(Handle& h) {
    auto victim = [](int32_t prio, int64_t start) {
        Pod p = MakePod({MakeCtr({{0, 600}})}, 0, -1, prio);
        p.node_name = true;
        p.start = start;
        return p;
    };
    std::vector<Node> nodes = {makeNode(1000, 1 << 30), makeNode(1000, 1 << 30), makeNode(1000, 1 << 30), makeNode(1000, 1 << 30), makeNode(500, 1 << 30)};
    nodes[0].on = {victim(50, 0)};
    nodes[1].on = {victim(40, 5)};
    nodes[2].on = {victim(40, 9)}; // as node 1, but its victim started later: the one to evict (preemption.go:393)
    nodes[3].on = {victim(2000, 0)}; // more important than the preemptor
    Snapshot snap(nodes);
    h.UpdateSnapshot(snap.get());
    Profile prof;
    prof.Filter = {Fit::Name};
    prof.Score = {{Fit::Name, 1}};
    Framework fw(h, prof);
    CapacityScheduling cs(h);
    Podset pods({MakePod({MakeCtr({{0, 900}})}, 0, -1, 1000), MakePod({MakeCtr({{0, 900}})}, 0, -1, 10)}, 2);
    CycleState st;
    CHECK(fw.RunPreFilterPlugins(st, pods.get(), 0).IsSuccess());
    CHECK(st.Read(Framework::kStateKey)->feasible_count == 0); // nothing fits: PostFilter's turn
    CHECK(fw.FitError() == "0/5 nodes are available: 5 Insufficient cpu."); // framework/types.go:221-250
    auto r = cs.PostFilter(pods.get(), 0, nullptr, st.Read(Framework::kStateKey).get());
    CHECK(r.second.IsSuccess());
    CHECK(r.first.Candidates == 3 && r.first.NominatedNode == 2);
    CHECK(r.first.Victims.size() == 1 && r.first.Victims[0] == snap.get()->pod_off[2]);
    // a preemptor less important than everybody: no candidates, the nomination is cleared (preemption.go:162-173)
    CycleState st2;
    CHECK(fw.RunPreFilterPlugins(st2, pods.get(), 1).IsSuccess());
    auto r2 = cs.PostFilter(pods.get(), 1, nullptr, st2.Read(Framework::kStateKey).get());
    CHECK(r2.second.GetCode() == Code::Unschedulable && r2.first.NominatedNode == -1 && r2.first.Candidates == 0);
    CHECK(r2.second.Message() == "preemption: 0/5 nodes are available");
}

// ---- Coscheduling PreFilter / Permit (core_test.go) ------------------------------------------------------------------------------
static void TestCoscheduling(Handle& h) {
    // 3 nodes of 1 cpu each; PodGroup 0: MinMember 2, MinResources cpu 2 (enough) — PodGroup 1: MinMember 2,
    // MinResources cpu 40 (resource gap) — PodGroup 2: MinMember 3 with 2 pods only (not enough siblings)
    std::vector<Node> nodes = {makeNode(1000, 1 << 30), makeNode(1000, 1 << 30), makeNode(1000, 1 << 30)};
    Pod assigned = MakePod({MakeCtr({{0, 100}})}, 0, 0);
    assigned.node_name = true;
    nodes[0].on = {assigned};
    Snapshot snap(nodes);
    h.UpdateSnapshot(snap.get());
    std::vector<uint8_t> exists = {1, 1, 1}, has_min = {1, 1, 0}, denied = {0, 0, 0}, permitted = {0, 0, 0};
    std::vector<int32_t> ns = {0, 0, 0}, min_member = {2, 2, 3};
    std::vector<int64_t> min_res = {2000, 0, 0, 0, 0, 0, 40000, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
    std::vector<uint32_t> min_present = {1, 1, 0};
    fgs_podgroups pgs{};
    pgs.n_groups = 3;
    pgs.n_scalars = 2;
    pgs.exists = exists.data(), pgs.ns = ns.data(), pgs.min_member = min_member.data(), pgs.has_min_resources = has_min.data();
    pgs.min_res = min_res.data(), pgs.min_res_present = min_present.data(), pgs.denied = denied.data(), pgs.permitted = permitted.data();
    Podset pending({MakePod({MakeCtr({{0, 100}})}, 0, 0), MakePod({MakeCtr({{0, 100}})}, 0, 0), MakePod({MakeCtr({{0, 100}})}, 0, 1),
                    MakePod({MakeCtr({{0, 100}})}, 0, 1), MakePod({MakeCtr({{0, 100}})}, 0, 2), MakePod({MakeCtr({{0, 100}})}, 0, 2),
                    MakePod({MakeCtr({{0, 100}})}, 0, -1)},
                   2);
    Coscheduling co(h);
    co.Evaluate(pending.get(), &pgs);
    CHECK(co.PreFilter(0).IsSuccess());                                          // enough siblings, enough cluster resource
    CHECK(co.PreFilter(2).GetCode() == Code::UnschedulableAndUnresolvable && co.PreFilter(2).Message() == "resource gap");
    CHECK(co.PreFilter(4).Message() == "pre-filter pod cannot find enough sibling pods");
    CHECK(co.PreFilter(6).IsSuccess());                                          // no PodGroup label
    CHECK(co.Permit(0).GetCode() == Code::Success);                              // 1 assigned + this one >= MinMember 2
    CHECK(co.Permit(2).GetCode() == Code::Wait);                                 // 0 assigned + 1 < 2
    CHECK(co.Permit(6).GetCode() == Code::Success);                              // PodGroupNotSpecified
    // PostFilter (coscheduling.go:140-176): group 0 has 1 of its 2 members assigned — a gap of 50 % — so an unschedulable
    // member takes the whole group down; a pod without a group only reports that
    bool reject = false;
    Status pf = co.PostFilter(0, "default/pg0", "pod-0", &reject);
    CHECK(reject && pf.GetCode() == Code::Unschedulable &&
          pf.Message() == "PodGroup default/pg0 gets rejected due to Pod pod-0 is unschedulable even after PostFilter");
    pf = co.PostFilter(6, "", "pod-6", &reject);
    CHECK(!reject && pf.GetCode() == Code::Unschedulable && pf.Message() == "can not find pod group");
}

int main() {
    try {
        Handle h(0, 2);
        TestNodeResourcesAllocatable(h);
        TestFlexGPU(h);
        TestCapacityPreFilter(h);
        TestCapacityPostFilter(h);
        TestCoscheduling(h);
    } catch (const std::exception& e) {
        std::fprintf(stderr, "exception: %s\n", e.what());
        return 2;
    }
    std::printf("%d checks, %d failed\n", g_checks, g_failed);
    return g_failed ? 1 : 0;
}
