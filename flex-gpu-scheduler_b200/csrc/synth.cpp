// synth.cpp — deterministic synthetic clusters for the configs of BASELINE.json.
// Product-side host code: bench.py and the parity tests both build their inputs here, so the
// GPU path and the CPU oracle always see the same object snapshot.
//
// Generator = SplitMix64. All quantities are integers. The NodeInfo aggregates (Requested,
// NonZeroRequested) are accumulated with the AddPod rule of
// vendor/k8s.io/kubernetes/pkg/scheduler/framework/types.go:583-609.
#include <cstring>
#include <new>
#include <vector>

#include "../../include/fgs.h"
#include "podmath.h"

namespace {

struct Rng {
    uint64_t s;
    explicit Rng(uint64_t seed) : s(seed) {}
    uint64_t next() {
        uint64_t z = (s += 0x9E3779B97F4A7C15ull);
        z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
        z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
        return z ^ (z >> 31);
    }
    int64_t uni(int64_t lo, int64_t hi) { return lo + (int64_t)(next() % (uint64_t)(hi - lo + 1)); } // inclusive
    bool pm(int per_mille) { return (int)(next() % 1000) < per_mille; }
    template <typename T, size_t K>
    T pick(const T (&a)[K]) { return a[next() % K]; }
};

struct PodsetStore {
    int S = 2;
    std::vector<int32_t> ns, pg, priority, index_annot, n_init;
    std::vector<uint8_t> has_node_name;
    std::vector<int64_t> ctr_off{0};
    std::vector<uint32_t> overhead_present;
    std::vector<int64_t> overhead;
    std::vector<int64_t> c_req, c_lim_flex;
    std::vector<uint32_t> c_req_present, c_lim_present;
    bool any_overhead = false;

    int W() const { return 3 + S; }
    void add_ctr(const int64_t* req, uint32_t req_present, int64_t gpu_lim, int64_t mem_lim, uint32_t lim_present) {
        c_req.insert(c_req.end(), req, req + W());
        c_req_present.push_back(req_present);
        c_lim_flex.push_back(gpu_lim);
        c_lim_flex.push_back(mem_lim);
        c_lim_present.push_back(lim_present);
    }
    // call after the pod's containers were added (regular first, then init)
    void end_pod(int32_t ns_, int32_t pg_, int32_t prio, bool node_name, int32_t annot, int32_t ninit,
                 const int64_t* ovh, uint32_t ovh_present) {
        ns.push_back(ns_);
        pg.push_back(pg_);
        priority.push_back(prio);
        has_node_name.push_back(node_name ? 1 : 0);
        index_annot.push_back(annot);
        n_init.push_back(ninit);
        ctr_off.push_back((int64_t)c_req_present.size());
        overhead_present.push_back(ovh_present);
        if (ovh) {
            overhead.insert(overhead.end(), ovh, ovh + W());
            any_overhead = true;
        } else {
            overhead.insert(overhead.end(), (size_t)W(), 0);
        }
    }
    int64_t size() const { return (int64_t)ns.size(); }
    void view(fgs_podset* v) const {
        std::memset(v, 0, sizeof(*v));
        v->n_pods = size();
        v->n_scalars = S;
        v->ns = ns.data();
        v->pg = pg.data();
        v->priority = priority.data();
        v->has_node_name = has_node_name.data();
        v->index_annot = index_annot.data();
        v->ctr_off = ctr_off.data();
        v->n_init = n_init.data();
        v->overhead_present = overhead_present.data();
        v->overhead = overhead.data();
        v->n_ctrs = (int64_t)c_req_present.size();
        v->c_req = c_req.data();
        v->c_req_present = c_req_present.data();
        v->c_lim_flex = c_lim_flex.data();
        v->c_lim_present = c_lim_present.data();
    }
};

const int64_t kCpuReq[] = {100, 250, 500, 1000, 2000, 4000};
const int64_t kMi = 1024 * 1024, kGi = 1024 * kMi;
const int64_t kMemReq[] = {256 * kMi, 512 * kMi, 1 * kGi, 2 * kGi, 4 * kGi, 8 * kGi, 16 * kGi};
const int64_t kNodeCpu[] = {32000, 64000, 96000, 128000};
const int64_t kNodeMem[] = {256 * kGi, 512 * kGi, 1024 * kGi, 2048 * kGi};
const int64_t kGpuMemMiB[] = {24576, 40960, 81920, 143360, 196608};

} // namespace

struct fgs_synth {
    fgs_synth_params p;
    // nodes
    std::vector<uint8_t> node_valid;
    std::vector<int64_t> alloc, requested, nonzero, pod_off;
    std::vector<uint32_t> alloc_scalar_present;
    PodsetStore existing, pending, nominated;
    // groups
    std::vector<uint8_t> pg_exists, pg_has_min, pg_denied, pg_permitted;
    std::vector<int32_t> pg_ns, pg_min_member;
    std::vector<int64_t> pg_min_res;
    std::vector<uint32_t> pg_min_present;
    // quotas
    std::vector<uint8_t> q_has;
    std::vector<int64_t> q_min, q_max, q_used;
    std::vector<uint32_t> q_used_present;
    // views
    fgs_snapshot snap_v;
    fgs_podset pending_v, nominated_v;
    fgs_podgroups pgs_v;
    fgs_quotas quotas_v;
};

namespace {

// a CPU-only / flex pod with 1–2 containers; returns through `store`
struct PodSpec {
    int kind; // 0 none, 1 whole gpu, 2 fractional, 3 both (degenerate)
    int64_t gpu_lim, mem_lim;
};

void emit_pod(PodsetStore& st, Rng& r, const fgs_synth_params& p, const PodSpec& spec, int32_t ns, int32_t pg,
              int32_t prio, bool node_name, int32_t annot) {
    const int S = st.S, W = 3 + S;
    int n_reg = r.pm(p.pm_two_ctrs) ? 2 : 1;
    for (int c = 0; c < n_reg; ++c) {
        std::vector<int64_t> req((size_t)W, 0);
        uint32_t present = 0;
        if (!r.pm(p.pm_missing_req)) {
            req[0] = r.pick(kCpuReq);
            present |= 1u;
        }
        if (!r.pm(p.pm_missing_req)) {
            req[1] = r.pick(kMemReq);
            present |= 2u;
        }
        if (r.pm(100)) {
            req[2] = r.uni(1, 64) * kGi;
            present |= 4u;
        }
        int64_t gl = 0, ml = 0;
        uint32_t lp = 0;
        if (c == 0) { // the flex limits sit on the first container; requests == limits for extended resources
            if (spec.kind == 1 || spec.kind == 3) {
                gl = spec.gpu_lim;
                lp |= 1u;
                req[3 + FGS_SCALAR_FLEX_GPU] = gl;
                present |= 8u << FGS_SCALAR_FLEX_GPU;
            }
            if (spec.kind == 2 || spec.kind == 3) {
                ml = spec.mem_lim;
                lp |= 2u;
                req[3 + FGS_SCALAR_FLEX_MEM] = ml;
                present |= 8u << FGS_SCALAR_FLEX_MEM;
            }
            if (S > 2 && r.pm(p.pm_extra_scalar)) {
                int s = 2 + (int)(r.next() % (uint64_t)(S - 2));
                req[3 + s] = r.uni(0, 4);
                present |= 8u << s;
            }
        }
        st.add_ctr(req.data(), present, gl, ml, lp);
    }
    int ninit = 0;
    if (r.pm(p.pm_init_ctr)) {
        std::vector<int64_t> req((size_t)W, 0);
        uint32_t present = 3u;
        req[0] = r.pick(kCpuReq) * 2;
        req[1] = r.pick(kMemReq);
        st.add_ctr(req.data(), present, 0, 0, 0);
        ninit = 1;
    }
    if (r.pm(p.pm_overhead)) {
        std::vector<int64_t> ovh((size_t)W, 0);
        ovh[0] = r.uni(1, 8) * 125; // 125m .. 1000m
        ovh[1] = r.uni(1, 4) * 64 * kMi;
        st.end_pod(ns, pg, prio, node_name, annot, ninit, ovh.data(), 0x80000000u | 3u);
    } else {
        st.end_pod(ns, pg, prio, node_name, annot, ninit, nullptr, 0);
    }
}

} // namespace

extern "C" int fgs_synth_create(const fgs_synth_params* pp, fgs_synth** out) {
    if (!pp || !out) return FGS_E_INVALID;
    const fgs_synth_params& p = *pp;
    if (p.n_scalars < 2 || p.n_scalars > FGS_MAX_SCALARS || p.n_devices < 1 || p.n_devices > FGS_MAX_DEVICES ||
        p.n_nodes < 0 || p.n_pending < 0)
        return FGS_E_INVALID;
    fgs_synth* sy = new (std::nothrow) fgs_synth();
    if (!sy) return FGS_E_INVALID;
    sy->p = p;
    const int S = p.n_scalars, AW = 4 + S, RW = 3 + S, D = p.n_devices;
    const int64_t N = p.n_nodes;
    const int64_t NS = p.n_namespaces > 0 ? p.n_namespaces : 1;
    const int64_t G = p.n_groups;
    sy->existing.S = sy->pending.S = sy->nominated.S = S;

    Rng rn(p.seed_nodes);
    Rng rp(p.seed_pods);

    // ---- pod groups (members are created below) ------------------------------------------
    std::vector<int64_t> pg_unit_cpu((size_t)G), pg_unit_mem((size_t)G), pg_unit_gpu((size_t)G);
    std::vector<int32_t> pg_placed((size_t)G, 0), pg_total((size_t)G, 0);
    for (int64_t g = 0; g < G; ++g) {
        int32_t mm = (int32_t)rp.uni(p.min_member_lo, p.min_member_hi);
        sy->pg_exists.push_back(rp.pm(980) ? 1 : 0);
        sy->pg_ns.push_back((int32_t)rp.uni(0, NS - 1));
        sy->pg_min_member.push_back(mm);
        bool has_min = rp.pm(850);
        sy->pg_has_min.push_back(has_min ? 1 : 0);
        sy->pg_denied.push_back(rp.pm(20) ? 1 : 0);
        sy->pg_permitted.push_back(rp.pm(20) ? 1 : 0);
        pg_unit_cpu[(size_t)g] = rp.uni(1, 8) * 1000;
        pg_unit_mem[(size_t)g] = rp.uni(1, 32) * kGi;
        pg_unit_gpu[(size_t)g] = rp.uni(0, 1);
        std::vector<int64_t> mr((size_t)AW, 0);
        uint32_t present = 0;
        if (has_min) {
            // a few groups ask for far more than the cluster holds → resource gap
            int64_t scale = rp.pm(60) ? 4000 : 1;
            mr[0] = mm * pg_unit_cpu[(size_t)g] * scale;
            present |= 1u;
            mr[1] = mm * pg_unit_mem[(size_t)g];
            present |= 2u;
            if (pg_unit_gpu[(size_t)g]) {
                mr[4 + FGS_SCALAR_FLEX_GPU] = mm;
                present |= 1u << (4 + FGS_SCALAR_FLEX_GPU);
            }
            if (rp.pm(50)) { // a zero / negative-proof key
                mr[2] = 0;
                present |= 4u;
            }
        }
        sy->pg_min_res.insert(sy->pg_min_res.end(), mr.begin(), mr.end());
        sy->pg_min_present.push_back(present);
        if (rp.pm(p.pm_groups_placed)) pg_placed[(size_t)g] = (int32_t)rp.uni(1, mm > 1 ? mm - 1 : 1);
        pg_total[(size_t)g] = rp.pm(p.pm_groups_short) ? (int32_t)rp.uni(1, mm > 1 ? mm - 1 : 1) : mm;
        if (pg_total[(size_t)g] < pg_placed[(size_t)g]) pg_total[(size_t)g] = pg_placed[(size_t)g];
    }
    // placed group members by node
    std::vector<std::vector<int32_t>> placed_on((size_t)N);
    if (N > 0)
        for (int64_t g = 0; g < G; ++g)
            for (int k = 0; k < pg_placed[(size_t)g]; ++k) placed_on[(size_t)rn.uni(0, N - 1)].push_back((int32_t)g);

    // ---- nodes + existing pods --------------------------------------------------------------
    sy->alloc.assign((size_t)(N * AW), 0);
    sy->requested.assign((size_t)(N * RW), 0);
    sy->nonzero.assign((size_t)(N * 2), 0);
    sy->alloc_scalar_present.assign((size_t)N, 0);
    sy->node_valid.assign((size_t)N, 1);
    sy->pod_off.assign((size_t)(N + 1), 0);
    for (int64_t n = 0; n < N; ++n) {
        int64_t* al = &sy->alloc[(size_t)(n * AW)];
        al[0] = rn.pick(kNodeCpu);
        al[1] = rn.pick(kNodeMem);
        al[2] = rn.uni(1, 4) * 1024 * kGi;
        al[3] = 110;
        const int64_t M = rn.pick(kGpuMemMiB);
        const bool nongpu = rn.pm(p.pm_nongpu_nodes);
        const bool panic = rn.pm(p.pm_panic_nodes);
        const bool quirk = rn.pm(p.pm_quirk_nodes);
        if (rn.pm(p.pm_invalid_nodes)) sy->node_valid[(size_t)n] = 0;
        int gpus = D;
        if (!nongpu) {
            if (rn.pm(150) && D > 4) gpus = D / 2; // some nodes carry fewer devices
            al[4 + FGS_SCALAR_FLEX_GPU] = gpus;
            al[4 + FGS_SCALAR_FLEX_MEM] = gpus * M;
            sy->alloc_scalar_present[(size_t)n] |= (1u << FGS_SCALAR_FLEX_GPU) | (1u << FGS_SCALAR_FLEX_MEM);
            if (panic && rn.pm(500)) al[4 + FGS_SCALAR_FLEX_GPU] = 0; // gpu count 0 with the key present
        } else if (rn.pm(300)) {
            // only one of the two keys
            al[4 + FGS_SCALAR_FLEX_GPU] = gpus;
            sy->alloc_scalar_present[(size_t)n] |= 1u << FGS_SCALAR_FLEX_GPU;
        }
        for (int s = 2; s < S; ++s)
            if (rn.pm(500)) {
                al[4 + s] = rn.uni(0, 8);
                sy->alloc_scalar_present[(size_t)n] |= 1u << s;
            }
        // device occupancy while generating, so that annotations point at sensible devices
        std::vector<int64_t> used((size_t)D, 0);
        std::vector<uint8_t> mono((size_t)D, 0);
        int64_t n_pods = rn.uni(0, p.max_existing_pods);
        const int64_t first = sy->existing.size();
        auto account = [&](int64_t q) {
            fgs_podset v;
            sy->existing.view(&v);
            fgs::PodRequest pr = fgs::pod_request(&v, q);
            int64_t* rq = &sy->requested[(size_t)(n * RW)];
            for (int k = 0; k < RW; ++k) rq[k] += pr.v[k];
            sy->nonzero[(size_t)(n * 2 + 0)] += pr.nz_mcpu;
            sy->nonzero[(size_t)(n * 2 + 1)] += pr.nz_mem;
        };
        for (int64_t k = 0; k < n_pods; ++k) {
            PodSpec spec{0, 0, 0};
            int32_t annot = FGS_ANNOT_NONE;
            int roll = (int)(rn.next() % 100);
            if (!nongpu && roll < 25) { // whole GPU on a free device
                int dev = -1;
                for (int d = 0; d < gpus; ++d)
                    if (!mono[(size_t)d] && used[(size_t)d] == 0) {
                        dev = d;
                        break;
                    }
                if (dev >= 0) {
                    spec = PodSpec{1, 1, 0};
                    annot = dev;
                    mono[(size_t)dev] = 1;
                }
            } else if (!nongpu && roll < 75) { // fractional, first-fit
                int64_t ml = M * rn.uni(1, 8) / 16;
                int dev = -1;
                for (int d = 0; d < gpus; ++d)
                    if (!mono[(size_t)d] && used[(size_t)d] + ml <= M) {
                        dev = d;
                        break;
                    }
                if (dev >= 0) {
                    spec = PodSpec{2, 0, ml};
                    annot = dev;
                    used[(size_t)dev] += ml;
                }
            }
            if (quirk && spec.kind != 0) {
                int q = (int)(rn.next() % 4);
                if (q == 0) annot = FGS_ANNOT_NONE;                                // missing annotation
                else if (q == 1 && spec.kind == 2) spec.mem_lim *= 6;               // over-committed device
                else if (q == 2) annot = (int32_t)rn.uni(0, gpus - 1);              // monopoly + used conflicts
                else if (q == 3 && spec.kind == 1) spec.gpu_lim = 2;                // invalid gpu limit (warned only)
            }
            if (panic && spec.kind != 0 && rn.pm(300)) annot = gpus + (int32_t)rn.uni(0, 3); // index out of range
            emit_pod(sy->existing, rn, p, spec, (int32_t)rn.uni(0, NS - 1), -1, (int32_t)rn.uni(0, 3) * 100, true, annot);
            account(sy->existing.size() - 1);
        }
        for (int32_t g : placed_on[(size_t)n]) { // already-placed gang members (CPU pods)
            PodSpec spec{0, 0, 0};
            emit_pod(sy->existing, rn, p, spec, sy->pg_ns[(size_t)g], g, 100, true, FGS_ANNOT_NONE);
            account(sy->existing.size() - 1);
        }
        if (rn.pm(p.pm_neg_left_nodes)) { // over-committed node: negative left-over cpu / memory
            sy->requested[(size_t)(n * RW + 0)] = al[0] + rn.uni(1, 16) * 1000;
            sy->requested[(size_t)(n * RW + 1)] = al[1] + rn.uni(1, 64) * kGi;
        }
        (void)first;
        sy->pod_off[(size_t)(n + 1)] = sy->existing.size();
    }

    // ---- pending pods ---------------------------------------------------------------------------
    // gang members first (C3), then free pods up to n_pending
    std::vector<int32_t> gang_member_of;
    for (int64_t g = 0; g < G; ++g)
        for (int k = pg_placed[(size_t)g]; k < pg_total[(size_t)g]; ++k) gang_member_of.push_back((int32_t)g);
    const int64_t P = p.n_pending;
    for (int64_t i = 0; i < P; ++i) {
        PodSpec spec{0, 0, 0};
        int roll = (int)(rp.next() % 1000);
        if (roll < p.pm_fractional) spec = PodSpec{2, 0, rp.uni(1, 64) * 1024};
        else if (roll < p.pm_fractional + p.pm_whole) spec = PodSpec{1, rp.pm(950) ? 1 : rp.uni(0, 3), 0};
        else if (roll < p.pm_fractional + p.pm_whole + p.pm_nogpu) spec = PodSpec{0, 0, 0};
        else if (roll < p.pm_fractional + p.pm_whole + p.pm_nogpu + p.pm_degenerate) {
            if (rp.pm(500)) spec = PodSpec{3, 1, 4096};                     // both limits
            else spec = PodSpec{2, 0, 400000 + rp.uni(0, 1000)};            // larger than any device
        } else spec = PodSpec{2, 0, rp.uni(1, 64) * 1024};
        int32_t pg = -1;
        int32_t ns;
        if ((size_t)i < gang_member_of.size()) {
            pg = gang_member_of[(size_t)i];
            ns = sy->pg_ns[(size_t)pg];
        } else {
            ns = (int32_t)rp.uni(0, NS - 1);
            if (G > 0 && rp.pm(20)) pg = (int32_t)rp.uni(0, G - 1), ns = sy->pg_ns[(size_t)pg];
        }
        emit_pod(sy->pending, rp, p, spec, ns, pg, (int32_t)rp.uni(0, 3) * 100, false, FGS_ANNOT_NONE);
    }
    for (int i = 0; i < p.n_nominated; ++i) {
        PodSpec spec{0, 0, 0};
        emit_pod(sy->nominated, rp, p, spec, (int32_t)rp.uni(0, NS - 1), -1, (int32_t)rp.uni(0, 3) * 100, false, FGS_ANNOT_NONE);
    }

    // ---- elastic quotas -----------------------------------------------------------------------------
    {
        const int QW = 2 + S;
        sy->q_has.assign((size_t)NS, 0);
        sy->q_min.assign((size_t)(NS * QW), 0);
        sy->q_max.assign((size_t)(NS * QW), 0);
        sy->q_used.assign((size_t)(NS * QW), 0);
        sy->q_used_present.assign((size_t)NS, 0);
        // Σmin ≈ 0.9 × cluster
        int64_t tot_cpu = 0, tot_mem = 0, tot_gpu = 0, tot_gmem = 0;
        for (int64_t n = 0; n < N; ++n) {
            tot_cpu += sy->alloc[(size_t)(n * AW + 0)];
            tot_mem += sy->alloc[(size_t)(n * AW + 1)];
            tot_gpu += sy->alloc[(size_t)(n * AW + 4 + FGS_SCALAR_FLEX_GPU)];
            tot_gmem += sy->alloc[(size_t)(n * AW + 4 + FGS_SCALAR_FLEX_MEM)];
        }
        Rng rq(p.seed_pods ^ 0xE1A571C0ull);
        for (int64_t k = 0; k < NS; ++k) {
            bool has = p.n_namespaces > 0 && !rq.pm(p.pm_pods_without_eq);
            sy->q_has[(size_t)k] = has ? 1 : 0;
            if (!has) continue;
            int64_t base[4] = {tot_cpu * 9 / 10 / NS, tot_mem * 9 / 10 / NS, tot_gpu * 9 / 10 / NS, tot_gmem * 9 / 10 / NS};
            int cols[4] = {0, 1, 2 + FGS_SCALAR_FLEX_GPU, 2 + FGS_SCALAR_FLEX_MEM};
            for (int c = 0; c < 4; ++c) {
                int64_t mn = base[c] * rq.uni(50, 150) / 100;
                sy->q_min[(size_t)(k * QW + cols[c])] = mn;
                sy->q_max[(size_t)(k * QW + cols[c])] = mn * rq.uni(100, 200) / 100;
                sy->q_used[(size_t)(k * QW + cols[c])] = mn * rq.uni(0, 120) / 100;
            }
            if (rq.pm(30)) sy->q_max[(size_t)(k * QW + 2 + FGS_SCALAR_FLEX_GPU)] = 0; // Max lacks the key
            sy->q_used_present[(size_t)k] = (uint32_t)rq.uni(0, 3);
            for (int sc = 0; sc < S; ++sc)
                if (!((sy->q_used_present[(size_t)k] >> sc) & 1u)) sy->q_used[(size_t)(k * QW + 2 + sc)] = 0;
        }
    }

    // ---- views -------------------------------------------------------------------------------------------
    std::memset(&sy->snap_v, 0, sizeof(sy->snap_v));
    sy->snap_v.n_nodes = N;
    sy->snap_v.n_scalars = S;
    sy->snap_v.node_valid = sy->node_valid.data();
    sy->snap_v.alloc = sy->alloc.data();
    sy->snap_v.alloc_scalar_present = sy->alloc_scalar_present.data();
    sy->snap_v.requested = sy->requested.data();
    sy->snap_v.nonzero = sy->nonzero.data();
    sy->snap_v.pod_off = sy->pod_off.data();
    sy->existing.view(&sy->snap_v.pods);
    sy->pending.view(&sy->pending_v);
    sy->nominated.view(&sy->nominated_v);
    std::memset(&sy->pgs_v, 0, sizeof(sy->pgs_v));
    sy->pgs_v.n_groups = G;
    sy->pgs_v.n_scalars = S;
    sy->pgs_v.exists = sy->pg_exists.data();
    sy->pgs_v.ns = sy->pg_ns.data();
    sy->pgs_v.min_member = sy->pg_min_member.data();
    sy->pgs_v.has_min_resources = sy->pg_has_min.data();
    sy->pgs_v.min_res = sy->pg_min_res.data();
    sy->pgs_v.min_res_present = sy->pg_min_present.data();
    sy->pgs_v.denied = sy->pg_denied.data();
    sy->pgs_v.permitted = sy->pg_permitted.data();
    std::memset(&sy->quotas_v, 0, sizeof(sy->quotas_v));
    sy->quotas_v.n_ns = NS;
    sy->quotas_v.n_scalars = S;
    sy->quotas_v.has_eq = sy->q_has.data();
    sy->quotas_v.min = sy->q_min.data();
    sy->quotas_v.max = sy->q_max.data();
    sy->quotas_v.used = sy->q_used.data();
    sy->quotas_v.used_scalar_present = sy->q_used_present.data();
    *out = sy;
    return FGS_OK;
}

extern "C" void fgs_synth_destroy(fgs_synth* s) { delete s; }
extern "C" const fgs_snapshot* fgs_synth_snapshot(const fgs_synth* s) { return &s->snap_v; }
extern "C" const fgs_podset* fgs_synth_pending(const fgs_synth* s) { return &s->pending_v; }
extern "C" const fgs_podset* fgs_synth_nominated(const fgs_synth* s) { return &s->nominated_v; }
extern "C" const fgs_podgroups* fgs_synth_podgroups(const fgs_synth* s) { return &s->pgs_v; }
extern "C" const fgs_quotas* fgs_synth_quotas(const fgs_synth* s) { return &s->quotas_v; }
