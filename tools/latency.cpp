// latency.cpp — what one scheduling cycle costs through the C ABI, without an interpreter in the way: fgs_evaluate for
// P = 1 … 256 (one launch) and 257, 4096 (general path) pods on the 5 000-node C2 cluster, host pod table in, per-pod
// results out. Build: g++ -O2 -std=c++17 -I include tools/latency.cpp -L flex-gpu-scheduler_b200 -lfgs -Wl,-rpath,$PWD/flex-gpu-scheduler_b200
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "fgs.h"

int main(int argc, char** argv) {
    const int64_t n_nodes = argc > 1 ? atoll(argv[1]) : 5000;
    fgs_synth_params sp;
    std::memset(&sp, 0, sizeof(sp));
    sp.seed_nodes = 0xF6500002, sp.seed_pods = 0xF6510002;
    sp.n_nodes = n_nodes, sp.n_pending = 4096, sp.n_scalars = 2, sp.n_devices = 8, sp.max_existing_pods = 24;
    sp.pm_fractional = 600, sp.pm_whole = 300, sp.pm_nogpu = 50, sp.pm_degenerate = 50, sp.pm_missing_req = 100, sp.pm_init_ctr = 100;
    sp.pm_overhead = 30, sp.pm_two_ctrs = 200, sp.pm_quirk_nodes = 20;
    fgs_synth* sy = nullptr;
    if (fgs_synth_create(&sp, &sy)) return 1;
    fgs_config c;
    std::memset(&c, 0, sizeof(c));
    c.n_scalars = 2, c.total_nodes = n_nodes, c.select_seed = 1;
    fgs_engine* e = nullptr;
    if (fgs_create(&c, &e)) {
        std::fprintf(stderr, "%s\n", fgs_last_error(nullptr));
        return 1;
    }
    fgs_packer *pn = nullptr, *pp = nullptr;
    fgs_packer_create(2, &pn);
    fgs_packer_create(2, &pp);
    const fgs_node_table* nt = nullptr;
    fgs_packer_nodes(pn, fgs_synth_snapshot(sy), 0, n_nodes, &nt);
    if (fgs_upload_nodes(e, nt)) {
        std::fprintf(stderr, "%s\n", fgs_last_error(e));
        return 1;
    }
    std::printf("{\"nodes\": %lld", (long long)n_nodes);
    for (int P : {1, 2, 16, 64, 256, 257, 4096}) {
        const fgs_pod_table* t = nullptr;
        fgs_packer_pods(pp, fgs_synth_pending(sy), 0, P, &t);
        std::vector<int32_t> cnt(P), sel(P);
        std::vector<int64_t> sc(P);
        std::vector<uint32_t> err(P);
        fgs_results r;
        std::memset(&r, 0, sizeof(r));
        r.feasible_count = cnt.data(), r.selected_node = sel.data(), r.selected_score = sc.data(), r.pod_err = err.data();
        for (int k = 0; k < 20; ++k)
            if (fgs_evaluate(e, t, nullptr, &r)) {
                std::fprintf(stderr, "%s\n", fgs_last_error(e));
                return 1;
            }
        const int reps = 2000;
        const auto t0 = std::chrono::steady_clock::now();
        for (int k = 0; k < reps; ++k) fgs_evaluate(e, t, nullptr, &r);
        const double us = std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now() - t0).count() / reps;
        fgs_timings tm;
        fgs_get_timings(e, &tm);
        std::printf(", \"evaluate_%d_pods_us\": %.1f, \"one_launch_%d\": %d", P, us, P, tm.small);
    }
    // The reference's real loop: between two cycles a NodeInfo changed (the pod just bound; cache.go:197-276 bumps its
    // Generation). fgs_sync_nodes repacks and patches that row, then the next pod is evaluated — the level tables are
    // stale every time, so the one-launch kernel filters by comparing.
    {
        fgs_snapshot snap = *fgs_synth_snapshot(sy);
        std::vector<int64_t> gen(n_nodes);
        for (int64_t i = 0; i < n_nodes; ++i) gen[i] = i + 1;
        snap.generation = gen.data();
        int64_t g = n_nodes + 1, nrep = 0;
        if (fgs_sync_nodes(e, pn, &snap, 0, n_nodes, &nrep)) {
            std::fprintf(stderr, "%s\n", fgs_last_error(e));
            return 1;
        }
        std::vector<const fgs_pod_table*> one(64);
        std::vector<fgs_packer*> pks(64);
        for (int k = 0; k < 64; ++k) {
            fgs_packer_create(2, &pks[k]);
            fgs_packer_pods(pks[k], fgs_synth_pending(sy), k, k + 1, &one[k]);
        }
        int32_t cnt = 0, sel = 0;
        int64_t sc = 0;
        uint32_t err = 0;
        fgs_results r;
        std::memset(&r, 0, sizeof(r));
        r.feasible_count = &cnt, r.selected_node = &sel, r.selected_score = &sc, r.pod_err = &err;
        const int reps = 2000;
        double sync_us = 0, eval_us = 0;
        int direct = 0;
        for (int k = -50; k < reps; ++k) {
            const auto t0 = std::chrono::steady_clock::now();
            gen[(int64_t)(k + 50) * 37 % n_nodes] = g++;
            if (fgs_sync_nodes(e, pn, &snap, 0, n_nodes, &nrep) || nrep != 1) {
                std::fprintf(stderr, "sync: %s (%lld rows)\n", fgs_last_error(e), (long long)nrep);
                return 1;
            }
            const auto t1 = std::chrono::steady_clock::now();
            if (fgs_evaluate(e, one[(k + 50) & 63], nullptr, &r)) {
                std::fprintf(stderr, "%s\n", fgs_last_error(e));
                return 1;
            }
            const auto t2 = std::chrono::steady_clock::now();
            if (k >= 0) {
                sync_us += std::chrono::duration<double, std::micro>(t1 - t0).count();
                eval_us += std::chrono::duration<double, std::micro>(t2 - t1).count();
                fgs_timings tm;
                fgs_get_timings(e, &tm);
                direct += tm.small && !tm.reduce_lookup;
            }
        }
        std::printf(", \"cycle_sync_one_row_us\": %.1f, \"cycle_evaluate_1_pod_us\": %.1f, \"cycle_total_us\": %.1f, \"cycles_without_tables\": %d, \"cycles\": %d",
                    sync_us / reps, eval_us / reps, (sync_us + eval_us) / reps, direct, reps);
    }
    std::printf("}\n");
    fgs_destroy(e);
    return 0;
}
