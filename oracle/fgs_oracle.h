/*
 * fgs_oracle.h — C interface of the CPU oracle (TEST INFRASTRUCTURE, not product code).
 *
 * The oracle is a restatement, on plain structs, of the reference's Go plugins for the hot
 * path named in BASELINE.json (see ref_walk.cpp for the file:line each function follows).
 * It walks the OBJECT-level snapshot (include/fgs_types.h) the way the Go code walks
 * NodeInfo/Pod objects — per (pod,node): re-sum the limits of every pod on the node, rebuild
 * the per-device table, then Filter/Score — so that it doubles as the timed CPU baseline.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * load this library. Nothing under flex-gpu-scheduler_b200/ links, imports or calls it.
 *
 * Parity pinning: the reference cannot be compiled in this image (no Go toolchain), so the
 * oracle is pinned against the golden vectors transcribed from the reference's own unit tests
 * (tests/golden/, see tests/test_oracle_golden.py):
 *   pkg/noderesources/allocatable_test.go:112-236, pkg/coscheduling/core/core_test.go:42-234,
 *   303-355, pkg/capacityscheduling/capacity_scheduling_test.go:52-164.
 * pkg/flexgpu has no tests in the reference: for FlexGPU, and for the vendored in-tree
 * NodeResourcesFit / DefaultNormalizeScore / selectHost, parity is UNPINNED by reference
 * vectors; hand-derived known-answer cases stand in (tests/test_oracle_flexgpu.py).
 */
#ifndef FGS_ORACLE_H
#define FGS_ORACLE_H

#include "../include/fgs_types.h"

#ifdef __cplusplus
extern "C" {
#endif

/* status byte per (pod,node): low nibble = framework.Code, high nibble = failing plugin id
 * (FGS_PL_*), 0xF0|code when the reference would panic inside Filter on that node. */
typedef struct fgs_oracle_out {
    uint8_t* status;        /* [P][N] or NULL */
    int64_t* raw_flex;      /* [P][N] raw Score per plugin, valid on feasible nodes; or NULL */
    int64_t* raw_fit;
    int64_t* raw_alloc;
    int64_t* norm_flex;     /* [P][N] after NormalizeScore, before weight; or NULL */
    int64_t* norm_fit;
    int64_t* norm_alloc;
    int64_t* total;         /* [P][N] Σ weight·normalised over score plugins; or NULL */
    int32_t* feasible_count;/* [P] */
    uint32_t* pod_err;      /* [P] FGS_PODERR_* */
    int64_t* max_total;     /* [P] selectHost's maxScore (INT64_MIN if no feasible node) */
    int32_t* tie_count;     /* [P] size of the max-score tie set */
    int32_t* first_max;     /* [P] lowest node index in the tie set, -1 if none */
} fgs_oracle_out;

/* Evaluate pending pods [p0,p1) against every node (percentageOfNodesToScore = 100).
 * Output rows are indexed by (p - p0). workers: goroutine-count analogue (reference default 16,
 * parallelize/parallelism.go:27); nodes are chunked with chunkSizeFor (:42-51).
 * Scores are computed on every feasible node even when only one node is feasible (the reference
 * skips scoring then, generic_scheduler.go:120-127); status / counts are unaffected by that. */
int fgs_oracle_evaluate(const fgs_snapshot* snap, const fgs_podset* pending,
                        const fgs_plugin_cfg* cfg, int64_t p0, int64_t p1, int workers,
                        fgs_oracle_out* out);

/* The reason inside each plugin's Filter status of pod p, per node: flex[n] = FGS_FLEX_REASON_* (the message
 * FlexGPU.Filter returns, flex_gpu.go:60-114), fit[n] = FGS_FIT_REASON_* bits (fitsRequest's InsufficientResource
 * list, fit.go:256-329). Either may be NULL. */
int fgs_oracle_filter_reasons(const fgs_snapshot* snap, const fgs_podset* pending, const fgs_plugin_cfg* cfg,
                              int64_t p, uint16_t* flex, uint16_t* fit);

/* fgs_oracle_evaluate with nominated pods: every node is filtered as RunFilterPluginsWithNominatedPods does
 * (runtime/framework.go:752-823) — first with the nominated pods of equal or higher priority added to the node
 * (addNominatedPods), then, if that passed, once more without them. nom_node[q] = snapshot index of pod q's
 * NominatedNodeName (< 0: none); nom_uid / pending_uid (optional) keep a pod from yielding to itself (:812). */
int fgs_oracle_evaluate_nominated(const fgs_snapshot* snap, const fgs_podset* pending, const fgs_plugin_cfg* cfg,
                                  int64_t p0, int64_t p1, int workers, const fgs_podset* nominated,
                                  const int32_t* nom_node, const int64_t* nom_uid, const int64_t* pending_uid,
                                  fgs_oracle_out* out);

/* fgs_oracle_evaluate when the profile has filter plugins OUTSIDE this path (the in-tree defaults the reference's chart
 * keeps enabled): ext_mask[(p - p0) * N + n] == 0 means one of them rejected node n for pod p, so the node is missing
 * from the feasible list that prioritizeNodes / RunScorePlugins are handed (generic_scheduler.go:420-426,
 * runtime/framework.go:861-933) and NormalizeScore / selectHost run over the rest. Such a node's status reads 0xE2. */
int fgs_oracle_evaluate_masked(const fgs_snapshot* snap, const fgs_podset* pending, const fgs_plugin_cfg* cfg,
                               int64_t p0, int64_t p1, int workers, const uint8_t* ext_mask, fgs_oracle_out* out);

/* numFeasibleNodesToFind (generic_scheduler.go:171-191): percentage 0 = adaptive, >= 100 = all nodes. */
int32_t fgs_oracle_num_feasible_nodes_to_find(int32_t numAllNodes, int32_t percentageOfNodesToScore);

/* fgs_oracle_evaluate with the reference's node sampling as it runs with ONE worker (findNodesThatPassFilters,
 * generic_scheduler.go:266-338): pod p visits nodes[(start+i) % N], i = 0,1,…, keeps the first
 * K = numFeasibleNodesToFind(N) feasible ones and stops at the node that would be the (K+1)-th; then
 * start = (start + processed) % N with processed = kept + failed (:329-330). *start_index carries
 * g.nextStartNodeIndex in and out; processed[p - p0] may be NULL. status of a node the walk did not reach (or
 * reached last and dropped) is 0xEE. Scores / NormalizeScore / selectHost run on the kept nodes only. */
int fgs_oracle_evaluate_sampled(const fgs_snapshot* snap, const fgs_podset* pending,
                                const fgs_plugin_cfg* cfg, int64_t p0, int64_t p1, int percentage,
                                int64_t* start_index, int32_t* processed, fgs_oracle_out* out);

/* FlexGPU.Reserve's device choice (flex_gpu.go:178-223) for pod p on node n:
 * returns the index written into the annotation, -1 when no index fits (Unschedulable),
 * -2 when the pod has neither limit (nothing to reserve), -3 both limits, -4 reference panic. */
int fgs_oracle_reserve_index(const fgs_snapshot* snap, const fgs_podset* pending, int64_t p,
                             int64_t n);

/* Coscheduling PodGroupManager.PreFilter (core/core.go:149-196) per pending pod.
 * code[p]: 0 pass; 1 denied-cache; 2 not enough sibling pods; 3 resource gap. */
int fgs_oracle_gang_prefilter(const fgs_snapshot* snap, const fgs_podset* pending,
                              const fgs_podgroups* pgs, int64_t p0, int64_t p1, uint8_t* code);

/* Same, with `unlisted[q] != 0` marking pending pods the pod lister does not know (the reference's
 * unit test evaluates a pod it never added to the informer store, core_test.go:159-164). */
int fgs_oracle_gang_prefilter_listed(const fgs_snapshot* snap, const fgs_podset* pending,
                                     const fgs_podgroups* pgs, int64_t p0, int64_t p1, uint8_t* code,
                                     const uint8_t* unlisted);

/* CheckClusterResource on an explicit request map: request[4+S] = {mcpu, mem, eph, pods, scalars},
 * bit k of `keys` = column k is a key of the map; desired_pg = group id whose pods are removed. */
int fgs_oracle_check_cluster_resource_raw(const fgs_snapshot* snap, const int64_t* request, uint32_t keys,
                                          int32_t desired_pg);

/* CheckClusterResource alone (core/core.go:322-342) for group g: 1 enough, 0 gap. */
int fgs_oracle_check_cluster_resource(const fgs_snapshot* snap, const fgs_podgroups* pgs,
                                      int64_t g);

/* PodGroupManager.Permit (core/core.go:199-216): status[p] = 0 Success, 1 Wait,
 * 2 PodGroupNotSpecified, 3 PodGroupNotFound; assigned[g] = CalculateAssignedPods (:301-318). */
int fgs_oracle_gang_permit(const fgs_snapshot* snap, const fgs_podset* pending,
                           const fgs_podgroups* pgs, int64_t p0, int64_t p1, uint8_t* status,
                           int32_t* assigned);

/* CapacityScheduling.PreFilter (capacity_scheduling.go:201-275) per pending pod.
 * `nominated` = pods nominated to nodes (may be NULL); nominated_uid / pending_uid let the walk
 * skip the pod itself (:237). code[p]: 0 Success, 1 rejected: over Max, 2 rejected: aggregated
 * used over Σmin. */
int fgs_oracle_quota_prefilter(const fgs_podset* pending, const fgs_quotas* q,
                               const fgs_podset* nominated, const int64_t* pending_uid,
                               const int64_t* nominated_uid, int64_t p0, int64_t p1,
                               uint8_t* code);

/* ElasticQuotaInfo.reserveResource (op = +1) / unreserveResource (op = -1) with pod p's request
 * (elasticquota.go:74-88,126-147) on one namespace's Used row {mcpu, mem, scalars} and its scalar keys. */
int fgs_oracle_quota_reserve(const fgs_podset* pods, int64_t p, int op, int64_t* used, uint32_t* used_present);

/* preemptor.SelectVictimsOnNode (capacity_scheduling.go:465-644) for pending pod p on EVERY node, as
 * DryRunPreemption runs it (per-node clones; no PodDisruptionBudgets; empty PodNominator). q may be NULL (no
 * ElasticQuota anywhere). tracks_pods: 1 = assigned pods are members of their quota's pod set, so removing / re-adding
 * one moves its request out of / into Used (production); 0 = they are not (the reference's unit-test fixtures).
 * node_status[n] = FGS_PREEMPT_*; victim[a] = 1 for rows of snap->pods chosen as victims on a CANDIDATE node;
 * n_victims[n] (may be NULL). */
int fgs_oracle_select_victims(const fgs_snapshot* snap, const fgs_podset* pending, int64_t p, const fgs_plugin_cfg* cfg,
                              const fgs_quotas* q, int tracks_pods, uint8_t* node_status, uint8_t* victim,
                              int32_t* n_victims);

/* pickOneNodeForPreemption (vendor/k8s.io/kubernetes/pkg/scheduler/framework/preemption/preemption.go:386-505) over
 * the candidates a dry run produced (node_status / victim as fgs_oracle_select_victims wrote them; no
 * PodDisruptionBudgets, so NumPDBViolations is 0 everywhere). The reference's last tie-break is the iteration order of
 * a Go map: tie_set[n] = 1 for EVERY node it may return (the nodes still level after criterion 5). Returns their
 * number (0: no candidate). */
int64_t fgs_oracle_pick_preemption_node(const fgs_snapshot* snap, const uint8_t* node_status, const uint8_t* victim,
                                        uint8_t* tie_set);

#ifdef __cplusplus
}
#endif
#endif
