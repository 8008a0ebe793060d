/*
 * fgs.h — C ABI of libfgs, the B200-native batched Filter+Score engine.
 *
 * libfgs replaces ONE path of the reference scheduler (WLBF/flex-gpu-scheduler, a fork of
 * kubernetes-sigs/scheduler-plugins): the per-cycle evaluation of pending pods against every
 * NodeInfo of the snapshot — Filter, Score, NormalizeScore, weighting and selectHost for the
 * FlexGPU, NodeResourcesFit (Least/MostAllocated) and NodeResourcesAllocatable plugins, plus the
 * PreFilter arithmetic of Coscheduling (PodGroup minMember / MinResources) and CapacityScheduling
 * (ElasticQuota). The Go plugins keep their framework interfaces; a PreFilter/PreScore stage
 * fills the flat tables of fgs_types.h, calls these entry points through cgo and
 * answers Filter/Score from the returned bitmap / score matrix.
 *
 * The reference has no FFI of its own (it is pure Go, CGO_ENABLED=0, Makefile:17); each entry
 * point therefore cites the Go function(s) whose work it takes over. Paths are relative to the
 * reference repo; "sched/" abbreviates vendor/k8s.io/kubernetes/pkg/scheduler/.
 *
 * Conventions: every function returns 0 on success or a negative FGS_E_* code;
 * fgs_last_error(engine) gives the message. All pointers are plain host pointers unless the
 * name says `dev`. No function falls back to a CPU path: without a usable CUDA device
 * fgs_create fails with FGS_E_CUDA.
 *
 * Threading: one caller at a time per engine (the scheduling goroutine); result buffers handed
 * back by fgs_evaluate are read-only afterwards and may be read concurrently (the 16 Filter /
 * Score goroutines of sched/framework/parallelize/parallelism.go:27).
 */
#ifndef FGS_H
#define FGS_H

#include "fgs_types.h"

#ifdef __cplusplus
extern "C" {
#endif

#define FGS_OK 0
#define FGS_E_INVALID (-1)   /* bad argument / table failed validation */
#define FGS_E_CUDA (-2)      /* CUDA runtime error (message has the cudaError string) */
#define FGS_E_RANGE (-3)     /* a quantity is outside the supported int64 range (|v| >= 2^53) */
#define FGS_E_STATE (-4)     /* call sequence error (e.g. evaluate before upload) */
#define FGS_E_TOO_LARGE (-5) /* exceeds a stated fixed maximum */

typedef struct fgs_engine fgs_engine;

typedef struct fgs_config {
    int32_t device;            /* CUDA device ordinal this engine owns */
    int32_t n_scalars;         /* S, number of scalar resource slots (2..FGS_MAX_SCALARS) */
    int64_t node_index_base;   /* global index of this shard's first node (multi-GPU sharding) */
    int64_t total_nodes;       /* nodes over all shards (== shard size on one GPU) */
    uint64_t select_seed;      /* seed of the selectHost tie-break hash */
    void* stream;              /* cudaStream_t to run on, NULL = a stream the engine creates */
} fgs_config;

/* Per-pod outputs of one evaluation (host buffers owned by the caller, each [P]). */
typedef struct fgs_results {
    int32_t* feasible_count;   /* len(feasibleNodes) — sched/generic_scheduler.go:106-127 */
    int32_t* selected_node;    /* global node index chosen by selectHost, -1 when none feasible */
    int64_t* selected_score;   /* its total score */
    uint32_t* pod_err;         /* FGS_PODERR_* */
    /* optional full planes; NULL = leave on the device */
    uint32_t* feasible_bitmap; /* [P][words_per_row]; bit n%32 of word n/32 = node n feasible */
    int64_t* scores;           /* [P][score_row_stride] total weighted score per node */
    int64_t words_per_row;     /* filled in by the engine */
    int64_t score_row_stride;  /* filled in by the engine */
    /* node sampling (fgs_eval_opts.sample): g.nextStartNodeIndex after the last pod, and per pod the nodes its
     * walk counted (optional, [P]); without sampling next_start_node_index is -1 */
    int64_t next_start_node_index;
    int32_t* processed_nodes;
} fgs_results;

typedef struct fgs_eval_opts {
    uint32_t emit_mask;   /* score plugins summed into `scores` (0 = all of score_mask); lets a
                             caller read ONE plugin's normalised scores (the Go Score/NormalizeScore
                             answers) */
    uint32_t emit_raw;    /* 1: emit the plugin's raw Score instead of the normalised×weight value */
    uint32_t skip_pass_b; /* 1: Filter only (bitmap + counts come from the reduce pass) */
    /* 1: the reference's node sampling (findNodesThatPassFilters, sched/generic_scheduler.go:266-338, as ONE
     * worker runs it — with 16 workers which nodes are found is a race, :302-308): pod p walks
     * nodes[(start + i) % N], keeps the first K = numFeasibleNodesToFind(N) feasible nodes (:171-191), stops at
     * the node that would be the (K+1)-th and moves `start` on by the nodes it counted (:329-330); the next pod
     * of the batch starts there. Bitmap, counts, NormalizeScore and selectHost then cover the kept nodes only.
     * 0 (default): every node is evaluated for every pod (percentageOfNodesToScore = 100). Single engine only. */
    uint32_t sample;
    int32_t percentage_of_nodes_to_score; /* with sample: KubeSchedulerConfiguration's field; 0 = adaptive */
    int32_t _pad;
    int64_t start_node_index;             /* with sample: g.nextStartNodeIndex before the first pod */
} fgs_eval_opts;

/* device-side buffers of the current batch, for callers that keep results in HBM or that run
 * the cross-GPU merges themselves (torch.distributed / NCCL on these pointers) */
typedef struct fgs_device_view {
    void* reduce_max;      /* int64 [FGS_REDUCE_ROWS][P_pad]: merge across shards with MAX */
    void* reduce_sum;      /* int64 [P_pad] feasible counts: merge with SUM */
    void* select_key;      /* int64 [3][P_pad]: row 0 packed (score,hash,node) keys, row 1 score-range
                              error flag, row 2 score-time panic flag: merge with MAX */
    void* feasible_bitmap; /* uint32 [P][words_per_row] */
    void* scores;          /* int64 [P][score_row_stride] */
    int64_t p_pad;
    int64_t words_per_row;
    int64_t score_row_stride;
    void* stream;          /* the cudaStream_t all engine work is ordered on */
} fgs_device_view;

#define FGS_REDUCE_ROWS 6 /* flex max, alloc hi, -alloc lo, filter-panic flag, (2 spare) */

/* ---- lifecycle ------------------------------------------------------------------------- */

/* Replaces plugin construction: flexgpu.New (pkg/flexgpu/flex_gpu.go:37-39), NewAllocatable
 * (pkg/noderesources/allocatable.go:79-118), NewFit (sched/framework/plugins/noderesources/
 * fit.go:104-130) — the engine is the shared state those plugins point at. */
int fgs_create(const fgs_config* cfg, fgs_engine** out);
void fgs_destroy(fgs_engine* e);
const char* fgs_last_error(const fgs_engine* e); /* e may be NULL: last create error */

/* Plugin args → engine configuration. Mirrors NodeResourcesAllocatableArgs / NodeResourcesFitArgs
 * decoding and validation (pkg/noderesources/allocatable.go:51-58,84-108; weights per
 * sched/framework/runtime/framework.go:395-415). */
int fgs_set_plugins(fgs_engine* e, const fgs_plugin_cfg* cfg);

/* ---- packing (host only: no CUDA call, usable before any engine exists) -------------------- */

typedef struct fgs_packer fgs_packer; /* owns the tables it returns */

int fgs_packer_create(int32_t n_scalars, fgs_packer** out);
void fgs_packer_destroy(fgs_packer* p);
const char* fgs_packer_error(const fgs_packer* p);

/* Object snapshot → SoA node table for nodes [n0, n1) of the snapshot (a shard; 0..n_nodes = all).
 * Takes over the per-call object walks of nodeResourceLimitSum (pkg/flexgpu/flex_gpu.go:133-140)
 * and NewGPUNode/constructGPUs (pkg/flexgpu/gpu_node.go:30-120), and reads the NodeInfo aggregates
 * (sched/framework/types.go:359-419). Valid until the next fgs_packer_nodes on this packer. */
int fgs_packer_nodes(fgs_packer* p, const fgs_snapshot* snap, int64_t n0, int64_t n1,
                     const fgs_node_table** out);

/* The same for an arbitrary list of snapshot rows: table row i = snapshot row rows[i] (what a generation diff found
 * stale; the table then goes to fgs_update_nodes with `rows` as the dirty indices — fgs_sync_nodes does both). */
int fgs_packer_node_rows(fgs_packer* p, const fgs_snapshot* snap, const int64_t* rows, int64_t n_rows,
                         const fgs_node_table** out);

/* Pods [p0, p1) of the podset → SoA pod table (a batch; 0..n_pods = all). Takes over podResourceLimit (flex_gpu.go:121-131),
 * computePodResourceRequest (fit.go:162-180) and calculatePodResourceRequest
 * (sched/framework/plugins/noderesources/resource_allocation.go:119-143).
 * Valid until the next fgs_packer_pods on this packer. */
int fgs_packer_pods(fgs_packer* p, const fgs_podset* pods, int64_t p0, int64_t p1,
                    const fgs_pod_table** out);

/* Page-locks the packer's pod-table columns (cudaHostRegister, now and after every later fgs_packer_pods), so that the
 * engine can DMA them to the GPU without a staging copy (fgs_pod_stats). The one packer call that needs a CUDA device.
 * A pinned table that was handed to fgs_submit must not be packed over before that batch's fgs_collect. */
int fgs_packer_pin(fgs_packer* p);

/* ---- upload + evaluate ----------------------------------------------------------------- */

/* H2D of the node table and derivation of the pod-independent columns on the GPU
 * (memEach, per-device max free memory, GPUScore/MemScore — gpu_node.go:55,179-199 —
 * Allocatable raw score — allocatable.go:120-143). Replaces Cache.UpdateSnapshot's role as the
 * per-cycle input (sched/internal/cache/cache.go:197-276). */
int fgs_upload_nodes(fgs_engine* e, const fgs_node_table* t);

/* Same for a subset of rows: `rows` has n_dirty rows, row r replaces node dirty_idx[r]. Up to 4 096 rows are stored,
 * re-derived and checked by one kernel and the call returns without waiting for it (the next evaluation is ordered behind it
 * on the engine's stream). A batch staged before the call no longer matches the node table: fgs_pass_reduce / fgs_pass_emit /
 * fgs_score_masked / fgs_resubmit on it return FGS_E_STATE until it is staged again. */
int fgs_update_nodes(fgs_engine* e, const fgs_node_table* rows, const int64_t* dirty_idx);

/* The incremental form of the two calls above, keyed by NodeInfo.Generation as the scheduler cache's own snapshot update
 * is (sched/internal/cache/cache.go:197-276: UpdateSnapshot walks the NodeInfos whose Generation is newer than the
 * snapshot's and re-clones only those): the engine remembers the generation of every row of its shard [n0, n1) as of
 * the last call; rows whose snap->generation differs are packed (fgs_packer_node_rows) and patched (fgs_update_nodes),
 * the rest stays on the GPU. The first call, a shard of another size, a stale row with more device slots than the
 * uploaded table, or more than half of the rows stale fall back to a full pack + fgs_upload_nodes. Needs
 * snap->generation. n_repacked (optional) <- rows packed and copied by this call. */
int fgs_sync_nodes(fgs_engine* e, fgs_packer* p, const fgs_snapshot* snap, int64_t n0, int64_t n1, int64_t* n_repacked);

/* One batched scheduling evaluation = for every pod: findNodesThatPassFilters over all nodes
 * (sched/generic_scheduler.go:266-338 with percentageOfNodesToScore=100; FlexGPU.Filter
 * flex_gpu.go:41-119; Fit.Filter fit.go:221-329), RunScorePlugins (runtime/framework.go:861-933;
 * FlexGPU.Score/NormalizeScore flex_gpu.go:142-176; Fit.Score fit.go:332-339 with
 * least_allocated.go:29-57 / most_allocated.go:29-61; Allocatable.Score/NormalizeScore
 * allocatable.go:61-72,146-171), the per-node sum (generic_scheduler.go:440-447) and selectHost
 * (:146-167). Host pod table in, per-pod results out (H2D + kernels + D2H). */
int fgs_evaluate(fgs_engine* e, const fgs_pod_table* pods, const fgs_eval_opts* opts,
                 fgs_results* out);
/* The reference's own call shape is ONE pod per cycle (sched/scheduler.go:425-455). For up to 256 pods, default options, a
 * single engine and values in the int32 range, fgs_evaluate runs as a single launch — one CTA per pod reads the pod's
 * columns from mapped host memory, ANDs the level rows, reduces, scores every node and writes the result record back to
 * mapped host memory — and leaves the same device state behind as the general path (planes, per-pod rows, staged batch),
 * so that fgs_status_plane, fgs_pass_emit, fgs_score_masked … work after it. fgs_timings.small tells. */

/* Score + NormalizeScore + selectHost under the framework's OWN feasible list — the PreScore / Score half of the drop-in
 * when the profile keeps filter plugins outside this path (the reference's chart only ADDS FlexGPU to the in-tree
 * defaults, manifests/flexgpu/templates/configmap.yaml:12-28). prioritizeNodes hands RunScorePlugins the nodes that
 * passed EVERY filter plugin (sched/generic_scheduler.go:420-426), and each NormalizeScore runs over that list
 * (sched/framework/runtime/framework.go:861-933: DefaultNormalizeScore's maxCount, Allocatable's hi / lo) — a shorter list
 * than the engine's own as soon as another plugin rejected a node. Sequence: fgs_evaluate with skip_pass_b = 1 (or
 * fgs_stage_pods + fgs_pass_reduce) answers Filter; the Go side runs its other filters; at PreScore it calls this with
 *   mask[P][words_per_row]   bit n%32 of word n/32 of row p = node n is in pod p's feasible list (host memory, or
 *                            device memory with mask_on_device = 1; on a node-sharded engine: this shard's nodes)
 * The engine ANDs its bitmap with the mask (a node its own filters rejected stays out), recounts, redoes the
 * NormalizeScore reductions over what is left, merges across shards when a peer group is attached (every rank calls),
 * scores, selects and fills `out` as fgs_evaluate does; the planes then describe the masked list. Calling it again for
 * the same batch can only shrink the list further. */
int fgs_score_masked(fgs_engine* e, const uint32_t* mask, int32_t mask_on_device, const fgs_eval_opts* opts, fgs_results* out);

/* The PodNominator's content: pods an earlier preemption nominated to nodes (Status.NominatedNodeName).
 * While it is set, every evaluation filters a node that has such pods the way RunFilterPluginsWithNominatedPods
 * does (sched/framework/runtime/framework.go:752-823): once with the nominated pods whose priority is >= the
 * evaluated pod's added to the node (addNominatedPods: Requested, pod count, flex limit sums, devices by index
 * annotation), and — only if that passed — once without them; feasible = both pass. Scores are not affected
 * (Score plugins read the snapshot's NodeInfo). node[q] = snapshot index of pod q's nominated node (global index;
 * < 0 or outside this engine's shard: ignored); uid (optional) with fgs_pod_table.uid of the evaluated batch keeps
 * a pod from yielding to itself. pods == NULL or empty clears. The tables are copied. */
int fgs_set_nominated(fgs_engine* e, const fgs_pod_table* pods, const int32_t* node, const int64_t* uid);

/* fgs_evaluate for a STREAM of batches, two in flight: fgs_submit stages the batch (into the buffers the
 * previous-but-one batch has released), copies it on a copy stream and enqueues both passes and the D2H of the
 * per-pod results, then returns; fgs_collect blocks until the OLDEST submitted batch is done and unpacks its
 * per-pod results. While the kernels of one batch run, the host stages the next and unpacks the previous one.
 * At most two batches may be in flight (a third fgs_submit before fgs_collect → FGS_E_STATE). The bitmap / score
 * planes are shared by consecutive batches, so `out` receives per-pod results only; fgs_evaluate, the staged
 * calls and the plane getters are refused while batches are in flight. No sampling / skip_pass_b here. */
int fgs_submit(fgs_engine* e, const fgs_pod_table* pods, const fgs_eval_opts* opts);
int fgs_collect(fgs_engine* e, fgs_results* out);
/* Node-sharded engines (fgs_peer_attach): fgs_evaluate, fgs_submit and fgs_collect include the two cross-shard
 * merges; every rank must make the same sequence of calls. In the submit / collect pipeline the merges run on their
 * own stream and overlap a pass of the OTHER batch in flight: the emit pass of a batch is issued by the next
 * fgs_submit, or by fgs_collect when nothing follows. */
/* fgs_resubmit evaluates again the batch that an earlier fgs_submit left resident in the next pipeline slot (no staging,
 * no H2D): steady-state measurements with the tables already in HBM, on single and on node-sharded engines alike. The
 * slots are walked round-robin. By default there are as many slots as batches in flight (2, or 3 with a peer group);
 * fgs_set_resident_batches(e, n), 3 <= n <= 32, makes the ring n slots long, so that a job of n batches — BASELINE's
 * config 5: 1 000 000 pods in 65 536-pod batches — stays resident and n fgs_resubmit calls walk it once; a longer ring
 * also admits ONE more batch in flight (3, or 4 with a peer group), so that fgs_submit issues a batch's H2D copy a whole
 * pass earlier; n = 0 restores the default. A batch staged before the node table last changed is
 * refused (FGS_E_STATE). */
int fgs_resubmit(fgs_engine* e, const fgs_eval_opts* opts);
int fgs_set_resident_batches(fgs_engine* e, int32_t n);
int fgs_max_in_flight(fgs_engine* e); /* batches fgs_submit / fgs_resubmit accept before an fgs_collect is due (2..4) */

/* The same evaluation in stages, for node-sharded multi-GPU runs (one engine per GPU/process):
 *   fgs_stage_pods → fgs_pass_reduce → [merge reduce_max / reduce_sum across shards]
 *   → fgs_pass_emit → [merge select_key across shards] → fgs_fetch_results.
 * All stages are asynchronous on the engine stream except fgs_fetch_results. */
int fgs_stage_pods(fgs_engine* e, const fgs_pod_table* pods);
/* Filter AND Score in one pass over pod × node (opt-in, FGS_FUSED=1 in the environment when the engine is created:
 * fgs_evaluate / fgs_submit then take it by themselves when they can; measured slower than the two passes on config 2,
 * DESIGN.md §3): a
 * per-pod prepass finds DefaultNormalizeScore's maxCount — the raw FlexGPU score depends on the node and the pod's kind
 * only, so it is the score of the first feasible node in score order — then one kernel runs the filters, writes the
 * bitmap and the counts, scores, normalises, sums and keeps the selectHost key. Needs: one engine holding the whole
 * cluster (nothing to merge between Filter and Score), no sampling, no nominated pods, default outputs, no
 * NodeResourcesAllocatable among the score plugins, the int32 specialisation. fgs_fused_eligible: 1 / 0 for the staged
 * batch; fgs_pass_fused replaces fgs_pass_reduce + fgs_pass_emit (FGS_E_STATE when not eligible). */
int fgs_fused_eligible(fgs_engine* e, const fgs_eval_opts* opts);
int fgs_pass_fused(fgs_engine* e, const fgs_eval_opts* opts);
/* the same in its two steps (for callers that time them apart): per-pod prepass, then the pass over pod × node */
int fgs_pass_fused_prepass(fgs_engine* e, const fgs_eval_opts* opts);
int fgs_pass_fused_main(fgs_engine* e, const fgs_eval_opts* opts);
int fgs_pass_reduce(fgs_engine* e);
int fgs_pass_emit(fgs_engine* e, const fgs_eval_opts* opts);
int fgs_fetch_results(fgs_engine* e, fgs_results* out);
int fgs_get_device_view(fgs_engine* e, fgs_device_view* v);

/* ---- the cross-shard merges over NVLink peer memory --------------------------------------- */

/* The two merges of the sharded sequence above as one kernel each, over exchange buffers every rank maps
 * (CUDA IPC / fabric handles / torch symmetric memory — the caller's plumbing) instead of NCCL all-reduces:
 * DefaultNormalizeScore's maxCount and Allocatable's hi / lo are maxima over ALL nodes (vendor/…/plugins/helper/
 * normalize_score.go:27-35, pkg/noderesources/allocatable.go:121-153), len(feasibleNodes) a sum, selectHost a
 * maximum of keys (schedule_one.go:783-806). With `multicast` set the NVSwitch reduces and replicates
 * (multimem.ld_reduce / multimem.st); without, slices are pushed to their owner and the results pushed back.
 * Every rank must call the same merges in the same order; a peer that does not show up within 10 s raises an
 * error (fgs_peer_status) instead of hanging the GPU. */
#define FGS_PEER_MAX_RANKS 8
typedef struct fgs_peer_group {
    int32_t rank, world;               /* 2 ≤ world ≤ FGS_PEER_MAX_RANKS */
    void* buffers[FGS_PEER_MAX_RANKS]; /* every rank's exchange buffer as mapped in THIS process; [rank] = own */
    void* multicast;                   /* multicast mapping of the same buffers, or NULL */
    uint64_t bytes;                    /* size of each buffer, ≥ fgs_peer_bytes(max pods per batch) */
} fgs_peer_group;
uint64_t fgs_peer_bytes(int64_t max_pods);
/* Zeroes the own buffer's header; synchronise the ranks (a host barrier) between attach and the first merge.
 * g == NULL detaches. */
int fgs_peer_attach(fgs_engine* e, const fgs_peer_group* g);
int fgs_peer_merge_reduce(fgs_engine* e); /* after fgs_pass_reduce: reduce_max (MAX) and reduce_sum (SUM) */
int fgs_peer_merge_select(fgs_engine* e); /* after fgs_pass_emit: select_key (MAX) */
/* Waits for the engine stream; FGS_E_STATE if a merge timed out. uses_multicast (optional) ← 1 / 0. */
int fgs_peer_status(fgs_engine* e, int32_t* uses_multicast);
int fgs_synchronize(fgs_engine* e);

/* Filter verdicts of the last staged batch as framework.Code bytes, out[P][n_nodes]:
 *   plugin == FGS_PL_FLEXGPU  what FlexGPU.Filter returns (flex_gpu.go:41-119): 0 Success,
 *                             2 Unschedulable, 3 UnschedulableAndUnresolvable
 *   plugin == FGS_PL_FIT      what Fit.Filter returns (fit.go:221-238)
 *   plugin == 0               RunFilterPlugins' merged result (runtime/framework.go:667-693):
 *                             low nibble = code, high nibble = id of the plugin that failed
 * 0xF1 marks a (pod,node) on which the reference would panic inside NewGPUNode. This is the
 * diagnosis plane the Go Filter methods answer from (NodeToStatusMap, generic_scheduler.go:309-313). */
int fgs_status_plane(fgs_engine* e, uint32_t plugin, uint8_t* out);

/* The REASON behind each non-success verdict of the last staged batch, out[P][n_nodes] — what the reference
 * writes into the Status message that ends up in the pod's FitError / events (sched/framework/types.go:221-250):
 *   plugin == FGS_PL_FLEXGPU  one of FGS_FLEX_REASON_* (FlexGPU.Filter returns at its first failing step)
 *   plugin == FGS_PL_FIT      OR of FGS_FIT_REASON_* (fitsRequest reports every insufficient resource) */
int fgs_reason_plane(fgs_engine* e, uint32_t plugin, uint16_t* out);

/* FitError.Error() (sched/framework/types.go:221-250) without moving the planes to the host: for every pod of the
 * last evaluated batch that NO node took, counts[p][FGS_DIAG_*] = how many nodes carry each reason in the pod's
 * NodeToStatusMap — a node's status being that of the first filter plugin (profile order) that rejected it
 * (runtime/framework.go:667-693), with every insufficient resource Fit lists counted. Rows of pods with a feasible
 * node stay 0. counts is int32 [n_pods][FGS_DIAG_SLOTS]. The host sorts "<count> <reason>" and joins:
 * "0/<N> nodes are available: 3 Insufficient cpu, 2 insufficient resource nvidia.flex.com/gpu." Call after the
 * evaluation (on a sharded engine: after the merges; each shard reports its own nodes, the host adds the rows up). */
int fgs_fit_error_histogram(fgs_engine* e, int32_t* counts);

/* ---- sequential scheduling cycles: assume + Reserve between pods ---------------------------- */

/* device_index values that are not an index (FlexGPU.Reserve, pkg/flexgpu/flex_gpu.go:178-223) */
#define FGS_RESERVE_FAIL (-1)     /* "allocate index fail" (:198-200,208-210): the cycle fails, the pod is forgotten */
#define FGS_RESERVE_NONE (-2)     /* the pod has neither flex limit: nothing to reserve (:187-189) */
#define FGS_RESERVE_CONFLICT (-3) /* both limits: UnschedulableAndUnresolvable (:191-193) */
#define FGS_RESERVE_PANIC (-4)    /* NewGPUNode panics on the chosen node (gpu_node.go:40-58,109,115) */
#define FGS_RESERVE_SKIPPED (-5)  /* Reserve did not run: no node chosen, cycle error, or reserve_flex == 0 */

typedef struct fgs_seq_opts {
    int32_t reserve_flex; /* 1: FlexGPU is a Reserve plugin of the profile (it is in the reference's chart) */
    int32_t _pad;
    /* non-NULL: CapacityScheduling is in the profile. Every cycle starts with its PreFilter against the CURRENT
     * ElasticQuotaInfos (pkg/capacityscheduling/capacity_scheduling.go:201-275, no nominated pods) and a committed
     * pod is added to its namespace's Used (Reserve :340-353 → reserveResource, elasticquota.go:74-80). */
    const fgs_quotas* quotas;
} fgs_seq_opts;

/* Per-pod outputs of fgs_schedule_sequential (host buffers owned by the caller, each [P]; any may be NULL). */
typedef struct fgs_seq_results {
    int32_t* feasible_count;
    int32_t* selected_node;  /* -1 when no node is feasible */
    int64_t* selected_score;
    uint32_t* pod_err;       /* FGS_PODERR_* */
    int32_t* device_index;   /* the nvidia.flex.com/index annotation Reserve sets, or FGS_RESERVE_* */
    uint8_t* committed;      /* 1: the pod was assumed on selected_node and the node row was updated */
    uint8_t* quota_code;     /* CapacityScheduling.PreFilter: 0 Success, 1 over Max, 2 Σused over Σmin (then no node is tried) */
    int64_t* quota_used;     /* [n_ns][2+S] ElasticQuotaInfo.Used after the queue (opts->quotas given) */
    uint32_t* quota_used_present; /* [n_ns] its scalar keys */
} fgs_seq_results;

/* What the reference does with a queue of pods, one scheduling cycle each, in `pods` order
 * (vendor/k8s.io/kubernetes/pkg/scheduler/scheduler.go:441-540): Filter + Score + selectHost against the
 * CURRENT node table, then `assume` — NodeInfo.AddPod on the chosen node (scheduler.go:353-370,
 * framework/types.go:583-609: Requested, NonZeroRequested, len(Pods)) — and FlexGPU.Reserve, which picks the
 * device index (whole GPU: first idle non-monopoly device, gpu_node.go:163-177; fractional: first device that
 * still fits, :122-161) that the next cycle's NewGPUNode sees as used. A pod without a feasible node, with a
 * cycle error, or whose Reserve fails leaves the table unchanged (ForgetPod, scheduler.go:512-521).
 * One persistent kernel runs all cycles; the node rows stay on the GPU and only the chosen row is re-derived
 * between cycles. Afterwards the engine's node table reflects every commit (fgs_evaluate sees it).
 * Single GPU: the engine must hold the whole cluster (node_index_base 0). */
int fgs_schedule_sequential(fgs_engine* e, const fgs_pod_table* pods, const fgs_seq_opts* opts,
                            fgs_seq_results* out);

/* FlexGPU.Reserve alone, for callers that keep `assume` on the Go side: the device index each pod would get on
 * node[p] of the current table (no commit). device_index[p] as above; node[p] < 0 gives FGS_RESERVE_SKIPPED. */
int fgs_reserve(fgs_engine* e, const fgs_pod_table* pods, const int32_t* node, int32_t* device_index);

/* ---- coscheduling / capacity scheduling --------------------------------------------------- */

/* PodGroupManager.PreFilter for every pending pod (pkg/coscheduling/core/core.go:149-196 incl.
 * CheckClusterResource :322-342 and getNodeResource :349-382) and CalculateAssignedPods / Permit
 * (:199-216,301-318). Needs fgs_upload_nodes first; `snap` supplies the pods already on nodes.
 * prefilter_code[p]: 0 pass, 1 denied, 2 too few siblings, 3 resource gap.
 * permit_status[p]: 0 Success, 1 Wait, 2 PodGroupNotSpecified, 3 PodGroupNotFound.
 * assigned[g]: CalculateAssignedPods. Any output may be NULL. */
int fgs_gang_check(fgs_engine* e, const fgs_snapshot* snap, const fgs_podset* pending,
                   const fgs_podgroups* pgs, uint8_t* prefilter_code, uint8_t* permit_status,
                   int32_t* assigned);

/* Coscheduling.PostFilter's decision for every pending pod, should the pod turn out unschedulable
 * (pkg/coscheduling/coscheduling.go:140-176): CalculateAssignedPods of the pod's group (core.go:301-318) against
 * Spec.MinMember, then the gap test notAssignedPercentage = float32(MinMember - assigned) / float32(MinMember) <= 0.1 in
 * IEEE single precision as Go evaluates it. postfilter_code[p] = FGS_GANG_POSTFILTER_*; only REJECT_GROUP makes the plugin
 * reject the group's waiting pods, add the group to lastDeniedPG and drop it from permittedPG (:168-172) — the Go side
 * does that with the code in hand; every code returns Unschedulable to the framework. assigned[g] (optional) as in
 * fgs_gang_check. Same inputs and preconditions as fgs_gang_check. */
#define FGS_GANG_POSTFILTER_NO_GROUP 1        /* "can not find pod group" (:143-147): no label, or no PodGroup object */
#define FGS_GANG_POSTFILTER_ENOUGH_ASSIGNED 2 /* assigned >= MinMember (:151-155): the group stands */
#define FGS_GANG_POSTFILTER_SMALL_GAP 3       /* gap <= 10 % (:159-163): later pods may still complete the group */
#define FGS_GANG_POSTFILTER_REJECT_GROUP 4    /* "PodGroup … gets rejected due to Pod … is unschedulable even after PostFilter" */
int fgs_gang_postfilter(fgs_engine* e, const fgs_snapshot* snap, const fgs_podset* pending, const fgs_podgroups* pgs,
                        uint8_t* postfilter_code, int32_t* assigned);

/* CapacityScheduling.PreFilter for every pending pod
 * (pkg/capacityscheduling/capacity_scheduling.go:201-275; elasticquota.go:40-51,90-100,161-181).
 * code[p]: 0 Success, 1 over Max, 2 aggregated used over Σmin. `nominated` may be NULL. */
int fgs_quota_check(fgs_engine* e, const fgs_pod_table* pods, const fgs_quotas* q,
                    const fgs_pod_table* nominated, const int64_t* pod_uid,
                    const int64_t* nominated_uid, uint8_t* code);

/* Batched dry-run preemption: preemptor.SelectVictimsOnNode (pkg/capacityscheduling/capacity_scheduling.go:465-644)
 * for pending pod p on EVERY node of the uploaded snapshot at once — what preemption.Evaluator.DryRunPreemption
 * (sched/framework/preemption/preemption.go:546-597) runs node by node on clones. Per node: the potential victims by the
 * ElasticQuota rules, the filters with all of them removed, the quota bounds, then the reprieve loop from the most
 * important victim down. No PodDisruptionBudgets, empty PodNominator. q may be NULL (no ElasticQuota anywhere).
 * tracks_pods: 1 = assigned pods are members of their quota's pod set, so removing / re-adding one moves its request
 * out of / into Used (production); 0 = they are not (the reference's unit-test fixtures).
 *   node_status[n_nodes]  FGS_PREEMPT_*
 *   victim[snap->pods.n_pods]  1 for the pods chosen as victims on a CANDIDATE node (may be NULL)
 *   n_victims[n_nodes]    (may be NULL)
 *   picked_node           (may be NULL) the candidate pickOneNodeForPreemption (preemption.go:386-505) chooses among
 *                         ALL candidate nodes, -1 if there is none: lowest highest-victim-priority, then smallest sum
 *                         of victim priorities, then fewest victims, then the latest earliest-start-time of the
 *                         highest-priority victims (a pod without a start time counts as 0; the reference takes
 *                         time.Now()); what is left the reference breaks by Go map order, here the lowest node index
 *                         wins. The reference dry-runs a random window of the nodes (GetOffsetAndNumCandidates) and
 *                         picks among those; this is the pick over the whole cluster.
 * `snap` must be the snapshot the node table was packed from; single engine holding the whole cluster; at most 128
 * pods per node. The engine keeps the snapshot's pods on the device between calls and packs them again only when it
 * is handed another fgs_snapshot object or the node table has been uploaded / patched since: a snapshot edited in
 * place goes through fgs_upload_nodes before the next call. The caller filters the nodes by nodesWherePreemptionMightHelp (status plane code != 3) — a node it
 * drops must not be picked: pass picked_node = NULL then and choose among node_status / n_victims on the host. */
int fgs_select_victims(fgs_engine* e, const fgs_snapshot* snap, const fgs_podset* pending, int64_t p,
                       const fgs_quotas* q, int32_t tracks_pods, uint8_t* node_status, uint8_t* victim,
                       int32_t* n_victims, int32_t* picked_node);

/* The same ElasticQuota check for callers that must not allocate. This entry point has no counterpart
 * in the reference; it exists for hosts that keep device memory in a fixed arena of their own, for code on a
 * latency-critical path that may never wait on the CUDA allocator, and for callers that are recording work
 * and cannot touch the allocator at that moment. It needs no engine and calls no cudaMalloc / cudaFree (or
 * anything else that allocates): every table and the result are DEVICE pointers the caller owns, and the
 * kernels' working memory is a fixed __device__ array inside libfgs (about 4.6 KB per GPU, held until exit).
 *   dev_has_eq[n_ns], dev_min/max/used[n_ns][2+S] ({mcpu, mem, scalars}), dev_used_present[n_ns]: as fgs_quotas
 *   dev_pod_req[2+S][n_pods] (column-major {mcpu, mem, scalars}), dev_pod_present[n_pods], dev_pod_ns[n_pods]
 *   dev_code[n_pods]: 0 Success, 1 over Max, 2 aggregated used over the sum of Min (no nominated pods)
 * Limits: n_ns <= 4096 and 2 <= n_scalars <= 8, otherwise FGS_E_TOO_LARGE / FGS_E_INVALID with the reason in
 * errbuf (may be NULL). The first call on a device resolves the array's device address: make it at start-up,
 * before any stream capture on that device. All calls share the array (each clears what it reads first), so
 * two calls that can overlap on the GPU race: keep them on one stream or order them with events.
 * `stream` must belong to the device that is current when the call is made (the array that is used is that device's).
 * The resolved address does not survive cudaDeviceReset. Asynchronous on `stream`; the result is ready when the stream
 * reaches that point. */
int fgs_quota_check_static(void* stream, int32_t n_scalars, int64_t n_ns, const uint8_t* dev_has_eq,
                           const int64_t* dev_min, const int64_t* dev_max, const int64_t* dev_used,
                           const uint32_t* dev_used_present, int64_t n_pods, const int64_t* dev_pod_req,
                           const uint32_t* dev_pod_present, const int32_t* dev_pod_ns, uint8_t* dev_code,
                           char* errbuf, int64_t errbuf_len);

/* ---- timings ---------------------------------------------------------------------------- */

typedef struct fgs_timings {
    float h2d_ms, reduce_ms, emit_ms, d2h_ms; /* last fgs_evaluate, CUDA events on the stream */
    int64_t kernel_launches;                  /* kernels launched by this engine so far */
    int32_t reduce_int32, emit_int32;         /* 1: the last pass ran the int32 specialisation */
    float seq_ms;                             /* last fgs_schedule_sequential: the persistent kernel alone */
    int32_t seq_ctas;                         /* CTAs it ran on */
    int32_t fused;                            /* 1: the last evaluation ran Filter and Score in one pass (then reduce_ms =
                                                 the per-pod prepass, emit_ms = the fused kernel) */
    int32_t reduce_lookup;                    /* 1: the last reduce pass ran Filter as table look-ups (level tables of the node
                                                 side) instead of the compare kernels (then reduce_int32 = 0) */
    int32_t small;                            /* 1: the last fgs_evaluate ran as ONE launch (at most 256 pods, a CTA per pod: Filter,
                                                 reductions, Score, selectHost and the result record; the *_ms fields are 0 then) */
    int32_t _pad;
    int64_t node_patches;                     /* fgs_update_nodes calls served by the row-patch kernel (no host round trip), and … */
    int64_t node_patch_redos;                 /* … how many times a patched row left the node analysis, so that it was redone */
} fgs_timings;
int fgs_get_timings(fgs_engine* e, fgs_timings* t);

/* ---- synthetic clusters (deterministic) ------------------------------------ */

typedef struct fgs_synth_params {
    uint64_t seed_nodes, seed_pods;
    int64_t n_nodes, n_pending;
    int32_t n_scalars, n_devices;
    int32_t max_existing_pods;     /* pods per node ~ U{0..max} */
    /* pending mix, per mille (rest = whole-GPU) */
    int32_t pm_fractional, pm_whole, pm_nogpu, pm_degenerate;
    int32_t pm_missing_req, pm_init_ctr, pm_overhead, pm_two_ctrs, pm_extra_scalar;
    /* node quirks, per mille */
    int32_t pm_quirk_nodes, pm_nongpu_nodes, pm_panic_nodes, pm_invalid_nodes;
    /* coscheduling */
    int64_t n_groups;
    int32_t min_member_lo, min_member_hi, pm_groups_placed, pm_groups_short, pm_neg_left_nodes;
    /* capacity scheduling */
    int64_t n_namespaces;
    int32_t pm_pods_without_eq, n_nominated;
} fgs_synth_params;

typedef struct fgs_synth fgs_synth;
int fgs_synth_create(const fgs_synth_params* p, fgs_synth** out);
void fgs_synth_destroy(fgs_synth* s);
const fgs_snapshot* fgs_synth_snapshot(const fgs_synth* s);
const fgs_podset* fgs_synth_pending(const fgs_synth* s);
const fgs_podset* fgs_synth_nominated(const fgs_synth* s);
const fgs_podgroups* fgs_synth_podgroups(const fgs_synth* s);
const fgs_quotas* fgs_synth_quotas(const fgs_synth* s);

#ifdef __cplusplus
}
#endif
#endif /* FGS_H */
