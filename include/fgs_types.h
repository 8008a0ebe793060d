/*
 * fgs_types.h — flat, pointer+size data formats that cross the libfgs C ABI.
 *
 * Two levels of input exist on purpose:
 *
 *  (1) the OBJECT level (fgs_snapshot / fgs_podset / fgs_podgroups / fgs_quotas): a flattened
 *      picture of what the reference's plugins walk — NodeInfo rows with their pod lists,
 *      containers with request/limit maps, annotations, PodGroup and ElasticQuota objects.
 *      The Go shim fills these straight from the scheduler snapshot; the CPU oracle under
 *      oracle/ walks them the way the Go plugins do.
 *
 *  (2) the TABLE level (fgs_node_table / fgs_pod_table): SoA int64 columns produced from (1)
 *      by the packer (fgs_packer_nodes / fgs_packer_pods); this is what is uploaded to HBM and
 *      what the CUDA kernels read.
 *
 * Quantities are int64 in the units the reference's framework.Resource uses
 * (vendor/k8s.io/kubernetes/pkg/scheduler/framework/types.go:410-419): cpu in milli-cores
 * (Quantity.MilliValue(), rounds up), everything else in base units (Quantity.Value()).
 * The packer rejects |v| >= 2^53 so that no int64 sum the reference forms can wrap.
 *
 * Resource slot numbering used by every per-resource array:
 *      0 cpu (milli)   1 memory   2 ephemeral-storage   3+s scalar resource s (0 <= s < n_scalars)
 * Node "alloc" rows carry one more column for the pod-count limit:
 *      alloc row = { mcpu, mem, eph, pods, scalar_0 .. scalar_{S-1} }            (4+S columns)
 *      request row = { mcpu, mem, eph, scalar_0 .. scalar_{S-1} }                (3+S columns)
 * Scalar slot 0 is nvidia.flex.com/gpu and slot 1 is nvidia.flex.com/memory
 * (pkg/flexgpu/flex_gpu.go:18-19); further slots are whatever extended / hugepage resources the
 * cluster uses.
 */
#ifndef FGS_TYPES_H
#define FGS_TYPES_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define FGS_MAX_SCALARS 8
#define FGS_MAX_DEVICES 16
#define FGS_MAX_SCORE_RES 8

#define FGS_RES_CPU 0
#define FGS_RES_MEM 1
#define FGS_RES_EPH 2
#define FGS_RES_SCALAR0 3
#define FGS_RES_UNKNOWN (-1) /* a name the scorer does not consider: contributes (0,0) */

#define FGS_SCALAR_FLEX_GPU 0
#define FGS_SCALAR_FLEX_MEM 1

/* index annotation missing or not parseable by strconv.Atoi (pkg/flexgpu/gpu_node.go:91-96) */
#define FGS_ANNOT_NONE INT32_MIN

/* plugin ids / mask bits */
#define FGS_PL_FLEXGPU 1u     /* "FlexGPU"                  pkg/flexgpu/flex_gpu.go:17 */
#define FGS_PL_FIT 2u         /* "NodeResourcesFit"         vendor/.../noderesources/fit.go:42 */
#define FGS_PL_ALLOCATABLE 4u /* "NodeResourcesAllocatable" pkg/noderesources/allocatable.go:43 */

/* framework.Code values (vendor/.../framework/interface.go:60-80) */
#define FGS_CODE_SUCCESS 0
#define FGS_CODE_ERROR 1
#define FGS_CODE_UNSCHEDULABLE 2
#define FGS_CODE_UNRESOLVABLE 3
#define FGS_CODE_WAIT 4

/* per-pod cycle error bits (fgs_results.pod_err) */
#define FGS_PODERR_SCORE_RANGE 1u /* a normalised score left [0,100]: runtime/framework.go:920-923 */
#define FGS_PODERR_REF_PANIC 2u   /* the reference would panic here (gpu_node.go:40-58,109,115) */

/* fgs_reason_plane: why FlexGPU.Filter rejected the node = which of its messages the Status carries */
#define FGS_FLEX_REASON_NONE 0
#define FGS_FLEX_REASON_CONFLICT 1         /* "pod conflict resources"                         flex_gpu.go:60 */
#define FGS_FLEX_REASON_UNKNOWN_RESOURCE 2 /* "unknown resource type"                          :67,73 */
#define FGS_FLEX_REASON_INSUFFICIENT_GPU 3 /* "insufficient resource nvidia.flex.com/gpu"      :92 */
#define FGS_FLEX_REASON_INSUFFICIENT_MEM 4 /* "insufficient resource nvidia.flex.com/memory"   :97 */
#define FGS_FLEX_REASON_NO_FIT_INDEX 5     /* "no fit indexes resource nvidia.flex.com/memory" :114 */
#define FGS_FLEX_REASON_PANIC 0xF1         /* NewGPUNode panics on the node (gpu_node.go:40-58,109,115) */
/* … and Fit.Filter's InsufficientResource list as bits (fit.go:256-329; every failing dimension is listed) */
#define FGS_FIT_REASON_PODS 1u     /* "Too many pods" */
#define FGS_FIT_REASON_CPU 2u      /* "Insufficient cpu" */
#define FGS_FIT_REASON_MEM 4u      /* "Insufficient memory" */
#define FGS_FIT_REASON_EPH 8u      /* "Insufficient ephemeral-storage" */
#define FGS_FIT_REASON_SCALAR0 16u /* << s: "Insufficient <scalar resource s>" */

/* fgs_fit_error_histogram: slots of the per-pod reason histogram (FitError, framework/types.go:221-250) */
#define FGS_DIAG_FLEX0 0       /* + (FGS_FLEX_REASON_x - 1): the five FlexGPU messages */
#define FGS_DIAG_FLEX_PANIC 5  /* nodes on which NewGPUNode panics */
#define FGS_DIAG_FIT0 6        /* + bit index of FGS_FIT_REASON_x: Too many pods, cpu, memory, ephemeral-storage, scalar s */
#define FGS_DIAG_NOT_A_NODE 30 /* rows that are not nodes of the list (NodeInfo without a Node) */
#define FGS_DIAG_SLOTS 32

/* fgs_select_victims: what SelectVictimsOnNode returns for a node (pkg/capacityscheduling/capacity_scheduling.go:465-644) */
#define FGS_PREEMPT_CANDIDATE 0   /* Success with at least one victim: the node is a preemption candidate */
#define FGS_PREEMPT_NO_VICTIMS 1  /* "No victims found on node …", UnschedulableAndUnresolvable (:560-565) */
#define FGS_PREEMPT_STILL_UNFIT 2 /* the pod fails the filters even with every potential victim removed (:573-575) */
#define FGS_PREEMPT_QUOTA 3       /* "global quota max exceeded" (:580-585) */
#define FGS_PREEMPT_ERROR 4       /* reprieve removes a pod it has just removed: RemovePod errors (:598-612) */
#define FGS_PREEMPT_NOT_A_NODE 5  /* a row without a Node is not in the list DryRunPreemption walks */
#define FGS_PREEMPT_PANIC 6       /* NewGPUNode panics in one of the filter runs */

/* ---------------------------------------------------------------- object level ---------- */

typedef struct fgs_podset {
    int64_t n_pods;
    int32_t n_scalars;
    int32_t _pad0;
    const int32_t* ns;               /* [n] namespace id >= 0 */
    const int32_t* pg;               /* [n] PodGroup full-name id ("ns/label"), -1 = no label */
    const int32_t* priority;         /* [n] corev1helpers.PodPriority */
    const uint8_t* has_node_name;    /* [n] Spec.NodeName != "" */
    const int32_t* index_annot;      /* [n] Atoi(nvidia.flex.com/index) or FGS_ANNOT_NONE */
    const int64_t* ctr_off;          /* [n+1] container rows of pod i: [ctr_off[i], ctr_off[i+1]) */
    const int32_t* n_init;           /* [n] the LAST n_init rows of that range are init containers */
    const uint32_t* overhead_present;/* [n] bit31: Spec.Overhead != nil; bit r: key of slot r present */
    const int64_t* overhead;         /* [n][3+S] (may be NULL when no pod has overhead) */
    int64_t n_ctrs;
    const int64_t* c_req;            /* [C][3+S] Resources.Requests */
    const uint32_t* c_req_present;   /* [C] bit r: Requests has the key of slot r */
    const int64_t* c_lim_flex;       /* [C][2] Resources.Limits of flex gpu / flex memory */
    const uint32_t* c_lim_present;   /* [C] bit0 flex gpu limit key present, bit1 flex memory */
    /* optional (NULL = all equal): Status.StartTime, any monotone int64 — the tie-break of util.MoreImportantPod
     * (same priority: the pod that started earlier is more important), read by the dry-run preemption only */
    const int64_t* start_time;
} fgs_podset;

typedef struct fgs_snapshot {
    int64_t n_nodes;
    int32_t n_scalars;
    int32_t _pad0;
    const uint8_t* node_valid;            /* [N] NodeInfo.Node() != nil */
    const int64_t* alloc;                 /* [N][4+S] NodeInfo.Allocatable */
    const uint32_t* alloc_scalar_present; /* [N] bit s: Status.Allocatable has scalar s */
    const int64_t* requested;             /* [N][3+S] NodeInfo.Requested */
    const int64_t* nonzero;               /* [N][2]   NodeInfo.NonZeroRequested {mcpu, mem} */
    const int64_t* pod_off;               /* [N+1] rows of `pods` on node n (snapshot list order) */
    fgs_podset pods;                      /* NodeInfo.Pods of all nodes, grouped by node */
    /* optional (NULL = not tracked): NodeInfo.Generation (sched/framework/types.go:396-398), the counter the scheduler
     * cache bumps whenever a NodeInfo changes; fgs_sync_nodes repacks and re-uploads only the rows whose generation moved
     * since its last call, as Cache.UpdateSnapshot does for the Go snapshot (sched/internal/cache/cache.go:197-276) */
    const int64_t* generation;            /* [N] */
} fgs_snapshot;

typedef struct fgs_podgroups {
    int64_t n_groups;
    int32_t n_scalars;
    int32_t _pad0;
    const uint8_t* exists;           /* [G] a PodGroup object with that full name was found */
    const int32_t* ns;               /* [G] namespace id of the group */
    const int32_t* min_member;       /* [G] Spec.MinMember */
    const uint8_t* has_min_resources;/* [G] Spec.MinResources != nil */
    const int64_t* min_res;          /* [G][4+S] {mcpu, mem, eph, pods(unused), scalars} */
    const uint32_t* min_res_present; /* [G] bit k: column k of min_res is a key of MinResources */
    const uint8_t* denied;           /* [G] in lastDeniedPG (core.go:155) */
    const uint8_t* permitted;        /* [G] in permittedPG  (core.go:176) */
} fgs_podgroups;

typedef struct fgs_quotas {
    int64_t n_ns;
    int32_t n_scalars;
    int32_t _pad0;
    const uint8_t* has_eq;               /* [NS] namespace has an ElasticQuotaInfo */
    const int64_t* min;                  /* [NS][2+S] {mcpu, mem, scalars}; missing key = 0 */
    const int64_t* max;                  /* [NS][2+S] */
    const int64_t* used;                 /* [NS][2+S] */
    const uint32_t* used_scalar_present; /* [NS] bit s: Used.ScalarResources has key s */
} fgs_quotas;

/* ---------------------------------------------------------------- plugin configuration -- */

typedef struct fgs_plugin_cfg {
    uint32_t filter_mask;     /* FGS_PL_FLEXGPU | FGS_PL_FIT */
    uint32_t score_mask;      /* FGS_PL_FLEXGPU | FGS_PL_FIT | FGS_PL_ALLOCATABLE */
    int32_t filter_order[2];  /* plugin ids in the order the profile lists them; 0 = end */
    int64_t weight_flex;      /* score-plugin weights (runtime/framework.go:405-408: 0 means 1) */
    int64_t weight_fit;
    int64_t weight_alloc;
    /* NodeResourcesFitArgs: ScoringStrategy {LeastAllocated|MostAllocated, resources}, ignored */
    int32_t fit_strategy;     /* 0 LeastAllocated, 1 MostAllocated */
    int32_t fit_n_res;
    int32_t fit_res[FGS_MAX_SCORE_RES];    /* resource slot ids */
    int64_t fit_res_w[FGS_MAX_SCORE_RES];
    uint32_t fit_ignored_scalar_mask;      /* IgnoredResources / IgnoredResourceGroups → slots */
    /* NodeResourcesAllocatableArgs {Resources, Mode} (apis/config/types.go:50-62) */
    int32_t alloc_mode;       /* 0 Least, 1 Most */
    int32_t alloc_n_res;
    int32_t _pad0;
    int32_t alloc_res[FGS_MAX_SCORE_RES];  /* slot ids or FGS_RES_UNKNOWN */
    int64_t alloc_res_w[FGS_MAX_SCORE_RES];
} fgs_plugin_cfg;

/* ---------------------------------------------------------------- table level ----------- */

/* node flags column */
#define FGS_NF_VALID 1u      /* Node() != nil */
#define FGS_NF_HAS_FLEX 2u   /* both flex resources are keys of Status.Allocatable */
#define FGS_NF_REF_PANIC 4u  /* NewGPUNode panics on this node (gpu count <= 0 or bad index) */
#define FGS_NF_HAS_GPU_RES 8u
#define FGS_NF_HAS_MEM_RES 16u

typedef struct fgs_node_table {
    int64_t n_nodes;
    int32_t n_scalars;
    int32_t n_devices;               /* D: device slots per node in used_mem (max gpu count) */
    /* columns, each [N] int64 unless noted */
    const int64_t* alloc_mcpu;
    const int64_t* alloc_mem;
    const int64_t* alloc_eph;
    const int64_t* alloc_pods;
    const int64_t* req_mcpu;
    const int64_t* req_mem;
    const int64_t* req_eph;
    const int64_t* num_pods;         /* len(NodeInfo.Pods) */
    const int64_t* nz_mcpu;
    const int64_t* nz_mem;
    const int64_t* alloc_scalar;     /* [S][N] */
    const int64_t* req_scalar;       /* [S][N] */
    const uint32_t* alloc_scalar_present; /* [N] */
    const int64_t* flex_gpu_sum;     /* Σ flex gpu limits of pods on node (flex_gpu.go:133-140) */
    const int64_t* flex_mem_sum;
    const int64_t* used_mem;         /* [D][N] per-device used memory (gpu_node.go:113-116) */
    const uint32_t* monopoly_mask;   /* [N] bit i: device i held by a whole-GPU pod */
    const uint32_t* flags;           /* [N] FGS_NF_* */
} fgs_node_table;

/* pod flags column */
#define FGS_PF_GPU_LIMIT 1u  /* a container has a flex gpu limit key */
#define FGS_PF_MEM_LIMIT 2u  /* a container has a flex memory limit key */
#define FGS_PF_REQ_NONE 4u   /* fit.go:270-275: request is all zero and has no scalar keys */

/* Optional companion of a pod table: what the engine's staging pass would otherwise compute by reading every value —
 * the OR and the MAX of each int64 column (they pick the exact power-of-two shifts of the int32 kernels) — and whether
 * the columns lie in page-locked host memory. With both, fgs_submit / fgs_evaluate copy the columns to the GPU by DMA
 * straight from the table (no host pass, no staging copy): the table must then stay untouched until the batch has been
 * collected. The packer fills it (fgs_packer_pods; FGS_POD_STATS_PINNED after fgs_packer_pin). A caller that builds
 * tables itself may pass supersets: or_bits with extra bits and max_val too high are safe, missing bits are not.
 * Column index: 0 flex_gpu_limit, 1 flex_mem_limit, 2 req_mcpu, 3 req_mem, 4 req_eph, 5 nz_mcpu, 6 nz_mem, 7 unused,
 * 8+s req_scalar[s]. */
#define FGS_POD_STATS_PINNED 1u
typedef struct fgs_pod_stats {
    uint64_t or_bits[8 + FGS_MAX_SCALARS];
    int64_t max_val[8 + FGS_MAX_SCALARS];
    uint32_t present_or; /* OR of req_scalar_present */
    uint32_t flags;      /* FGS_POD_STATS_* */
} fgs_pod_stats;

typedef struct fgs_pod_table {
    int64_t n_pods;
    int32_t n_scalars;
    int32_t _pad0;
    const int64_t* flex_gpu_limit;   /* Σ containers' limits (flex_gpu.go:121-131) */
    const int64_t* flex_mem_limit;
    const uint32_t* flags;           /* FGS_PF_* */
    const int64_t* req_mcpu;         /* computePodResourceRequest (fit.go:162-180) */
    const int64_t* req_mem;
    const int64_t* req_eph;
    const int64_t* req_scalar;       /* [S][P] */
    const uint32_t* req_scalar_present; /* [P] bit s: key present in ScalarResources */
    const int64_t* nz_mcpu;          /* calculatePodResourceRequest with non-zero defaults */
    const int64_t* nz_mem;
    const int32_t* ns;               /* namespace id */
    const int32_t* pg;               /* PodGroup id or -1 */
    const int32_t* priority;
    /* read only by the sequential-commit path (fgs_schedule_sequential): what NodeInfo.AddPod adds to
     * NonZeroRequested (calculateResource's non0CPU / non0Mem, framework/types.go:715-745 — unlike the
     * scorer's figure above, the overhead is added in milli-cores) and the pod's own index annotation */
    const int64_t* acct_nz_mcpu;
    const int64_t* acct_nz_mem;
    const int32_t* index_annot;      /* Atoi(nvidia.flex.com/index) or FGS_ANNOT_NONE */
    /* optional (the packer leaves it NULL): pod UIDs, so that a pod that is itself nominated does not yield to
     * itself in the nominated-pod filter pass (runtime/framework.go:812) */
    const int64_t* uid;
    /* optional (NULL = the engine reads the columns itself): see fgs_pod_stats */
    const fgs_pod_stats* stats;
} fgs_pod_table;

#ifdef __cplusplus
}
#endif
#endif /* FGS_TYPES_H */
