"""Thin Python host layer over the libfgs C ABI.

Everything here is plumbing for tests and bench.py: it builds ctypes views of numpy arrays, calls the
``extern "C"`` entry points of include/fgs.h and hands numpy arrays back. No scheduling arithmetic
lives in Python; if libfgs.so or a CUDA device is missing the calls raise.
"""
import ctypes as C

import numpy as np

from . import _lib as L


class FgsError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"libfgs error {code}: {msg}")
        self.code = code
        self.msg = msg


def _ptr(a, ctype):
    return a.ctypes.data_as(C.POINTER(ctype))


def plugin_cfg(filter_plugins=("NodeResourcesFit", "FlexGPU"), score_plugins=(("FlexGPU", 1), ("NodeResourcesFit", 1)),
               fit_strategy="LeastAllocated", fit_resources=(("cpu", 1), ("memory", 1)), fit_ignored_scalar_mask=0,
               alloc_mode="Least", alloc_resources=(("cpu", 1 << 20), ("memory", 1))):
    """Build an fgs_plugin_cfg the way a KubeSchedulerConfiguration profile would.

    Plugin names are the reference's: "FlexGPU" (pkg/flexgpu/flex_gpu.go:17), "NodeResourcesFit",
    "NodeResourcesAllocatable" (pkg/noderesources/allocatable.go:43). Resources are "cpu", "memory",
    "ephemeral-storage", ("scalar", s) for slot s, or any other string for a name the scorer ignores.
    """
    ids = {"FlexGPU": L.FGS_PL_FLEXGPU, "NodeResourcesFit": L.FGS_PL_FIT, "NodeResourcesAllocatable": L.FGS_PL_ALLOCATABLE}

    def res_id(r):
        if r == "cpu":
            return L.FGS_RES_CPU
        if r == "memory":
            return L.FGS_RES_MEM
        if r == "ephemeral-storage":
            return L.FGS_RES_EPH
        if isinstance(r, tuple) and r[0] == "scalar":
            return L.FGS_RES_SCALAR0 + int(r[1])
        return L.FGS_RES_UNKNOWN

    c = L.PluginCfg()
    for k, name in enumerate(filter_plugins):
        c.filter_mask |= ids[name]
        c.filter_order[k] = ids[name]
    c.weight_flex = c.weight_fit = c.weight_alloc = 1
    for name, w in score_plugins:
        c.score_mask |= ids[name]
        if name == "FlexGPU":
            c.weight_flex = w
        elif name == "NodeResourcesFit":
            c.weight_fit = w
        else:
            c.weight_alloc = w
    c.fit_strategy = {"LeastAllocated": 0, "MostAllocated": 1}[fit_strategy]
    c.fit_n_res = len(fit_resources)
    for k, (r, w) in enumerate(fit_resources):
        c.fit_res[k] = res_id(r)
        c.fit_res_w[k] = w
    c.fit_ignored_scalar_mask = fit_ignored_scalar_mask
    c.alloc_mode = {"Least": 0, "Most": 1}.get(alloc_mode, 2)
    c.alloc_n_res = len(alloc_resources)
    for k, (r, w) in enumerate(alloc_resources):
        c.alloc_res[k] = res_id(r)
        c.alloc_res_w[k] = w
    return c


class Synth:
    """Deterministic synthetic cluster; owns the object-level snapshot."""

    DEFAULTS = dict(seed_nodes=0xF6500000, seed_pods=0xF6510000, n_nodes=1000, n_pending=1000, n_scalars=2, n_devices=8,
                    max_existing_pods=24, pm_fractional=600, pm_whole=300, pm_nogpu=50, pm_degenerate=50, pm_missing_req=100,
                    pm_init_ctr=100, pm_overhead=30, pm_two_ctrs=200, pm_extra_scalar=0, pm_quirk_nodes=20, pm_nongpu_nodes=0,
                    pm_panic_nodes=0, pm_invalid_nodes=0, n_groups=0, min_member_lo=4, min_member_hi=64, pm_groups_placed=400,
                    pm_groups_short=50, pm_neg_left_nodes=10, n_namespaces=0, pm_pods_without_eq=50, n_nominated=0)

    def __init__(self, **kw):
        lib = L.load()
        p = L.SynthParams()
        vals = dict(self.DEFAULTS)
        vals.update(kw)
        for k, v in vals.items():
            setattr(p, k, v)
        self.params = vals
        h = C.c_void_p()
        rc = lib.fgs_synth_create(C.byref(p), C.byref(h))
        if rc != 0:
            raise FgsError(rc, "fgs_synth_create: invalid parameters")
        self._h = h
        self._lib = lib
        self.snapshot = lib.fgs_synth_snapshot(h)
        self.pending = lib.fgs_synth_pending(h)
        self.nominated = lib.fgs_synth_nominated(h)
        self.podgroups = lib.fgs_synth_podgroups(h)
        self.quotas = lib.fgs_synth_quotas(h)

    @property
    def n_nodes(self):
        return int(self.snapshot.contents.n_nodes)

    @property
    def n_pending(self):
        return int(self.pending.contents.n_pods)

    def close(self):
        if self._h:
            self._lib.fgs_synth_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class Packer:
    """Host-side packer (no GPU needed): object snapshot → SoA tables. A table stays valid until the
    next call of the same kind on the same Packer, so use one Packer per table you keep."""

    def __init__(self, n_scalars=2, pin=False):
        """pin=True: page-lock the pod tables this packer makes (fgs_packer_pin), so that the engine copies them to the
        GPU by DMA without a staging pass; needs a CUDA device"""
        self._lib = L.load()
        h = C.c_void_p()
        rc = self._lib.fgs_packer_create(n_scalars, C.byref(h))
        if rc != 0:
            raise FgsError(rc, "fgs_packer_create")
        self._h = h
        if pin:
            self._check(self._lib.fgs_packer_pin(self._h))

    def _check(self, rc):
        if rc != 0:
            raise FgsError(rc, self._lib.fgs_packer_error(self._h).decode())

    def nodes(self, snapshot, n0=0, n1=None):
        if n1 is None:
            n1 = int(snapshot.contents.n_nodes)
        t = C.POINTER(L.NodeTable)()
        self._check(self._lib.fgs_packer_nodes(self._h, snapshot, n0, n1, C.byref(t)))
        t._owner = self  # the table lives inside the packer: keep it alive with the pointer
        return t

    def node_rows(self, snapshot, rows):
        """fgs_packer_node_rows: table row i = snapshot row rows[i] (for fgs_update_nodes with the same list)"""
        idx = np.ascontiguousarray(rows, dtype=np.int64)
        t = C.POINTER(L.NodeTable)()
        self._check(self._lib.fgs_packer_node_rows(self._h, snapshot, _ptr(idx, C.c_int64), len(idx), C.byref(t)))
        t._owner = self
        return t

    def pods(self, podset, p0=0, p1=None):
        if p1 is None:
            p1 = int(podset.contents.n_pods)
        t = C.POINTER(L.PodTable)()
        self._check(self._lib.fgs_packer_pods(self._h, podset, p0, p1, C.byref(t)))
        t._owner = self
        return t

    def close(self):
        if self._h:
            self._lib.fgs_packer_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class Results:
    def __init__(self, P):
        self.feasible_count = np.zeros(P, np.int32)
        self.selected_node = np.zeros(P, np.int32)
        self.selected_score = np.zeros(P, np.int64)
        self.pod_err = np.zeros(P, np.uint32)
        self.processed_nodes = np.zeros(max(P, 1), np.int32)[:P]  # node sampling: nodes each pod's walk counted
        self.next_start_node_index = -1
        self.feasible_bitmap = None
        self.scores = None
        self.words_per_row = 0
        self.score_row_stride = 0

    def feasible(self, n_nodes):
        """bool [P][n_nodes] from the bitmap"""
        bits = np.unpackbits(self.feasible_bitmap.view(np.uint8), axis=1, bitorder="little")
        return bits[:, :n_nodes].astype(bool)


class SeqResults:
    def __init__(self, P):
        n = max(P, 1)
        self.feasible_count = np.zeros(n, np.int32)[:P]
        self.selected_node = np.zeros(n, np.int32)[:P]
        self.selected_score = np.zeros(n, np.int64)[:P]
        self.pod_err = np.zeros(n, np.uint32)[:P]
        self.device_index = np.zeros(n, np.int32)[:P]
        self.committed = np.zeros(n, np.uint8)[:P]
        self.quota_code = np.zeros(n, np.uint8)[:P]
        self.quota_used = None
        self.quota_used_present = None


class Engine:
    """One libfgs engine = one GPU's shard of the node table."""

    def __init__(self, device=0, n_scalars=2, node_index_base=0, total_nodes=0, seed=0x5EED, stream=None):
        self._lib = L.load()
        cfg = L.Config(device=device, n_scalars=n_scalars, node_index_base=node_index_base, total_nodes=total_nodes,
                       select_seed=seed, stream=stream)
        h = C.c_void_p()
        rc = self._lib.fgs_create(C.byref(cfg), C.byref(h))
        if rc != 0:
            raise FgsError(rc, self._lib.fgs_last_error(None).decode())
        self._h = h
        self.n_nodes = 0
        self.n_pods = 0
        self._keep = []

    def _check(self, rc):
        if rc != 0:
            raise FgsError(rc, self._lib.fgs_last_error(self._h).decode())

    def close(self):
        if self._h:
            self._lib.fgs_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- configuration / upload ------------------------------------------------------------------
    def set_plugins(self, cfg):
        self._check(self._lib.fgs_set_plugins(self._h, C.byref(cfg)))

    def upload_nodes(self, table):
        self._check(self._lib.fgs_upload_nodes(self._h, table))
        self.n_nodes = int(table.contents.n_nodes)

    def sync_nodes(self, packer, snapshot, n0=0, n1=None):
        """fgs_sync_nodes: repack and re-upload only the rows whose NodeInfo.Generation moved; returns how many"""
        if n1 is None:
            n1 = int(snapshot.contents.n_nodes)
        n = C.c_int64(0)
        self._check(self._lib.fgs_sync_nodes(self._h, packer._h, snapshot, n0, n1, C.byref(n)))
        self.n_nodes = n1 - n0  # the result planes are sized by it
        return int(n.value)

    def update_nodes(self, rows, dirty_idx):
        idx = np.ascontiguousarray(dirty_idx, dtype=np.int64)
        self._check(self._lib.fgs_update_nodes(self._h, rows, _ptr(idx, C.c_int64)))

    # -- evaluation ----------------------------------------------------------------------------------
    def _results(self, P, want_planes):
        r = Results(P)
        cr = L.Results()
        cr.feasible_count = _ptr(r.feasible_count, C.c_int32)
        cr.selected_node = _ptr(r.selected_node, C.c_int32)
        cr.selected_score = _ptr(r.selected_score, C.c_int64)
        cr.pod_err = _ptr(r.pod_err, C.c_uint32)
        cr.processed_nodes = _ptr(r.processed_nodes, C.c_int32)
        if want_planes:
            n_pad = (max(self.n_nodes, 1) + 1023) // 1024 * 1024  # kNodePad of the engine
            r.feasible_bitmap = np.zeros((P, n_pad // 32), np.uint32)
            r.scores = np.zeros((P, n_pad), np.int64)
            cr.feasible_bitmap = _ptr(r.feasible_bitmap, C.c_uint32)
            cr.scores = _ptr(r.scores, C.c_int64)
        return r, cr

    @staticmethod
    def _opts(emit_mask=0, emit_raw=0, skip_pass_b=0, sample=0, percentage_of_nodes_to_score=0, start_node_index=0):
        return L.EvalOpts(emit_mask=emit_mask, emit_raw=emit_raw, skip_pass_b=skip_pass_b, sample=1 if sample else 0,
                          percentage_of_nodes_to_score=percentage_of_nodes_to_score, start_node_index=start_node_index)

    def evaluate(self, pod_table, want_planes=False, **opts):
        P = int(pod_table.contents.n_pods)
        self.n_pods = P
        r, cr = self._results(P, want_planes)
        o = self._opts(**opts)
        self._check(self._lib.fgs_evaluate(self._h, pod_table, C.byref(o), C.byref(cr)))
        r.words_per_row, r.score_row_stride = int(cr.words_per_row), int(cr.score_row_stride)
        r.next_start_node_index = int(cr.next_start_node_index)
        return r

    def score_masked(self, mask_bool, want_planes=False, **opts):
        """fgs_score_masked: Score / NormalizeScore / selectHost of the filtered batch under the framework's feasible list.
        mask_bool: bool [P][n_nodes of this engine] — True = the node passed every filter plugin of the profile."""
        P = self.n_pods
        m = np.asarray(mask_bool, bool)
        W = (m.shape[1] + 1023) // 1024 * 1024 // 32
        bits = np.zeros((P, W * 32), np.uint8)
        bits[:, :m.shape[1]] = m
        words = np.packbits(bits.reshape(P, W, 32), axis=2, bitorder="little").view(np.uint32).reshape(P, W)
        words = np.ascontiguousarray(words)
        r, cr = self._results(P, want_planes)
        o = self._opts(**opts)
        self._check(self._lib.fgs_score_masked(self._h, _ptr(words, C.c_uint32), 0, C.byref(o), C.byref(cr)))
        r.words_per_row, r.score_row_stride = int(cr.words_per_row), int(cr.score_row_stride)
        return r

    def select_victims(self, snapshot, pending, p, quotas=None, tracks_pods=True, pick=False):
        """Dry-run preemption of pending pod p on every node (fgs_select_victims):
        (status [N] FGS_PREEMPT_*, victim flags [pods in the snapshot], victims per node [N]);
        pick=True appends the node pickOneNodeForPreemption chooses among the candidates (-1: none)"""
        N = int(snapshot.contents.n_nodes)
        T = int(snapshot.contents.pods.n_pods)
        st, vic, cnt = np.zeros(max(N, 1), np.uint8), np.zeros(max(T, 1), np.uint8), np.zeros(max(N, 1), np.int32)
        picked = C.c_int32(-1)
        self._check(self._lib.fgs_select_victims(self._h, snapshot, pending, p, quotas, 1 if tracks_pods else 0, _ptr(st, C.c_uint8),
                                                 _ptr(vic, C.c_uint8), _ptr(cnt, C.c_int32), C.byref(picked) if pick else None))
        if pick:
            return st[:N], vic[:T], cnt[:N], int(picked.value)
        return st[:N], vic[:T], cnt[:N]

    def set_nominated(self, pod_table=None, nodes=None, uid=None):
        """The PodNominator's content (fgs_set_nominated): pod_table = packed nominated pods, nodes[q] = snapshot index
        of pod q's nominated node. No arguments: clear."""
        if pod_table is None:
            self._check(self._lib.fgs_set_nominated(self._h, None, None, None))
            return
        nodes = np.ascontiguousarray(nodes, np.int32)
        u = _ptr(np.ascontiguousarray(uid, np.int64), C.c_int64) if uid is not None else None
        self._check(self._lib.fgs_set_nominated(self._h, pod_table, _ptr(nodes, C.c_int32), u))

    @staticmethod
    def with_uids(pod_table, uid):
        """Attach pod UIDs to a packed pod table (fgs_pod_table.uid); returns the table."""
        arr = np.ascontiguousarray(uid, np.int64)
        pod_table.contents.uid = _ptr(arr, C.c_int64)
        pod_table._uid_keep = arr
        return pod_table

    def submit(self, pod_table, **opts):
        """Enqueue one batch (fgs_submit); at most two may be in flight. Pair every submit with a collect()."""
        P = int(pod_table.contents.n_pods)
        o = self._opts(**opts)
        self._check(self._lib.fgs_submit(self._h, pod_table, C.byref(o)))
        self._pending = getattr(self, "_pending", []) + [P]

    def resubmit(self, n_pods, **opts):
        """fgs_resubmit: the batch already resident in the next pipeline slot again (no staging, no H2D)."""
        o = self._opts(**opts)
        self._check(self._lib.fgs_resubmit(self._h, C.byref(o)))
        self._pending = getattr(self, "_pending", []) + [int(n_pods)]

    def set_resident_batches(self, n):
        """fgs_set_resident_batches: a ring of n pipeline slots, so that a job of n batches stays resident for resubmit()"""
        self._check(self._lib.fgs_set_resident_batches(self._h, int(n)))

    def max_in_flight(self):
        return int(self._lib.fgs_max_in_flight(self._h))

    def collect(self):
        """Per-pod results of the oldest submitted batch (blocks until it is done)."""
        P = self._pending.pop(0)
        r, cr = self._results(P, False)
        self._check(self._lib.fgs_collect(self._h, C.byref(cr)))
        return r

    def stage_pods(self, pod_table):
        self.n_pods = int(pod_table.contents.n_pods)
        self._check(self._lib.fgs_stage_pods(self._h, pod_table))

    def pass_reduce(self):
        self._check(self._lib.fgs_pass_reduce(self._h))

    def fused_eligible(self, **opts):
        o = self._opts(**opts)
        return bool(self._lib.fgs_fused_eligible(self._h, C.byref(o)))

    def pass_fused(self, **opts):
        """Filter and Score in one pass over pod × node (fgs_pass_fused) instead of pass_reduce + pass_emit"""
        o = self._opts(**opts)
        self._check(self._lib.fgs_pass_fused(self._h, C.byref(o)))

    def pass_fused_prepass(self, **opts):
        o = self._opts(**opts)
        self._check(self._lib.fgs_pass_fused_prepass(self._h, C.byref(o)))

    def pass_fused_main(self, **opts):
        o = self._opts(**opts)
        self._check(self._lib.fgs_pass_fused_main(self._h, C.byref(o)))

    def pass_emit(self, **opts):
        o = self._opts(**opts)
        self._check(self._lib.fgs_pass_emit(self._h, C.byref(o)))

    def fetch_results(self, want_planes=False):
        r, cr = self._results(self.n_pods, want_planes)
        self._check(self._lib.fgs_fetch_results(self._h, C.byref(cr)))
        r.words_per_row, r.score_row_stride = int(cr.words_per_row), int(cr.score_row_stride)
        r.next_start_node_index = int(cr.next_start_node_index)
        return r

    def device_view(self):
        v = L.DeviceView()
        self._check(self._lib.fgs_get_device_view(self._h, C.byref(v)))
        return v

    def peer_bytes(self, max_pods):
        return int(self._lib.fgs_peer_bytes(int(max_pods)))

    def peer_attach(self, rank, world, buffer_ptrs, multicast_ptr, nbytes):
        """fgs_peer_attach: the exchange buffers of all shards (device addresses as mapped in this process)."""
        g = L.PeerGroup()
        g.rank, g.world, g.bytes = int(rank), int(world), int(nbytes)
        for r, p in enumerate(buffer_ptrs):
            g.buffers[r] = int(p)
        g.multicast = int(multicast_ptr) if multicast_ptr else None
        self._check(self._lib.fgs_peer_attach(self._h, C.byref(g)))

    def peer_detach(self):
        self._check(self._lib.fgs_peer_attach(self._h, None))

    def peer_merge_reduce(self):
        self._check(self._lib.fgs_peer_merge_reduce(self._h))

    def peer_merge_select(self):
        self._check(self._lib.fgs_peer_merge_select(self._h))

    def peer_status(self):
        """waits for the engine stream; raises if a merge timed out; True when the multicast path is in use"""
        mc = C.c_int32(0)
        self._check(self._lib.fgs_peer_status(self._h, C.byref(mc)))
        return bool(mc.value)

    def synchronize(self):
        self._check(self._lib.fgs_synchronize(self._h))

    def status_plane(self, plugin=0):
        out = np.zeros((self.n_pods, self.n_nodes), np.uint8)
        self._check(self._lib.fgs_status_plane(self._h, plugin, _ptr(out, C.c_uint8)))
        return out

    def reason_plane(self, plugin):
        """uint16 [P][N]: FGS_FLEX_REASON_* ids (plugin FlexGPU) or FGS_FIT_REASON_* bit masks (plugin Fit)."""
        out = np.zeros((self.n_pods, self.n_nodes), np.uint16)
        self._check(self._lib.fgs_reason_plane(self._h, plugin, _ptr(out, C.c_uint16)))
        return out

    def fit_error_histogram(self):
        """int32 [P][FGS_DIAG_SLOTS]: per pod without a feasible node, nodes per reason of its NodeToStatusMap
        (fgs_fit_error_histogram — the numbers in FitError.Error())"""
        out = np.zeros((max(self.n_pods, 1), L.FGS_DIAG_SLOTS), np.int32)
        self._check(self._lib.fgs_fit_error_histogram(self._h, _ptr(out, C.c_int32)))
        return out[:self.n_pods]

    def timings(self):
        t = L.Timings()
        self._check(self._lib.fgs_get_timings(self._h, C.byref(t)))
        return dict(h2d_ms=t.h2d_ms, reduce_ms=t.reduce_ms, emit_ms=t.emit_ms, d2h_ms=t.d2h_ms, kernel_launches=t.kernel_launches,
                    reduce_int32=bool(t.reduce_int32), emit_int32=bool(t.emit_int32), seq_ms=t.seq_ms, seq_ctas=int(t.seq_ctas),
                    fused=bool(t.fused), reduce_lookup=bool(t.reduce_lookup), small=bool(t.small),
                    node_patches=int(t.node_patches), node_patch_redos=int(t.node_patch_redos))

    # -- sequential cycles: assume + Reserve between pods -------------------------------------------------
    def schedule_sequential(self, pod_table, reserve_flex=True, quotas=None):
        """One scheduling cycle per pod of the table, in order, committing each placement (fgs_schedule_sequential).
        quotas: POINTER(Quotas) when CapacityScheduling is in the profile (PreFilter per cycle, Reserve per commit)."""
        P = int(pod_table.contents.n_pods)
        r = SeqResults(P)
        cr = L.SeqResults()
        cr.feasible_count = _ptr(r.feasible_count, C.c_int32)
        cr.selected_node = _ptr(r.selected_node, C.c_int32)
        cr.selected_score = _ptr(r.selected_score, C.c_int64)
        cr.pod_err = _ptr(r.pod_err, C.c_uint32)
        cr.device_index = _ptr(r.device_index, C.c_int32)
        cr.committed = _ptr(r.committed, C.c_uint8)
        cr.quota_code = _ptr(r.quota_code, C.c_uint8)
        o = L.SeqOpts(reserve_flex=1 if reserve_flex else 0)
        if quotas is not None:
            q = quotas.contents
            NS, W = int(q.n_ns), 2 + int(q.n_scalars)
            r.quota_used = np.zeros((max(NS, 1), W), np.int64)[:NS]
            r.quota_used_present = np.zeros(max(NS, 1), np.uint32)[:NS]
            cr.quota_used = _ptr(r.quota_used, C.c_int64)
            cr.quota_used_present = _ptr(r.quota_used_present, C.c_uint32)
            o.quotas = C.cast(quotas, C.c_void_p)
        self._check(self._lib.fgs_schedule_sequential(self._h, pod_table, C.byref(o), C.byref(cr)))
        return r

    def reserve(self, pod_table, nodes):
        """FlexGPU.Reserve's device index for pod p on nodes[p] of the current table, without committing."""
        P = int(pod_table.contents.n_pods)
        nodes = np.ascontiguousarray(nodes, np.int32)
        assert nodes.shape == (P,)
        dev = np.zeros(max(P, 1), np.int32)
        self._check(self._lib.fgs_reserve(self._h, pod_table, _ptr(nodes, C.c_int32), _ptr(dev, C.c_int32)))
        return dev[:P]

    # -- coscheduling / capacity scheduling ----------------------------------------------------------
    def gang_check(self, snapshot, pending, podgroups):
        P = int(pending.contents.n_pods)
        G = int(podgroups.contents.n_groups)
        pre = np.zeros(max(P, 1), np.uint8)
        permit = np.zeros(max(P, 1), np.uint8)
        assigned = np.zeros(max(G, 1), np.int32)
        self._check(self._lib.fgs_gang_check(self._h, snapshot, pending, podgroups, _ptr(pre, C.c_uint8), _ptr(permit, C.c_uint8),
                                             _ptr(assigned, C.c_int32)))
        return pre[:P], permit[:P], assigned[:G]

    def gang_postfilter(self, snapshot, pending, podgroups):
        """Coscheduling.PostFilter's decision per pending pod (fgs_gang_postfilter): FGS_GANG_POSTFILTER_* codes"""
        P = int(pending.contents.n_pods)
        code = np.zeros(max(P, 1), np.uint8)
        self._check(self._lib.fgs_gang_postfilter(self._h, snapshot, pending, podgroups, _ptr(code, C.c_uint8), None))
        return code[:P]

    def quota_check(self, pod_table, quotas, nominated_table=None, pod_uid=None, nominated_uid=None):
        P = int(pod_table.contents.n_pods)
        code = np.zeros(max(P, 1), np.uint8)
        pu = _ptr(np.ascontiguousarray(pod_uid, np.int64), C.c_int64) if pod_uid is not None else None
        nu = _ptr(np.ascontiguousarray(nominated_uid, np.int64), C.c_int64) if nominated_uid is not None else None
        self._check(self._lib.fgs_quota_check(self._h, pod_table, quotas, nominated_table, pu, nu, _ptr(code, C.c_uint8)))
        return code[:P]
