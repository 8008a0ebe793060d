"""One engine, a long random sequence of calls: node rows change (patched on the device or uploaded again), batches of three
granularities are evaluated alone, in flight together, resident and resubmitted, filtered and scored under an outside
feasible list, sampled — and after every call the answers are the oracle's on the cluster as it is at that moment. What the
single-feature tests cannot see is the state the features leave for each other (node analysis under patches, the ring of
int32 node tables, pipeline slots, staged-batch generation)."""
import random

import numpy as np
import pytest

import builders as B
import mixed_shifts as M
import flex_gpu_scheduler_b200 as fgs
from flex_gpu_scheduler_b200 import _lib as L

pytestmark = pytest.mark.gpu


class World:
    def __init__(self, oracle, seed, n_nodes, cfg=None):
        self.oracle, self.rng = oracle, random.Random(seed)
        self.N = n_nodes
        self.nodes = M.cluster(n_nodes, seed=seed)
        self.spare = M.cluster(n_nodes, seed=seed + 1000)  # rows to swap in
        self.gen = np.arange(1, n_nodes + 1, dtype=np.int64)
        self.next_gen = n_nodes + 1
        self.version = 0
        self.cfg = cfg if cfg is not None else fgs.config_plugins(2)
        sizes = {"A": 700, "B": 520, "C": 900, "a": 1, "b": 3, "c": 8, "d": 40}
        kinds = {"A": "A", "B": "B", "C": "C", "a": "A", "b": "B", "c": "C", "d": "A"}
        self.sets = {k: B.build_podset(M.batch(kinds[k], n, seed * 7 + i)) for i, (k, n) in enumerate(sizes.items())}
        self.sizes = sizes
        self._snap = None
        self._refs = {}

    def snapshot(self):
        if self._snap is None:
            s = B.build_snapshot(self.nodes)
            s.struct.generation = self.gen.ctypes.data_as(L.p_i64)
            self._snap = s
        return self._snap

    def change_nodes(self, k):
        for i in self.rng.sample(range(self.N), k):
            self.nodes[i], self.spare[i] = self.spare[i], self.nodes[i]
            self.gen[i] = self.next_gen
            self.next_gen += 1
        self.version += 1
        self._snap = None
        self._refs = {}

    def ref(self, key, **kw):
        tag = (key, tuple(sorted((k, v if not isinstance(v, np.ndarray) else v.tobytes()) for k, v in kw.items())))
        if tag not in self._refs:
            self._refs[tag] = self.oracle.evaluate(self.snapshot().ptr, self.sets[key].ptr, self.cfg, workers=4, **kw)
        return self._refs[tag]


def _same(got, ref, N, planes):
    assert np.array_equal(got.feasible_count, ref.feasible_count), "counts"
    assert np.array_equal(got.pod_err, ref.pod_err), "error bits"
    ok = (ref.pod_err == 0) & (ref.feasible_count >= 1)
    assert np.array_equal(got.selected_score[ok], ref.max_total[ok]), "selected score"
    idx = np.nonzero(ok)[0]
    assert np.all(ref.status[idx, got.selected_node[idx]] == 0), "selected node infeasible"
    assert np.array_equal(ref.total[idx, got.selected_node[idx]], ref.max_total[idx]), "selected node outside the tie set"
    assert np.all(got.selected_node[ref.feasible_count == 0] == -1)
    if planes:
        assert np.array_equal(got.feasible(N), ref.status == 0), "bitmap"
        m = (ref.status == 0) & ok[:, None]
        assert np.array_equal(got.scores[:, :N][m], ref.total[m]), "scores"


CFGS = {
    "c2": None,
    "c1": lambda: fgs.config_plugins(1),
    "all-three-flex-first": lambda: fgs.plugin_cfg(filter_plugins=("FlexGPU", "NodeResourcesFit"),
                                                  score_plugins=(("FlexGPU", 2), ("NodeResourcesFit", 1), ("NodeResourcesAllocatable", 3)),
                                                  fit_strategy="MostAllocated", fit_resources=(("cpu", 2), ("memory", 3)), alloc_mode="Most"),
}


@pytest.mark.parametrize("seed,n_nodes,cfg_name", [(11, 1500, "c2"), (12, 2600, "c2"), (13, 1200, "c1"), (14, 900, "all-three-flex-first")])
def test_random_call_sequences_match_the_oracle(gpu_ok, oracle, seed, n_nodes, cfg_name):
    w = World(oracle, seed, n_nodes, CFGS[cfg_name]() if CFGS[cfg_name] else None)
    rng, N = w.rng, w.N
    eng = fgs.Engine(device=0)
    eng.set_plugins(w.cfg)
    pk = fgs.Packer(2)
    packers = {k: fgs.Packer(2, pin=(i % 2 == 0)) for i, k in enumerate(w.sets)}
    tabs = {k: packers[k].pods(v.ptr) for k, v in w.sets.items()}
    eng.sync_nodes(pk, w.snapshot().ptr)
    done = {}
    for step in range(70):
        op = rng.choice(["nodes", "nodes", "eval", "eval", "eval", "pipe", "resident", "masked", "sample"])
        done[op] = done.get(op, 0) + 1
        if op == "nodes":
            w.change_nodes(rng.choice([1, 1, 1, 2, 4, 5, 30]))
            n = eng.sync_nodes(pk, w.snapshot().ptr)
            assert 1 <= n <= 30
        elif op == "eval":
            k = rng.choice(list(w.sets))
            planes = rng.random() < 0.5
            _same(eng.evaluate(tabs[k], want_planes=planes), w.ref(k), N, planes)
        elif op == "pipe":
            order = [rng.choice("ABCd") for _ in range(rng.randint(2, 6))]
            depth, q = eng.max_in_flight(), []
            for k in order:
                eng.submit(tabs[k])
                q.append(k)
                if len(q) == depth:
                    _same(eng.collect(), w.ref(q.pop(0)), N, False)
            while q:
                _same(eng.collect(), w.ref(q.pop(0)), N, False)
        elif op == "resident":
            eng.set_resident_batches(3)
            for k in "ACB":
                eng.submit(tabs[k])
                _same(eng.collect(), w.ref(k), N, False)
            for _ in range(2):
                q = []
                for k in "ACB":
                    eng.resubmit(w.sizes[k])
                    q.append(k)
                while q:
                    _same(eng.collect(), w.ref(q.pop(0)), N, False)
            eng.set_resident_batches(0)
        elif op == "masked":
            k = rng.choice("ABCcd")
            P = w.sizes[k]
            mask = np.random.default_rng(step).random((P, N)) < 0.7
            eng.evaluate(tabs[k], skip_pass_b=1)
            got = eng.score_masked(mask, want_planes=True)
            _same(got, w.ref(k, ext_mask=mask.astype(np.uint8)), N, True)
        elif op == "sample":
            k = rng.choice("ABC")
            pct, start = rng.choice([0, 10, 30]), rng.randrange(N)
            got = eng.evaluate(tabs[k], want_planes=True, sample=1, percentage_of_nodes_to_score=pct, start_node_index=start)
            ref = w.ref(k, sample=(pct, start))
            assert np.array_equal(got.processed_nodes, ref.processed) and got.next_start_node_index == ref.next_start
            assert np.array_equal(got.feasible(N), ref.status == 0) and np.array_equal(got.feasible_count, ref.feasible_count)
            ok = (ref.pod_err == 0) & (ref.feasible_count > 0)
            assert np.array_equal(got.selected_score[ok], ref.max_total[ok])
    tm = eng.timings()
    assert tm["node_patches"] >= 1 and len(done) >= 6, (tm, done)
    eng.close()
