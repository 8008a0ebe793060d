"""The two CPU restatements of the reference's plugins against each other (CPU only): oracle/ref_walk.cpp, which the GPU
parity tests check against, and oracle/literal_walk.py, written from the Go sources independently. pkg/flexgpu and the
vendored NodeResourcesFit / DefaultNormalizeScore / selectHost ship without tests, so neither restatement can be pinned on
reference vectors for them; agreement of two separate readings on random snapshots (quirk nodes, annotations out of
range, overheads, init containers, every plugin order / strategy / weight) is the evidence there is. The 10^5-snapshot run
of the same harness is logged in profiles/r02_cross_fuzz.log (python -m oracle.cross_fuzz --n 100000)."""
import random

import numpy as np

import builders as B
import flex_gpu_scheduler_b200 as fgs
from oracle import cross_fuzz as F
from oracle import literal_walk as LW


def test_two_restatements_agree_on_random_snapshots(oracle):
    n_bad, n_run, evals = F.run(1500, seed=20261107, quiet=True)
    assert n_run == 1500 and evals > 10000
    assert n_bad == 0


def test_literal_walk_on_the_hand_derived_flexgpu_cases(oracle):
    """the aliasing in MemAssumeFitIndexes (gpu_node.go:131-144): `remain := u.memory` shares ONE Quantity between the
    devices, so every fit shrinks the memory the later devices are compared with; Filter only asks whether ANY device
    fits, Reserve takes the first"""
    GPU, MEM = ("scalar", 0), ("scalar", 1)
    node = B.node({"cpu": 8000, "memory": 1 << 34, "pods": 10, GPU: 2, MEM: 2000},
                  [B.pod([B.ctr({MEM: 600}, mem_limit=600)], node_name=True, index=0)])
    snap = B.build_snapshot([node])
    pods = [B.pod([B.ctr({"cpu": 100, MEM: 400}, mem_limit=400)]),   # fits device 0 (600 + 400 <= 1000) → index 0
            B.pod([B.ctr({"cpu": 100, MEM: 401}, mem_limit=401)]),   # device 0 full, device 1 takes it → index 1
            B.pod([B.ctr({"cpu": 100, MEM: 1001}, mem_limit=1001)])]  # no device: "no fit indexes"
    pend = B.build_podset(pods)
    cfg = fgs.plugin_cfg(filter_plugins=("FlexGPU",), score_plugins=(("FlexGPU", 1),))
    res = LW.evaluate(snap.ptr, pend.ptr, cfg)
    assert [len(r.feasible) for r in res] == [1, 1, 0]
    assert int(res[2].status[0]) == (LW.Unschedulable | (LW.PL_FLEXGPU << 4))
    assert [LW.reserve_index(snap.ptr, pend.ptr, p, 0) for p in range(3)] == [0, 1, -1]
    assert [oracle.reserve_index(snap.ptr, pend.ptr, p, 0) for p in range(3)] == [0, 1, -1]
    # MemScore sums memory − used over ALL devices, monopolised ones included (gpu_node.go:189-199)
    assert int(res[0].raw[LW.PL_FLEXGPU][0]) == 2 * 1000 - 600
