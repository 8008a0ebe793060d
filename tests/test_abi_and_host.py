"""CPU tests: the C-ABI library loads and exports every symbol include/fgs.h declares (no compute calls —
there is no GPU here), and the host-side logic (synthetic generator, packer, validation) behaves."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import builders as B
import flex_gpu_scheduler_b200 as fgs
from flex_gpu_scheduler_b200 import _lib as L

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_functions():
    names = set()
    for hdr in ("fgs.h",):
        text = open(os.path.join(ROOT, "include", hdr)).read()
        text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
        for m in re.finditer(r"\b(fgs_[a-z_0-9]+)\s*\(", text):
            names.add(m.group(1))
    return names


def test_library_exports_every_declared_symbol():
    lib = L.load()
    decl = declared_functions()
    assert len(decl) >= 25
    for name in sorted(decl):
        assert hasattr(lib, name), f"libfgs.so does not export {name}"
    # and the Python binding table covers exactly the header
    assert set(L.ABI) == decl


def test_no_cpu_fallback_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(fgs.FgsError) as ei:
        fgs.Engine(device=0)
    assert ei.value.code == L.FGS_E_CUDA
    assert "no CPU fallback" in str(ei.value)


def test_product_does_not_touch_the_oracle():
    """only tests/, __graft_entry__.smoke() and bench.py's baseline legs may use oracle/"""
    pkg = os.path.join(ROOT, "flex-gpu-scheduler_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h", "Makefile")):
                text = open(os.path.join(dirpath, f), errors="ignore").read()
                assert "fgs_oracle" not in text and "ref_walk" not in text and "from oracle" not in text and "import oracle" not in text, f


def test_synth_is_deterministic_and_consistent():
    a = fgs.config_synth(2, n_nodes=200, n_pending=300)
    b = fgs.config_synth(2, n_nodes=200, n_pending=300)
    sa, sb = a.snapshot.contents, b.snapshot.contents
    N, S = 200, 2
    al_a = np.ctypeslib.as_array(sa.alloc, shape=(N, 4 + S))
    al_b = np.ctypeslib.as_array(sb.alloc, shape=(N, 4 + S))
    assert np.array_equal(al_a, al_b)
    rq = np.ctypeslib.as_array(sa.requested, shape=(N, 3 + S))
    off = np.ctypeslib.as_array(sa.pod_off, shape=(N + 1,))
    assert off[0] == 0 and np.all(np.diff(off) >= 0) and off[-1] == sa.pods.n_pods
    # node aggregates equal the sum over the node's pods (AddPod, framework/types.go:583-609) except on the
    # deliberately over-committed nodes
    pk = fgs.Packer(2)
    pt = pk.pods(C.pointer(sa.pods)).contents
    req_cpu = np.ctypeslib.as_array(pt.req_mcpu, shape=(int(sa.pods.n_pods),))
    sums = np.array([req_cpu[off[n]:off[n + 1]].sum() for n in range(N)])
    same = sums == rq[:, 0]
    assert same.mean() > 0.9
    assert np.all(rq[~same, 0] > al_a[~same, 0])  # the over-committed ones


def test_packer_flex_columns_by_hand():
    # gpu_node.go:67-120 on a hand-made node: used per device, monopoly mask, limit sums
    pods = [
        B.pod(containers=[B.ctr({}, mem_limit=300)], node_name=True, index=1),
        B.pod(containers=[B.ctr({}, mem_limit=200), B.ctr({}, mem_limit=50)], node_name=True, index=1),  # limits add up per pod
        B.pod(containers=[B.ctr({}, gpu_limit=1)], node_name=True, index=0),
        B.pod(containers=[B.ctr({}, mem_limit=999)], node_name=True, index=None),                       # no annotation: sum only
        B.pod(containers=[B.ctr({"cpu": 500})], node_name=True, index=3),                                # no limits: ignored
        B.pod(containers=[B.ctr({})], init_containers=[B.ctr({}, mem_limit=7)], node_name=True, index=2), # init containers do not count
    ]
    snap = B.build_snapshot([B.node(alloc={"cpu": 8000, "memory": 1 << 36, "pods": 110, ("scalar", 0): 4, ("scalar", 1): 4000}, pods=pods),
                             B.node(alloc={"cpu": 8000, "memory": 1 << 36, "pods": 110})])
    t = fgs.Packer(2).nodes(snap.ptr).contents
    assert t.n_nodes == 2 and t.n_devices == 4
    used = np.ctypeslib.as_array(t.used_mem, shape=(4, 2))
    assert used[:, 0].tolist() == [0, 550, 0, 0]
    assert t.monopoly_mask[0] == 0b0001
    assert t.flex_gpu_sum[0] == 1 and t.flex_mem_sum[0] == 300 + 250 + 999
    assert t.num_pods[0] == 6 and t.num_pods[1] == 0
    assert t.flags[0] & L.FGS_PL_FLEXGPU  # FGS_NF_VALID == 1
    NF_HAS_FLEX, NF_REF_PANIC = 2, 4
    assert t.flags[0] & NF_HAS_FLEX and not (t.flags[0] & NF_REF_PANIC)
    assert not (t.flags[1] & NF_HAS_FLEX) and (t.flags[1] & NF_REF_PANIC)  # no gpus → NewGPUNode divides by zero


def test_packer_ranges_and_validation():
    s = fgs.config_synth(2, n_nodes=50, n_pending=40)
    pk = fgs.Packer(2)
    full = pk.nodes(s.snapshot)
    alloc_full = np.ctypeslib.as_array(full.contents.alloc_mcpu, shape=(50,)).copy()
    part = fgs.Packer(2).nodes(s.snapshot, 10, 25)
    assert part.contents.n_nodes == 15
    assert np.array_equal(np.ctypeslib.as_array(part.contents.alloc_mcpu, shape=(15,)), alloc_full[10:25])
    pp = fgs.Packer(2).pods(s.pending, 5, 17)
    assert pp.contents.n_pods == 12
    with pytest.raises(fgs.FgsError):
        fgs.Packer(2).nodes(s.snapshot, 10, 51)
    with pytest.raises(fgs.FgsError):
        fgs.Packer(3).nodes(s.snapshot)  # scalar-slot count mismatch
    # quantities at or beyond 2^53 are rejected (FGS_E_RANGE), like negative pod requests
    big = B.build_snapshot([B.node(alloc={"cpu": 1 << 53, "memory": 1, "pods": 1})])
    with pytest.raises(fgs.FgsError) as ei:
        fgs.Packer(2).nodes(big.ptr)
    assert ei.value.code == L.FGS_E_RANGE
    neg = B.build_podset([B.pod(containers=[B.ctr({"cpu": -5})])])
    with pytest.raises(fgs.FgsError) as ei:
        fgs.Packer(2).pods(neg.ptr)
    assert ei.value.code == L.FGS_E_RANGE
    # more devices than the engine supports
    many = B.build_snapshot([B.node(alloc={"cpu": 1, "memory": 1, "pods": 1, ("scalar", 0): 17, ("scalar", 1): 17})])
    with pytest.raises(fgs.FgsError) as ei:
        fgs.Packer(2).nodes(many.ptr)
    assert ei.value.code == L.FGS_E_TOO_LARGE


def test_pod_table_matches_reference_request_rules():
    # computePodResourceRequest (fit.go:162-180): max(Σ containers, each init container) + overhead;
    # scoring request (resource_allocation.go:119-143): non-zero defaults per container, overhead cpu added as
    # whole cores (the reference adds Quantity.Value() to a milli-core sum)
    Mi = 1 << 20
    pods = [
        B.pod(containers=[B.ctr({"cpu": 300, "memory": 100 * Mi}), B.ctr({"memory": 50 * Mi})],
              init_containers=[B.ctr({"cpu": 1000})], overhead={"cpu": 250, "memory": 10 * Mi}),
        B.pod(containers=[B.ctr({})]),
        B.pod(containers=[B.ctr({("scalar", 0): 0}, gpu_limit=0)]),
    ]
    t = fgs.Packer(2).pods(B.build_podset(pods).ptr).contents
    assert t.req_mcpu[0] == max(300, 1000) + 250
    assert t.req_mem[0] == 150 * Mi + 10 * Mi
    assert t.nz_mcpu[0] == max(300 + 100, 1000) + 1        # container 2 has no cpu request → 100m default; +ceil(0.25 cores)
    assert t.nz_mem[0] == max(150 * Mi, 200 * Mi) + 10 * Mi   # the init container has no memory request → 200Mi default
    PF_GPU, PF_MEM, PF_NONE = 1, 2, 4
    assert t.flags[1] == PF_NONE and t.nz_mcpu[1] == 100 and t.nz_mem[1] == 200 * Mi
    assert t.flags[2] == PF_GPU                              # a zero-valued scalar key still makes the request non-empty
    assert t.req_scalar_present[2] == 1


def test_division_by_invariant_multipliers_is_exact(tmp_path):
    """The int32 emit pass divides by multiplication (kernels_narrow.cuh: narrow_magic): the host-side function that fills
    the device tables, checked against '/' over every small divisor, the neighbours of every power of two and random
    31-bit divisors (tests/cpp/magic_test.cpp, ~7.5e7 quotients)."""
    import subprocess
    exe = str(tmp_path / "magic_test")
    cmd = ["g++", "-O2", "-std=c++17", "-I", os.path.join(ROOT, "flex-gpu-scheduler_b200", "csrc"), "-I", os.path.join(ROOT, "include"),
           os.path.join(ROOT, "tests", "cpp", "magic_test.cpp"), "-o", exe]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
    r = subprocess.run([exe], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and "0 failed" in r.stdout, r.stdout[-2000:]


def test_ctypes_mirrors_match_the_c_headers(tmp_path):
    """Every struct that crosses the C ABI exists twice: in include/*.h and as a ctypes.Structure in _lib.py. A C program
    prints sizeof and the offset of every field of the header's version; both must agree with the Python mirror."""
    import subprocess
    from flex_gpu_scheduler_b200 import _lib as L
    pairs = {"fgs_podset": L.Podset, "fgs_snapshot": L.Snapshot, "fgs_podgroups": L.Podgroups, "fgs_quotas": L.Quotas,
             "fgs_plugin_cfg": L.PluginCfg, "fgs_node_table": L.NodeTable, "fgs_pod_table": L.PodTable, "fgs_config": L.Config,
             "fgs_results": L.Results, "fgs_eval_opts": L.EvalOpts, "fgs_device_view": L.DeviceView, "fgs_peer_group": L.PeerGroup,
             "fgs_timings": L.Timings, "fgs_seq_opts": L.SeqOpts, "fgs_seq_results": L.SeqResults, "fgs_synth_params": L.SynthParams}
    lines = ['#include <stdio.h>', '#include <stddef.h>', '#include "fgs.h"', 'int main(void) {']
    for cname, py in pairs.items():
        lines.append(f'  printf("{cname} size %zu\\n", sizeof({cname}));')
        for fname, *_ in py._fields_:
            lines.append(f'  printf("{cname} {fname} %zu\\n", offsetof({cname}, {fname}));')
    lines += ['  return 0;', '}']
    src = tmp_path / "layout.c"
    src.write_text("\n".join(lines))
    exe = str(tmp_path / "layout")
    r = subprocess.run(["gcc", "-std=c11", "-I", os.path.join(ROOT, "include"), str(src), "-o", exe], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stderr[-3000:]  # a field the header does not have fails here
    out = subprocess.run([exe], capture_output=True, text=True, timeout=60).stdout
    for line in out.splitlines():
        cname, what, val = line.split()
        py = pairs[cname]
        if what == "size":
            assert C.sizeof(py) == int(val), (cname, C.sizeof(py), val)
        else:
            assert getattr(py, what).offset == int(val), (cname, what, getattr(py, what).offset, val)
    # … and the header has no field the mirror lacks (sizes equal and every mirrored offset matches ⇒ same layout)


def test_cpp_host_layer_compiles():
    """include/fgs_plugins.hpp (the C++ mirror of the Go plugin interface) and its test program compile without a GPU;
    tests/test_gpu_cpp_plugins.py builds and RUNS them on one."""
    import subprocess
    cmd = ["g++", "-std=c++17", "-fsyntax-only", "-Wall", "-Wextra", "-I", os.path.join(ROOT, "include"),
           "-I", os.path.join(ROOT, "flex-gpu-scheduler_b200", "csrc"), "-I", os.path.join(ROOT, "tests", "cpp"),
           os.path.join(ROOT, "tests", "cpp", "plugin_test.cpp")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and "warning" not in r.stderr, r.stderr[-3000:]
