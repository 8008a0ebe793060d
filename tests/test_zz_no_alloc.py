"""GPU test of fgs_quota_check_static, the ElasticQuota check for callers that cannot allocate device memory.

Some users of libfgs are not free to call cudaMalloc: a host process that gives the library a fixed device
arena, a plugin on a path that must never block on the CUDA allocator, or code that is recording work at the
time. For them the library offers one entry point that takes every table and the result as device pointers
the caller already owns and keeps its own working memory in a fixed array inside the library (at most 4096
namespaces). This test plays such a caller: torch owns all device buffers, the call goes straight through the
C ABI, and the answer must be identical to fgs_quota_check's — twice in a row, the second time on a smaller
and different quota table, because the entry point's scratch is shared between calls and still holds what the
first call left there.

The file name sorts last on purpose: a run that stops at the first failure has then already run every other test.
"""
import ctypes as C

import numpy as np
import pytest

import flex_gpu_scheduler_b200 as fgs
from flex_gpu_scheduler_b200 import _lib as L

pytestmark = pytest.mark.gpu

S = 3


def _device_tables(torch, quotas, pod_table):
    """Copy the quota table and the pods' request columns into torch-owned device memory."""
    q = quotas.contents
    NS, W = int(q.n_ns), 2 + S
    P = int(pod_table.contents.n_pods)
    pt = pod_table.contents

    def dev(ptr, shape, dtype):
        a = np.ctypeslib.as_array(ptr, shape=shape).astype(dtype, copy=True)
        if dtype == np.uint32:
            a = a.view(np.int32)  # same bits; torch's uint32 support is partial
        return torch.from_numpy(a).cuda()

    t = dict(
        has_eq=dev(q.has_eq, (NS,), np.uint8), qmin=dev(q.min, (NS, W), np.int64), qmax=dev(q.max, (NS, W), np.int64),
        qused=dev(q.used, (NS, W), np.int64), used_present=dev(q.used_scalar_present, (NS,), np.uint32),
        pod_present=dev(pt.req_scalar_present, (P,), np.uint32), pod_ns=dev(pt.ns, (P,), np.int32),
    )
    req = np.empty((W, P), np.int64)
    req[0] = np.ctypeslib.as_array(pt.req_mcpu, shape=(P,))
    req[1] = np.ctypeslib.as_array(pt.req_mem, shape=(P,))
    req[2:] = np.ctypeslib.as_array(pt.req_scalar, shape=(S, P))
    t["pod_req"] = torch.from_numpy(req).cuda()
    t["code"] = torch.full((P,), 0xEE, dtype=torch.uint8, device="cuda")  # poisoned: every byte must be overwritten
    return t, NS, P


def _call_static(lib, torch, t, NS, P):
    err = C.create_string_buffer(256)
    stream = torch.cuda.current_stream().cuda_stream
    rc = lib.fgs_quota_check_static(stream, S, NS, t["has_eq"].data_ptr(), t["qmin"].data_ptr(), t["qmax"].data_ptr(), t["qused"].data_ptr(),
                                    t["used_present"].data_ptr(), P, t["pod_req"].data_ptr(), t["pod_present"].data_ptr(),
                                    t["pod_ns"].data_ptr(), t["code"].data_ptr(), err, 256)
    assert rc == 0, err.value.decode()
    torch.cuda.synchronize()
    return t["code"].cpu().numpy()


def test_static_quota_check_matches_engine_twice(gpu_ok, oracle):
    import torch
    lib = L.load()
    eng = fgs.Engine(device=0, n_scalars=S)
    # two different problems, the second with fewer namespaces than the first: whatever the first call leaves in
    # the library's fixed scratch must not leak into the second answer
    cases = [dict(n_namespaces=300, n_pending=20000, seed_pods=0xF6510004), dict(n_namespaces=37, n_pending=6000, seed_pods=0xABCD)]
    for kw in cases:
        s = fgs.config_synth(4, n_nodes=400, n_scalars=S, pm_extra_scalar=300, **kw)
        pk = fgs.Packer(S)
        table = pk.pods(s.pending)
        want = eng.quota_check(table, s.quotas)  # the existing entry point (allocates its own arena)
        t, NS, P = _device_tables(torch, s.quotas, table)
        got = _call_static(lib, torch, t, NS, P)
        assert np.array_equal(got, want)  # integer verdicts: exact equality is the tolerance
        assert np.array_equal(got, oracle.quota_prefilter(s.pending, s.quotas))  # … and the CPU restatement of the Go plugin agrees
        assert set(np.unique(want)) >= {0, 1}
    # a third call on a hand-made table that reaches every verdict, the "aggregated used over the sum of Min" one
    # included (capacity_scheduling.go:262-272), which the synthetic quotas above never produce
    import builders as B
    GPU = ("scalar", 0)
    quotas = B.build_quotas({0: dict(min={"cpu": 1000, GPU: 2}, max={"cpu": 5000, GPU: 8}, used={"cpu": 900, GPU: 2}),
                             1: dict(min={"cpu": 1000, GPU: 2}, max={"cpu": 1200, GPU: 8}, used={"cpu": 900, GPU: 1})}, n_ns=4, S=S)
    pods = [B.pod([B.ctr({"cpu": 100})], ns=0),            # stays within its Min: Success
            B.pod([B.ctr({"cpu": 300})], ns=0),            # over its Min and the cluster's Min is spent: code 2
            B.pod([B.ctr({"cpu": 400})], ns=1),            # over its Max: code 1
            B.pod([B.ctr({"cpu": 50, GPU: 1})], ns=0),     # the scalar key: used 2 + 1 over Min 2, aggregate 4 > 4 is false … cpu 950 <= 1000
            B.pod([B.ctr({"cpu": 50, GPU: 2})], ns=0),     # aggregate gpu 3 + 2 > 4: code 2
            B.pod([B.ctr({"cpu": 10 ** 6})], ns=3)]        # a namespace without an ElasticQuota: Success
    pend = B.build_podset(pods, S)
    table = fgs.Packer(S).pods(pend.ptr)
    want = eng.quota_check(table, quotas.ptr)
    ref = oracle.quota_prefilter(pend.ptr, quotas.ptr)
    assert np.array_equal(want, ref)
    assert 2 in set(ref) and 1 in set(ref) and 0 in set(ref), list(ref)
    t, NS, P = _device_tables(torch, quotas.ptr, table)
    got = _call_static(lib, torch, t, NS, P)
    assert np.array_equal(got, ref)
    eng.close()


def test_static_quota_check_refuses_oversized_input(gpu_ok):
    import torch
    lib = L.load()
    err = C.create_string_buffer(256)
    dummy = torch.zeros(64, dtype=torch.int64, device="cuda")
    p = dummy.data_ptr()
    rc = lib.fgs_quota_check_static(None, S, 4097, p, p, p, p, p, 1, p, p, p, p, err, 256)
    assert rc == L.FGS_E_TOO_LARGE and b"4096" in err.value
    rc = lib.fgs_quota_check_static(None, 9, 10, p, p, p, p, p, 1, p, p, p, p, err, 256)
    assert rc == L.FGS_E_INVALID
