"""literal_walk.py — a SECOND, independent restatement of the reference's plugins on this path. TEST INFRASTRUCTURE ONLY.

Why it exists: pkg/flexgpu ships no tests and the in-tree NodeResourcesFit / DefaultNormalizeScore / selectHost are
vendored without theirs, so nothing in the reference can pin oracle/ref_walk.cpp for them. This file was written from
the Go sources alone, WITHOUT consulting ref_walk.cpp, as literally as Python allows: objects with identity where Go has
pointers (so `assumed := u.usedMemory; assumed.Add(...)` mutates the device and `remain := u.memory` mutates the
Quantity all devices of a node share), dicts where Go has maps, arbitrary-precision ints standing in for
resource.Quantity (no quantity on this path leaves int64), and Go's truncating '/'. tests/test_oracle_cross.py fuzzes
the two restatements against each other. Two readings that agree are still not a pin: PARITY UNPINNED, by definition.

Only tests/ and oracle/cross_fuzz.py import this module; nothing under flex-gpu-scheduler_b200/ does.

Go sources followed (paths relative to the reference repo; sched/ = vendor/k8s.io/kubernetes/pkg/scheduler/):
  pkg/flexgpu/flex_gpu.go:41-223            Filter, podResourceLimit, nodeResourceLimitSum, Score, NormalizeScore, Reserve
  pkg/flexgpu/gpu_node.go:30-199            NewGPUNode, constructGPUs, MemAssumeFitIndexes, GPUAssumeFitIndexes, GPUScore, MemScore
  sched/framework/plugins/noderesources/fit.go:162-180,221-238,256-329,332-339
  sched/framework/plugins/noderesources/resource_allocation.go:49-143
  sched/framework/plugins/noderesources/least_allocated.go:29-57, most_allocated.go:29-61
  sched/framework/plugins/helper/normalize_score.go:26-54
  sched/util/pod_resources.go:25-85
  sched/framework/types.go:429-454,487-509  Resource.Add / SetMaxResource
  sched/framework/runtime/framework.go:405-408,667-693,861-933
  sched/generic_scheduler.go:93-167,414-447
  pkg/noderesources/allocatable.go:61-72,120-171, pkg/noderesources/resource_allocation.go:49-120
"""
import numpy as np

GPUResourceName = "nvidia.flex.com/gpu"
MemResourceName = "nvidia.flex.com/memory"
GPUIndexAnnotationKey = "nvidia.flex.com/index"
CPU, MEMORY, EPHEMERAL, PODS = "cpu", "memory", "ephemeral-storage", "pods"
MaxNodeScore, MinNodeScore = 100, 0
DefaultMilliCPURequest, DefaultMemoryRequest = 100, 200 * 1024 * 1024
Success, Error, Unschedulable, UnschedulableAndUnresolvable = 0, 1, 2, 3
PL_FLEXGPU, PL_FIT, PL_ALLOCATABLE = 1, 2, 4
ANNOT_NONE = -(2 ** 31)
PODERR_SCORE_RANGE, PODERR_REF_PANIC = 1, 2
INT64_MIN, INT64_MAX = -(2 ** 63), 2 ** 63 - 1


class GoPanic(Exception):
    """a run-time panic of the Go process (integer divide by zero, index out of range, makeslice: len out of range)"""


def wrap64(v):
    """int64 arithmetic wraps in Go"""
    v &= (1 << 64) - 1
    return v - (1 << 64) if v >= (1 << 63) else v


def go_div(a, b):
    """Go's '/' on integers truncates toward zero and panics on a zero divisor"""
    if b == 0:
        raise GoPanic("integer divide by zero")
    q = abs(a) // abs(b)
    return q if (a < 0) == (b < 0) else -q


def scalar_name(s):
    return GPUResourceName if s == 0 else MemResourceName if s == 1 else "example.com/scalar-%d" % s


def slot_name(slot, S):
    if slot == 0:
        return CPU
    if slot == 1:
        return MEMORY
    if slot == 2:
        return EPHEMERAL
    if 3 <= slot < 3 + S:
        return scalar_name(slot - 3)
    return "unknown-%d" % slot


def is_scalar_resource_name(name):
    """schedutil.IsScalarResourceName: extended resources, hugepages, native prefixed, attachable volumes — of the
    names this format can carry, exactly the scalar slots (all extended: they have a '/' and are not kubernetes.io/)"""
    return "/" in name


# quantities: a ResourceList maps name → int; cpu is kept in milli-units (what MilliValue() returns), the rest in base
# units. Value() of a cpu quantity rounds UP to whole cores.
def milli_value(name, v):
    assert name == CPU
    return v


def value(name, v):
    return -((-v) // 1000) if name == CPU else v


class Quantity:
    """*resource.Quantity: an object with identity; Add / Sub mutate the receiver"""

    def __init__(self, v=0):
        self.v = v

    def Add(self, o):
        self.v += o.v

    def Sub(self, o):
        self.v -= o.v

    def Cmp(self, o):
        return (self.v > o.v) - (self.v < o.v)

    def CmpInt64(self, x):
        return (self.v > x) - (self.v < x)

    def IsZero(self):
        return self.v == 0

    def AsInt64(self):
        return self.v, True


class Container:
    def __init__(self, requests, limits):
        self.requests = requests  # name → int (cpu in milli)
        self.limits = limits      # flex names only → int


class Pod:
    def __init__(self):
        self.containers, self.init_containers = [], []
        self.overhead = None  # None or name → int
        self.annotations = {}
        self.namespace = 0
        self.priority = 0


class Resource:
    """framework.Resource"""

    def __init__(self):
        self.MilliCPU = self.Memory = self.EphemeralStorage = 0
        self.AllowedPodNumber = 0
        self.ScalarResources = None

    def AddScalar(self, name, q):
        self.SetScalar(name, (self.ScalarResources or {}).get(name, 0) + q)

    def SetScalar(self, name, q):
        if self.ScalarResources is None:
            self.ScalarResources = {}
        self.ScalarResources[name] = q

    def Add(self, rl):  # types.go:429-454
        for rName, rQuant in rl.items():
            if rName == CPU:
                self.MilliCPU += milli_value(rName, rQuant)
            elif rName == MEMORY:
                self.Memory += value(rName, rQuant)
            elif rName == PODS:
                self.AllowedPodNumber += value(rName, rQuant)
            elif rName == EPHEMERAL:
                self.EphemeralStorage += value(rName, rQuant)
            elif is_scalar_resource_name(rName):
                self.AddScalar(rName, value(rName, rQuant))

    def SetMaxResource(self, rl):  # types.go:487-509
        for rName, rQuantity in rl.items():
            if rName == MEMORY:
                self.Memory = max(self.Memory, value(rName, rQuantity))
            elif rName == CPU:
                self.MilliCPU = max(self.MilliCPU, milli_value(rName, rQuantity))
            elif rName == EPHEMERAL:
                self.EphemeralStorage = max(self.EphemeralStorage, value(rName, rQuantity))
            elif is_scalar_resource_name(rName):
                self.SetScalar(rName, max((self.ScalarResources or {}).get(rName, 0), value(rName, rQuantity)))

    def scalar(self, name):
        return (self.ScalarResources or {}).get(name, 0)


class Node:
    def __init__(self, status_allocatable):
        self.status_allocatable = status_allocatable  # name → int


class NodeInfo:
    def __init__(self):
        self.node = None
        self.Pods = []
        self.Allocatable, self.Requested, self.NonZeroRequested = Resource(), Resource(), Resource()


# ---------------------------------------------------------------- the flat snapshot → objects ------------------------

def _arr(ptr, n, dtype):
    if n <= 0 or not ptr:
        return np.zeros(0, dtype)
    return np.ctypeslib.as_array(ptr, shape=(n,)).astype(dtype, copy=True)


def podset_to_pods(ps):
    S, n = int(ps.n_scalars), int(ps.n_pods)
    W = 3 + S
    C = int(ps.n_ctrs)
    ctr_off = _arr(ps.ctr_off, n + 1, np.int64)
    if n:
        C = max(C, int(ctr_off[n]))
    ns, prio, annot, n_init = (_arr(ps.ns, n, np.int64), _arr(ps.priority, n, np.int64), _arr(ps.index_annot, n, np.int64),
                               _arr(ps.n_init, n, np.int64))
    ovh_p = _arr(ps.overhead_present, n, np.uint32)
    ovh = _arr(ps.overhead, n * W, np.int64).reshape(-1, W) if ps.overhead else np.zeros((n, W), np.int64)
    c_req = _arr(ps.c_req, C * W, np.int64).reshape(-1, W)
    c_pres = _arr(ps.c_req_present, C, np.uint32)
    c_lim = _arr(ps.c_lim_flex, C * 2, np.int64).reshape(-1, 2)
    c_lp = _arr(ps.c_lim_present, C, np.uint32)
    out = []
    for i in range(n):
        p = Pod()
        p.namespace, p.priority = int(ns[i]), int(prio[i])
        if int(annot[i]) != ANNOT_NONE:
            p.annotations[GPUIndexAnnotationKey] = str(int(annot[i]))
        lo, hi = int(ctr_off[i]), int(ctr_off[i + 1])
        first_init = hi - int(n_init[i])
        for c in range(lo, hi):
            req = {slot_name(r, S): int(c_req[c, r]) for r in range(W) if (int(c_pres[c]) >> r) & 1}
            lim = {}
            if int(c_lp[c]) & 1:
                lim[GPUResourceName] = int(c_lim[c, 0])
            if int(c_lp[c]) & 2:
                lim[MemResourceName] = int(c_lim[c, 1])
            (p.containers if c < first_init else p.init_containers).append(Container(req, lim))
        if int(ovh_p[i]) & 0x80000000:
            p.overhead = {slot_name(r, S): int(ovh[i, r]) for r in range(W) if (int(ovh_p[i]) >> r) & 1}
        out.append(p)
    return out


def snapshot_to_nodeinfos(sn):
    S, N = int(sn.n_scalars), int(sn.n_nodes)
    AW, RW = 4 + S, 3 + S
    valid = _arr(sn.node_valid, N, np.uint8)
    alloc = _arr(sn.alloc, N * AW, np.int64).reshape(-1, AW)
    present = _arr(sn.alloc_scalar_present, N, np.uint32)
    req = _arr(sn.requested, N * RW, np.int64).reshape(-1, RW)
    nz = _arr(sn.nonzero, N * 2, np.int64).reshape(-1, 2)
    pod_off = _arr(sn.pod_off, N + 1, np.int64)
    pods = podset_to_pods(sn.pods)
    out = []
    for n in range(N):
        ni = NodeInfo()
        ni.Allocatable.MilliCPU, ni.Allocatable.Memory, ni.Allocatable.EphemeralStorage = int(alloc[n, 0]), int(alloc[n, 1]), int(alloc[n, 2])
        ni.Allocatable.AllowedPodNumber = int(alloc[n, 3])
        status_alloc = {CPU: int(alloc[n, 0]), MEMORY: int(alloc[n, 1]), EPHEMERAL: int(alloc[n, 2]), PODS: int(alloc[n, 3])}
        for s in range(S):
            if (int(present[n]) >> s) & 1:
                ni.Allocatable.SetScalar(scalar_name(s), int(alloc[n, 4 + s]))
                status_alloc[scalar_name(s)] = int(alloc[n, 4 + s])
        ni.Requested.MilliCPU, ni.Requested.Memory, ni.Requested.EphemeralStorage = int(req[n, 0]), int(req[n, 1]), int(req[n, 2])
        for s in range(S):
            if int(req[n, 3 + s]) != 0:
                ni.Requested.SetScalar(scalar_name(s), int(req[n, 3 + s]))
        ni.NonZeroRequested.MilliCPU, ni.NonZeroRequested.Memory = int(nz[n, 0]), int(nz[n, 1])
        ni.Pods = pods[int(pod_off[n]):int(pod_off[n + 1])]
        if int(valid[n]):
            ni.node = Node(status_alloc)
        out.append(ni)
    return out


# ---------------------------------------------------------------- pkg/flexgpu ------------------------------------------------

def podResourceLimit(name, pod):  # flex_gpu.go:121-131
    exist = False
    resourceLimitSum = Quantity(0)
    for container in pod.containers:
        if name in container.limits:
            exist = True
            resourceLimitSum.Add(Quantity(container.limits[name]))
    return resourceLimitSum, exist


def nodeResourceLimitSum(name, nodeInfo):  # flex_gpu.go:133-140
    resourceLimitSum = Quantity(0)
    for podInfo in nodeInfo.Pods:
        limit, _ = podResourceLimit(name, podInfo)
        resourceLimitSum.Add(limit)
    return resourceLimitSum


class Gpu:
    def __init__(self, index, memory):
        self.index = index
        self.monopoly = False
        self.memory = memory            # the SAME Quantity object for every device of the node (gpu_node.go:66)
        self.usedMemory = Quantity(0)


class GpuNode:
    def __init__(self, gpus):
        self.gpus = gpus

    def MemAssumeFitIndexes(self, memLimit):  # gpu_node.go:122-161
        fits = []
        for u in self.gpus:
            assumed = u.usedMemory          # a pointer copy: the Add below changes the device's usedMemory
            assumed.Add(memLimit)
            if not u.monopoly and u.memory.Cmp(assumed) >= 0:
                remain = u.memory           # a pointer copy of the Quantity ALL devices share: Sub changes it for the rest
                remain.Sub(assumed)
                fits.append((u.index, remain))
        # sort.Slice by remain: every element holds the same Quantity, so `less` is false for every pair; up to 12
        # elements Go 1.17 runs insertion sort, which then moves nothing
        if len(fits) > 12:
            raise NotImplementedError("sort.Slice beyond its insertion-sort range with all keys equal")
        return [f[0] for f in fits]

    def GPUAssumeFitIndexes(self, gpuLimit):  # gpu_node.go:163-177
        return [u.index for u in self.gpus if not u.monopoly and u.usedMemory.IsZero()]

    def GPUScore(self):  # gpu_node.go:179-187
        cnt = 0
        for u in self.gpus:
            if not u.monopoly and u.usedMemory.IsZero():
                cnt += 1
        return cnt

    def MemScore(self):  # gpu_node.go:189-199
        remainSum = Quantity(0)
        for u in self.gpus:
            remainSum.Add(u.memory)
            remainSum.Sub(u.usedMemory)
        return remainSum.AsInt64()[0]


def go_atoi(s):
    """strconv.Atoi: optional sign, decimal digits, nothing else; the empty string is an error"""
    t = s[1:] if s[:1] in "+-" else s
    if not t or not t.isdigit() or not t.isascii():
        return 0, True
    return int(s), False


def constructGPUs(cnt, memEachGPU, pods):  # gpu_node.go:70-120
    if cnt < 0:
        raise GoPanic("makeslice: len out of range")
    gpus = [Gpu(i, memEachGPU) for i in range(cnt)]
    for pod in pods:
        gpuLimit, gpuExist = podResourceLimit(GPUResourceName, pod)
        memLimit, memExist = podResourceLimit(MemResourceName, pod)
        if not gpuExist and not memExist:
            continue
        val = pod.annotations.get(GPUIndexAnnotationKey, "")
        hasAnnotation = GPUIndexAnnotationKey in pod.annotations
        index, err = go_atoi(val)
        if err:
            continue
        if not hasAnnotation:
            continue
        if gpuExist:
            if not 0 <= index < len(gpus):
                raise GoPanic("index out of range")
            gpus[index].monopoly = True
        if memExist:
            if not 0 <= index < len(gpus):
                raise GoPanic("index out of range")
            gpus[index].usedMemory.Add(memLimit)
    return gpus


def NewGPUNode(nodeInfo):  # gpu_node.go:30-68
    pods = list(nodeInfo.Pods)
    gpuCnt = nodeInfo.node.status_allocatable.get(GPUResourceName, 0)  # a missing key reads as the zero Quantity
    memCnt = nodeInfo.node.status_allocatable.get(MemResourceName, 0)
    memEachGPU = Quantity(go_div(memCnt, gpuCnt))
    return GpuNode(constructGPUs(int(gpuCnt), memEachGPU, pods))


def FlexGPU_Filter(pod, nodeInfo):  # flex_gpu.go:41-119 → None (Success) or (code, message)
    gpuLimit, gpuLimitExist = podResourceLimit(GPUResourceName, pod)
    memLimit, memLimitExist = podResourceLimit(MemResourceName, pod)
    if not gpuLimitExist and not memLimitExist:
        return None
    if gpuLimitExist and memLimitExist:
        return UnschedulableAndUnresolvable, "pod conflict resources"
    alloc = nodeInfo.node.status_allocatable
    if GPUResourceName not in alloc:
        return UnschedulableAndUnresolvable, "unknown resource type"
    gpuAllocatable = Quantity(alloc[GPUResourceName])
    if MemResourceName not in alloc:
        return UnschedulableAndUnresolvable, "unknown resource type"
    memAllocatable = Quantity(alloc[MemResourceName])
    gpuCurrLimitSum = nodeResourceLimitSum(GPUResourceName, nodeInfo)
    memCurrLimitSum = nodeResourceLimitSum(MemResourceName, nodeInfo)
    gpuAssumeLimitSum = gpuCurrLimitSum
    memAssumeLimitSum = memCurrLimitSum
    gpuAssumeLimitSum.Add(gpuLimit)
    memAssumeLimitSum.Add(memLimit)
    if gpuAssumeLimitSum.Cmp(gpuAllocatable) > 0:
        return Unschedulable, "insufficient resource %s" % GPUResourceName
    if memAssumeLimitSum.Cmp(memAllocatable) > 0:
        return Unschedulable, "insufficient resource %s" % MemResourceName
    no = NewGPUNode(nodeInfo)
    if memLimitExist:
        indexes = no.MemAssumeFitIndexes(memLimit)
        if len(indexes) == 0:
            return Unschedulable, "no fit indexes resource %s" % MemResourceName
    return None


def FlexGPU_Score(pod, nodeInfo):  # flex_gpu.go:142-166
    no = NewGPUNode(nodeInfo)
    _, gpuLimitExist = podResourceLimit(GPUResourceName, pod)
    _, memLimitExist = podResourceLimit(MemResourceName, pod)
    if not gpuLimitExist and not memLimitExist:
        return 0
    if gpuLimitExist:
        return no.GPUScore()
    if memLimitExist:
        return no.MemScore()
    return 0


def FlexGPU_Reserve(pod, nodeInfo):
    """flex_gpu.go:178-223 → ("none" | "conflict" | "fail" | "ok", index)"""
    gpuLimit, gpuLimitExist = podResourceLimit(GPUResourceName, pod)
    memLimit, memLimitExist = podResourceLimit(MemResourceName, pod)
    if not gpuLimitExist and not memLimitExist:
        return "none", None
    if gpuLimitExist and memLimitExist:
        return "conflict", None
    no = NewGPUNode(nodeInfo)
    index = None
    if gpuLimitExist:
        indexes = no.GPUAssumeFitIndexes(gpuLimit)
        if len(indexes) == 0:
            return "fail", None
        index = indexes[0]
    if memLimitExist:
        indexes = no.MemAssumeFitIndexes(memLimit)
        if len(indexes) == 0:
            return "fail", None
        index = indexes[0]
    return "ok", index


def DefaultNormalizeScore(maxPriority, reverse, scores):  # normalize_score.go:26-54, in place
    maxCount = 0
    for s in scores:
        if s > maxCount:
            maxCount = s
    if maxCount == 0:
        if reverse:
            for i in range(len(scores)):
                scores[i] = maxPriority
        return
    for i in range(len(scores)):
        score = scores[i]
        score = go_div(wrap64(maxPriority * score), maxCount)
        if reverse:
            score = maxPriority - score
        scores[i] = score


# ---------------------------------------------------------------- NodeResourcesFit --------------------------------------

def computePodResourceRequest(pod):  # fit.go:162-180 (PodOverhead is on by default in this release)
    result = Resource()
    for container in pod.containers:
        result.Add(container.requests)
    for container in pod.init_containers:
        result.SetMaxResource(container.requests)
    if pod.overhead is not None:
        result.Add(pod.overhead)
    return result


def fitsRequest(podRequest, nodeInfo, ignoredExtendedResources):  # fit.go:256-329 → list of reasons
    insufficient = []
    allowedPodNumber = nodeInfo.Allocatable.AllowedPodNumber
    if len(nodeInfo.Pods) + 1 > allowedPodNumber:
        insufficient.append("Too many pods")
    if (podRequest.MilliCPU == 0 and podRequest.Memory == 0 and podRequest.EphemeralStorage == 0 and
            len(podRequest.ScalarResources or {}) == 0):
        return insufficient
    if podRequest.MilliCPU > nodeInfo.Allocatable.MilliCPU - nodeInfo.Requested.MilliCPU:
        insufficient.append("Insufficient cpu")
    if podRequest.Memory > nodeInfo.Allocatable.Memory - nodeInfo.Requested.Memory:
        insufficient.append("Insufficient memory")
    if podRequest.EphemeralStorage > nodeInfo.Allocatable.EphemeralStorage - nodeInfo.Requested.EphemeralStorage:
        insufficient.append("Insufficient ephemeral-storage")
    for rName, rQuant in (podRequest.ScalarResources or {}).items():
        if rName in ignoredExtendedResources:  # every scalar of this format is an extended resource name
            continue
        if rQuant > nodeInfo.Allocatable.scalar(rName) - nodeInfo.Requested.scalar(rName):
            insufficient.append("Insufficient %s" % rName)
    return insufficient


def GetRequestForResource(resource, requests, nonZero):  # util/pod_resources.go:53-85
    if resource == CPU:
        if CPU not in requests and nonZero:
            return DefaultMilliCPURequest
        return milli_value(CPU, requests.get(CPU, 0))
    if resource == MEMORY:
        if MEMORY not in requests and nonZero:
            return DefaultMemoryRequest
        return value(MEMORY, requests.get(MEMORY, 0))
    if resource not in requests:
        return 0
    return value(resource, requests[resource])


class FitScorer:
    """resourceAllocationScorer of NodeResourcesFit (resource_allocation.go:33-143) with the Least / MostAllocated
    scorer; useRequested is false for both strategies"""

    def __init__(self, strategy, resourceToWeightMap):
        self.strategy = strategy
        self.resourceToWeightMap = resourceToWeightMap
        self.useRequested = False

    def calculatePodResourceRequest(self, pod, resource):  # :119-143
        podRequest = 0
        for container in pod.containers:
            podRequest += GetRequestForResource(resource, container.requests, not self.useRequested)
        for initContainer in pod.init_containers:
            v = GetRequestForResource(resource, initContainer.requests, not self.useRequested)
            if podRequest < v:
                podRequest = v
        if pod.overhead is not None:
            if resource in pod.overhead:
                podRequest += value(resource, pod.overhead[resource])  # quantity.Value(), also for cpu: whole cores
        return podRequest

    def calculateResourceAllocatableRequest(self, nodeInfo, pod, resource):  # :84-114
        requested = nodeInfo.Requested if self.useRequested else nodeInfo.NonZeroRequested
        podRequest = self.calculatePodResourceRequest(pod, resource)
        if podRequest == 0 and is_scalar_resource_name(resource):
            return 0, 0
        if resource == CPU:
            return nodeInfo.Allocatable.MilliCPU, requested.MilliCPU + podRequest
        if resource == MEMORY:
            return nodeInfo.Allocatable.Memory, requested.Memory + podRequest
        if resource == EPHEMERAL:
            return nodeInfo.Allocatable.EphemeralStorage, nodeInfo.Requested.EphemeralStorage + podRequest
        if resource in (nodeInfo.Allocatable.ScalarResources or {}):
            return nodeInfo.Allocatable.ScalarResources[resource], nodeInfo.Requested.scalar(resource) + podRequest
        return 0, 0

    def score(self, pod, nodeInfo):  # :49-81
        requested, allocatable = {}, {}
        for resource in self.resourceToWeightMap:
            alloc, req = self.calculateResourceAllocatableRequest(nodeInfo, pod, resource)
            if alloc != 0:
                allocatable[resource], requested[resource] = alloc, req
        nodeScore = weightSum = 0
        for resource in requested:
            weight = self.resourceToWeightMap[resource]
            if self.strategy == 0:
                resourceScore = leastRequestedScore(requested[resource], allocatable[resource])
            else:
                resourceScore = mostRequestedScore(requested[resource], allocatable[resource])
            nodeScore += resourceScore * weight
            weightSum += weight
        if weightSum == 0:
            return 0
        return go_div(nodeScore, weightSum)


def leastRequestedScore(requested, capacity):  # least_allocated.go:49-57
    if capacity == 0:
        return 0
    if requested > capacity:
        return 0
    return go_div(wrap64((capacity - requested) * MaxNodeScore), capacity)


def mostRequestedScore(requested, capacity):  # most_allocated.go:50-61
    if capacity == 0:
        return 0
    if requested > capacity:
        requested = capacity
    return go_div(wrap64(requested * MaxNodeScore), capacity)


# ---------------------------------------------------------------- NodeResourcesAllocatable -------------------------------

class AllocatableScorer:
    """pkg/noderesources/allocatable.go + resource_allocation.go"""

    def __init__(self, mode, resourceToWeightMap):
        self.mode = mode
        self.resourceToWeightMap = resourceToWeightMap

    def score(self, pod, nodeInfo):
        allocatable = {}
        for resource in self.resourceToWeightMap:
            if resource == CPU:
                allocatable[resource] = nodeInfo.Allocatable.MilliCPU
            elif resource == MEMORY:
                allocatable[resource] = nodeInfo.Allocatable.Memory
            elif resource == EPHEMERAL:
                allocatable[resource] = nodeInfo.Allocatable.EphemeralStorage
            elif is_scalar_resource_name(resource):
                allocatable[resource] = nodeInfo.Allocatable.scalar(resource)
            else:
                allocatable[resource] = 0
        nodeScore = weightSum = 0
        for resource, weight in self.resourceToWeightMap.items():
            cap = allocatable[resource]
            resourceScore = wrap64(-1 * cap) if self.mode == 0 else cap
            nodeScore = wrap64(nodeScore + wrap64(resourceScore * weight))
            weightSum = wrap64(weightSum + weight)
        return go_div(nodeScore, weightSum)

    @staticmethod
    def normalize(scores):  # allocatable.go:146-171, in place
        highest, lowest = -INT64_MAX, INT64_MAX
        for s in scores:
            if s > highest:
                highest = s
            if s < lowest:
                lowest = s
        oldRange = wrap64(highest - lowest)
        newRange = MaxNodeScore - MinNodeScore
        for i, s in enumerate(scores):
            if oldRange == 0:
                scores[i] = MinNodeScore
            else:
                scores[i] = wrap64(go_div(wrap64(wrap64(s - lowest) * newRange), oldRange) + MinNodeScore)


# ---------------------------------------------------------------- the framework around the plugins ------------------------

class Profile:
    """what fgs_plugin_cfg says, in the reference's terms"""

    def __init__(self, cfg, S):
        self.filter_plugins = []
        for k in range(2):
            pid = int(cfg.filter_order[k])
            if pid == 0:
                break
            if int(cfg.filter_mask) & pid and pid not in self.filter_plugins:
                self.filter_plugins.append(pid)
        for pid in (PL_FIT, PL_FLEXGPU):  # enabled but not listed: after the listed ones
            if int(cfg.filter_mask) & pid and pid not in self.filter_plugins:
                self.filter_plugins.append(pid)
        sm = int(cfg.score_mask)
        self.score_plugins = [pid for pid in (PL_FLEXGPU, PL_FIT, PL_ALLOCATABLE) if sm & pid]
        w = {PL_FLEXGPU: int(cfg.weight_flex), PL_FIT: int(cfg.weight_fit), PL_ALLOCATABLE: int(cfg.weight_alloc)}
        self.weight = {pid: (w[pid] if w[pid] != 0 else 1) for pid in w}  # runtime/framework.go:405-408
        fit_map = {}
        for k in range(int(cfg.fit_n_res)):
            fit_map[slot_name(int(cfg.fit_res[k]), S)] = int(cfg.fit_res_w[k])
        self.fit = FitScorer(int(cfg.fit_strategy), fit_map)
        self.ignored = {scalar_name(s) for s in range(S) if (int(cfg.fit_ignored_scalar_mask) >> s) & 1}
        alloc_map = {}
        for k in range(int(cfg.alloc_n_res)):
            slot = int(cfg.alloc_res[k])
            alloc_map["unknown-%d" % k if slot < 0 else slot_name(slot, S)] = int(cfg.alloc_res_w[k])
        self.alloc = AllocatableScorer(int(cfg.alloc_mode), alloc_map)


def RunFilterPlugins(profile, pod, podRequest, nodeInfo):
    """runtime/framework.go:667-693 with runAllFilters = false → (code, plugin id) of the merged status"""
    for pid in profile.filter_plugins:
        if pid == PL_FLEXGPU:
            st = FlexGPU_Filter(pod, nodeInfo)
            code = Success if st is None else st[0]
        else:
            reasons = fitsRequest(podRequest, nodeInfo, profile.ignored)
            code = Unschedulable if reasons else Success
        if code != Success:
            return code, pid
    return Success, 0


class CycleResult:
    pass


def schedule_pod(profile, pod, nodeInfos, score_single=True):
    """generic_scheduler.go:93-167 for one pod against the whole list (percentageOfNodesToScore = 100), without the
    random tie-break: reports the maximum and its tie set. score_single: also score when only one node is feasible
    (the reference returns that node unscored, :120-127) — the convention of the first oracle's planes."""
    N = len(nodeInfos)
    r = CycleResult()
    r.status = np.zeros(N, np.uint8)
    r.filter_panic = np.zeros(N, bool)
    r.pod_err = 0
    podRequest = computePodResourceRequest(pod)  # Fit.PreFilter
    feasible = []
    for n, ni in enumerate(nodeInfos):
        if ni.node is None:  # not a member of the scheduler's node list
            r.status[n] = 0xFF
            continue
        try:
            code, pid = RunFilterPlugins(profile, pod, podRequest, ni)
        except GoPanic:
            r.filter_panic[n] = True
            r.pod_err |= PODERR_REF_PANIC
            r.status[n] = 0xF0
            continue
        r.status[n] = code | (pid << 4)
        if code == Success:
            feasible.append(n)
    r.feasible = feasible
    r.raw = {pid: np.zeros(N, np.int64) for pid in profile.score_plugins}
    r.norm = {pid: np.zeros(N, np.int64) for pid in profile.score_plugins}
    r.total = np.zeros(N, np.int64)
    r.max_total, r.tie_set = INT64_MIN, []
    if not feasible or (len(feasible) == 1 and not score_single):
        return r
    scored = len(feasible) >= 2
    # RunScorePlugins (runtime/framework.go:861-933)
    lists = {}
    for pid in profile.score_plugins:
        sc = []
        for n in feasible:
            ni = nodeInfos[n]
            try:
                if pid == PL_FLEXGPU:
                    s = FlexGPU_Score(pod, ni)
                elif pid == PL_FIT:
                    s = profile.fit.score(pod, ni)
                else:
                    s = profile.alloc.score(pod, ni)
            except GoPanic:
                if scored:
                    r.pod_err |= PODERR_REF_PANIC
                s = 0
            sc.append(s)
        for n, s in zip(feasible, sc):
            r.raw[pid][n] = s
        if pid == PL_FLEXGPU:
            DefaultNormalizeScore(MaxNodeScore, True, sc)
        elif pid == PL_ALLOCATABLE:
            AllocatableScorer.normalize(sc)
        for n, s in zip(feasible, sc):
            r.norm[pid][n] = s
            if (s > MaxNodeScore or s < MinNodeScore) and scored:
                r.pod_err |= PODERR_SCORE_RANGE
        lists[pid] = [wrap64(s * profile.weight[pid]) for s in sc]
    for k, n in enumerate(feasible):  # prioritizeNodes: the per-node sum (generic_scheduler.go:440-447)
        t = 0
        for pid in profile.score_plugins:
            t = wrap64(t + lists[pid][k])
        r.total[n] = t
    # selectHost (:146-167)
    maxScore = r.total[feasible[0]]
    for n in feasible[1:]:
        if r.total[n] > maxScore:
            maxScore = r.total[n]
    r.max_total = int(maxScore)
    r.tie_set = [n for n in feasible if r.total[n] == maxScore]
    return r


def evaluate(snapshot, pending, cfg, p0=0, p1=None):
    """every pending pod in [p0, p1) against every node; snapshot / pending: ctypes pointers to fgs_snapshot / fgs_podset"""
    sn, ps = snapshot.contents, pending.contents
    S = int(sn.n_scalars)
    nodeInfos = snapshot_to_nodeinfos(sn)
    pods = podset_to_pods(ps)
    if p1 is None:
        p1 = len(pods)
    profile = Profile(cfg, S)
    return [schedule_pod(profile, pods[p], nodeInfos) for p in range(p0, p1)]


def reserve_index(snapshot, pending, p, n):
    """FlexGPU.Reserve's outcome for pending pod p on node n, in fgs_oracle_reserve_index's encoding"""
    nodeInfos = snapshot_to_nodeinfos(snapshot.contents)
    pod = podset_to_pods(pending.contents)[p]
    try:
        what, index = FlexGPU_Reserve(pod, nodeInfos[n])
    except GoPanic:
        return -4
    return {"none": -2, "conflict": -3, "fail": -1}.get(what, index)
