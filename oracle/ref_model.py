"""oracle/ref_model.py -- clarity-first CPU restatement of the reference's hot path.

TEST INFRASTRUCTURE ONLY. Imported by tests/, __graft_entry__.smoke() and the
cpu_baseline leg of bench.py -- never by the product package.

PARITY UNPINNED: the reference (cjg/aliyun-gpushare-device-plugin) cannot be built or
run here (Go, no toolchain on either machine, ~50 un-vendored modules) and its only
test, pkg/kubelet/client/client_test.go:19-57, asserts nothing. There are no golden
vectors, known-answer tests or fixtures in the reference for this path. This model is
a line-by-line restatement; the golden vectors in tests/golden/ were derived from it
and by hand from the cited lines (tests/golden/make_golden.py), not from a run of
the reference.

Every function cites the reference file:line it follows (paths relative to the
reference tree). Third-party semantics the path depends on, restated from their
published behaviour (sources are not under /root/reference):
  - Go 1.21 stdlib (go.mod:3): strconv.Atoi / ParseUint, sort.Sort (pdqsort with
    insertion sort for n <= 12), fmt %d/%v/%s, encoding/json map-key ordering;
  - k8s.io/apimachinery v0.28.4 (go.mod:12): resource.Quantity parsing and Value();
  - k8s.io/kubelet v0.28.3 (go.mod:14): deviceplugin/v1beta1 message shapes.

Pods are plain dicts shaped like the Kubernetes JSON the reference decodes:
  {"metadata": {"uid", "name", "namespace", "annotations": {..}},
   "spec": {"nodeName", "containers": [{"resources": {"limits": {..}}}]},
   "status": {"phase"}}
"""
from __future__ import annotations

import copy
import json
from fractions import Fraction

# --- const.go:11-35 -----------------------------------------------------------
resourceName = "aliyun.com/gpu-mem"
resourceCount = "aliyun.com/gpu-count"
serverSock = "/var/lib/kubelet/device-plugins/" + "aliyungpushare.sock"
OptimisticLockErrorMsg = ("the object has been modified; please apply your changes to the "
                          "latest version and try again")
envNVGPU = "NVIDIA_VISIBLE_DEVICES"
EnvResourceIndex = "ALIYUN_COM_GPU_MEM_IDX"
EnvResourceByPod = "ALIYUN_COM_GPU_MEM_POD"
EnvResourceByContainer = "ALIYUN_COM_GPU_MEM_CONTAINER"
EnvResourceByDev = "ALIYUN_COM_GPU_MEM_DEV"
EnvAssignedFlag = "ALIYUN_COM_GPU_MEM_ASSIGNED"
EnvResourceAssumeTime = "ALIYUN_COM_GPU_MEM_ASSUME_TIME"
EnvResourceAssignTime = "ALIYUN_COM_GPU_MEM_ASSIGN_TIME"  # const.go:31, never used
EnvNodeLabelForDisableCGPU = "cgpu.disable.isolation"
GiBPrefix = "GiB"
MiBPrefix = "MiB"
Healthy = "Healthy"      # pluginapi.Healthy
Unhealthy = "Unhealthy"  # pluginapi.Unhealthy

UINT64 = (1 << 64) - 1
INT64_MAX = (1 << 63) - 1
INT64_MIN = -(1 << 63)


# --- Go stdlib semantics ---------------------------------------------------------
class GoSyntaxError(ValueError):
    pass


class GoRangeError(ValueError):
    pass


def go_parse_uint(s: str, bit_size: int = 64) -> int:
    """strconv.ParseUint(s, 10, 64) as called at podutils.go:66: base 10 given
    explicitly, so no sign, no underscores, no prefixes; empty is a syntax error;
    overflow is a range error."""
    if s == "":
        raise GoSyntaxError(s)
    n = 0
    for ch in s:
        if not ("0" <= ch <= "9"):
            raise GoSyntaxError(s)
        n = n * 10 + (ord(ch) - 48)
        if n > (1 << bit_size) - 1:
            raise GoRangeError(s)
    return n


def go_atoi(s: str) -> int:
    """strconv.Atoi as called at podutils.go:44 (64-bit int): one optional '+'/'-',
    then one or more ASCII digits; the value must fit int64."""
    if s == "":
        raise GoSyntaxError(s)
    neg = False
    body = s
    if s[0] in "+-":
        neg = s[0] == "-"
        body = s[1:]
    if body == "":
        raise GoSyntaxError(s)
    try:
        un = go_parse_uint(body, 64)
    except GoRangeError:
        raise GoRangeError(s)
    if not neg and un > INT64_MAX:
        raise GoRangeError(s)
    if neg and un > (1 << 63):
        raise GoRangeError(s)
    return -un if neg else un


def go_uint(v: int) -> int:
    """Go's uint(int64) conversion on a 64-bit platform (podutils.go:127)."""
    return v & UINT64


_DEC_SUFFIX = {"n": -9, "u": -6, "m": -3, "": 0, "k": 3, "M": 6, "G": 9, "T": 12, "P": 15,
               "E": 18}
_BIN_SUFFIX = {"Ki": 10, "Mi": 20, "Gi": 30, "Ti": 40, "Pi": 50, "Ei": 60}


class QuantityError(ValueError):
    pass


def quantity_value(s) -> int:
    """resource.MustParse(s).Value() (apimachinery v0.28.4 pkg/api/resource): the
    integer value rounded away from zero. Accepts what parseQuantityString accepts:
    sign, digits with an optional '.', then a decimal-SI / binary-SI suffix or a
    decimal exponent (e3, E-2). JSON numbers arrive as int/float and are
    stringified first, as the apiserver does. Values beyond int64 raise."""
    if isinstance(s, bool):
        raise QuantityError(repr(s))
    if isinstance(s, int):
        s = str(s)
    elif isinstance(s, float):
        s = repr(s)
    if not isinstance(s, str) or s == "":
        raise QuantityError(repr(s))
    i = 0
    positive = True
    if s[0] in "+-":
        positive = s[0] == "+"
        i = 1
    j = i
    while j < len(s) and s[j].isdigit() and s[j].isascii():
        j += 1
    num = s[i:j]
    denom = ""
    if j < len(s) and s[j] == ".":
        k = j + 1
        while k < len(s) and s[k].isdigit() and s[k].isascii():
            k += 1
        denom = s[j + 1:k]
        j = k
    # parseQuantityString turns an empty numerator into "0", so ".", "+" and "-" parse as 0
    suffix = s[j:]
    mant = int((num or "0") + denom)
    exp10 = -len(denom)
    exp2 = 0
    if suffix in _DEC_SUFFIX:
        exp10 += _DEC_SUFFIX[suffix]
    elif suffix in _BIN_SUFFIX:
        exp2 = _BIN_SUFFIX[suffix]
    elif len(suffix) > 1 and suffix[0] in "eE":
        body = suffix[1:]
        sign = 1
        if body[0] in "+-":
            sign = -1 if body[0] == "-" else 1
            body = body[1:]
        if body == "" or not all("0" <= c <= "9" for c in body):
            raise QuantityError(s)
        exp10 += sign * int(body)
    else:
        raise QuantityError(s)
    v = Fraction(mant) * (Fraction(10) ** exp10) * (1 << exp2)
    n = -(-v.numerator // v.denominator)  # ceil: magnitude rounds away from zero
    if not positive:
        n = -n
    if n > INT64_MAX or n < INT64_MIN:
        raise QuantityError("out of int64 range: " + s)
    return n


def go_sort_less_swap(n: int, less, swap) -> None:
    """sort.Sort(data) of Go 1.21 (go.mod:3) over the index interface (Len = n).
    n <= 12 is a plain insertion sort (exact); above that this follows pdqsort as
    published in sort/zsortinterface.go. With the reference's non-strict Less (`<=`,
    podmanager.go:342) the order of EQUAL keys for n > 12 depends on those internals
    and is unpinned; with distinct keys any correct sort gives the same
    result."""
    if n <= 1:
        return
    _pdqsort(less, swap, 0, n, n.bit_length())


def _insertion_sort(less, swap, a, b):
    for i in range(a + 1, b):
        j = i
        while j > a and less(j, j - 1):
            swap(j, j - 1)
            j -= 1


def _sift_down(less, swap, lo, hi, first):
    root = lo
    while True:
        child = 2 * root + 1
        if child >= hi:
            return
        if child + 1 < hi and less(first + child, first + child + 1):
            child += 1
        if not less(first + root, first + child):
            return
        swap(first + root, first + child)
        root = child


def _heap_sort(less, swap, a, b):
    first, lo, hi = a, 0, b - a
    for i in range((hi - 1) // 2, -1, -1):
        _sift_down(less, swap, i, hi, first)
    for i in range(hi - 1, -1, -1):
        swap(first, first + i)
        _sift_down(less, swap, lo, i, first)


def _partition(less, swap, a, b, pivot):
    swap(a, pivot)
    i, j = a + 1, b - 1
    while i <= j and less(i, a):
        i += 1
    while i <= j and not less(j, a):
        j -= 1
    if i > j:
        swap(j, a)
        return j, True
    swap(i, j)
    i += 1
    j -= 1
    while True:
        while i <= j and less(i, a):
            i += 1
        while i <= j and not less(j, a):
            j -= 1
        if i > j:
            break
        swap(i, j)
        i += 1
        j -= 1
    swap(j, a)
    return j, False


def _partition_equal(less, swap, a, b, pivot):
    swap(a, pivot)
    i, j = a + 1, b - 1
    while True:
        while i <= j and not less(a, i):
            i += 1
        while i <= j and less(a, j):
            j -= 1
        if i > j:
            break
        swap(i, j)
        i += 1
        j -= 1
    return i


def _partial_insertion_sort(less, swap, a, b):
    max_steps, shortest_shifting = 5, 50
    i = a + 1
    for _ in range(max_steps):
        while i < b and not less(i, i - 1):
            i += 1
        if i == b:
            return True
        if b - a < shortest_shifting:
            return False
        swap(i, i - 1)
        if i - a >= 2:
            for j in range(i - 1, 0, -1):
                if not less(j, j - 1):
                    break
                swap(j, j - 1)
        if b - i >= 2:
            for j in range(i + 1, b):
                if not less(j, j - 1):
                    break
                swap(j, j - 1)
    return False


def _break_patterns(swap, a, b):
    length = b - a
    if length >= 8:
        random = length & UINT64
        modulus = 1 << length.bit_length()
        idx = a + (length // 4) * 2 - 1
        for i in range(3):
            random ^= (random << 13) & UINT64
            random ^= random >> 7
            random ^= (random << 17) & UINT64
            other = int(random & (modulus - 1))
            if other >= length:
                other -= length
            swap(idx - 1 + i, a + other)


def _order2(less, a, b, swaps):
    if less(b, a):
        swaps[0] += 1
        return b, a
    return a, b


def _median(less, a, b, c, swaps):
    a, b = _order2(less, a, b, swaps)
    b, c = _order2(less, b, c, swaps)
    a, b = _order2(less, a, b, swaps)
    return b


def _median_adjacent(less, a, swaps):
    return _median(less, a - 1, a, a + 1, swaps)


_INCREASING, _DECREASING, _UNKNOWN = 1, 2, 0


def _choose_pivot(less, a, b):
    shortest_ninther, max_swaps = 50, 4 * 3
    length = b - a
    swaps = [0]
    i = a + length // 4 * 1
    j = a + length // 4 * 2
    k = a + length // 4 * 3
    if length >= 8:
        if length >= shortest_ninther:
            i = _median_adjacent(less, i, swaps)
            j = _median_adjacent(less, j, swaps)
            k = _median_adjacent(less, k, swaps)
        j = _median(less, i, j, k, swaps)
    if swaps[0] == 0:
        return j, _INCREASING
    if swaps[0] == max_swaps:
        return j, _DECREASING
    return j, _UNKNOWN


def _reverse_range(swap, a, b):
    i, j = a, b - 1
    while i < j:
        swap(i, j)
        i += 1
        j -= 1


def _pdqsort(less, swap, a, b, limit):
    max_insertion = 12
    was_balanced, was_partitioned = True, True
    while True:
        length = b - a
        if length <= max_insertion:
            _insertion_sort(less, swap, a, b)
            return
        if limit == 0:
            _heap_sort(less, swap, a, b)
            return
        if not was_balanced:
            _break_patterns(swap, a, b)
            limit -= 1
        pivot, hint = _choose_pivot(less, a, b)
        if hint == _DECREASING:
            _reverse_range(swap, a, b)
            pivot = (b - 1) - (pivot - a)
            hint = _INCREASING
        if was_balanced and was_partitioned and hint == _INCREASING:
            if _partial_insertion_sort(less, swap, a, b):
                return
        if a > 0 and not less(a - 1, pivot):
            a = _partition_equal(less, swap, a, b, pivot)
            continue
        mid, already = _partition(less, swap, a, b, pivot)
        was_partitioned = already
        left_len, right_len = mid - a, b - mid
        balance_threshold = length // 8
        if left_len < right_len:
            was_balanced = left_len >= balance_threshold
            _pdqsort(less, swap, a, mid, limit)
            a = mid + 1
        else:
            was_balanced = right_len >= balance_threshold
            _pdqsort(less, swap, mid + 1, b, limit)
            b = mid


# --- nvidia.go -----------------------------------------------------------------
class Globals:
    """Package-level state of pkg/gpu/nvidia (nvidia.go:12-15): survives plugin
    restarts inside one process."""

    def __init__(self, metric=GiBPrefix):
        self.gpuMemory = 0
        self.metric = metric


def translatememoryUnits(value: str):
    """cmd/nvidia/main.go:67-78. Returns (unit, warned)."""
    if value in (MiBPrefix, GiBPrefix):
        return value, False
    return GiBPrefix, True


def generateFakeDeviceID(realID: str, fakeCounter: int) -> str:
    """nvidia.go:23-25."""
    return "%s-_-%d" % (realID, fakeCounter)


def extractRealDeviceID(fakeDeviceID: str) -> str:
    """nvidia.go:27-29."""
    return fakeDeviceID.split("-_-")[0]


def setGPUMemory(g: Globals, raw: int) -> None:
    """nvidia.go:31-38."""
    v = raw
    if g.metric == GiBPrefix:
        v = raw // 1024
    g.gpuMemory = v


def getGPUMemory(g: Globals) -> int:
    """nvidia.go:40-42."""
    return g.gpuMemory


def getDevices(g: Globals, gpus):
    """nvidia.go:50-86. `gpus` = [(uuid, raw_total)] in NVML index order, where
    raw_total is the number the fork feeds to setGPUMemory. Decision H1: callers
    pass the MiB-valued total (bytes >> 20), which is what the division by 1024 was
    written for; the fork's literal bytes-as-MiB behaviour is a recorded divergence.
    Returns (devs, realDevNames): devs = [[id, health]] (mutable health)."""
    devs = []
    realDevNames = {}
    for i, (uuid, total) in enumerate(gpus):
        realDevNames[uuid] = i
        if getGPUMemory(g) == 0:
            setGPUMemory(g, total)
        for j in range(getGPUMemory(g)):
            devs.append([generateFakeDeviceID(uuid, j), Healthy])
    return devs, realDevNames


XID_APPLICATION_ERRORS = (31, 43, 45)


def watchXIDs_event(devs, event_uuid, xid):
    """The rule inside the commented-out watchXIDs (nvidia.go:97-153): which fake
    devices one XID critical-error event marks unhealthy. Returns indices into devs."""
    if xid in XID_APPLICATION_ERRORS:            # nvidia.go:134-136
        return []
    if event_uuid is None or len(event_uuid) == 0:  # nvidia.go:138-143
        return list(range(len(devs)))
    return [k for k, d in enumerate(devs)          # nvidia.go:146-150
            if extractRealDeviceID(d[0]) == event_uuid]


# --- podutils.go ------------------------------------------------------------------
def _annotations(pod):
    return (pod.get("metadata") or {}).get("annotations") or {}


def patchPodAnnotationSpecAssigned(now_unix_nano: int) -> bytes:
    """podutils.go:27-35: json.Marshal of nested maps => keys sorted, no spaces."""
    patch = {"metadata": {"annotations": {
        EnvAssignedFlag: "true",
        EnvResourceAssumeTime: "%d" % now_unix_nano,
    }}}
    return json.dumps(patch, sort_keys=True, separators=(",", ":")).encode()


def getGPUIDFromPodAnnotation(pod) -> int:
    """podutils.go:37-61."""
    id_ = -1
    ann = _annotations(pod)
    if len(ann) > 0:
        if EnvResourceIndex in ann:
            try:
                id_ = go_atoi(ann[EnvResourceIndex])
            except ValueError:
                id_ = -1
    return id_


def getAssumeTimeFromPodAnnotation(pod) -> int:
    """podutils.go:64-75."""
    assumeTime = 0
    ann = _annotations(pod)
    if EnvResourceAssumeTime in ann:
        try:
            assumeTime = go_parse_uint(ann[EnvResourceAssumeTime], 64)
        except ValueError:
            pass
    return assumeTime


def getGPUMemoryFromPodResource(pod) -> int:
    """podutils.go:122-131: Spec.Containers only (no init containers)."""
    total = 0
    for container in (pod.get("spec") or {}).get("containers") or []:
        limits = ((container.get("resources") or {}).get("limits")) or {}
        if resourceName in limits:
            total = (total + go_uint(quantity_value(limits[resourceName]))) & UINT64
    return total


def isGPUMemoryAssumedPod(pod) -> bool:
    """podutils.go:78-119."""
    if getGPUMemoryFromPodResource(pod) <= 0:
        return False
    ann = _annotations(pod)
    if EnvResourceAssumeTime not in ann:
        return False
    if EnvAssignedFlag in ann:
        return ann[EnvAssignedFlag] == "false"
    return False


# --- podmanager.go:247-347 -----------------------------------------------------------
class PodSource:
    """In-memory stand-in for the lister (kubelet /pods or apiserver LIST,
    podmanager.go:187-245) and the patcher (allocate.go:136). Network I/O is out of
    scope; both arms of every comparison use the same source."""

    def __init__(self, pods, fail_list=False, fail_patch=None):
        self.pods = copy.deepcopy(pods)
        self.fail_list = fail_list
        self.fail_patch = fail_patch  # None | "always" | "conflict-once"
        self.patches = []             # [(namespace, name, body bytes)]

    def list_pending(self):
        if self.fail_list:
            raise RuntimeError("failed to get Pods assigned to node")
        return copy.deepcopy(self.pods)

    def patch(self, namespace, name, body: bytes):
        if self.fail_patch == "always":
            raise RuntimeError("patch failed")
        if self.fail_patch == "conflict-once":
            self.fail_patch = None
            raise RuntimeError(OptimisticLockErrorMsg)
        self.patches.append((namespace, name, body))
        ann = json.loads(body)["metadata"]["annotations"]
        for pod in self.pods:
            md = pod["metadata"]
            if md.get("namespace") == namespace and md.get("name") == name:
                md.setdefault("annotations", {}).update(ann)


def getPendingPodsInNode(source: PodSource, nodeName: str):
    """podmanager.go:247-297."""
    pods = []
    podIDMap = {}
    for pod in source.list_pending():
        if (pod.get("spec") or {}).get("nodeName", "") != nodeName:   # :274
            continue
        uid = (pod.get("metadata") or {}).get("uid", "")
        if uid not in podIDMap:                                        # :287-290
            pods.append(pod)
            podIDMap[uid] = True
    return pods


def makePodOrderdByAge(pods):
    """podmanager.go:326-347: sort.Sort with Less = `<=` on the assume time."""
    lst = list(pods)

    def less(i, j):
        return getAssumeTimeFromPodAnnotation(lst[i]) <= getAssumeTimeFromPodAnnotation(lst[j])

    def swap(i, j):
        lst[i], lst[j] = lst[j], lst[i]

    go_sort_less_swap(len(lst), less, swap)
    return lst


def getCandidatePods(source: PodSource, nodeName: str):
    """podmanager.go:300-323."""
    candidatePods = [pod for pod in getPendingPodsInNode(source, nodeName)
                     if isGPUMemoryAssumedPod(pod)]
    return makePodOrderdByAge(candidatePods)


# --- server.go + allocate.go ----------------------------------------------------------
class NvidiaDevicePlugin:
    """server.go:19-35 with the methods on the hot path."""

    def __init__(self, g: Globals, gpus, nodeName, source: PodSource,
                 disableCGPUIsolation=False, clock=lambda: 0):
        self.g = g
        self.devs, self.devNameMap = getDevices(g, gpus)      # server.go:44
        self.devIndxMap = {}
        self.nodeName = nodeName
        self.source = source
        self.disableCGPUIsolation = disableCGPUIsolation
        self.clock = clock

    def GetDeviceNameByIndex(self, index: int):
        """server.go:80-91."""
        if len(self.devIndxMap) == 0:
            self.devIndxMap = {v: k for k, v in self.devNameMap.items()}
        if index in self.devIndxMap:
            return self.devIndxMap[index], True
        return "", False

    def ListAndWatchResponse(self):
        """server.go:181,190: the full device list, every send."""
        return [(d[0], d[1]) for d in self.devs]

    def unhealthy_event(self, event_uuid, xid):
        """healthcheck pump (server.go:211-229) + ListAndWatch body (server.go:187-190)
        driven by one XID event. Returns how many full-list sends the reference design
        would make (one per fake device)."""
        idxs = watchXIDs_event(self.devs, event_uuid, xid)
        for k in idxs:
            self.devs[k][1] = Unhealthy
        return len(idxs)

    def buildErrResponse(self, reqs, podReqGPU):
        """allocate.go:25-40."""
        out = []
        for req in reqs:
            out.append({
                envNVGPU: "no-gpu-has-%d%s-to-run" % (podReqGPU, self.g.metric),
                EnvResourceIndex: "-1",
                EnvResourceByPod: "%d" % podReqGPU,
                EnvResourceByContainer: "%d" % len(req),
                EnvResourceByDev: "%d" % getGPUMemory(self.g),
            })
        return out

    def Allocate(self, reqs):
        """allocate.go:43-201. `reqs` = list (one per container) of device-id lists.
        Returns the list of env maps, one per container request (every failure is
        still a normal response, H8)."""
        responses = []
        podReqGPU = 0
        for req in reqs:                                  # :55-57
            podReqGPU += len(req)
        try:
            pods = getCandidatePods(self.source, self.nodeName)   # :63
        except Exception:
            return self.buildErrResponse(reqs, podReqGPU)          # :64-67

        found, assumePod = False, None
        for pod in pods:                                  # :79-89
            if getGPUMemoryFromPodResource(pod) == podReqGPU:
                assumePod, found = pod, True
                break

        if found:
            id_ = getGPUIDFromPodAnnotation(assumePod)   # :92
            if id_ >= 0:
                _, ok = self.GetDeviceNameByIndex(go_uint(id_))   # :100
                if not ok:
                    id_ = -1
            if id_ < 0:
                return self.buildErrResponse(reqs, podReqGPU)      # :109-111
            for req in reqs:                              # :114-129
                envs = {
                    envNVGPU: "%d" % id_,
                    EnvResourceIndex: "%d" % id_,
                    EnvResourceByPod: "%d" % podReqGPU,
                    EnvResourceByContainer: "%d" % len(req),
                    EnvResourceByDev: "%d" % getGPUMemory(self.g),
                }
                if self.disableCGPUIsolation:
                    envs["CGPU_DISABLE"] = "true"
                responses.append(envs)
            body = patchPodAnnotationSpecAssigned(self.clock())    # :132
            md = assumePod["metadata"]
            try:
                self.source.patch(md.get("namespace", ""), md.get("name", ""), body)  # :136
            except Exception as e:                        # :138-152
                if str(e) == OptimisticLockErrorMsg:
                    try:
                        self.source.patch(md.get("namespace", ""), md.get("name", ""), body)
                    except Exception:
                        return self.buildErrResponse(reqs, podReqGPU)
                else:
                    return self.buildErrResponse(reqs, podReqGPU)
        elif len(self.devNameMap) == 1:                   # :154-180
            devName, devIndex = next(iter(self.devNameMap.items()))
            for req in reqs:
                envs = {
                    envNVGPU: devName,
                    EnvResourceIndex: "%d" % devIndex,
                    EnvResourceByPod: "%d" % podReqGPU,
                    EnvResourceByContainer: "%d" % len(req),
                    EnvResourceByDev: "%d" % getGPUMemory(self.g),
                }
                if self.disableCGPUIsolation:
                    envs["CGPU_DISABLE"] = "true"
                responses.append(envs)
            return responses
        else:                                             # :182-187
            return self.buildErrResponse(reqs, podReqGPU)
        return responses                                  # :200


# --- deviceplugin/v1beta1 wire format (k8s.io/kubelet v0.28.3 api.proto) --------------
def _varint(n: int) -> bytes:
    out = bytearray()
    while True:
        b = n & 0x7F
        n >>= 7
        if n:
            out.append(b | 0x80)
        else:
            out.append(b)
            return bytes(out)


def _ld(field: int, payload: bytes) -> bytes:
    return _varint((field << 3) | 2) + _varint(len(payload)) + payload


def encode_allocate_request(reqs) -> bytes:
    """AllocateRequest{repeated ContainerAllocateRequest container_requests = 1},
    ContainerAllocateRequest{repeated string devices_ids = 1}."""
    return b"".join(_ld(1, b"".join(_ld(1, i.encode()) for i in req)) for req in reqs)


def encode_list_and_watch(devs) -> bytes:
    """ListAndWatchResponse{repeated Device devices = 1}, Device{ID=1, health=2}."""
    return b"".join(_ld(1, _ld(1, i.encode()) + _ld(2, h.encode())) for i, h in devs)


def _read_varint(buf, pos):
    shift = n = 0
    while True:
        if pos >= len(buf):
            raise ValueError("truncated varint")
        b = buf[pos]
        pos += 1
        n |= (b & 0x7F) << shift
        if not b & 0x80:
            return n, pos
        shift += 7
        if shift > 63:
            raise ValueError("varint too long")


def _fields(buf):
    pos = 0
    while pos < len(buf):
        key, pos = _read_varint(buf, pos)
        field, wt = key >> 3, key & 7
        if wt == 0:
            v, pos = _read_varint(buf, pos)
        elif wt == 2:
            ln, pos = _read_varint(buf, pos)
            if pos + ln > len(buf):
                raise ValueError("truncated field")
            v = bytes(buf[pos:pos + ln])
            pos += ln
        elif wt == 1:
            if pos + 8 > len(buf):
                raise ValueError("truncated fixed64")
            v = bytes(buf[pos:pos + 8])
            pos += 8
        elif wt == 5:
            if pos + 4 > len(buf):
                raise ValueError("truncated fixed32")
            v = bytes(buf[pos:pos + 4])
            pos += 4
        else:
            raise ValueError("unsupported wire type %d" % wt)
        yield field, wt, v


def decode_allocate_request(buf: bytes):
    return [[v.decode() for f2, w2, v in _fields(c) if f2 == 1 and w2 == 2]
            for f, w, c in _fields(buf) if f == 1 and w == 2]


def decode_allocate_response(buf: bytes):
    """AllocateResponse{repeated ContainerAllocateResponse container_responses = 1},
    ContainerAllocateResponse{map<string,string> envs = 1, ...}. Returns the list of
    env dicts; other fields (mounts, devices, annotations, cdi) must be absent."""
    out = []
    for f, w, c in _fields(buf):
        assert f == 1 and w == 2, (f, w)
        envs = {}
        for f2, w2, entry in _fields(c):
            assert f2 == 1 and w2 == 2, "only envs are ever populated (allocate.go:116-124)"
            k = v = ""
            for f3, w3, x in _fields(entry):
                if f3 == 1:
                    k = x.decode()
                elif f3 == 2:
                    v = x.decode()
            assert k not in envs, "duplicate env key"
            envs[k] = v
        out.append(envs)
    return out


def decode_list_and_watch(buf: bytes):
    out = []
    for f, w, d in _fields(buf):
        assert f == 1 and w == 2
        i = h = ""
        for f2, w2, x in _fields(d):
            if f2 == 1:
                i = x.decode()
            elif f2 == 2:
                h = x.decode()
        out.append((i, h))
    return out


# --- N1: the lister's retry ladders (podmanager.go:187-245) -------------------------------
retries = 8  # podmanager.go:29


class LadderPodSource(PodSource):
    """The reference's real pod source with its I/O abstracted: `transport` supplies
    kubelet_pods() -> PodList dict (pkg/kubelet/client/client.go:119-134),
    api_list(field_selector, label_selector) -> PodList dict (podmanager.go:229) and
    api_patch(namespace, name, body) (allocate.go:136); `sleep(seconds)` is time.Sleep.
    Every transport call may raise. Drop-in for PodSource under NvidiaDevicePlugin."""

    def __init__(self, transport, sleep, nodeName, queryKubelet):
        self.t = transport
        self.sleep = sleep
        self.nodeName = nodeName
        self.queryKubelet = queryKubelet
        self.patches = []

    def getPodList(self):
        """podmanager.go:187-208."""
        podList = self.t.kubelet_pods()
        items = [p for p in (podList.get("items") or [])
                 if ((p.get("status") or {}).get("phase")) == "Pending"]
        if len(items) == 0:
            raise RuntimeError("not found pending pod")
        return {"items": items}

    def getPodListsByQueryKubelet(self):
        """podmanager.go:210-225: first try, then up to `retries` more, each followed by a
        100 ms sleep (the sleep also follows the try that finally succeeds)."""
        podList, err = None, None
        try:
            podList = self.getPodList()
        except Exception as e:
            err = e
        i = 0
        while i < retries and err is not None:
            try:
                podList, err = self.getPodList(), None
            except Exception as e:
                err = e
            self.sleep(0.1)
            i += 1
        if err is not None:
            podList = self.getPodListsByListAPIServer()
        return podList

    def getPodListsByListAPIServer(self):
        """podmanager.go:227-245: fields.SelectorFromSet sorts its keys; first try plus up to
        three more, each followed by a 1 s sleep."""
        selector = "spec.nodeName=%s,status.phase=Pending" % self.nodeName
        podList, err = None, None
        try:
            podList = self.t.api_list(selector, "")
        except Exception as e:
            err = e
        i = 0
        while i < 3 and err is not None:
            try:
                podList, err = self.t.api_list(selector, ""), None
            except Exception as e:
                err = e
            self.sleep(1.0)
            i += 1
        if err is not None:
            raise RuntimeError("failed to get Pods assigned to node %s" % self.nodeName)
        return podList

    def list_pending(self):
        """podmanager.go:256-267."""
        if self.queryKubelet:
            podList = self.getPodListsByQueryKubelet()
        else:
            podList = self.getPodListsByListAPIServer()
        return copy.deepcopy(podList.get("items") or [])

    def patch(self, namespace, name, body: bytes):
        self.t.api_patch(namespace, name, body)
        self.patches.append((namespace, name, body))
