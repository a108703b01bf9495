"""Host-side objects over the C ABI, named after the reference's own types.

    DeviceTable          getDevices / generateFakeDeviceID / GetDeviceNameByIndex
                         (pkg/gpu/nvidia/nvidia.go:23-86, server.go:80-91)
    NvidiaDevicePlugin   the plugin struct and its RPC bodies (server.go:19-35,
                         93-95,180-201; allocate.go:43-201)
    Nvml                 the go-nvml call sites (nvidia.go:45-64, gpumanager.go:36,41)

Everything here forwards to libgpushare_core.so; there is no Python implementation of
any decision, encoding or probe.
"""
import ctypes as C
import threading
from dataclasses import dataclass

from . import _lib as L

GiB, MiB = "GiB", "MiB"
_UNIT = {GiB: L.GS_UNIT_GIB, MiB: L.GS_UNIT_MIB}
ALLOC_KIND = {L.GS_ALLOC_MATCHED: "matched", L.GS_ALLOC_SINGLE_GPU: "single-gpu",
              L.GS_ALLOC_ERR_NO_MATCH: "err-no-match", L.GS_ALLOC_ERR_BAD_INDEX: "err-bad-index",
              L.GS_ALLOC_ERR_PODS: "err-pods", L.GS_ALLOC_ERR_PATCH: "err-patch"}


def translate_memory_units(value: str):
    """cmd/nvidia/main.go:67-78 -> (unit string, warned)."""
    lib = L.load()
    u, w = C.c_int(), C.c_int()
    b = value.encode()
    L.check(lib.gs_translate_memory_unit(b, len(b), C.byref(u), C.byref(w)))
    return (MiB if u.value == L.GS_UNIT_MIB else GiB), bool(w.value)


def quantity_value(s: str) -> int:
    lib = L.load()
    v = C.c_int64()
    b = s.encode()
    L.check(lib.gs_quantity_value(b, len(b), C.byref(v)), "gs_quantity_value(%r)" % s)
    return v.value


def go_atoi(s: str) -> int:
    lib = L.load()
    v = C.c_int64()
    b = s.encode()
    L.check(lib.gs_go_atoi(b, len(b), C.byref(v)), "gs_go_atoi(%r)" % s)
    return v.value


def go_parse_uint(s: str) -> int:
    lib = L.load()
    v = C.c_uint64()
    b = s.encode()
    L.check(lib.gs_go_parse_uint64(b, len(b), C.byref(v)), "gs_go_parse_uint64(%r)" % s)
    return v.value


class Nvml:
    """NVML through dlopen; `path` may name the fake library used by tests."""

    def __init__(self, path=None):
        self.lib = L.load()
        L.check(self.lib.gs_nvml_open(path.encode() if path else None), "gs_nvml_open")
        self._open = True
        self._health = False

    def close(self):
        if self._open:
            self.lib.gs_nvml_close()
            self._open = False

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def device_count(self):
        n = C.c_uint32()
        L.check(self.lib.gs_nvml_device_count(C.byref(n)), "gs_nvml_device_count")
        return n.value

    def device_info(self, i):
        info = L.GpuInfo()
        L.check(self.lib.gs_nvml_device_info(i, C.byref(info)), "gs_nvml_device_info")
        return info

    def devices(self):
        n = self.device_count()
        arr = (L.GpuInfo * n)()
        for i in range(n):
            L.check(self.lib.gs_nvml_device_info(i, C.byref(arr[i])), "gs_nvml_device_info")
        return arr

    def map_cuda_ordinals(self, arr):
        L.check(self.lib.gs_cuda_map_ordinals(arr, len(arr)), "gs_cuda_map_ordinals")
        return [g.cuda_ordinal for g in arr]

    def health_open(self):
        m = C.c_uint64()
        L.check(self.lib.gs_health_open(C.byref(m)), "gs_health_open")
        self._health = True
        return m.value

    def health_wait(self, timeout_ms=5000):
        """Returns (gpu_index | -1 for all, xid, ignored) or None on timeout."""
        ev = L.HealthEvent()
        rc = self.lib.gs_health_wait(timeout_ms, C.byref(ev))
        if rc == L.GS_E_TIMEOUT:
            return None
        L.check(rc, "gs_health_wait")
        return ev.gpu_index, ev.xid, bool(ev.ignored)

    def health_close(self):
        if self._health:
            self.lib.gs_health_close()
            self._health = False


class DeviceTable:
    """The fake-device model: ids, O(1) lookup, ListAndWatch payload, health."""

    def __init__(self, uuids, units_per_gpu, unit=GiB, handle=None):
        self.lib = L.load()
        if handle is not None:
            self.h = handle
        else:
            enc = [u.encode() for u in uuids]
            arr = (C.c_char_p * len(enc))(*enc)
            lens = (C.c_uint32 * len(enc))(*[len(e) for e in enc])
            h = C.c_void_p()
            L.check(self.lib.gs_table_build(arr, lens, len(enc), units_per_gpu, _UNIT[unit],
                                            C.byref(h)), "gs_table_build")
            self.h = h

    @classmethod
    def from_gpus(cls, gpu_infos, unit=GiB):
        """getDevices over NVML-reported GPUs (units from GPU 0, MiB-valued: H1/H7)."""
        lib = L.load()
        h = C.c_void_p()
        L.check(lib.gs_table_build_from_gpus(gpu_infos, len(gpu_infos), _UNIT[unit], C.byref(h)),
                "gs_table_build_from_gpus")
        return cls(None, 0, unit, handle=h)

    @classmethod
    def from_raw(cls, uuids, raw_mib_gpu0, unit=GiB):
        """setGPUMemory(raw) semantics (nvidia.go:31-38) for a MiB-valued raw total."""
        lib = L.load()
        return cls(uuids, lib.gs_units_from_raw(raw_mib_gpu0, _UNIT[unit]), unit)

    def close(self):
        if self.h:
            self.lib.gs_table_free(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def gpu_count(self):
        return self.lib.gs_table_gpu_count(self.h)

    @property
    def units_per_gpu(self):
        return self.lib.gs_table_units_per_gpu(self.h)

    @property
    def device_count(self):
        return self.lib.gs_table_device_count(self.h)

    @property
    def unit(self):
        return MiB if self.lib.gs_table_unit(self.h) == L.GS_UNIT_MIB else GiB

    def fake_id(self, gpu, j):
        buf = C.create_string_buffer(160)
        n = C.c_size_t()
        L.check(self.lib.gs_fake_id_format(self.h, gpu, j, buf, len(buf), C.byref(n)),
                "gs_fake_id_format")
        return buf.raw[:n.value].decode()

    def extract_real_device_id(self, fake_id):
        b = fake_id.encode()
        return b[:self.lib.gs_extract_real_device_id(b, len(b))].decode()

    def lookup(self, fake_id):
        """-> (gpu, unit_index) or None."""
        b = fake_id.encode()
        g, j = C.c_uint32(), C.c_uint64()
        rc = self.lib.gs_id_lookup(self.h, b, len(b), C.byref(g), C.byref(j))
        if rc == L.GS_E_NOTFOUND:
            return None
        L.check(rc, "gs_id_lookup")
        return g.value, j.value

    def device_name_by_index(self, index):
        """GetDeviceNameByIndex -> (name, found)."""
        p, n = C.c_void_p(), C.c_size_t()
        rc = self.lib.gs_device_name_by_index(self.h, index, C.byref(p), C.byref(n))
        if rc == L.GS_E_NOTFOUND:
            return "", False
        L.check(rc, "gs_device_name_by_index")
        return C.string_at(p.value, n.value).decode(), True

    def device_index_by_name(self, uuid):
        b = uuid.encode()
        i = C.c_uint32()
        rc = self.lib.gs_device_index_by_name(self.h, b, len(b), C.byref(i))
        if rc == L.GS_E_NOTFOUND:
            return None
        L.check(rc)
        return i.value

    # health
    def mark_unhealthy_device(self, gpu, j):
        L.check(self.lib.gs_law_mark_unhealthy_device(self.h, gpu, j))

    def mark_unhealthy_id(self, fake_id):
        b = fake_id.encode()
        L.check(self.lib.gs_law_mark_unhealthy_id(self.h, b, len(b)), "gs_law_mark_unhealthy_id")

    def mark_unhealthy_gpu(self, gpu):
        L.check(self.lib.gs_law_mark_unhealthy_gpu(self.h, gpu), "gs_law_mark_unhealthy_gpu")

    def reset_health(self):
        L.check(self.lib.gs_law_reset_health(self.h))

    def device_healthy(self, gpu, j):
        rc = self.lib.gs_law_device_healthy(self.h, gpu, j)
        if rc < 0:
            L.check(rc, "gs_law_device_healthy")
        return bool(rc)

    @property
    def generation(self):
        return self.lib.gs_law_generation(self.h)

    # ListAndWatch payload
    def payload_size(self):
        return self.lib.gs_law_payload_size(self.h)

    def list_and_watch_bytes(self):
        """The serialized v1beta1.ListAndWatchResponse as Python bytes. The bytes object is kept
        per health generation: a second stream (kubelet restart) or a second subscriber gets
        the same object without another 90 MB copy."""
        gen = self.generation
        cached = getattr(self, "_law_cache", None)
        if cached is not None and cached[0] == gen:
            return cached[1]
        out = self._encode_law()
        if self.generation == gen:           # no health change raced with the encode
            self._law_cache = (gen, out)
        return out

    def _encode_law(self):
        n = self.payload_size()
        buf = C.create_string_buffer(max(n, 1))
        out = C.c_size_t()
        L.check(self.lib.gs_law_encode(self.h, buf, n, C.byref(out)), "gs_law_encode")
        return buf.raw[:out.value]

    def payload_segments(self):
        """Borrowed (address, length) pairs, one per GPU (gs_law_payload)."""
        cap = max(self.gpu_count, 1)
        iov = (L.IoVec * cap)()
        n = C.c_uint32()
        L.check(self.lib.gs_law_payload(self.h, iov, cap, C.byref(n)), "gs_law_payload")
        return [(iov[i].base, iov[i].len) for i in range(n.value)]


@dataclass
class Pod:
    """What the lister has to hand over for one pod (gs_pod_view)."""
    uid: str
    namespace: str
    name: str
    node_name: str
    gpu_mem_limit: int
    ann_idx: str = None          # ALIYUN_COM_GPU_MEM_IDX
    ann_assume_time: str = None  # ALIYUN_COM_GPU_MEM_ASSUME_TIME
    ann_assigned: str = None     # ALIYUN_COM_GPU_MEM_ASSIGNED

    @classmethod
    def from_k8s(cls, pod: dict):
        """From the Kubernetes JSON shape; limits go through gs_quantity_value and are
        summed with Go's wrapping uint add (podutils.go:122-131)."""
        md = pod.get("metadata") or {}
        ann = md.get("annotations") or {}
        total = 0
        for c in (pod.get("spec") or {}).get("containers") or []:
            limits = (c.get("resources") or {}).get("limits") or {}
            if "aliyun.com/gpu-mem" in limits:
                q = limits["aliyun.com/gpu-mem"]
                total = (total + (quantity_value(str(q)) & ((1 << 64) - 1))) & ((1 << 64) - 1)
        return cls(md.get("uid", ""), md.get("namespace", ""), md.get("name", ""),
                   (pod.get("spec") or {}).get("nodeName", ""), total,
                   ann.get("ALIYUN_COM_GPU_MEM_IDX"), ann.get("ALIYUN_COM_GPU_MEM_ASSUME_TIME"),
                   ann.get("ALIYUN_COM_GPU_MEM_ASSIGNED"))


def _str(s, keep):
    if s is None:
        return L.Str(None, 0)
    b = s.encode() if isinstance(s, str) else s
    keep.append(b)
    return L.Str(b, len(b))


def pod_views(pods):
    """-> (ctypes array of gs_pod_view, keep-alive list)."""
    keep = []
    arr = (L.PodView * max(len(pods), 1))()
    for i, p in enumerate(pods):
        arr[i] = L.PodView(_str(p.uid, keep), _str(p.namespace, keep), _str(p.name, keep),
                           _str(p.node_name, keep), p.gpu_mem_limit & ((1 << 64) - 1),
                           _str(p.ann_idx, keep), _str(p.ann_assume_time, keep),
                           _str(p.ann_assigned, keep))
    return arr, keep


class PodList:
    """Pods already laid out as the C ABI's gs_pod_view array (what a lister hands over)."""

    def __init__(self, pods):
        self.n = len(pods)
        self.views, self._keep = pod_views(pods)


@dataclass
class AllocDecision:
    kind: str
    pod_index: int
    gpu_index: int
    n_containers: int
    pod_req_units: int
    response: bytes
    patch: bytes = b""


def _decision(res, resp, patch=b""):
    return AllocDecision(ALLOC_KIND.get(res.kind, str(res.kind)), res.pod_index, res.gpu_index,
                         res.n_containers, res.pod_req_units, resp, patch)


def allocate(table: DeviceTable, node_name: str, request: bytes, pods, now_unix_nano=0,
             disable_cgpu_isolation=False) -> AllocDecision:
    """gs_allocate: one decision against an explicit pod list (stateless form)."""
    lib = L.load()
    keep = []
    ctx = L.AllocCtx(table.h, _str(node_name, keep), int(disable_cgpu_isolation), now_unix_nano)
    views, keep2 = pod_views(pods)
    cap = 512 + 256 * 64
    resp = C.create_string_buffer(cap)
    patch = C.create_string_buffer(256)
    res = L.AllocResult()
    rc = lib.gs_allocate(C.byref(ctx), request, len(request), views, len(pods), resp, cap, patch,
                         len(patch), C.byref(res))
    if rc == L.GS_E_NOSPACE:
        cap = res.resp_len
        resp = C.create_string_buffer(cap)
        rc = lib.gs_allocate(C.byref(ctx), request, len(request), views, len(pods), resp, cap,
                             patch, len(patch), C.byref(res))
    L.check(rc, "gs_allocate")
    return _decision(res, resp.raw[:res.resp_len], patch.raw[:res.patch_len])


def allocate_error_response(table, node_name, request, kind=L.GS_ALLOC_ERR_PATCH):
    lib = L.load()
    keep = []
    ctx = L.AllocCtx(table.h, _str(node_name, keep), 0, 0)
    cap = 512 + 256 * 64
    resp = C.create_string_buffer(cap)
    res = L.AllocResult()
    L.check(lib.gs_allocate_error_response(C.byref(ctx), request, len(request), kind, resp, cap,
                                           C.byref(res)), "gs_allocate_error_response")
    return _decision(res, resp.raw[:res.resp_len])


def request_scan(request: bytes):
    """-> (per-container id counts, total)."""
    lib = L.load()
    counts = (C.c_uint32 * 1024)()
    n, tot = C.c_uint32(), C.c_uint64()
    L.check(lib.gs_request_scan(request, len(request), counts, 1024, C.byref(n), C.byref(tot)),
            "gs_request_scan")
    return list(counts[:n.value]), tot.value


class NvidiaDevicePlugin:
    """server.go:19-35. Holds the device table, the Allocate lock and an in-memory pod
    source standing in for the lister/patcher."""

    def __init__(self, table: DeviceTable, node_name: str, disable_cgpu_isolation=False):
        self.lib = L.load()
        self.table = table
        b = node_name.encode()
        h = C.c_void_p()
        L.check(self.lib.gs_plugin_new(table.h, b, len(b), int(disable_cgpu_isolation),
                                       C.byref(h)), "gs_plugin_new")
        self.h = h
        self._tls = threading.local()   # one response buffer per calling thread

    def close(self):
        if self.h:
            self.lib.gs_plugin_free(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_pods(self, pods):
        """Replace the pod source. `pods` is a list of Pod, or a PodList prepared once with
        PodList(pods) when the same list is installed repeatedly."""
        if not isinstance(pods, PodList):
            pods = PodList(pods)
        L.check(self.lib.gs_plugin_set_pods(self.h, pods.views, pods.n), "gs_plugin_set_pods")

    def set_clock(self, fixed_unix_nano):
        L.check(self.lib.gs_plugin_set_clock(self.h, fixed_unix_nano))

    def get_pod(self, i) -> Pod:
        v = L.PodView()
        L.check(self.lib.gs_plugin_get_pod(self.h, i, C.byref(v)), "gs_plugin_get_pod")

        def s(x):
            # absent (NULL) and present-but-empty differ: an empty ASSUME_TIME annotation still
            # makes a pod "assumed" (podutils.go:92-96 tests presence, not content)
            addr = C.cast(C.byref(x), C.POINTER(C.c_void_p))[0]
            return None if not addr else C.string_at(addr, x.n).decode()

        return Pod(s(v.uid) or "", s(v.ns) or "", s(v.name) or "", s(v.node_name) or "",
                   v.gpu_mem_limit, s(v.ann_idx), s(v.ann_assume_time), s(v.ann_assigned))

    def pod_count(self):
        n = C.c_size_t()
        L.check(self.lib.gs_plugin_pod_count(self.h, C.byref(n)))
        return n.value

    # --- the five RPC bodies, bytes in / bytes out ---------------------------------
    def Allocate(self, request: bytes) -> AllocDecision:
        res = L.AllocResult()
        buf = getattr(self._tls, "resp", None)
        if buf is None:
            buf = self._tls.resp = C.create_string_buffer(1 << 16)
        rc = self.lib.gs_plugin_allocate(self.h, request, len(request), buf, len(buf),
                                         C.byref(res))
        L.check(rc, "gs_plugin_allocate")
        return _decision(res, C.string_at(buf, res.resp_len))

    def GetDevicePluginOptions(self, request: bytes = b"") -> bytes:
        n = C.c_size_t()
        L.check(self.lib.gs_plugin_get_device_plugin_options(self.h, None, 0, C.byref(n)))
        return b""

    def PreStartContainer(self, request: bytes = b"") -> bytes:
        n = C.c_size_t()
        L.check(self.lib.gs_plugin_pre_start_container(self.h, request, len(request), None, 0,
                                                       C.byref(n)))
        return b""

    def GetPreferredAllocation(self, request: bytes = b""):
        # server.go:37-40 panics ("implement me"); never advertised (server.go:93-95)
        raise NotImplementedError("implement me")

    def ListAndWatch(self, poll_ms=200):
        """Generator of serialized ListAndWatchResponse messages: the full list at once,
        then once per health change, until Stop (server.go:180-193)."""
        seen = C.c_uint64(self.table.generation)
        yield self.table.list_and_watch_bytes()
        while True:
            rc = self.lib.gs_plugin_wait_health_change(self.h, C.byref(seen), poll_ms)
            if rc == L.GS_E_STATE:
                return
            if rc == L.GS_E_TIMEOUT:
                continue
            L.check(rc, "gs_plugin_wait_health_change")
            yield self.table.list_and_watch_bytes()

    def unhealthy(self, gpu):
        """healthcheck pump (server.go:211-229) for one XID event: every fake device of
        `gpu` (or all GPUs for -1), one coalesced re-advertise (H13)."""
        self.table.mark_unhealthy_gpu(gpu)
        L.check(self.lib.gs_plugin_notify_health(self.h))

    def notify_health(self):
        """Wake ListAndWatch senders after a health change made on the table directly."""
        L.check(self.lib.gs_plugin_notify_health(self.h))

    def unhealthy_device(self, fake_id):
        """m.unhealthy(dev) for one fake device, the reference's granularity
        (server.go:195-197)."""
        self.table.mark_unhealthy_id(fake_id)
        L.check(self.lib.gs_plugin_notify_health(self.h))

    def Stop(self):
        L.check(self.lib.gs_plugin_stop(self.h))

    @property
    def stopped(self):
        return bool(self.lib.gs_plugin_stopped(self.h))
