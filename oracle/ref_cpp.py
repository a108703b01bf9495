"""ctypes wrapper of oracle/ref_model.cpp (the cost-faithful CPU baseline).

TEST INFRASTRUCTURE / CPU BASELINE ONLY -- see the header of ref_model.cpp.
"""
import ctypes as C
import os

_SO = os.path.join(os.path.dirname(os.path.abspath(__file__)), "libgs_oracle.so")
_lib = None


def lib():
    global _lib
    if _lib is None:
        l = C.CDLL(_SO)
        l.ref_new.restype = C.c_void_p
        l.ref_new.argtypes = [C.POINTER(C.c_char_p), C.c_int, C.c_uint64, C.c_char_p, C.c_char_p, C.c_int]
        l.ref_free.argtypes = [C.c_void_p]
        l.ref_get_devices.restype = C.c_uint64
        l.ref_get_devices.argtypes = [C.c_void_p]
        l.ref_units.restype = C.c_uint64
        l.ref_units.argtypes = [C.c_void_p]
        l.ref_clear_pods.argtypes = [C.c_void_p]
        l.ref_add_pod.argtypes = [C.c_void_p, C.c_char_p, C.c_char_p, C.c_char_p, C.c_char_p, C.c_int,
                                  C.POINTER(C.c_char_p), C.c_int, C.POINTER(C.c_char_p),
                                  C.POINTER(C.c_char_p)]
        l.ref_save_pods.argtypes = [C.c_void_p]
        l.ref_restore_pods.argtypes = [C.c_void_p]
        l.ref_allocate.restype = C.c_int
        l.ref_allocate.argtypes = [C.c_void_p, C.c_char_p, C.c_size_t, C.c_int64, C.c_void_p,
                                   C.c_size_t, C.POINTER(C.c_size_t)]
        l.ref_last_patched.restype = C.c_int
        l.ref_last_patched.argtypes = [C.c_void_p]
        l.ref_last_patch.restype = C.c_size_t
        l.ref_last_patch.argtypes = [C.c_void_p, C.c_char_p, C.c_size_t]
        l.ref_list_and_watch.restype = C.c_size_t
        l.ref_list_and_watch.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t]
        l.ref_mark_unhealthy.argtypes = [C.c_void_p, C.c_uint64]
        _lib = l
    return _lib


KINDS = {0: "matched", 1: "single-gpu", 2: "err-no-match", 3: "err-bad-index"}


class RefPlugin:
    def __init__(self, uuids, raw_mib, unit, node, cgpu=False):
        self.l = lib()
        enc = [u.encode() for u in uuids]
        arr = (C.c_char_p * max(len(enc), 1))(*enc)
        self.h = self.l.ref_new(arr, len(enc), raw_mib, unit.encode(), node.encode(), int(cgpu))
        self.buf = C.create_string_buffer(1 << 16)

    def close(self):
        if self.h:
            self.l.ref_free(self.h)
            self.h = None

    def __del__(self):
        self.close()

    @property
    def units(self):
        return self.l.ref_units(self.h)

    def get_devices(self):
        return self.l.ref_get_devices(self.h)

    def set_pods_k8s(self, pods):
        """pods in the Kubernetes JSON shape (as oracle/ref_model.py takes them)."""
        self.l.ref_clear_pods(self.h)
        for p in pods:
            md = p.get("metadata") or {}
            ann = md.get("annotations") or {}
            cs = (p.get("spec") or {}).get("containers") or []
            lims = []
            for c in cs:
                v = ((c.get("resources") or {}).get("limits") or {}).get("aliyun.com/gpu-mem")
                lims.append(None if v is None else str(v).encode())
            larr = (C.c_char_p * max(len(lims), 1))(*lims)
            keys = (C.c_char_p * max(len(ann), 1))(*[k.encode() for k in ann])
            vals = (C.c_char_p * max(len(ann), 1))(*[v.encode() for v in ann.values()])
            self.l.ref_add_pod(self.h, md.get("uid", "").encode(), md.get("namespace", "").encode(),
                               md.get("name", "").encode(),
                               (p.get("spec") or {}).get("nodeName", "").encode(), len(lims), larr,
                               len(ann), keys, vals)

    def save_pods(self):
        self.l.ref_save_pods(self.h)

    def restore_pods(self):
        """Re-install the saved lister answer without going through Python marshalling."""
        self.l.ref_restore_pods(self.h)

    def allocate(self, wire: bytes, now=0):
        """-> (kind string, response bytes, patched pod index or -1)."""
        n = C.c_size_t()
        kind = self.l.ref_allocate(self.h, wire, len(wire), now, self.buf, len(self.buf), C.byref(n))
        if n.value > len(self.buf):
            self.buf = C.create_string_buffer(n.value)
            raise RuntimeError("response buffer grown; call again")
        return KINDS.get(kind, str(kind)), self.buf.raw[:n.value], self.l.ref_last_patched(self.h)

    def last_patch(self):
        b = C.create_string_buffer(256)
        n = self.l.ref_last_patch(self.h, b, 256)
        return b.raw[:n]

    def list_and_watch(self):
        n = self.l.ref_list_and_watch(self.h, None, 0)
        b = C.create_string_buffer(max(n, 1))
        self.l.ref_list_and_watch(self.h, b, n)
        return b.raw[:n]

    def mark_unhealthy(self, dev_index):
        self.l.ref_mark_unhealthy(self.h, dev_index)
