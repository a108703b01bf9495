"""ctypes binding of libgpushare_core.so (the C ABI declared in include/gpushare.h).

Fails loudly when the shared library has not been built: there is no Python or CPU
fallback for anything this library does.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libgpushare_core.so")

GS_OK = 0
GS_E_INVAL, GS_E_NOSPACE, GS_E_NOTFOUND, GS_E_WIRE = -1, -2, -3, -4
GS_E_NVML, GS_E_CUDA, GS_E_NOLIB, GS_E_STATE, GS_E_TIMEOUT, GS_E_LIMIT = -5, -6, -7, -8, -9, -10
GS_UNIT_GIB, GS_UNIT_MIB = 0, 1
GS_ALL_GPUS = -1
GS_UUID_MAX = 96
GS_PROBE_MAX_SLICES = 64
GS_PROBE_WRITE, GS_PROBE_VERIFY, GS_PROBE_SCRUB = 1, 2, 4
GS_PROBE_PATTERN_LOWTOGGLE, GS_PROBE_PATTERN_HASH = 0, 1
(GS_ALLOC_MATCHED, GS_ALLOC_SINGLE_GPU, GS_ALLOC_ERR_NO_MATCH, GS_ALLOC_ERR_BAD_INDEX,
 GS_ALLOC_ERR_PODS, GS_ALLOC_ERR_PATCH) = range(6)


class GpuInfo(C.Structure):
    _fields_ = [("uuid", C.c_char * GS_UUID_MAX), ("pci_bus_id", C.c_char * 32),
                ("total_bytes", C.c_uint64), ("nvml_index", C.c_int32),
                ("cuda_ordinal", C.c_int32)]


class HealthEvent(C.Structure):
    _fields_ = [("gpu_index", C.c_int32), ("ignored", C.c_int32), ("xid", C.c_uint64)]


class IoVec(C.Structure):
    _fields_ = [("base", C.c_void_p), ("len", C.c_size_t)]


class Str(C.Structure):
    _fields_ = [("p", C.c_char_p), ("n", C.c_size_t)]


class PodView(C.Structure):
    _fields_ = [("uid", Str), ("ns", Str), ("name", Str), ("node_name", Str),
                ("gpu_mem_limit", C.c_uint64),
                ("ann_idx", Str), ("ann_assume_time", Str), ("ann_assigned", Str)]


class AllocCtx(C.Structure):
    _fields_ = [("table", C.c_void_p), ("node_name", Str),
                ("disable_cgpu_isolation", C.c_int32), ("now_unix_nano", C.c_int64)]


class AllocResult(C.Structure):
    _fields_ = [("kind", C.c_int32), ("pod_index", C.c_int32), ("gpu_index", C.c_int32),
                ("n_containers", C.c_uint32), ("pod_req_units", C.c_uint64),
                ("resp_len", C.c_size_t), ("patch_len", C.c_size_t)]


class Slice(C.Structure):
    _fields_ = [("base", C.c_void_p), ("bytes", C.c_uint64), ("seed", C.c_uint32),
                ("tag", C.c_uint32), ("pattern", C.c_uint32), ("reserved", C.c_uint32)]


class SliceResult(C.Structure):
    _fields_ = [("mismatches", C.c_uint64), ("first_bad_byte", C.c_uint64),
                ("bit_flips", C.c_uint64), ("reserved", C.c_uint64)]


class ProbeTiming(C.Structure):
    _fields_ = [("write_ms", C.c_float), ("verify_ms", C.c_float), ("bytes", C.c_uint64),
                ("launches", C.c_uint32), ("scrub_ms", C.c_float)]


ServerAllocateFn = C.CFUNCTYPE(C.c_int, C.c_void_p, C.POINTER(C.c_uint8), C.c_size_t,
                               C.POINTER(C.c_uint8), C.c_size_t, C.POINTER(C.c_size_t))
ServerDecisionFn = C.CFUNCTYPE(None, C.c_void_p, C.POINTER(AllocResult))


class ServerConfig(C.Structure):
    _fields_ = [("socket_path", C.c_char_p), ("allocate", ServerAllocateFn),
                ("on_decision", ServerDecisionFn), ("user", C.c_void_p),
                ("max_request_bytes", C.c_size_t)]


class ServerStats(C.Structure):
    _fields_ = [("connections", C.c_uint64), ("requests", C.c_uint64),
                ("allocate_calls", C.c_uint64), ("law_messages", C.c_uint64),
                ("protocol_errors", C.c_uint64), ("bytes_in", C.c_uint64),
                ("bytes_out", C.c_uint64), ("recv_body_ns", C.c_uint64), ("allocate_ns", C.c_uint64)]


_P = C.POINTER
_vp, _sz, _u32, _u64, _i32, _i64, _int = (C.c_void_p, C.c_size_t, C.c_uint32, C.c_uint64,
                                          C.c_int32, C.c_int64, C.c_int)
_u8p = _P(C.c_uint8)

# name -> (restype, argtypes); mirrors include/gpushare.h one to one
SIGNATURES = {
    "gs_strerror": (C.c_char_p, [_int]),
    "gs_version": (C.c_char_p, []),
    "gs_translate_memory_unit": (_int, [C.c_char_p, _sz, _P(_int), _P(_int)]),
    "gs_unit_name": (C.c_char_p, [_int]),
    "gs_nvml_open": (_int, [C.c_char_p]),
    "gs_nvml_close": (_int, []),
    "gs_nvml_device_count": (_int, [_P(_u32)]),
    "gs_nvml_device_info": (_int, [_u32, _P(GpuInfo)]),
    "gs_nvml_last_error": (C.c_char_p, []),
    "gs_cuda_map_ordinals": (_int, [_P(GpuInfo), _u32]),
    "gs_cuda_last_error": (C.c_char_p, []),
    "gs_health_open": (_int, [_P(_u64)]),
    "gs_health_wait": (_int, [_u32, _P(HealthEvent)]),
    "gs_health_close": (_int, []),
    "gs_units_from_raw": (_u64, [_u64, _int]),
    "gs_table_build": (_int, [_P(C.c_char_p), _P(_u32), _u32, _u64, _int, _P(_vp)]),
    "gs_table_build_from_gpus": (_int, [_P(GpuInfo), _u32, _int, _P(_vp)]),
    "gs_table_free": (None, [_vp]),
    "gs_table_gpu_count": (_u32, [_vp]),
    "gs_table_units_per_gpu": (_u64, [_vp]),
    "gs_table_device_count": (_u64, [_vp]),
    "gs_table_unit": (_int, [_vp]),
    "gs_fake_id_format": (_int, [_vp, _u32, _u64, C.c_char_p, _sz, _P(_sz)]),
    "gs_extract_real_device_id": (_sz, [C.c_char_p, _sz]),
    "gs_id_lookup": (_int, [_vp, C.c_char_p, _sz, _P(_u32), _P(_u64)]),
    "gs_device_name_by_index": (_int, [_vp, _u64, _P(C.c_void_p), _P(_sz)]),
    "gs_device_index_by_name": (_int, [_vp, C.c_char_p, _sz, _P(_u32)]),
    "gs_law_mark_unhealthy_device": (_int, [_vp, _u32, _u64]),
    "gs_law_mark_unhealthy_id": (_int, [_vp, C.c_char_p, _sz]),
    "gs_law_mark_unhealthy_gpu": (_int, [_vp, _i32]),
    "gs_law_reset_health": (_int, [_vp]),
    "gs_law_device_healthy": (_int, [_vp, _u32, _u64]),
    "gs_law_generation": (_u64, [_vp]),
    "gs_law_payload_size": (_sz, [_vp]),
    "gs_law_encode": (_int, [_vp, _vp, _sz, _P(_sz)]),
    "gs_law_payload": (_int, [_vp, _P(IoVec), _u32, _P(_u32)]),
    "gs_allocate": (_int, [_P(AllocCtx), _vp, _sz, _P(PodView), _sz, _vp, _sz, _vp, _sz,
                           _P(AllocResult)]),
    "gs_allocate_error_response": (_int, [_P(AllocCtx), _vp, _sz, _i32, _vp, _sz,
                                          _P(AllocResult)]),
    "gs_request_scan": (_int, [_vp, _sz, _P(_u32), _u32, _P(_u32), _P(_u64)]),
    "gs_quantity_value": (_int, [C.c_char_p, _sz, _P(_i64)]),
    "gs_go_atoi": (_int, [C.c_char_p, _sz, _P(_i64)]),
    "gs_go_parse_uint64": (_int, [C.c_char_p, _sz, _P(_u64)]),
    "gs_plugin_new": (_int, [_vp, C.c_char_p, _sz, _i32, _P(_vp)]),
    "gs_plugin_free": (None, [_vp]),
    "gs_plugin_set_pods": (_int, [_vp, _P(PodView), _sz]),
    "gs_plugin_pod_count": (_int, [_vp, _P(_sz)]),
    "gs_plugin_get_pod": (_int, [_vp, _sz, _P(PodView)]),
    "gs_plugin_set_clock": (_int, [_vp, _i64]),
    "gs_plugin_allocate": (_int, [_vp, _vp, _sz, _vp, _sz, _P(AllocResult)]),
    "gs_plugin_get_device_plugin_options": (_int, [_vp, _vp, _sz, _P(_sz)]),
    "gs_plugin_pre_start_container": (_int, [_vp, _vp, _sz, _vp, _sz, _P(_sz)]),
    "gs_plugin_wait_health_change": (_int, [_vp, _P(_u64), _u32]),
    "gs_plugin_notify_health": (_int, [_vp]),
    "gs_plugin_stop": (_int, [_vp]),
    "gs_plugin_stopped": (_int, [_vp]),
    "gs_server_start": (_int, [_vp, _P(ServerConfig), _P(_vp)]),
    "gs_server_stop": (_int, [_vp]),
    "gs_server_get_stats": (_int, [_vp, _P(ServerStats)]),
    "gs_register": (_int, [C.c_char_p, C.c_char_p, C.c_char_p, _u32]),
    "gs_grpc_client_connect": (_int, [C.c_char_p, _u32, _P(_vp)]),
    "gs_grpc_client_close": (None, [_vp]),
    "gs_grpc_client_unary": (_int, [_vp, C.c_char_p, C.c_char_p, _sz, _vp, _sz, _P(_sz), _P(_i32),
                                    C.c_char_p, _sz]),
    "gs_grpc_client_stream_open": (_int, [_vp, C.c_char_p, C.c_char_p, _sz, _P(_u32)]),
    "gs_grpc_client_stream_next": (_int, [_vp, _u32, _vp, _sz, _P(_sz), _P(_i32), _u32]),
    "gs_grpc_client_stream_next_borrow": (_int, [_vp, _u32, _P(_vp), _P(_sz), _P(_i32), _u32]),
    "gs_grpc_client_stream_cancel": (_int, [_vp, _u32]),
    "gs_grpc_last_error": (C.c_char_p, []),
    "gs_probe_open": (_int, [_int, _P(_vp)]),
    "gs_probe_close": (_int, [_vp]),
    "gs_probe_sm_count": (_int, [_vp, _P(_int)]),
    "gs_probe_mem_alloc": (_int, [_vp, _u64, _P(_vp)]),
    "gs_probe_mem_free": (_int, [_vp, _vp]),
    "gs_probe_mem_info": (_int, [_vp, _P(_u64), _P(_u64)]),
    "gs_probe_test_skip_stores": (_int, [_vp, _int]),
    "gs_probe_next_seed": (_int, [_vp, _P(_u32)]),
    "gs_probe_submit": (_int, [_vp, _P(Slice), _u32, _u32]),
    "gs_probe_wait": (_int, [_vp, _P(SliceResult), _u32, _P(ProbeTiming)]),
    "gs_probe_poll": (_int, [_vp]),
    "gs_probe_run": (_int, [_vp, _P(Slice), _u32, _u32, _P(SliceResult), _P(ProbeTiming)]),
    "gs_probe_run_granted": (_int, [_int, _u64, _u32, _P(SliceResult), _P(ProbeTiming)]),
    "gs_probe_inject": (_int, [_vp, _vp, _u64, _u32]),
    "gs_probe_read": (_int, [_vp, _vp, _vp, _u64]),
    "gs_probe_set_variant": (_int, [_vp, _int, _int, _int]),
    "gs_probe_set_write_rate_auto": (_int, [_vp, _int]),
    "gs_probe_get_write_rate": (_int, [_vp, _P(_u32), _P(_int)]),
    "gs_probe_set_write_rate": (_int, [_vp, _u32]),
    "gs_probe_verify_noalloc": (_int, [_int, _vp, _P(Slice), _u32, _P(SliceResult)]),
}

_lib = None


class GpushareError(RuntimeError):
    def __init__(self, status, where=""):
        self.status = status
        detail = ""
        if _lib is not None:
            detail = _lib.gs_strerror(status).decode()
            if status == GS_E_CUDA:
                detail += ": " + _lib.gs_cuda_last_error().decode()
            elif status in (GS_E_NVML, GS_E_NOLIB):
                detail += ": " + _lib.gs_nvml_last_error().decode()
        super().__init__(f"{where} failed with status {status} ({detail})")


def load(path=None, strict=True):
    """Load the shared library and declare every prototype. Raises if it is missing.

    strict=False (tuning tools only) tolerates a library built from a subset of csrc/.
    """
    global _lib
    if _lib is not None and path is None:
        return _lib
    path = path or LIB_PATH
    if not os.path.exists(path):
        raise ImportError(
            f"{path} is not built: run `make` (or __graft_entry__.build()) first; "
            "gpushare has no CPU/Python fallback")
    lib = C.CDLL(path)
    for name, (res, args) in SIGNATURES.items():
        try:
            fn = getattr(lib, name)  # AttributeError = header/library mismatch: stay loud
        except AttributeError:
            if strict:
                raise
            continue
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(status, where=""):
    if status != GS_OK:
        raise GpushareError(status, where)
    return status
