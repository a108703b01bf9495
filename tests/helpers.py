"""Shared by the test modules: run one golden case through the product and the oracle."""
import json

from gpushare_b200 import api
from oracle import ref_model as R


def product_for_case(c):
    uuids = [g[0] for g in c["gpus"]]
    table = api.DeviceTable.from_raw(uuids, c["gpus"][0][1] if c["gpus"] else 0, c["unit"])
    plug = api.NvidiaDevicePlugin(table, c["node"], c["cgpu_disable"])
    plug.set_clock(c["clock"])
    plug.set_pods([api.Pod.from_k8s(p) for p in c["pods"]])
    return table, plug


def oracle_for_case(c):
    g = R.Globals(c["unit"])
    src = R.PodSource(c["pods"], fail_list=c["fail_list"], fail_patch=c["fail_patch"])
    plug = R.NvidiaDevicePlugin(g, [tuple(x) for x in c["gpus"]], c["node"], src,
                                c["cgpu_disable"], clock=lambda: c["clock"])
    return src, plug


def product_step(c, table, plug, request_lists):
    """One Allocate through the C ABI, with the caller-side failure handling the
    reference has around its lister and patcher (allocate.go:64-67,133-151).
    Returns (env dicts, kind, patched pod name or None)."""
    wire = R.encode_allocate_request(request_lists)
    if c["fail_list"]:
        d = api.allocate_error_response(table, c["node"], wire, api.L.GS_ALLOC_ERR_PODS)
        return R.decode_allocate_response(d.response), d.kind, None
    if c["fail_patch"] == "always":
        # decision is made statelessly, the PATCH fails, the caller asks for the error form
        pods = [plug.get_pod(i) for i in range(plug.pod_count())]
        d = api.allocate(table, c["node"], wire, pods, c["clock"], c["cgpu_disable"])
        if d.kind == "matched":
            d = api.allocate_error_response(table, c["node"], wire, api.L.GS_ALLOC_ERR_PATCH)
        return R.decode_allocate_response(d.response), d.kind, None
    d = plug.Allocate(wire)
    patched = plug.get_pod(d.pod_index).name if d.kind == "matched" else None
    return R.decode_allocate_response(d.response), d.kind, patched


def patch_body(clock):
    return json.dumps({"metadata": {"annotations": {
        "ALIYUN_COM_GPU_MEM_ASSIGNED": "true",
        "ALIYUN_COM_GPU_MEM_ASSUME_TIME": "%d" % clock}}}, sort_keys=True,
        separators=(",", ":")).encode()
