"""kubectl-inspect-gpushare-v2 -- per-node / per-GPU allocation report.

Restates cmd/inspect of the reference: main.go:33-79 (flags: -d, optional node name),
nodeinfo.go:47-272 (node/device bookkeeping, the allocation annotation), podinfo.go:58-135
(pod listing with 5 retries x 100 ms, active-pod filter), display.go:15-255 (the two tables,
laid out by Go's text/tabwriter with padding 2).

Deliberate differences, each recorded here:
  * memory unit: the reference guesses "MiB" when a GPU advertises more than 100 units
    (nodeinfo.go:228-244), which mislabels a 179-GiB B200 in GiB mode. Here the guess flips at
    1024 units per GPU (no GPU has more than 1 TiB, none less than 1 GiB), and --unit overrides.
  * Go iterates maps in random order (nodes: nodeinfo.go:131-133; devices: display.go:60). Here
    nodes keep the order the apiserver returned and devices go by index, pending (-1) last.
"""
import argparse
import json
import os
import sys
import time

from . import api, k8s_io

resourceName = "aliyun.com/gpu-mem"                      # main.go:14
countName = "aliyun.com/gpu-count"                       # main.go:15
envNVGPUID = "ALIYUN_COM_GPU_MEM_IDX"                    # main.go:22
gpushareAllocationFlag = "scheduler.framework.gpushare.allocation"   # main.go:25
RETRIES = 5                                              # podinfo.go:25


# ---------------------------------------------------------------- text/tabwriter --------
def tabwrite(text, padding=2):
    """Go's text/tabwriter with NewWriter(out, 0, 0, padding, ' ', 0): tab-terminated cells of
    consecutive lines form column blocks; a block is as wide as its widest cell plus the
    padding; the cell after a line's last tab is written as it is."""
    lines = [ln.split("\t") for ln in text.split("\n")]
    out = [None] * len(lines)

    def write_lines(widths, a, b):
        for i in range(a, b):
            cells, s = lines[i], ""
            for j, c in enumerate(cells):
                if j < len(widths) and j < len(cells) - 1:
                    s += c + " " * (widths[j] - len(c))
                else:
                    s += c
            out[i] = s

    def fmt(widths, line0, line1):
        column = len(widths)
        this = line0
        while this < line1:
            if column >= len(lines[this]) - 1:
                this += 1
                continue
            write_lines(widths, line0, this)
            line0 = this
            width = 0
            while this < line1 and column < len(lines[this]) - 1:
                width = max(width, len(lines[this][column]) + padding)
                this += 1
            fmt(widths + [width], line0, this)
            line0 = this
        write_lines(widths, line0, line1)

    fmt([], 0, len(lines))
    return "\n".join(out)


# ---------------------------------------------------------------- model -----------------
def _value(q):
    return api.quantity_value(str(q))


def gpu_memory_in_pod(pod):
    """podinfo.go:125-135 / display.go:247-255."""
    total = 0
    for c in (pod.get("spec") or {}).get("containers") or []:
        limits = (c.get("resources") or {}).get("limits") or {}
        if resourceName in limits:
            total += _value(limits[resourceName])
    return total


def get_allocation(pod):
    """nodeinfo.go:245-272: {gpu index: memory} from the scheduler-framework annotation."""
    ann = (pod.get("metadata") or {}).get("annotations")
    if not ann or gpushareAllocationFlag not in ann:
        return {}
    try:
        doc = json.loads(ann[gpushareAllocationFlag])
        out = {}
        for _container, per_gpu in doc.items():
            int(_container)                          # map[int]... keys must parse
            for gid, mem in per_gpu.items():
                out[api.go_atoi(gid)] = out.get(api.go_atoi(gid), 0) + int(mem)
        return out
    except Exception:
        return {}


class DeviceInfo:
    def __init__(self, idx, total):
        self.idx, self.pods, self.used, self.total = idx, [], 0, total

    def __str__(self):                               # nodeinfo.go:23-28
        return "%d" % self.used if self.idx == -1 else "%d/%d" % (self.used, self.total)


class NodeInfo:
    def __init__(self, node):
        alloc = (node.get("status") or {}).get("allocatable") or {}
        self.node, self.pods = node, []
        self.gpu_count = _value(alloc[countName]) if countName in alloc else 0
        self.gpu_total = _value(alloc[resourceName]) if resourceName in alloc else 0
        per = self.gpu_total // self.gpu_count if self.gpu_count else 0
        self.devs = {i: DeviceInfo(i, per) for i in range(self.gpu_count)}

    @property
    def name(self):
        return (self.node.get("metadata") or {}).get("name", "")

    def has_pending(self):
        return -1 in self.devs

    def device_info(self, pod):
        """nodeinfo.go:169-197."""
        allocation = get_allocation(pod)
        if allocation:
            return allocation
        dev_id = -1
        ann = (pod.get("metadata") or {}).get("annotations") or {}
        if ann and envNVGPUID in ann:
            try:
                dev_id = api.go_atoi(ann[envNVGPUID])
            except Exception:
                dev_id = -1
        return {dev_id: gpu_memory_in_pod(pod)}

    def build_device_info(self):
        """nodeinfo.go:143-167."""
        per = self.gpu_total // self.gpu_count if self.gpu_count > 0 else 0
        for pod in self.pods:
            if gpu_memory_in_pod(pod) <= 0:
                continue
            for dev_id, used in self.device_info(pod).items():
                if dev_id not in self.devs:
                    self.devs[dev_id] = DeviceInfo(dev_id, per)
                self.devs[dev_id].used += used
                self.devs[dev_id].pods.append(pod)


def guess_unit(gpu_memory, gpu_count):
    """setUnit (nodeinfo.go:228-244) with the threshold moved from 100 to 1024."""
    if gpu_count == 0:
        return ""
    return "MiB" if gpu_memory // gpu_count > 1024 else "GiB"


def build_all_node_infos(pods, nodes, unit=""):
    """nodeinfo.go:47-60,96-135. -> ([NodeInfo], unit string)."""
    infos, by_name = [], {}
    for node in nodes:
        name = (node.get("metadata") or {}).get("name", "")
        info = by_name.get(name)
        if info is None:
            info = by_name[name] = NodeInfo(node)
            infos.append(info)
        for pod in pods:
            if (pod.get("spec") or {}).get("nodeName") == name:
                info.pods.append(pod)
    for info in infos:
        if info.gpu_total > 0:
            if not unit:
                unit = guess_unit(info.gpu_total, info.gpu_count)
            info.build_device_info()
    return infos, unit


def _address(node):
    for addr in (node.get("status") or {}).get("addresses") or []:
        if addr.get("type") == "InternalIP":
            return addr.get("address", "unknown")
    return "unknown"


def _dev_order(info):
    return sorted(info.devs, key=lambda k: (k == -1, k))


# ---------------------------------------------------------------- display ---------------
def display_summary(infos, unit):
    """display.go:141-245."""
    has_pending = any(i.has_pending() for i in infos)
    max_gpus = max([i.gpu_count for i in infos] or [0])
    text = "NAME\tIPADDRESS\t" + "".join("GPU%d(Allocated/Total)\t" % i for i in range(max_gpus))
    if has_pending:
        text += "PENDING(Allocated)\t"
    text += "GPU Memory(%s)\n" % unit
    used_cluster = total_cluster = 0
    line_len = 0
    for info in infos:
        if info.gpu_total <= 0:
            continue
        used, cells = 0, []
        for i in range(max_gpus):
            if i in info.devs:
                cells.append(str(info.devs[i]))
                used += info.devs[i].used
            else:
                cells.append("0/0")
        pending = ""
        if -1 in info.devs:
            pending = "%d" % info.devs[-1].used
            used += info.devs[-1].used
        line = "%s\t%s\t" % (info.name, _address(info.node)) + "".join(c + "\t" for c in cells)
        if has_pending:
            line += pending + "\t"
        line += "%d/%d\n" % (used, info.gpu_total)
        text += line
        if line_len == 0:
            line_len = len(line) + 20
        used_cluster += used
        total_cluster += info.gpu_total
    text += "-" * line_len + "\n"
    text += "Allocated/Total GPU Memory In Cluster:\n"
    usage = int(used_cluster / total_cluster * 100) if total_cluster > 0 else 0
    text += "%d/%d (%d%%)\t\n" % (used_cluster, total_cluster, usage)
    return tabwrite(text)


def display_details(infos):
    """display.go:15-129."""
    text = ""
    used_cluster = total_cluster = 0
    line_len = 0
    for info in infos:
        if info.gpu_total <= 0:
            continue
        text += "\nNAME:\t%s\nIPADDRESS:\t%s\n\n" % (info.name, _address(info.node))
        text += "NAME\tNAMESPACE\t" + "".join("GPU%d(Allocated)\t" % i for i in range(info.gpu_count))
        if info.has_pending():
            text += "Pending(Allocated)\t"
        text += "\n"
        used, body, seen = 0, "", set()
        for i in _dev_order(info):
            dev = info.devs[i]
            used += dev.used
            for pod in dev.pods:
                md = pod.get("metadata") or {}
                if md.get("uid") in seen:
                    continue
                body += "%s\t%s\t" % (md.get("name", ""), md.get("namespace", ""))
                count = info.gpu_count + (1 if info.has_pending() else 0)
                allocation = get_allocation(pod)
                for k in range(count):
                    if allocation:
                        body += "%d\t" % allocation.get(k, 0)
                    elif k == i or (i == -1 and k == info.gpu_count):
                        body += "%d\t" % gpu_memory_in_pod(pod)
                    else:
                        body += "0\t"
                body += "\n"
                seen.add(md.get("uid"))
        if line_len == 0:
            line_len = len(body) + 10
        text += body
        usage = int(used / info.gpu_total * 100)
        text += "Allocated :\t%d (%d%%)\t\n" % (used, usage)
        text += "Total :\t%d \t\n" % info.gpu_total
        text += "-" * line_len + "\n"
        used_cluster += used
        total_cluster += info.gpu_total
    text += "\n\nAllocated/Total GPU Memory In Cluster:\t"
    usage = int(used_cluster / total_cluster * 100) if total_cluster > 0 else 0
    text += "%d/%d (%d%%)\t\n" % (used_cluster, total_cluster, usage)
    return tabwrite(text)


# ---------------------------------------------------------------- fetch -----------------
def _retry_list(fn, what):
    """podinfo.go:58-94: first try plus up to 5 more, 100 ms after each retry."""
    err = None
    try:
        return fn()
    except Exception as e:
        err = e
    for _ in range(RETRIES):
        try:
            out, err = fn(), None
        except Exception as e:
            err = e
        time.sleep(0.1)
        if err is None:
            return out
    raise RuntimeError(what)


def filter_active_pods(pods):
    """podinfo.go:96-107."""
    return [p for p in pods if (p.get("status") or {}).get("phase") not in ("Succeeded", "Failed")]


def fetch(transport, node_name=""):
    if node_name:
        nodes = [transport.get_node(node_name)]
        doc = _retry_list(lambda: transport._api(
            "GET", "/api/v1/pods?fieldSelector=spec.nodeName%3D" + node_name),
            "failed to get Pods in node %s" % node_name)
    else:
        all_nodes = transport._api("GET", "/api/v1/nodes").get("items") or []
        nodes = [n for n in all_nodes                       # isGPUSharingNode, nodeinfo.go:214-222
                 if resourceName in ((n.get("status") or {}).get("allocatable") or {})
                 and _value(n["status"]["allocatable"][resourceName]) > 0]
        doc = _retry_list(lambda: transport._api("GET", "/api/v1/pods"), "failed to get Pods")
    return filter_active_pods(doc.get("items") or []), nodes


def main(argv=None):
    ap = argparse.ArgumentParser(prog="kubectl-inspect-gpushare-v2")
    ap.add_argument("-d", action="store_true", help="details")
    ap.add_argument("--unit", default="", choices=["", "GiB", "MiB"], help="override the unit guess")
    ap.add_argument("node", nargs="?", default="")
    a = ap.parse_args(argv)
    path = os.environ.get("KUBECONFIG") or os.path.join(os.environ.get("HOME", ""), ".kube/config")
    if not os.path.exists(path):                          # podinfo.go:30-36
        sys.exit("kubeconfig %s failed to find, please set KUBECONFIG env" % path)
    transport = k8s_io.HttpTransport(k8s_io.ApiConfig.from_kubeconfig(path), timeout=15.0)
    try:
        pods, nodes = fetch(transport, a.node)
    except Exception as e:
        sys.stdout.write("Failed due to %s" % e)
        return 1
    infos, unit = build_all_node_infos(pods, nodes, a.unit)
    sys.stdout.write(display_details(infos) if a.d else display_summary(infos, unit))
    return 0


if __name__ == "__main__":
    sys.exit(main())
