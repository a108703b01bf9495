"""Renders the Kubernetes manifests of the plugin.

Drop-in names are kept from the reference's manifests so an operator can swap the image
only: DaemonSet `gpushare-device-plugin-ds` in kube-system (labels component/app/name,
nodeSelector gpushare=true, hostNetwork, serviceAccount `gpushare-device-plugin`, args
`-logtostderr --v=5 --memory-unit=GiB`, env KUBECONFIG / NODE_NAME, hostPath
/var/lib/kubelet/device-plugins) -- device-plugin-ds.yaml:1-58 -- and the ClusterRole /
ServiceAccount / ClusterRoleBinding `gpushare-device-plugin` -- device-plugin-rbac.yaml:1-59.
What differs is listed under DIFFERENCES and asserted by tests/test_packaging.py.

    python deploy/render.py            # rewrites deploy/*.yaml
"""
import os

import yaml

HERE = os.path.dirname(os.path.abspath(__file__))
NAME = "gpushare-device-plugin"
LABELS = {"component": NAME, "app": "gpushare", "name": NAME + "-ds"}
PLUGIN_DIR = "/var/lib/kubelet/device-plugins"

DIFFERENCES = [
    "image: a CUDA runtime image carrying libgpushare_core.so and the Python host (deploy/Dockerfile)",
    "memory limit 512Mi instead of 300Mi: the pre-encoded ListAndWatch payload of 8 B200s at MiB "
    "granularity alone is 90 MB",
    "NVIDIA_VISIBLE_DEVICES=all and NVIDIA_DRIVER_CAPABILITIES=compute,utility so the probe can run "
    "(the reference needs `utility` only: it never touches CUDA)",
    "optional args --health-check and --probe (the probe costs each GPU the plugin's CUDA context, about "
    "0.5 GiB, and a 64 MiB resident sentinel range; it sweeps HBM that is free when it runs, never a tenant's)",
]


def daemonset(memory_unit="GiB", probe=False, health_check=True, image="gpushare-b200:latest"):
    args = ["-logtostderr", "--v=5", "--memory-unit=%s" % memory_unit]
    if health_check:
        args.append("--health-check")
    if probe:
        args.append("--probe")
    container = {
        "name": "gpushare",
        "image": image,
        "command": ["gpushare-device-plugin-v2"] + args,
        "env": [
            {"name": "KUBECONFIG", "value": "/etc/kubernetes/kubelet.conf"},
            {"name": "NODE_NAME", "valueFrom": {"fieldRef": {"fieldPath": "spec.nodeName"}}},
            {"name": "NVIDIA_VISIBLE_DEVICES", "value": "all"},
            {"name": "NVIDIA_DRIVER_CAPABILITIES", "value": "compute,utility"},
        ],
        # requests == limits: a Guaranteed pod is never evicted for node pressure
        "resources": {"limits": {"cpu": "1", "memory": "512Mi"}, "requests": {"cpu": "1", "memory": "512Mi"}},
        "securityContext": {"allowPrivilegeEscalation": False, "capabilities": {"drop": ["ALL"]}},
        "volumeMounts": [{"name": "device-plugin", "mountPath": PLUGIN_DIR}],
    }
    return {
        "apiVersion": "apps/v1", "kind": "DaemonSet",
        "metadata": {"name": NAME + "-ds", "namespace": "kube-system"},
        "spec": {
            "selector": {"matchLabels": dict(LABELS)},
            "template": {
                "metadata": {"labels": dict(LABELS),
                             "annotations": {"scheduler.alpha.kubernetes.io/critical-pod": ""}},
                "spec": {
                    "serviceAccount": NAME, "hostNetwork": True, "nodeSelector": {"gpushare": "true"},
                    "containers": [container],
                    "volumes": [{"name": "device-plugin", "hostPath": {"path": PLUGIN_DIR}}],
                },
            },
        },
    }


def rbac():
    rule = lambda res, verbs: {"apiGroups": [""], "resources": res, "verbs": verbs}  # noqa: E731
    return [
        {"apiVersion": "rbac.authorization.k8s.io/v1", "kind": "ClusterRole", "metadata": {"name": NAME},
         "rules": [
             rule(["nodes", "nodes/proxy"], ["get", "list", "watch"]),          # GET node; kubelet /pods
             rule(["nodes/status"], ["patch", "update"]),                      # aliyun.com/gpu-count
             rule(["pods"], ["get", "list", "watch", "patch", "update"]),      # LIST pending, PATCH assigned
             rule(["events"], ["create", "patch"]),
         ]},
        {"apiVersion": "v1", "kind": "ServiceAccount", "metadata": {"name": NAME, "namespace": "kube-system"}},
        {"apiVersion": "rbac.authorization.k8s.io/v1", "kind": "ClusterRoleBinding",
         "metadata": {"name": NAME},
         "roleRef": {"apiGroup": "rbac.authorization.k8s.io", "kind": "ClusterRole", "name": NAME},
         "subjects": [{"kind": "ServiceAccount", "name": NAME, "namespace": "kube-system"}]},
    ]


def main():
    head = "# rendered by deploy/render.py -- edit that, not this\n"
    with open(os.path.join(HERE, "gpushare-b200-daemonset.yaml"), "w") as f:
        f.write(head)
        yaml.safe_dump(daemonset(), f, sort_keys=True, default_flow_style=False)
    with open(os.path.join(HERE, "gpushare-b200-daemonset-mib-probe.yaml"), "w") as f:
        f.write(head)
        yaml.safe_dump(daemonset("MiB", probe=True), f, sort_keys=True, default_flow_style=False)
    with open(os.path.join(HERE, "gpushare-b200-rbac.yaml"), "w") as f:
        f.write(head)
        yaml.safe_dump_all(rbac(), f, sort_keys=True, default_flow_style=False)


if __name__ == "__main__":
    main()
