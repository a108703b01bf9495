"""podgetter -- print what kubelet's /pods answers (cmd/podgetter/main.go:19-57).

A debugging aid: same four flags, kubelet at 127.0.0.1:10250, service-account token when no
credentials are given. Prints the decoded pod list as JSON (the reference prints Go's %v).
"""
import argparse
import json
import sys

from . import k8s_io


def main(argv=None):
    ap = argparse.ArgumentParser(prog="podgetter", allow_abbrev=False)
    for name, default in (("client-cert", ""), ("client-key", ""), ("token", "")):
        ap.add_argument("-" + name, "--" + name, default=default)
    ap.add_argument("-timeout", "--timeout", type=int, default=10)
    ap.add_argument("--kubelet-url", default="https://127.0.0.1:10250", help="(addition) where kubelet listens")
    a = ap.parse_args(argv)
    token = a.token
    if not (a.client_cert or a.client_key or token):
        try:
            token = open(k8s_io.SA_DIR + "/token").read()
        except OSError as e:
            sys.exit("in cluster mode, find token failed, error: %s" % e)
    t = k8s_io.HttpTransport(None, a.kubelet_url, token, a.client_cert or None, a.client_key or None,
                             timeout=float(a.timeout))
    try:
        print(json.dumps(t.kubelet_pods(), indent=1))
    except k8s_io.ApiError as e:
        print(e)
    return 0


if __name__ == "__main__":
    sys.exit(main())
