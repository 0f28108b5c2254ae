#!/usr/bin/env python3
"""Run a parameter-server job on Kubernetes: one scheduler pod, S server pods, W worker pods.

    docker/kube/submit.py up   --name l1lr --servers 2 --workers 2 --app-file config/batch_l1lr.conf \\
                               --data /mnt/data --output /mnt/model [--gpu] [--image IMAGE] [--apply]
    docker/kube/submit.py down --name l1lr [--apply]

The counterpart of the reference's docker/client/script/submit.py (Python 2, v1beta1 pods and
replicationControllers driven through a `kubecfg` binary that is a missing blob there) for a
current cluster: it WRITES manifests (apps/v1 StatefulSets behind a headless Service, a ConfigMap
with the app config) and, with --apply, hands them to `kubectl`.  What it keeps from the
reference: the node command lines (`build/ps -my_node ... -scheduler ... -num_servers ...
-app_file ...`, docker/client/script/submit.py:60-66), the scheduler on a fixed port, the data and
output directories mounted from the host, one container per node.  What changes: pods find the
scheduler through cluster DNS (<name>-scheduler-0.<name>) instead of a file with the minion's
address, node ids come from the pod ordinal, and --gpu asks for one `nvidia.com/gpu` per server /
worker and starts them with `-device gpu` (model shards in HBM, payloads over CUDA IPC when the
pods share a host with the NVIDIA device plugin's IPC settings).
ref: docker/client/script/submit.py:1-464, docker/server/script/*.sh, docker/client/config/*.conf
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys

SCHEDULER_PORT = 8000
NODE_PORT = 9000


def node_arg(role: str, host: str, port: int, node_id: str) -> str:
    """The text-format Node the binaries take on the command line (src/system/van.h:33-37)."""
    return f"role:{role},hostname:'{host}',port:{port},id:'{node_id}'"


def shell_command(role: str, a) -> str:
    """What a pod runs.  $(ORDINAL) is derived from the StatefulSet pod name at start."""
    sched = node_arg("SCHEDULER", f"{a.name}-scheduler-0.{a.name}", SCHEDULER_PORT, "H")
    common = (f"-num_servers {a.servers} -num_workers {a.workers} -app_file /config/app.conf "
              f"-scheduler \"{sched}\" -num_threads {a.threads}")
    if a.gpu:
        common += " -device gpu"
    if role == "scheduler":
        me = sched
    else:
        prefix = "S" if role == "server" else "W"
        me = node_arg(role.upper(), f"$(hostname).{a.name}", NODE_PORT, f"{prefix}$ORDINAL")
    return ("ORDINAL=${HOSTNAME##*-}; cd /work && exec /opt/psb200/build/ps "
            f"-my_node \"{me}\" {common}")


def stateful_set(role: str, replicas: int, a) -> dict:
    labels = {"app": a.name, "role": role}
    container = {
        "name": role,
        "image": a.image,
        "command": ["sh", "-c", shell_command(role, a)],
        "ports": [{"containerPort": SCHEDULER_PORT if role == "scheduler" else NODE_PORT}],
        "volumeMounts": [{"name": "config", "mountPath": "/config"},
                         {"name": "data", "mountPath": "/work/data"},
                         {"name": "output", "mountPath": "/work/model"}],
    }
    if a.gpu and role != "scheduler":
        container["resources"] = {"limits": {"nvidia.com/gpu": 1}}
    spec = {
        "serviceName": a.name, "replicas": replicas, "podManagementPolicy": "Parallel",
        "selector": {"matchLabels": labels},
        "template": {"metadata": {"labels": labels},
                     "spec": {"restartPolicy": "Always", "containers": [container],
                              # -device gpu moves payloads through CUDA IPC: same host, shared IPC namespace
                              **({"hostIPC": True} if a.gpu else {}),
                              "volumes": [{"name": "config", "configMap": {"name": f"{a.name}-config"}},
                                          {"name": "data", "hostPath": {"path": a.data}},
                                          {"name": "output", "hostPath": {"path": a.output}}]}},
    }
    return {"apiVersion": "apps/v1", "kind": "StatefulSet",
            "metadata": {"name": f"{a.name}-{role}", "labels": labels}, "spec": spec}


def manifests(a) -> list:
    conf = open(a.app_file).read() if a.app_file and os.path.exists(a.app_file) else (a.app_conf or "")
    out = [
        {"apiVersion": "v1", "kind": "ConfigMap", "metadata": {"name": f"{a.name}-config"},
         "data": {"app.conf": conf}},
        # headless: gives every pod the DNS name <pod>.<name>
        {"apiVersion": "v1", "kind": "Service", "metadata": {"name": a.name},
         "spec": {"clusterIP": "None", "selector": {"app": a.name},
                  "ports": [{"name": "sched", "port": SCHEDULER_PORT}, {"name": "node", "port": NODE_PORT}]}},
        stateful_set("scheduler", 1, a),
        stateful_set("server", a.servers, a),
        stateful_set("worker", a.workers, a),
    ]
    return out


def kubectl(args: list, doc: str | None = None) -> int:
    return subprocess.run(["kubectl", *args], input=doc, text=True).returncode


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("action", choices=["up", "down"])
    ap.add_argument("--name", default="psb200")
    ap.add_argument("--servers", type=int, default=2)
    ap.add_argument("--workers", type=int, default=2)
    ap.add_argument("--threads", type=int, default=2)
    ap.add_argument("--app-file", default="", help="AppConfig text file (e.g. example/linear/ctr/batch_l1lr.conf)")
    ap.add_argument("--app-conf", default="", help="the same as a string")
    ap.add_argument("--data", default="/mnt/psb200/data", help="host directory mounted at /work/data")
    ap.add_argument("--output", default="/mnt/psb200/model", help="host directory mounted at /work/model")
    ap.add_argument("--image", default="psb200:latest", help="built from docker/Dockerfile")
    ap.add_argument("--gpu", action="store_true", help="one GPU per server / worker, -device gpu")
    ap.add_argument("--apply", action="store_true", help="run kubectl instead of printing the manifests")
    a = ap.parse_args(argv)
    if a.action == "down":
        kinds = [("statefulset", f"{a.name}-{r}") for r in ("worker", "server", "scheduler")] + \
                [("service", a.name), ("configmap", f"{a.name}-config")]
        for kind, name in kinds:
            if a.apply:
                kubectl(["delete", kind, name, "--ignore-not-found"])
            else:
                print(f"kubectl delete {kind} {name} --ignore-not-found")
        return 0
    docs = manifests(a)
    text = "\n---\n".join(json.dumps(d, indent=2) for d in docs)  # JSON is YAML
    if a.apply:
        return kubectl(["apply", "-f", "-"], text)
    print(text)
    return 0


if __name__ == "__main__":
    sys.exit(main())
