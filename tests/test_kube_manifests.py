"""docker/kube/submit.py: the Kubernetes manifests of a job (no cluster needed to check them)."""
import json
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent


def test_manifests_describe_scheduler_servers_workers(tmp_path):
    conf = tmp_path / "app.conf"
    conf.write_text("linear_method { darlin { max_pass_of_data: 3 } }")
    out = subprocess.run([sys.executable, str(ROOT / "docker" / "kube" / "submit.py"), "up", "--name", "job", "--servers", "3",
                          "--workers", "5", "--app-file", str(conf), "--gpu"], capture_output=True, text=True, check=True).stdout
    docs = [json.loads(d) for d in out.split("\n---\n")]
    kinds = [(d["kind"], d["metadata"]["name"]) for d in docs]
    assert kinds == [("ConfigMap", "job-config"), ("Service", "job"), ("StatefulSet", "job-scheduler"),
                     ("StatefulSet", "job-server"), ("StatefulSet", "job-worker")]
    assert docs[0]["data"]["app.conf"].startswith("linear_method")
    assert docs[1]["spec"]["clusterIP"] == "None"
    sched, server, worker = docs[2:]
    assert (sched["spec"]["replicas"], server["spec"]["replicas"], worker["spec"]["replicas"]) == (1, 3, 5)
    cmd = worker["spec"]["template"]["spec"]["containers"][0]["command"][2]
    assert "-num_servers 3 -num_workers 5" in cmd and "-device gpu" in cmd and "id:'W$ORDINAL'" in cmd
    assert "role:SCHEDULER,hostname:'job-scheduler-0.job',port:8000,id:'H'" in cmd
    assert worker["spec"]["template"]["spec"]["containers"][0]["resources"]["limits"]["nvidia.com/gpu"] == 1
    assert "resources" not in sched["spec"]["template"]["spec"]["containers"][0]
    down = subprocess.run([sys.executable, str(ROOT / "docker" / "kube" / "submit.py"), "down", "--name", "job"],
                          capture_output=True, text=True, check=True).stdout
    assert "kubectl delete statefulset job-worker" in down
