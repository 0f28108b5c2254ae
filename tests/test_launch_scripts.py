"""The launch scripts start real processes that talk over the TCP van (ref script/local.sh,
script/mpi_node.sh and the expected output quoted in src/app/hello_world/main.cc)."""
import os
import random
import subprocess
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent
HELLO = ROOT / "build" / "hello"


@pytest.fixture(scope="module", autouse=True)
def _binaries():
    if not HELLO.exists():
        from parameter_server_b200 import build
        build.build_host()
    assert HELLO.exists()


def _check_hello(out: str):
    lines = [l for l in out.splitlines() if "pulled weight" in l]
    got = {l.split(":")[0]: " ".join(l.split(":", 1)[1].split()) for l in lines}
    assert got == {
        "W0": "pulled weight [4]: 0 0.2 0.4 0.5 at key [4]: 0 2 4 5",
        "W1": "pulled weight [4]: 0 0.1 0.3 0.4 at key [4]: 0 1 3 4",
    }, out


def test_local_sh_hello_world():
    env = dict(os.environ, PS_BASE_PORT=str(random.randint(20000, 40000)))
    r = subprocess.run([str(ROOT / "script" / "local.sh"), str(HELLO), "2", "2"], cwd=ROOT, env=env,
                       capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stdout + r.stderr
    _check_hello(r.stdout)


def test_mpi_node_sh_ranks_from_environment(tmp_path):
    """Every process gets the same command line; role, id, address and port come from its rank."""
    port = random.randint(20000, 40000)
    conf = tmp_path / "launch.conf"
    conf.write_text(f"bin={HELLO}\nnum_servers=1\nnum_workers=2\nnetwork_interface=\narg=\n")
    root = subprocess.run([str(ROOT / "script" / "get_root_node.sh"), "lo", str(port)],
                          capture_output=True, text=True, timeout=30).stdout.strip()
    assert root.startswith("role:SCHEDULER,hostname:'") and f"port:{port}" in root
    sched = f"role:SCHEDULER,hostname:'127.0.0.1',port:{port},id:'H'"
    procs = []
    for rank in range(4):
        env = dict(os.environ, RANK=str(rank), WORLD_SIZE="4")
        for k in ("PMI_RANK", "OMPI_COMM_WORLD_RANK", "SLURM_PROCID", "PMI_SIZE", "OMPI_COMM_WORLD_SIZE",
                  "SLURM_NTASKS"):
            env.pop(k, None)
        procs.append(subprocess.Popen([str(ROOT / "script" / "mpi_node.sh"), sched, str(conf)], cwd=ROOT,
                                      env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True))
    outs = []
    try:
        for p in procs:
            o, _ = p.communicate(timeout=120)
            outs.append(o)
    finally:
        for p in procs:
            if p.poll() is None:
                p.kill()
    assert all(p.returncode == 0 for p in procs), "\n".join(outs)
    _check_hello("\n".join(outs))
