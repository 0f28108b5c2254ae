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


# Base ports stay below 30000: the nodes listen on base, base + 1500 + i and base + 1600 + i, and from
# 32768 on a port may already belong to somebody's outgoing connection (the ephemeral range).


def test_local_sh_stops_when_a_node_cannot_start():
    """A node whose port is taken dies at start-up; the launcher must stop the other nodes and fail
    instead of leaving the scheduler waiting for the missing registration (it used to hang)."""
    import socket
    import time
    base = random.randint(10000, 30000)
    blocker = socket.socket()
    blocker.bind(("0.0.0.0", base + 1500))      # the port of worker W0
    blocker.listen(1)
    try:
        t0 = time.time()
        r = subprocess.run([str(ROOT / "script" / "local.sh"), str(HELLO), "1", "2"], cwd=ROOT,
                           env=dict(os.environ, PS_BASE_PORT=str(base)), capture_output=True, text=True, timeout=60)
        assert r.returncode != 0 and "cannot bind" in r.stderr, r.stdout + r.stderr
        assert time.time() - t0 < 30
    finally:
        blocker.close()


def test_local_sh_hello_world():
    env = dict(os.environ, PS_BASE_PORT=str(random.randint(10000, 30000)))
    r = subprocess.run([str(ROOT / "script" / "local.sh"), str(HELLO), "2", "2"], cwd=ROOT, env=env,
                       capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stdout + r.stderr
    _check_hello(r.stdout)


def test_mpi_node_sh_ranks_from_environment(tmp_path):
    """Every process gets the same command line; role, id, address and port come from its rank."""
    port = random.randint(10000, 30000)
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


def test_text2proto_and_show_example(tmp_path):
    """text -> RecordIO of Example protos (+ .info); a DataConfig{format: PROTO} reads it back."""
    from parameter_server_b200.utils.data_io import load_slot_data
    from parameter_server_b200.utils.synthetic import make_rcv1_like, write_libsvm
    shard = make_rcv1_like(200, 80, 8.0, seed=5)
    write_libsvm(shard, str(tmp_path / "a.txt"))
    r = subprocess.run([str(ROOT / "build" / "text2proto"), "-input", str(tmp_path / "a.txt"), "-format", "libsvm",
                        "-output", str(tmp_path / "a")], capture_output=True, text=True, timeout=60)
    assert r.returncode == 0 and "written 200 examples" in r.stderr, r.stderr
    info = (tmp_path / "a.info").read_text()
    assert "num_ex: 200" in info and f"nnz_ele: {shard.nnz}" in info
    r = subprocess.run([str(ROOT / "build" / "show_example"), "-input", str(tmp_path / "a.recordio"), "-n", "2"],
                       capture_output=True, text=True, timeout=60)
    assert r.returncode == 0 and r.stdout.count("example ") == 2 and "slot { id: 0 val:" in r.stdout
    g = load_slot_data(f'format: PROTO file: "{tmp_path}/a.recordio"', str(tmp_path / "cache_"))[1]
    assert g.n_rows == 200 and g.nnz == shard.nnz
    import torch
    assert torch.equal(g.keys, shard.keys)


def test_selftest_binary():
    """build/selftest: native unit checks, in-process clusters, node replacement, apps. The same
    binary is what tools/sanitize.py runs under ThreadSanitizer / AddressSanitizer (make tsan / asan)."""
    exe = ROOT / "build" / "selftest"
    assert exe.exists()
    r = subprocess.run([str(exe), "all"], capture_output=True, text=True, timeout=170)
    assert r.returncode == 0 and "0 failure(s)" in r.stderr, r.stderr[-3000:]
    assert r.stderr.count("\nok ") + r.stderr.startswith("ok ") >= 12


def test_gzip_text_input(tmp_path):
    """*.gz training files are read through zlib (ref src/util/file.cc gz support)"""
    import gzip
    import torch
    from parameter_server_b200.utils.data_io import load_slot_data
    from parameter_server_b200.utils.synthetic import make_rcv1_like, write_libsvm
    shard = make_rcv1_like(150, 60, 6.0, seed=8)
    write_libsvm(shard, str(tmp_path / "p.txt"))
    with open(tmp_path / "p.txt", "rb") as f, gzip.open(tmp_path / "p.txt.gz", "wb") as g:
        g.write(f.read())
    os.remove(tmp_path / "p.txt")
    g = load_slot_data(f'format: TEXT text: LIBSVM file: "{tmp_path}/p.txt.gz"', str(tmp_path / "cache_"))[1]
    assert g.n_rows == 150 and g.nnz == shard.nnz and torch.equal(g.keys, shard.keys)


def test_observability_flags(tmp_path):
    """-print_van (message trace per node), -traffic_statistics (bytes at exit), -log_to_file
    (ref SURVEY 5.1 / 5.5: van.cc, postoffice.cc)"""
    env = dict(os.environ, PS_BASE_PORT=str(random.randint(10000, 30000)))
    r = subprocess.run([str(ROOT / "script" / "local.sh"), str(HELLO), "1", "2", "-print_van", "-traffic_statistics",
                        "-log_to_file"], cwd=tmp_path, env=env, capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stdout + r.stderr
    names = sorted(os.listdir(tmp_path))
    for node in ("H", "S0", "W0", "W1"):
        assert f"van_{node}" in names and f"log_{node}" in names, names
    trace = (tmp_path / "van_W0").read_text()
    assert "SND" in trace and "RCV" in trace and "shared_para { cmd: PULL }" in trace
    logs = "".join((tmp_path / f"log_{n}").read_text() for n in ("H", "S0", "W0", "W1")) + r.stderr
    assert logs.count("Gbyte") >= 4            # one traffic line per node
    _check_hello(r.stdout)
