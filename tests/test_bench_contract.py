"""bench.py's command line as far as it can be checked without a GPU: the reference arm prints its one
JSON line and exits 0, unknown options are refused, the clock sampler degrades without nvidia-smi, and
every module the timed path imports lazily exists under the name bench.py uses."""
import ast
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _bench(*args):
    return subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), *args], capture_output=True, text=True,
                          timeout=300, cwd=ROOT)


def test_reference_arm_prints_one_json_line():
    r = _bench("--impl", "reference", "--gpus", "1", "--steps", "2", "--warmup", "3")
    assert r.returncode == 0, r.stderr
    d = json.loads(r.stdout.strip().splitlines()[-1])
    assert d["impl"] == "reference" and ("unavailable" in d or "value" in d)


def test_unknown_option_is_refused():
    r = _bench("--no-such-option")
    assert r.returncode != 0 and "unrecognized arguments" in r.stderr


def test_clock_sampler_without_nvidia_smi(monkeypatch):
    from parameter_server_b200.utils.clocks import ClockSampler
    monkeypatch.setenv("PATH", "/nonexistent")
    s = ClockSampler(0)
    s.start()
    out = s.stop()
    assert out["sm_mhz"] is None and out["reasons"] == ["nvidia-smi unavailable"]


def test_lazy_imports_of_bench_resolve():
    """bench.py imports its engine inside functions (the reference arm must not load it); a renamed
    symbol would only show on a GPU box.  Resolve every `from parameter_server_b200... import ...`."""
    import importlib
    tree = ast.parse(open(os.path.join(ROOT, "bench.py")).read())
    seen = 0
    for node in ast.walk(tree):
        if isinstance(node, ast.ImportFrom) and node.module and node.module.startswith("parameter_server_b200"):
            mod = importlib.import_module(node.module)
            for a in node.names:
                assert hasattr(mod, a.name), f"{node.module}.{a.name} (bench.py:{node.lineno})"
                seen += 1
    assert seen >= 4
    # every name used in bench.py is bound somewhere in it (catches a removed import that is still used)
    bound = set(dir(__builtins__)) if not isinstance(__builtins__, dict) else set(__builtins__)
    bound |= {"__file__", "__name__", "__doc__"}
    for node in ast.walk(tree):
        if isinstance(node, (ast.Import, ast.ImportFrom)):
            bound |= {(a.asname or a.name).split(".")[0] for a in node.names}
        elif isinstance(node, (ast.FunctionDef, ast.ClassDef)):
            bound.add(node.name)
            if isinstance(node, ast.FunctionDef):
                bound |= {a.arg for a in node.args.args + node.args.kwonlyargs}
                if node.args.vararg:
                    bound.add(node.args.vararg.arg)
                if node.args.kwarg:
                    bound.add(node.args.kwarg.arg)
        elif isinstance(node, ast.Name) and isinstance(node.ctx, (ast.Store, ast.Del)):
            bound.add(node.id)
        elif isinstance(node, ast.ExceptHandler) and node.name:
            bound.add(node.name)
        elif isinstance(node, ast.Lambda):
            bound |= {a.arg for a in node.args.args}
        elif isinstance(node, ast.comprehension):
            bound |= {n.id for n in ast.walk(node.target) if isinstance(n, ast.Name)}
    used = {n.id for n in ast.walk(tree) if isinstance(n, ast.Name) and isinstance(n.ctx, ast.Load)}
    assert not (used - bound), sorted(used - bound)
