"""Build the host runtime + build/selftest with a sanitizer and run it.

    python tools/sanitize.py thread      # ThreadSanitizer
    python tools/sanitize.py address     # AddressSanitizer + UBSan

Objects go to $TMPDIR/psb200-san-<kind>/ (the normal build is untouched). Exit code = the binary's.
"""
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
HOST = ROOT / "csrc" / "host"


def main() -> int:
    kind = sys.argv[1] if len(sys.argv) > 1 else "thread"
    flags = {"thread": ["-fsanitize=thread"], "address": ["-fsanitize=address,undefined", "-fno-omit-frame-pointer"]}[kind]
    # outside the tree: these objects are large and must not travel with repository snapshots
    out = Path(os.environ.get("TMPDIR", "/tmp")) / f"psb200-san-{kind}"
    out.mkdir(parents=True, exist_ok=True)
    # the sanitizer runtimes ship with the distribution compiler, not with every custom toolchain
    cxx = os.environ.get("SAN_CXX") or ("/usr/bin/g++" if os.path.exists("/usr/bin/g++") else os.environ.get("CXX", "g++"))
    srcs = [s for s in sorted(HOST.rglob("*.cc")) if s.parent.name != "bin"] + [HOST / "bin" / "selftest.cc"]
    base = [cxx, "-std=c++17", "-O1", "-g", "-fPIC", "-pthread", "-I", str(HOST), "-I", str(ROOT / "include"), *flags]

    def compile_one(s: Path) -> Path:
        o = out / ("_".join(s.relative_to(HOST).with_suffix("").parts) + ".o")
        if not o.exists() or o.stat().st_mtime < max(p.stat().st_mtime for p in [s, *HOST.rglob("*.h")]):
            subprocess.run([*base, "-c", str(s), "-o", str(o)], check=True)
        return o
    with ThreadPoolExecutor(8) as ex:
        objs = list(ex.map(compile_one, srcs))
    exe = out / "selftest"
    subprocess.run([*base, "-o", str(exe), *map(str, objs), "-lz", "-ldl"], check=True)
    env = dict(os.environ, TSAN_OPTIONS="halt_on_error=0 second_deadlock_stack=1 history_size=4",
               ASAN_OPTIONS="detect_leaks=0", UBSAN_OPTIONS="print_stacktrace=1")
    return subprocess.run([str(exe), *sys.argv[2:]], env=env).returncode


if __name__ == "__main__":
    sys.exit(main())
