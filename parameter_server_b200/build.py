"""In-tree build of the native code: the sm_100a CUDA library and the C++ host runtime.

    python -m parameter_server_b200.build [--force] [--verbose]

Artifacts (git-ignored, but they travel to the GPU box with the snapshot):
    parameter_server_b200/lib/libpsb200_cuda.so   nvcc, -gencode arch=compute_100a,code=sm_100a
    parameter_server_b200/lib/libpsb200_host.so   g++ -std=c++17, no CUDA dependency
    build/ps, build/hello                         launchable node binaries (host runtime)
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
CSRC = ROOT / "csrc"
LIBDIR = ROOT / "parameter_server_b200" / "lib"
OBJDIR = ROOT / "build" / "obj"
BINDIR = ROOT / "build"

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC,-fvisibility=hidden", "--expt-relaxed-constexpr",
    "-Xptxas", "-v", "-I", str(ROOT / "include"),
]
CXX_FLAGS = ["-std=c++17", "-O2", "-g1", "-fPIC", "-Wall", "-Wno-sign-compare", "-pthread",
             "-I", str(ROOT / "include"), "-I", str(CSRC / "host")]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found (set NVCC=...)")


def _newer(src: Path, dst: Path, extra: list[Path]) -> bool:
    if not dst.exists():
        return True
    t = dst.stat().st_mtime
    return any(p.stat().st_mtime > t for p in [src, *extra] if p.exists())


def _run(cmd: list[str], log: Path | None, verbose: bool) -> None:
    if verbose:
        print(" ".join(cmd), flush=True)
    res = subprocess.run(cmd, capture_output=True, text=True)
    if log is not None:
        log.write_text(res.stdout + res.stderr)
    if res.returncode != 0:
        sys.stderr.write(res.stdout + res.stderr)
        raise RuntimeError(f"build step failed: {' '.join(cmd[:3])} ... ({cmd[-1]})")


def build_cuda(force: bool = False, verbose: bool = False) -> Path:
    srcs = sorted((CSRC / "cuda").glob("*.cu"))
    hdrs = sorted((CSRC / "cuda").glob("*.cuh")) + sorted((ROOT / "include").glob("*.h"))
    OBJDIR.mkdir(parents=True, exist_ok=True)
    LIBDIR.mkdir(parents=True, exist_ok=True)
    nvcc = _nvcc()
    jobs = []
    objs = []
    for s in srcs:
        o = OBJDIR / (s.stem + ".cu.o")
        objs.append(o)
        if force or _newer(s, o, hdrs):
            jobs.append(([nvcc, *NVCC_FLAGS, "-c", str(s), "-o", str(o)],
                         OBJDIR / (s.stem + ".ptxas.log")))
    with ThreadPoolExecutor(max_workers=min(8, max(1, len(jobs)))) as ex:
        list(ex.map(lambda j: _run(j[0], j[1], verbose), jobs))
    out = LIBDIR / "libpsb200_cuda.so"
    if force or jobs or not out.exists():
        _run([nvcc, "-shared", "-o", str(out), *map(str, objs), "-lcudart",
              "-Xlinker", "--exclude-libs,ALL", "-Xlinker", "-Bsymbolic"], None, verbose)
    return out


def build_host(force: bool = False, verbose: bool = False) -> Path | None:
    srcs = sorted((CSRC / "host").rglob("*.cc"))
    if not srcs:
        return None
    hdrs = sorted((CSRC / "host").rglob("*.h")) + sorted((ROOT / "include").glob("*.h"))
    OBJDIR.mkdir(parents=True, exist_ok=True)
    LIBDIR.mkdir(parents=True, exist_ok=True)
    cxx = os.environ.get("CXX", "g++")
    jobs, lib_objs, mains = [], [], []
    for s in srcs:
        rel = "_".join(s.relative_to(CSRC / "host").with_suffix("").parts)
        o = OBJDIR / (rel + ".o")
        if s.parent.name == "bin":
            mains.append((s.stem, o))
        else:
            lib_objs.append(o)
        if force or _newer(s, o, hdrs):
            jobs.append(([cxx, *CXX_FLAGS, "-c", str(s), "-o", str(o)], None))
    with ThreadPoolExecutor(max_workers=min(8, max(1, len(jobs)))) as ex:
        list(ex.map(lambda j: _run(j[0], j[1], verbose), jobs))
    out = LIBDIR / "libpsb200_host.so"
    if force or jobs or not out.exists():
        _run([cxx, "-shared", "-o", str(out), *map(str, lib_objs), "-lz", "-pthread", "-ldl",
              "-Wl,--exclude-libs,ALL", "-Wl,-Bsymbolic"],
             None, verbose)
    for name, o in mains:
        exe = BINDIR / name
        if force or jobs or not exe.exists():
            _run([cxx, "-o", str(exe), str(o), *map(str, lib_objs), "-lz", "-pthread", "-ldl"],
                 None, verbose)
    return out


def build_micro(force: bool = False, verbose: bool = False) -> list:
    """Stand-alone micro-benchmarks (benchmarks/micro/*.cu -> build/micro/<name>), sm_100a."""
    out = []
    (BINDIR / "micro").mkdir(parents=True, exist_ok=True)
    for s in sorted((ROOT / "benchmarks" / "micro").glob("*.cu")):
        exe = BINDIR / "micro" / s.stem
        if force or _newer(s, exe, []):
            _run([_nvcc(), "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
                  "-o", str(exe), str(s)], None, verbose)
        out.append(exe)
    return out


def build_all(force: bool = False, verbose: bool = False) -> dict:
    return {"cuda": build_cuda(force, verbose), "host": build_host(force, verbose),
            "micro": build_micro(force, verbose)}


if __name__ == "__main__":
    res = build_all(force="--force" in sys.argv, verbose="--verbose" in sys.argv)
    for k, v in res.items():
        print(f"{k}: {v}")
