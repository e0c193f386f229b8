"""In-tree build of libsdb200.so (sm_100a only) with plain nvcc commands.

The library is a handful of translation units (``csrc/unit_*.cu``, each a short include list) that are compiled in
parallel to ``csrc/_obj/*.o`` and linked into one shared object; a unit is recompiled only when one of the files it
includes changed."""
from __future__ import annotations

import glob
import os
import re
import shutil
import subprocess
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(CSRC, "_obj")
OUT = os.path.join(HERE, "libsdb200.so")
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-Xcompiler", "-fPIC"]


def _nvcc() -> str:
    for cand in (shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found")


def units():
    return sorted(glob.glob(os.path.join(CSRC, "unit_*.cu")))


def _deps(path: str, seen=None):
    """The file and everything it #include "..."s, recursively."""
    seen = set() if seen is None else seen
    path = os.path.normpath(path)
    if path in seen or not os.path.exists(path):
        return seen
    seen.add(path)
    for inc in re.findall(r'^\s*#\s*include\s+"([^"]+)"', open(path).read(), flags=re.M):
        _deps(os.path.join(os.path.dirname(path), inc), seen)
    return seen


def _obj(unit: str) -> str:
    return os.path.join(OBJ, os.path.basename(unit)[:-3] + ".o")


def _stale(unit: str) -> bool:
    o = _obj(unit)
    return not os.path.exists(o) or any(os.path.getmtime(d) > os.path.getmtime(o) for d in _deps(unit))


def needs_build() -> bool:
    if not os.path.exists(OUT):
        return True
    t = os.path.getmtime(OUT)
    return any(os.path.getmtime(d) > t for u in units() for d in _deps(u))


def build(force: bool = False, verbose: bool = False) -> str:
    if not force and not needs_build():
        return OUT
    os.makedirs(OBJ, exist_ok=True)
    nvcc = _nvcc()
    todo = [u for u in units() if force or _stale(u)]

    def compile_unit(u):
        cmd = [nvcc, *NVCC_FLAGS, "-c", u, "-o", _obj(u) + ".tmp"]
        if verbose:
            cmd.insert(1, "-Xptxas=-v")
        res = subprocess.run(cmd, capture_output=True, text=True, cwd=CSRC)
        if res.returncode == 0:
            os.replace(_obj(u) + ".tmp", _obj(u))
        return u, res

    with ThreadPoolExecutor(max_workers=min(len(todo), os.cpu_count() or 1) or 1) as ex:
        results = list(ex.map(compile_unit, todo))
    failed = [(u, r) for u, r in results if r.returncode != 0]
    if failed:
        raise RuntimeError("nvcc failed:\n" + "\n".join(f"[{os.path.basename(u)}]\n{r.stdout}{r.stderr}" for u, r in failed))
    if verbose:
        for u, r in results:
            print(f"==== {os.path.basename(u)}\n{r.stderr}")
    res = subprocess.run([nvcc, "-gencode", "arch=compute_100a,code=sm_100a", "-shared", "-o", OUT + ".tmp",
                          *[_obj(u) for u in units()]], capture_output=True, text=True, cwd=CSRC)
    if res.returncode != 0:
        raise RuntimeError("link failed:\n" + res.stdout + res.stderr)
    os.replace(OUT + ".tmp", OUT)
    return OUT


if __name__ == "__main__":
    import sys
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
