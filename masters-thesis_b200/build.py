"""Builds libb200lstm.so in-tree with nvcc for sm_100a (no JIT cache, so the .so travels with the repo).

    python masters-thesis_b200/build.py [--force]
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

PKG = Path(__file__).resolve().parent
CSRC = PKG / "csrc"
OBJ = PKG / "build"
LIB = PKG / "libb200lstm.so"
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
              "-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=hidden", "--expt-relaxed-constexpr"]


def _nvcc() -> str:
    exe = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(exe):
        raise RuntimeError("nvcc not found: cannot build libb200lstm.so")
    return exe


def _stale(target: Path, deps) -> bool:
    if not target.exists():
        return True
    t = target.stat().st_mtime
    return any(Path(d).stat().st_mtime > t for d in deps)


def source_hash(extra_flags=()) -> str:
    """sha256 over csrc/*, include/b200_lstm.h and the compile flags: compiled into the library (b200_build_hash) so that
    a stale prebuilt .so is detected by _lib.load() instead of passing silently."""
    import hashlib
    h = hashlib.sha256()
    for f in sorted(CSRC.glob("*.cu")) + sorted(CSRC.glob("*.cuh")) + [PKG.parent / "include" / "b200_lstm.h"]:
        h.update(f.name.encode()); h.update(f.read_bytes())
    h.update(" ".join(list(NVCC_FLAGS) + list(extra_flags)).encode())
    return h.hexdigest()[:16]


def _extra_flags():
    flags = []
    if os.environ.get("B200_TRACE") == "1":
        flags.append("-DB200_TRACE")
    flags += os.environ.get("B200_DEFS", "").split()
    if os.environ.get("B200_ABLATE") == "1":
        flags.append("-DB200_ABLATE")
    return flags


def build(force: bool = False, verbose: bool = False) -> Path:
    flags = list(NVCC_FLAGS)
    digest = source_hash(_extra_flags())
    stamp = OBJ / "source_hash.txt"
    old = stamp.read_text().split() if stamp.exists() else []
    flag_sig = "|".join(_extra_flags()) or "-"
    if not force and (len(old) != 2 or old[1] != flag_sig):
        force = True                     # B200_TRACE / B200_DEFS / B200_ABLATE switches changed: rebuild every object
    hash_changed = len(old) != 2 or old[0] != digest
    flags += _extra_flags()
    srcs = sorted(CSRC.glob("*.cu"))
    hdrs = sorted(CSRC.glob("*.cuh")) + [PKG.parent / "include" / "b200_lstm.h"]
    OBJ.mkdir(exist_ok=True)
    nvcc = _nvcc()
    jobs = []
    for s in srcs:
        o = OBJ / (s.stem + ".o")
        if force or _stale(o, [s] + hdrs) or (s.stem == "api" and hash_changed):
            cmd = [nvcc, *flags, "-c", str(s), "-o", str(o)]
            if s.stem == "api":           # the hash lives in api.cu only, so that one edited file recompiles two, not all
                cmd.insert(1, f'-DB200_SRC_HASH="{digest}"')
            if verbose:
                cmd.insert(1, "-Xptxas=-v")
            jobs.append(cmd)

    def run(cmd):
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("nvcc failed:\n" + " ".join(cmd) + "\n" + r.stdout + r.stderr)
        return r.stderr

    with ThreadPoolExecutor(max_workers=min(8, max(1, len(jobs)))) as ex:
        for out in ex.map(run, jobs):
            if verbose and out:
                print(out)
    objs = [OBJ / (s.stem + ".o") for s in srcs]
    if force or jobs or _stale(LIB, objs):
        cmd = [nvcc, "-shared", "-o", str(LIB), *map(str, objs), "-gencode", "arch=compute_100a,code=sm_100a",
               "-Xcompiler", "-fPIC", "-lcudart", "-lcuda", "-ldl"]
        run(cmd)
    stamp.write_text(f"{digest} {flag_sig}")
    return LIB


if __name__ == "__main__":
    p = build(force="--force" in sys.argv, verbose="-v" in sys.argv)
    print(p)
