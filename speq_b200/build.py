"""Build libspeq_cuda.so (and the `speq` host binary) in-tree with nvcc for sm_100a.

    python -m speq_b200.build [--force]

The library is the product; it links the CUDA runtime statically and nothing from oracle/.
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent
CSRC = ROOT / "csrc"
LIB_DIR = ROOT / "lib"
LIB = LIB_DIR / "libspeq_cuda.so"
OBJ_DIR = ROOT / "build"
CLI_SRC = ROOT / "host"
CLI_BIN = ROOT / "bin" / "speq"

LIB_SOURCES = ["scanh.cu", "scan32.cu", "fastq.cu", "capi.cu", "capi_scan.cu", "capi_refine.cu", "kernels.cu", "sa_device.cu", "probe.cu", "noalloc.cu", "index_build.cpp", "pack.cpp", "refidx.cpp"]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "-Xcompiler", "-fPIC,-Wall,-pthread"]


def nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and Path(cand).exists():
            return cand
    raise RuntimeError("nvcc not found: libspeq_cuda.so cannot be built (there is no CPU fallback)")


def _stale(target: Path, deps: list[Path]) -> bool:
    if not target.exists():
        return True
    t = target.stat().st_mtime
    return any(d.stat().st_mtime > t for d in deps if d.exists())


def build_lib(force: bool = False, verbose: bool = False) -> Path:
    deps = [p for p in CSRC.iterdir() if p.suffix in (".cu", ".cpp", ".h", ".hpp", ".cuh")] + [ROOT.parent / "include" / "speq_cuda.h"]
    if not force and not _stale(LIB, deps):
        return LIB
    LIB_DIR.mkdir(exist_ok=True)
    OBJ_DIR.mkdir(exist_ok=True)
    headers = [d for d in deps if d.suffix in (".h", ".hpp", ".cuh")]

    def compile_one(src: str) -> Path:
        obj = OBJ_DIR / (src + ".o")
        if not force and not _stale(obj, [CSRC / src] + headers):
            return obj
        cmd = [nvcc(), *NVCC_FLAGS, *os.environ.get("SPEQB_NVCC_EXTRA", "").split(), "-c", str(CSRC / src), "-o", str(obj)]
        if verbose:
            cmd.insert(1, "-Xptxas=-v")
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("nvcc failed:\n" + " ".join(cmd) + "\n" + r.stdout + r.stderr)
        if verbose:
            print(r.stderr)
        return obj

    from concurrent.futures import ThreadPoolExecutor
    with ThreadPoolExecutor(max_workers=min(len(LIB_SOURCES), os.cpu_count() or 2)) as ex:
        objs = list(ex.map(compile_one, LIB_SOURCES))
    cmd = [nvcc(), "-gencode", "arch=compute_100a,code=sm_100a", "-shared", "-Xcompiler", "-fPIC", "-o", str(LIB)] + [str(o) for o in objs]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("link failed:\n" + " ".join(cmd) + "\n" + r.stdout + r.stderr)
    return LIB


def build_cli(force: bool = False) -> Path | None:
    if not CLI_SRC.exists():
        return None
    srcs = sorted(CLI_SRC.glob("*.cpp"))
    deps = srcs + list(CLI_SRC.glob("*.h")) + [LIB]
    if not force and not _stale(CLI_BIN, deps):
        return CLI_BIN
    CLI_BIN.parent.mkdir(exist_ok=True)
    cmd = ["g++", "-std=c++17", "-O2", "-Wall", "-pthread", "-I", str(ROOT.parent / "include"), "-o", str(CLI_BIN)] + \
          [str(s) for s in srcs] + ["-L", str(LIB_DIR), "-lspeq_cuda", "-lz", "-Wl,-rpath,$ORIGIN/../lib"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("g++ failed:\n" + " ".join(cmd) + "\n" + r.stdout + r.stderr)
    return CLI_BIN


def build_all(force: bool = False, verbose: bool = False) -> None:
    build_lib(force, verbose)
    build_cli(force)


if __name__ == "__main__":
    build_all(force="--force" in sys.argv, verbose="-v" in sys.argv)
    print(LIB)
