"""Build the native library in-tree with nvcc (sm_100a only) and g++.

``python -m determined_b200._native.build`` compiles every ``csrc/*.cu|*.cpp`` into
``determined_b200/_native/libdetb200.so`` and the C++ launcher into
``determined_b200/_native/detb200-launch``.  Objects are cached by source mtime so rebuilds are
incremental.  nvcc cross-compiles without a GPU.
"""
import concurrent.futures
import os
import pathlib
import shutil
import subprocess
import sys
from typing import List

HERE = pathlib.Path(__file__).resolve().parent
ROOT = HERE.parent.parent
CSRC = ROOT / "csrc"
OBJ = HERE / "obj"
LIB = HERE / "libdetb200.so"
LAUNCHER = HERE / "detb200-launch"
HOSTLIB = HERE / "libdetb200_host.so"

NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "--use_fast_math", "-Xcompiler", "-fPIC,-fvisibility=hidden", "-I", str(ROOT / "include"),
              "-I", str(CSRC)]
CXX_FLAGS = ["-O2", "-std=c++17", "-fPIC", "-fvisibility=hidden", "-Wall", "-I", str(ROOT / "include"), "-I", str(CSRC)]


def _nvcc() -> str:
    for c in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if c and os.path.exists(c):
            return c
    raise RuntimeError("nvcc not found")


def _stale(src: pathlib.Path, obj: pathlib.Path, deps: List[pathlib.Path]) -> bool:
    if not obj.exists():
        return True
    t = obj.stat().st_mtime
    return any(p.stat().st_mtime > t for p in [src, *deps])


def _compile(src: pathlib.Path, verbose: bool) -> pathlib.Path:
    obj = OBJ / (src.name + ".o")
    headers = list(CSRC.glob("*.cuh")) + list(CSRC.glob("*.h")) + list((ROOT / "include").glob("*.h"))
    if not _stale(src, obj, headers):
        return obj
    if src.suffix == ".cu":
        cmd = [_nvcc(), *NVCC_FLAGS, "-c", str(src), "-o", str(obj)]
    else:
        cmd = ["g++", *CXX_FLAGS, "-I", "/usr/local/cuda/include", "-c", str(src), "-o", str(obj)]
    if verbose:
        print(" ".join(cmd), flush=True)
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"compile failed: {src.name}\n{r.stdout}")
    if verbose and r.stdout.strip():
        print(r.stdout)
    return obj


def build(verbose: bool = False) -> pathlib.Path:
    OBJ.mkdir(parents=True, exist_ok=True)
    lib_srcs = sorted(p for p in CSRC.iterdir() if p.suffix in (".cu", ".cpp"))
    with concurrent.futures.ThreadPoolExecutor(max_workers=min(8, os.cpu_count() or 2)) as ex:
        objs = list(ex.map(lambda s: _compile(s, verbose), lib_srcs))
    if not LIB.exists() or any(o.stat().st_mtime > LIB.stat().st_mtime for o in objs):
        cmd = [_nvcc(), "-shared", "-o", str(LIB), *map(str, objs), "-gencode", "arch=compute_100a,code=sm_100a",
               "-lcudart", "-Xlinker", "-rpath,/usr/local/cuda/lib64"]
        if verbose:
            print(" ".join(cmd), flush=True)
        r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"link failed\n{r.stdout}")
    # C++ launcher (process spawner + pidfd watchdog); sources under csrc/launcher/
    ldir = CSRC / "launcher"
    if ldir.is_dir():
        srcs = sorted(ldir.glob("*.cpp"))
        if srcs and (not LAUNCHER.exists() or any(s.stat().st_mtime > LAUNCHER.stat().st_mtime for s in srcs)):
            cmd = ["g++", "-O2", "-std=c++17", "-Wall", "-pthread", "-o", str(LAUNCHER), *map(str, srcs)]
            if verbose:
                print(" ".join(cmd), flush=True)
            r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
            if r.returncode != 0:
                raise RuntimeError(f"launcher build failed\n{r.stdout}")
    # host-only runtime pieces (no CUDA): shared-memory object collectives; sources under csrc/host/
    hdir = CSRC / "host"
    if hdir.is_dir():
        srcs = sorted(p for p in hdir.glob("*.cpp") if not p.name.endswith("_test.cpp"))
        if srcs and (not HOSTLIB.exists() or any(s.stat().st_mtime > HOSTLIB.stat().st_mtime for s in srcs)):
            cmd = ["g++", "-O2", "-std=c++17", "-Wall", "-fPIC", "-shared", "-pthread", "-fvisibility=hidden", "-o", str(HOSTLIB), *map(str, srcs), "-lrt"]
            if verbose:
                print(" ".join(cmd), flush=True)
            r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
            if r.returncode != 0:
                raise RuntimeError(f"host library build failed\n{r.stdout}")
    return LIB


if __name__ == "__main__":
    print(build(verbose="-q" not in sys.argv))
