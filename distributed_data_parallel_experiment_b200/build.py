"""In-tree build of the native libraries (nvcc, sm_100a only).

``python -m distributed_data_parallel_experiment_b200.build`` compiles every ``csrc/kernels/*.cu`` and
``csrc/runtime/*.cu|*.cpp`` into ``distributed_data_parallel_experiment_b200/lib/libddpx_kernels.so``.
The library has a plain C ABI (``include/ddpx_kernels.h``) and is loaded with ctypes, so the build needs
neither torch headers nor a GPU; nvcc cross-compiles here and the ``.so`` travels to the GPU box in-tree.
"""
from __future__ import annotations

import hashlib
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

PKG_DIR = Path(__file__).resolve().parent
REPO = PKG_DIR.parent
LIB_DIR = PKG_DIR / "lib"
OBJ_DIR = REPO / "build" / "obj"
LIB_PATH = LIB_DIR / "libddpx_kernels.so"

NVCC = os.environ.get("DDPX_NVCC", "/usr/local/cuda/bin/nvcc")
NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=default",
    "-I", str(REPO / "include"),
]


def sources() -> list[Path]:
    out = sorted((REPO / "csrc" / "kernels").glob("*.cu"))
    out += sorted((REPO / "csrc" / "runtime").glob("*.cu"))
    out += sorted((REPO / "csrc" / "runtime").glob("*.cpp"))
    return out


def _digest(src: Path) -> str:
    h = hashlib.sha1()
    h.update(src.read_bytes())
    for hdr in sorted(list((REPO / "include").glob("*.h")) + list(src.parent.glob("*.cuh")) + list(src.parent.glob("*.hpp"))):
        h.update(hdr.read_bytes())
    h.update(" ".join(NVCC_FLAGS).encode())
    return h.hexdigest()


def _compile_one(src: Path, verbose: bool) -> Path:
    OBJ_DIR.mkdir(parents=True, exist_ok=True)
    obj = OBJ_DIR / (src.stem + ".o")
    stamp = OBJ_DIR / (src.stem + ".sha1")
    dig = _digest(src)
    if obj.exists() and stamp.exists() and stamp.read_text() == dig:
        return obj
    cmd = [NVCC, *NVCC_FLAGS, "-c", str(src), "-o", str(obj)]
    if src.suffix == ".cpp":
        cmd = [NVCC, *NVCC_FLAGS, "-x", "cu", "-c", str(src), "-o", str(obj)]
    if verbose:
        print("[ddpx.build]", " ".join(cmd), flush=True)
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"nvcc failed for {src}:\n{r.stdout}\n{r.stderr}")
    stamp.write_text(dig)
    return obj


def build(verbose: bool = True, force: bool = False) -> Path:
    """Compile (incrementally) and link the kernel library; returns its path."""
    srcs = sources()
    if not srcs:
        raise RuntimeError("no CUDA sources found under csrc/")
    if force:
        for f in OBJ_DIR.glob("*.sha1"):
            f.unlink()
    with ThreadPoolExecutor(max_workers=min(8, len(srcs))) as ex:
        objs = list(ex.map(lambda s: _compile_one(s, verbose), srcs))
    LIB_DIR.mkdir(parents=True, exist_ok=True)
    newest = max(o.stat().st_mtime for o in objs)
    if (not LIB_PATH.exists()) or LIB_PATH.stat().st_mtime < newest or force:
        cmd = [NVCC, "-shared", "-o", str(LIB_PATH), *map(str, objs), "-lcudart", "-lpthread"]
        if verbose:
            print("[ddpx.build]", " ".join(cmd), flush=True)
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    return LIB_PATH


EXT_SRC = REPO / "csrc" / "torch_ext" / "autograd_hooks.cpp"
EXT_PATH = LIB_DIR / "ddpx_autograd.so"


def build_torch_ext(verbose: bool = True, force: bool = False) -> Path:
    """Compile the C++ autograd-hook extension (csrc/torch_ext) in-tree against the installed torch headers.
    No CUDA and no nvcc involved; ~25 s with g++."""
    import sysconfig
    import torch
    import logging
    logging.getLogger("torch.utils.cpp_extension").setLevel(logging.ERROR)
    from torch.utils import cpp_extension as ce
    OBJ_DIR.mkdir(parents=True, exist_ok=True)
    LIB_DIR.mkdir(parents=True, exist_ok=True)
    h = hashlib.sha1(EXT_SRC.read_bytes())
    h.update(torch.__version__.encode())
    stamp = OBJ_DIR / "ddpx_autograd.sha1"
    if EXT_PATH.exists() and stamp.exists() and stamp.read_text() == h.hexdigest() and not force:
        return EXT_PATH
    incs = [f"-I{p}" for p in ce.include_paths()] + [f"-I{sysconfig.get_paths()['include']}"]
    libdir = ce.library_paths()[0]
    abi = int(bool(torch._C._GLIBCXX_USE_CXX11_ABI))
    cmd = [os.environ.get("CXX", "g++"), "-O2", "-std=c++17", "-fPIC", "-shared", f"-D_GLIBCXX_USE_CXX11_ABI={abi}",
           "-DTORCH_EXTENSION_NAME=ddpx_autograd", "-DTORCH_API_INCLUDE_EXTENSION_H", *incs, str(EXT_SRC), "-o", str(EXT_PATH),
           f"-L{libdir}", "-ltorch_python", "-ltorch", "-ltorch_cpu", "-lc10", f"-Wl,-rpath,{libdir}"]
    if verbose:
        print("[ddpx.build]", " ".join(cmd), flush=True)
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"torch extension build failed:\n{r.stdout}\n{r.stderr}")
    stamp.write_text(h.hexdigest())
    return EXT_PATH


if __name__ == "__main__":
    p = build(verbose=True, force="--force" in sys.argv)
    print(p)
    print(build_torch_ext(verbose=True, force="--force" in sys.argv))
