"""In-tree build of libwxddp.so (nvcc, sm_100a only).  No JIT cache: the .so sits next to this file so it
travels with the repository snapshot to the GPU box."""
from __future__ import annotations

import hashlib
import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

PKG_DIR = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(PKG_DIR, "csrc")
INCLUDE = os.path.join(os.path.dirname(PKG_DIR), "include")
BUILD_DIR = os.path.join(CSRC, "build")
LIB_PATH = os.path.join(PKG_DIR, "libwxddp.so")

NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-Xcompiler", "-fPIC", f"-I{INCLUDE}", f"-I{CSRC}"]
# developer builds only (e.g. WX_NVCC_EXTRA=-DWX_TC_TIMERS=1 for the phase timers of the tensor-core step kernel)
NVCC_FLAGS += [fl for fl in os.environ.get("WX_NVCC_EXTRA", "").split() if fl]

# (source, object name, extra flags)
UNITS = [("ctx.cu", "ctx.o", []), ("preprocess.cu", "preprocess.o", []), ("mlp_host.cu", "mlp_host.o", []),
         ("tc_selftest.cu", "tc_selftest.o", []), ("tc_step.cu", "tc_step.o", []), ("wide.cu", "wide.o", [])]
UNITS += [("mlp_inst.cu", f"mlp_inst_g{g}.o", [f"-DWX_F_GROUP={g}"]) for g in range(4)]
UNITS += [("tc_step_inst.cu", f"tc_step_ft{ft}.o", [f"-DWX_TC_FT={ft}"]) for ft in (0, 5, 20)]
EXTRA_UNITS: list[tuple[str, str, list[str]]] = []


def _nvcc() -> str:
    for cand in (shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found: libwxddp.so cannot be built (there is no CPU fallback)")


def _source_digest() -> str:
    h = hashlib.sha256()
    files = sorted(os.listdir(CSRC)) + [os.path.join(INCLUDE, "wxddp.h")]
    for name in files:
        path = name if os.path.isabs(name) else os.path.join(CSRC, name)
        if os.path.isfile(path) and path.endswith((".cu", ".cuh", ".h")):
            h.update(os.path.basename(name).encode())
            with open(path, "rb") as fh:
                h.update(fh.read())
    # flags without the absolute -I paths, so the stamp stays valid when the tree is copied elsewhere
    h.update(" ".join(fl for fl in NVCC_FLAGS if not fl.startswith("-I")).encode())
    return h.hexdigest()


def is_current() -> bool:
    stamp = os.path.join(BUILD_DIR, "stamp")
    return os.path.exists(LIB_PATH) and os.path.exists(stamp) and open(stamp).read().strip() == _source_digest()


def build_library(force: bool = False, verbose: bool = True) -> str:
    """Compile csrc/*.cu for sm_100a and link libwxddp.so.  Returns its path."""
    if not force and is_current():
        return LIB_PATH
    nvcc = _nvcc()
    os.makedirs(BUILD_DIR, exist_ok=True)
    units = UNITS + EXTRA_UNITS

    def compile_one(unit):
        src, obj, extra = unit
        cmd = [nvcc] + NVCC_FLAGS + extra + ["-c", os.path.join(CSRC, src), "-o", os.path.join(BUILD_DIR, obj)]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"nvcc failed for {src} {extra}:\n{r.stdout}\n{r.stderr}")
        return obj

    with ThreadPoolExecutor(max_workers=min(8, os.cpu_count() or 1)) as ex:
        objs = list(ex.map(compile_one, units))
    cmd = [nvcc, "-gencode", "arch=compute_100a,code=sm_100a", "-shared", "-o", LIB_PATH] + \
          [os.path.join(BUILD_DIR, o) for o in objs]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    with open(os.path.join(BUILD_DIR, "stamp"), "w") as fh:
        fh.write(_source_digest())
    if verbose:
        print(f"[wxddp] built {LIB_PATH}", file=sys.stderr)
    return LIB_PATH


if __name__ == "__main__":
    build_library(force="--force" in sys.argv)
