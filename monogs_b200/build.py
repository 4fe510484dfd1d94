"""Build the sm_100a CUDA library (monogs_b200/libgsr_b200.so) with nvcc.

In-tree, explicit `nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo`; no
torch headers are involved (the library exposes the C ABI of include/gsr_b200.h
and is loaded with ctypes), so a full rebuild takes well under a minute.
"""
from __future__ import annotations

import concurrent.futures as cf
import hashlib
import os
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent
CSRC = ROOT / "csrc"
OBJ = ROOT / "build"
LIB = ROOT / "libgsr_b200.so"
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
SOURCES = ["preprocess.cu", "binning.cu", "radix_sort.cu", "render_fwd.cu", "render_bwd.cu",
           "preprocess_bwd.cu", "knn.cu", "loss.cu", "tile_scatter.cu", "tile_refine.cu", "stats.cu", "rows_exchange.cu", "insert.cu", "surgery.cu", "api.cu"]
FLAGS = ["-std=c++17", "-O3", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo",
         "-Xcompiler", "-fPIC", "--expt-relaxed-constexpr"]


def _stamp() -> str:
    h = hashlib.sha256()
    for p in sorted(list(CSRC.glob("*.cu")) + list(CSRC.glob("*.cuh")) + [ROOT.parent / "include" / "gsr_b200.h"]):
        h.update(p.name.encode())
        h.update(p.read_bytes())
    h.update(" ".join(FLAGS).encode())
    return h.hexdigest()


def build_variant(name: str, defines) -> Path:
    """Experimental variant monogs_b200/libgsr_b200_<name>.so with extra -D flags (selected with GSR_B200_LIB)."""
    obj = ROOT / f"build_{name}"
    obj.mkdir(exist_ok=True)
    lib = ROOT / f"libgsr_b200_{name}.so"
    objs = []
    for src in SOURCES:
        o = obj / (src + ".o")
        r = subprocess.run([NVCC, *FLAGS, *[f"-D{d}" for d in defines], "-c", str(CSRC / src), "-o", str(o)],
                           capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(r.stderr)
        objs.append(str(o))
    r = subprocess.run([NVCC, "-shared", "-o", str(lib), *objs, "-gencode", "arch=compute_100a,code=sm_100a"],
                       capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(r.stderr)
    return lib


def build(force: bool = False, verbose: bool = False) -> Path:
    OBJ.mkdir(exist_ok=True)
    stamp_file = OBJ / "stamp"
    stamp = _stamp()
    if not force and LIB.exists() and stamp_file.exists() and stamp_file.read_text() == stamp:
        return LIB
    if not Path(NVCC).exists():
        if LIB.exists():  # GPU box without a toolkit change: keep the shipped library
            return LIB
        raise RuntimeError(f"nvcc not found at {NVCC} and {LIB} is not built")

    def compile_one(src: str) -> str:
        obj = OBJ / (src + ".o")
        cmd = [NVCC, *FLAGS, "-c", str(CSRC / src), "-o", str(obj)]
        if verbose:
            cmd.insert(1, "-Xptxas=-v")
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"nvcc failed for {src}:\n{r.stdout}\n{r.stderr}")
        if verbose:
            sys.stderr.write(r.stderr)
        return str(obj)

    with cf.ThreadPoolExecutor(max_workers=min(8, len(SOURCES))) as ex:
        objs = list(ex.map(compile_one, SOURCES))
    r = subprocess.run([NVCC, "-shared", "-o", str(LIB), *objs, "-gencode", "arch=compute_100a,code=sm_100a"],
                       capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    stamp_file.write_text(stamp)
    return LIB


if __name__ == "__main__":
    if "--variant" in sys.argv:
        i = sys.argv.index("--variant")
        print(build_variant(sys.argv[i + 1], sys.argv[i + 2:]))
    else:
        print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
