"""In-tree build of libb200attn.so (sm_100a only) with plain nvcc.

The shared library is the product: a C-ABI (include/b200attn.h) over hand-written CUDA kernels.
There is no JIT and no torch C++ extension: the .so is built next to this file so that it ships
with a snapshot of the repo, and the ctypes loader (byox_vllm_b200/_lib.py) fails loudly when it
is missing.
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys
from pathlib import Path

PKG_DIR = Path(__file__).resolve().parent
CSRC = PKG_DIR / "csrc"
BUILD_DIR = PKG_DIR / "_build"
LIB_PATH = PKG_DIR / "libb200attn.so"

SOURCES = ["store_kv.cu", "decode_attn.cu", "prefill_attn.cu", "prefill_tc.cu", "cascade.cu",
           "qk_norm_rope.cu", "prepare_meta.cu", "api.cu"]
# every header under csrc/ (a new one is picked up without editing this file) + the C-ABI header
HEADERS = sorted(p.name for p in CSRC.glob("*.h")) + sorted(p.name for p in CSRC.glob("*.cuh")) + \
    ["../../include/b200attn.h"]

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC",
    "-Xptxas", "-v",
]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found (set NVCC=/path/to/nvcc)")


def _stale(target: Path, deps: list[Path]) -> bool:
    if not target.exists():
        return True
    t = target.stat().st_mtime
    return any(d.stat().st_mtime > t for d in deps if d.exists())


def build(force: bool = False, verbose: bool = False) -> Path:
    """Compile every .cu under csrc/ for sm_100a and link libb200attn.so. Returns its path."""
    nvcc = _nvcc()
    BUILD_DIR.mkdir(exist_ok=True)
    headers = [(CSRC / h).resolve() for h in HEADERS] + [Path(__file__)]
    objs = []
    log_lines = []
    for src in SOURCES:
        s = CSRC / src
        if not s.exists():
            continue
        o = BUILD_DIR / (s.stem + ".o")
        objs.append(o)
        if force or _stale(o, [s] + headers):
            cmd = [nvcc, *NVCC_FLAGS, "-c", str(s), "-o", str(o)]
            r = subprocess.run(cmd, capture_output=True, text=True)
            log_lines.append("$ " + " ".join(cmd))
            log_lines.append(r.stdout + r.stderr)
            if r.returncode != 0:
                sys.stderr.write(r.stdout + r.stderr)
                raise RuntimeError(f"nvcc failed on {src}")
            if verbose:
                sys.stderr.write(r.stderr)
    if force or _stale(LIB_PATH, objs):
        cmd = [nvcc, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", str(LIB_PATH),
               *map(str, objs)]
        r = subprocess.run(cmd, capture_output=True, text=True)
        log_lines.append("$ " + " ".join(cmd))
        log_lines.append(r.stdout + r.stderr)
        if r.returncode != 0:
            sys.stderr.write(r.stdout + r.stderr)
            raise RuntimeError("link of libb200attn.so failed")
    if log_lines:
        (BUILD_DIR / "build.log").write_text("\n".join(log_lines))
    return LIB_PATH


if __name__ == "__main__":
    p = build(force="--force" in sys.argv, verbose=True)
    print(p)
