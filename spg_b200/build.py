"""Build recipe for the spgx CUDA library (sm_100a only) and the test-only oracle libraries.

Everything is built in-tree so the shared objects travel with a `gpurun` snapshot:
  spg_b200/lib/libspgx.so        product: CUDA kernels + C ABI (include/spgx.h)
  oracle/libspgport.so           TEST INFRASTRUCTURE: independent CPU restatement
  oracle/_ref/libspgref.so       TEST INFRASTRUCTURE: the unmodified reference, only when /root/reference exists
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
CSRC = ROOT / "spg_b200" / "csrc"
LIBDIR = ROOT / "spg_b200" / "lib"
OBJDIR = ROOT / "build" / "obj"
LIB = LIBDIR / "libspgx.so"

CU_SOURCES = ["api_build.cu", "api_step.cu", "vbd.cu", "vbd_stream_cloth.cu", "vbd_stream_cloth_wedge.cu", "vbd_stream_tets.cu", "xpbd.cu", "state.cu", "comm.cu", "newton.cu", "api_newton.cu", "rigid.cu", "colour.cu"]
HEADERS = ["dual.cuh", "halo.cuh", "linalg.cuh", "energies.cuh", "streams.cuh", "stream_records.h", "vbd_stream.cuh", "energy_prepare.h", "host_build.h", "partition.h",
           "kernels.h", "group.h", "kinds.def", "xpbd_partition.h", "hinge_hessian.cuh", "membrane_hessian.cuh", "rigid.cuh"]

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-std=c++17", "-lineinfo",
    "-Xcompiler", "-fPIC",
    "--expt-relaxed-constexpr",
    "-Xptxas", "-v",
]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and Path(cand).exists():
            return cand
    raise RuntimeError("nvcc not found: the spgx library cannot be built (there is no CPU fallback)")


def _stale(target: Path, deps: list[Path]) -> bool:
    if not target.exists():
        return True
    t = target.stat().st_mtime
    return any(d.exists() and d.stat().st_mtime > t for d in deps)


EXTRA_DIR = ROOT / "tests" / "native"
LIB_EXTRA = LIBDIR / "libspgx_extra.so"


def build_library(force: bool = False, verbose: bool = False, extra_kind: bool = False) -> Path:
    """Compile every .cu for sm_100a and link spg_b200/lib/libspgx.so.

    extra_kind: the same sources with the throw-away 15th energy kind of tests/native/extra_kind.{def,cuh} registered
    (-DSPGX_EXTRA_KINDS_DEF / -DSPGX_EXTRA_KINDS_IMPL) -> spg_b200/lib/libspgx_extra.so, loaded only by tests/test_extra_kind.py:
    the proof that an energy plugs in without touching the library sources."""
    nvcc = _nvcc()
    LIBDIR.mkdir(parents=True, exist_ok=True)
    OBJDIR = ROOT / "build" / ("obj_extra" if extra_kind else "obj")
    LIB = LIB_EXTRA if extra_kind else globals()["LIB"]
    NVCC_FLAGS = list(globals()["NVCC_FLAGS"])
    OBJDIR.mkdir(parents=True, exist_ok=True)
    hdrs = [CSRC / h for h in HEADERS] + [ROOT / "include" / "spgx.h", Path(__file__)]
    if extra_kind:
        NVCC_FLAGS += [f'-DSPGX_EXTRA_KINDS_DEF="{EXTRA_DIR / "extra_kind.def"}"', f'-DSPGX_EXTRA_KINDS_IMPL="{EXTRA_DIR / "extra_kind.cuh"}"']
        hdrs += [EXTRA_DIR / "extra_kind.def", EXTRA_DIR / "extra_kind.cuh"]
    # a host compiler whose OpenMP / libstdc++ pieces are complete (see oracle/Makefile)
    ccbin = ["-ccbin", "/usr/bin/g++"] if Path("/usr/bin/g++").exists() else []

    def compile_one(src: str) -> Path:
        obj = OBJDIR / (src + ".o")
        if force or _stale(obj, [CSRC / src] + hdrs):
            cmd = [nvcc, *ccbin, *NVCC_FLAGS, "-c", str(CSRC / src), "-o", str(obj)]
            res = subprocess.run(cmd, capture_output=True, text=True)
            (OBJDIR / (src + ".log")).write_text(res.stdout + res.stderr)
            if res.returncode != 0:
                raise RuntimeError(f"nvcc failed on {src}:\n{res.stdout}\n{res.stderr}")
            if verbose:
                print(res.stderr)
        return obj

    with ThreadPoolExecutor(max_workers=len(CU_SOURCES)) as pool:
        objs = list(pool.map(compile_one, CU_SOURCES))
    if force or _stale(LIB, objs):
        cmd = [nvcc, *ccbin, "-shared", "-o", str(LIB), *map(str, objs), "-lcudart_static", "-ldl", "-lpthread", "-lrt"]
        res = subprocess.run(cmd, capture_output=True, text=True)
        if res.returncode != 0:
            raise RuntimeError(f"link failed:\n{res.stdout}\n{res.stderr}")
    return LIB


def build_oracle(verbose: bool = False) -> dict:
    """Build the test-only oracle libraries (building the checker is not using it)."""
    out = {}
    odir = ROOT / "oracle"
    targets = ["port"] if (odir / "spg_port.cpp").exists() else []
    if Path("/root/reference/spg/source").exists():
        targets += ["ref", "ref_extra"]
    for t in targets:
        res = subprocess.run(["make", "-C", str(odir), "-j8", t], capture_output=True, text=True)
        out[t] = res.returncode == 0
        if res.returncode != 0:
            raise RuntimeError(f"oracle target {t} failed:\n{res.stdout[-2000:]}\n{res.stderr[-4000:]}")
        if verbose:
            print(res.stdout[-500:])
    return out


if __name__ == "__main__":
    force = "--force" in sys.argv
    print(build_library(force=force, verbose="-v" in sys.argv))
    if "--extra" in sys.argv:
        print(build_library(force=force, verbose="-v" in sys.argv, extra_kind=True))
    if "--oracle" in sys.argv:
        print(build_oracle(verbose=True))
