"""In-tree build of ``librtdc_b200.so`` (sm_100a kernels + native runtime) with plain nvcc.

The library is built next to the package (``<pkg>/lib/librtdc_b200.so``) so that it travels with
the source tree; nothing is installed into site-packages and no JIT cache is used. nvcc
cross-compiles for ``sm_100a`` on a machine without a GPU.
"""
from __future__ import annotations

import concurrent.futures as _fut
import hashlib
import os
import shutil
import subprocess
import sys
from pathlib import Path

PKG_DIR = Path(__file__).resolve().parent.parent
CSRC = PKG_DIR / "csrc"
LIB_DIR = PKG_DIR / "lib"
LIB_PATH = LIB_DIR / "librtdc_b200.so"
LAUNCHER_PATH = LIB_DIR / "rtdc-launch"
STAMP = LIB_DIR / ".build_stamp"

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC,-fvisibility=hidden,-Wall,-Wno-unused-function",
    "--expt-relaxed-constexpr",
]


def _flags():
    """NVCC_FLAGS plus ``$RTDC_EXTRA_NVCC_FLAGS`` (e.g. ``-DRTDC_GEMM_DEBUG_HOOKS`` for the tools/gemm_*probe*.py scripts)."""
    return NVCC_FLAGS + os.environ.get("RTDC_EXTRA_NVCC_FLAGS", "").split()


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and Path(cand).exists():
            return cand
    raise RuntimeError("nvcc not found; set NVCC or add /usr/local/cuda/bin to PATH")


def _sources():
    cu = sorted(CSRC.glob("*.cu"))
    cpp = sorted(p for p in CSRC.glob("*.cpp") if p.name != "launcher_main.cpp")
    return cu + cpp


def _digest() -> str:
    h = hashlib.sha256()
    files = sorted(list(CSRC.glob("*")) + [PKG_DIR.parent / "include" / "rtdc_b200.h"])
    for f in files:
        if f.is_file():
            h.update(f.name.encode())
            h.update(f.read_bytes())
    h.update(" ".join(_flags()).encode())
    return h.hexdigest()


def is_current() -> bool:
    return LIB_PATH.exists() and STAMP.exists() and STAMP.read_text().strip() == _digest()


def _run(cmd, verbose):
    if verbose:
        print("[rtdc build]", " ".join(map(str, cmd)), flush=True)
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"build step failed: {' '.join(map(str, cmd))}\n{r.stdout}\n{r.stderr}")
    return r.stdout + r.stderr


def build(force: bool = False, verbose: bool = False, ptxas_info: bool = False) -> Path:
    """Compile every translation unit under csrc/ for sm_100a and link the shared library.

    Safe to call from many processes at once (every worker rank of a ``TorchTrainer.fit()`` calls it through the first
    op): an exclusive file lock serialises the check-and-build, objects and the library are produced in a private
    temporary directory and moved into place with ``os.replace``, so no process can ever map a half-written file."""
    if not force and is_current():
        return LIB_PATH
    import fcntl
    import tempfile
    LIB_DIR.mkdir(exist_ok=True)
    with open(LIB_DIR / ".build.lock", "w") as lock:
        fcntl.flock(lock, fcntl.LOCK_EX)
        try:
            if not force and is_current():          # somebody else built it while we waited
                return LIB_PATH
            return _build_locked(verbose, ptxas_info, Path(tempfile.mkdtemp(prefix=".build_", dir=LIB_DIR)))
        finally:
            fcntl.flock(lock, fcntl.LOCK_UN)


def _build_locked(verbose: bool, ptxas_info: bool, tmp: Path) -> Path:
    nvcc = _nvcc()
    obj_dir = LIB_DIR / "obj"
    obj_dir.mkdir(exist_ok=True)
    srcs = _sources()
    extra = ["-Xptxas", "-v"] if ptxas_info else []
    digest = _digest()

    def compile_one(src: Path):
        obj = tmp / (src.stem + ".o")
        log = _run([nvcc, *_flags(), *extra, "-I", str(PKG_DIR.parent / "include"), "-c", str(src), "-o", str(obj)], verbose)
        return obj, log

    try:
        logs = {}
        with _fut.ThreadPoolExecutor(max_workers=min(8, len(srcs))) as ex:
            objs = []
            for (obj, log), src in zip(ex.map(compile_one, srcs), srcs):
                objs.append(obj)
                logs[src.name] = log
        so = tmp / LIB_PATH.name
        _run([nvcc, "-shared", "-o", str(so), *map(str, objs), "-lcudart_static", "-ldl", "-lrt", "-lpthread"], verbose)
        # native worker-group launcher (plain C++ executable, no CUDA)
        main_cpp = CSRC / "launcher_main.cpp"
        launcher = None
        if main_cpp.exists():
            cxx = shutil.which("g++") or "g++"
            launcher = tmp / LAUNCHER_PATH.name
            _run([cxx, "-O2", "-std=c++17", "-Wall", "-o", str(launcher), str(main_cpp), "-lpthread"], verbose)
        for obj in objs:                             # kept next to the library for cuobjdump / incremental inspection
            os.replace(obj, obj_dir / obj.name)
        if launcher is not None:
            os.replace(launcher, LAUNCHER_PATH)
        os.replace(so, LIB_PATH)
        stamp_tmp = tmp / "stamp"
        stamp_tmp.write_text(digest)
        os.replace(stamp_tmp, STAMP)                 # last: a reader that sees the stamp sees the finished library
        if ptxas_info:
            (LIB_DIR / "ptxas_info.txt").write_text("\n".join(f"==== {k}\n{v}" for k, v in logs.items()))
    finally:
        shutil.rmtree(tmp, ignore_errors=True)
    return LIB_PATH


if __name__ == "__main__":
    p = build(force="--force" in sys.argv, verbose=True, ptxas_info="--ptxas" in sys.argv)
    print(p)
