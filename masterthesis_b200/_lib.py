"""Build + load the C-ABI shared library (csrc/*.cu -> libmt_b200.so) and bind it with ctypes.

The prototypes are parsed from include/mt_b200.h so that the header stays the single source of
truth for the boundary.  There is no fallback: if the library cannot be built or loaded, importing
the compute path raises.
"""
from __future__ import annotations

import ctypes
import os
import re
import subprocess
import threading
from pathlib import Path

PKG_DIR = Path(__file__).resolve().parent
REPO_DIR = PKG_DIR.parent
CSRC = PKG_DIR / "csrc"
HEADER = REPO_DIR / "include" / "mt_b200.h"
LIB_PATH = PKG_DIR / "libmt_b200.so"


class MtConv2d(ctypes.Structure):
    _fields_ = [(n, ctypes.c_int) for n in (
        "N", "Hin", "Win", "Cin", "Cin_s", "Hout", "Wout", "Cout", "Cout_s",
        "kh", "kw", "stride", "zpad", "out_H", "out_W", "out_pad", "act")] + [("slope", ctypes.c_float), ("window", ctypes.c_int),
                                                                               ("single", ctypes.c_int)]


def sources():
    return sorted(CSRC.glob("*.cu"))


def _stale() -> bool:
    if not LIB_PATH.exists():
        return True
    t = LIB_PATH.stat().st_mtime
    deps = list(CSRC.glob("*.cu")) + list(CSRC.glob("*.cuh")) + [HEADER]
    return any(d.stat().st_mtime > t for d in deps)


def build(force: bool = False, verbose: bool = False) -> Path:
    """Compile every CUDA source for sm_100a into one shared library, in-tree.  Safe across processes (N ranks
    of a torchrun job on a fresh checkout): an exclusive file lock serialises the builders, and a rank that
    waited finds the library fresh and returns."""
    if not force and not _stale():
        return LIB_PATH
    import fcntl
    build_dir = PKG_DIR / "build"
    build_dir.mkdir(exist_ok=True)
    with open(build_dir / ".lock", "w") as lock:
        fcntl.flock(lock, fcntl.LOCK_EX)
        try:
            if not force and not _stale():        # another process built it while this one waited
                return LIB_PATH
            return _build_locked(verbose)
        finally:
            fcntl.flock(lock, fcntl.LOCK_UN)


def _build_locked(verbose: bool) -> Path:
    nvcc = os.environ.get("NVCC", "nvcc")
    objs = []
    procs = []
    build_dir = PKG_DIR / "build"
    build_dir.mkdir(exist_ok=True)
    for src in sources():
        obj = build_dir / (src.stem + ".o")
        objs.append(obj)
        cmd = [nvcc, "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
               "-Xcompiler", "-fPIC", "-c", str(src), "-o", str(obj)]
        if verbose:
            cmd.insert(1, "-Xptxas=-v")
        procs.append((src, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT)))
    for src, p in procs:
        out, _ = p.communicate()
        if p.returncode != 0:
            raise RuntimeError(f"nvcc failed on {src.name}:\n{out.decode()}")
        if verbose and out:
            print(out.decode())
    tmp = LIB_PATH.with_suffix(f".so.tmp{os.getpid()}")
    cmd = [nvcc, "-shared", "-o", str(tmp)] + [str(o) for o in objs]
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT)
    if r.returncode != 0:
        raise RuntimeError(f"link failed:\n{r.stdout.decode()}")
    os.replace(tmp, LIB_PATH)
    return LIB_PATH


_CTYPE = {
    "int": ctypes.c_int, "float": ctypes.c_float, "long long": ctypes.c_longlong,
    "void*": ctypes.c_void_p, "const void*": ctypes.c_void_p,
    "float*": ctypes.c_void_p, "const float*": ctypes.c_void_p,
    "const char*": ctypes.c_char_p, "const MtConv2d*": ctypes.POINTER(MtConv2d),
    "const int*": ctypes.c_void_p, "int*": ctypes.c_void_p,
    "const float* const*": ctypes.c_void_p, "float* const*": ctypes.c_void_p,     # host arrays of device pointers
}


def header_prototypes() -> dict:
    """{name: (restype, [argtypes...])} for every function declared in include/mt_b200.h."""
    text = HEADER.read_text()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    protos = {}
    for m in re.finditer(r"^\s*((?:const\s+)?[A-Za-z_ ]+?\**)\s*\b(mt_\w+)\s*\(([^;{]*?)\)\s*;", text, flags=re.M):
        ret, name, args = m.group(1).strip(), m.group(2), m.group(3).strip()
        argtypes = []
        if args and args != "void":
            for a in args.split(","):
                a = " ".join(a.split())
                mm = re.match(r"^(.*?)(\w+)$", a)
                ty = mm.group(1).strip().replace(" *", "*")
                argtypes.append(ty)
        protos[name] = (ret.replace(" *", "*"), argtypes)
    return protos


_lib = None
_lock = threading.Lock()


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is not None:
            return _lib
        if _stale():
            build()
        l = ctypes.CDLL(str(LIB_PATH))
        for name, (ret, args) in header_prototypes().items():
            fn = getattr(l, name)  # AttributeError if the library lacks a declared symbol
            fn.restype = _CTYPE[ret]
            fn.argtypes = [_CTYPE[a] for a in args]
        _lib = l
    return _lib


class MtError(RuntimeError):
    pass


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        msg = lib().mt_error_string(rc)
        raise MtError(f"{what} failed with code {rc}: {msg.decode() if msg else '?'}")
