"""In-tree native builds.

Two artefacts, both written next to the Python sources so that they travel with a
snapshot of the repository (no JIT cache under ~/.cache is involved):

* ``_C.<abi>.so``            host runtime (C++): reducer, bucket assignment, parameter
                             verification, symmetric-arena handle manager.  Built with g++
                             against the torch headers.
* ``lib/libb200ddp_kernels.so``  every CUDA kernel, compiled by nvcc for **sm_100a only**
                             (``-gencode arch=compute_100a,code=sm_100a``), exported through the
                             C ABI declared in ``include/b200ddp.h``.  No torch headers in the
                             kernels: they compile in seconds and can be driven from any
                             host (ctypes here).

A content hash of the sources (+ flags) is stored beside each artefact; a rebuild happens
only when the hash changes, so a fresh checkout on a GPU box reuses the shipped binaries.
"""
from __future__ import annotations

import hashlib
import os
import shutil
import subprocess
import sys
import sysconfig
import time
from pathlib import Path
from typing import List, Optional

PKG = Path(__file__).resolve().parent
ROOT = PKG.parent
CSRC = ROOT / "csrc"
INCLUDE = ROOT / "include"
LIBDIR = PKG / "lib"
KERNEL_LIB = LIBDIR / "libb200ddp_kernels.so"
EXT_SUFFIX = sysconfig.get_config_var("EXT_SUFFIX") or ".so"
HOST_EXT = PKG / f"_C{EXT_SUFFIX}"

NVCC_ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
NVCC_FLAGS = ["-O3", "-std=c++17", "-lineinfo", "--use_fast_math", "-Xcompiler", "-fPIC",
              "--expt-relaxed-constexpr", "-Xptxas", "-v", "-ccbin", "/usr/bin/g++"]


def _nvcc() -> Optional[str]:
    cand = os.environ.get("NVCC") or shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    return cand if os.path.exists(cand) else None


def _host_cxx() -> str:
    # NOTE: $CXX on this image points at /opt/gcc/bin/g++, a wrapper that links libstdc++
    # statically; a second copy of libstdc++ inside a process that already loaded torch
    # corrupts iostream locale facets (segfault on `ostream << int`).  Use the distro g++.
    cand = os.environ.get("B200DDP_CXX") or ("/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++")
    return cand


def _digest(paths: List[Path], extra: str) -> str:
    h = hashlib.sha256(extra.encode())
    for p in sorted(paths):
        h.update(p.name.encode())
        h.update(p.read_bytes())
    return h.hexdigest()


def _stale(artefact: Path, digest: str) -> bool:
    stamp = artefact.with_suffix(artefact.suffix + ".hash")
    return not (artefact.exists() and stamp.exists() and stamp.read_text().strip() == digest)


def _stamp(artefact: Path, digest: str) -> None:
    artefact.with_suffix(artefact.suffix + ".hash").write_text(digest)


def _run(cmd: List[str], log: Path) -> None:
    t0 = time.time()
    proc = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    log.parent.mkdir(parents=True, exist_ok=True)
    log.write_text(" ".join(cmd) + "\n" + proc.stdout)
    if proc.returncode != 0:
        sys.stderr.write(proc.stdout[-8000:])
        raise RuntimeError(f"native build failed ({cmd[0]}), see {log}")
    if os.environ.get("B200DDP_BUILD_VERBOSE"):
        print(f"[build] {' '.join(cmd[:3])} ... {time.time() - t0:.1f}s", file=sys.stderr)


class _BuildLock:
    """Inter-process lock around a build.  When several ranks of a job find the shipped binary stale at the same time,
    one of them builds and the others wait and then re-check: without it a rank could dlopen a half-written library
    ("file too short") and leave its peers waiting in a collective."""

    def __init__(self, name: str) -> None:
        self.path = ROOT / "build" / f".{name}.lock"
        self.fh = None

    def __enter__(self):
        import fcntl

        self.path.parent.mkdir(parents=True, exist_ok=True)
        self.fh = open(self.path, "w")
        fcntl.flock(self.fh, fcntl.LOCK_EX)
        return self

    def __exit__(self, *exc):
        import fcntl

        fcntl.flock(self.fh, fcntl.LOCK_UN)
        self.fh.close()
        return False


def kernel_sources() -> List[Path]:
    return sorted((CSRC / "kernels").glob("*.cu"))


def build_kernels(force: bool = False, verbose: bool = False) -> Path:
    """nvcc-compile every ``csrc/kernels/*.cu`` into one shared library (sm_100a)."""
    srcs = kernel_sources()
    hdrs = sorted((CSRC / "kernels").glob("*.cuh")) + sorted(INCLUDE.glob("*.h"))
    digest = _digest(srcs + hdrs, " ".join(NVCC_ARCH + NVCC_FLAGS))
    if not force and not _stale(KERNEL_LIB, digest):
        return KERNEL_LIB
    with _BuildLock("kernels"):
        if not force and not _stale(KERNEL_LIB, digest):       # another process built it while we waited
            return KERNEL_LIB
        return _build_kernels_locked(srcs, hdrs, digest, force)


def _build_kernels_locked(srcs: List[Path], hdrs: List[Path], digest: str, force: bool) -> Path:
    nvcc = _nvcc()
    if nvcc is None:
        raise RuntimeError("nvcc not found: cannot build the sm_100a kernel library")
    LIBDIR.mkdir(parents=True, exist_ok=True)
    objdir = ROOT / "build" / "kernels"
    objdir.mkdir(parents=True, exist_ok=True)
    objs = []
    procs = []
    for s in srcs:
        o = objdir / (s.stem + ".o")
        sd = _digest([s] + hdrs, " ".join(NVCC_ARCH + NVCC_FLAGS))
        objs.append(o)
        if not force and not _stale(o, sd):
            continue
        cmd = [nvcc, *NVCC_ARCH, *NVCC_FLAGS, "-I", str(INCLUDE), "-I", str(CSRC / "kernels"),
               "-c", str(s), "-o", str(o)]
        procs.append((s, o, sd, cmd, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    failed = []
    for s, o, sd, cmd, p in procs:
        out, _ = p.communicate()
        (objdir / (s.stem + ".log")).write_text(" ".join(cmd) + "\n" + out)
        if p.returncode != 0:
            failed.append((s, out))
        else:
            _stamp(o, sd)
    if failed:
        for s, out in failed:
            sys.stderr.write(f"---- {s.name} ----\n{out[-6000:]}\n")
        raise RuntimeError("nvcc failed for: " + ", ".join(s.name for s, _ in failed))
    tmp = KERNEL_LIB.with_suffix(f".so.tmp{os.getpid()}")
    _run([nvcc, *NVCC_ARCH, "-ccbin", "/usr/bin/g++", "-shared", "-Xcompiler", "-fPIC", "-o", str(tmp), *map(str, objs),
          "-lcudart", "-lcuda"], objdir / "link.log")
    os.replace(tmp, KERNEL_LIB)                 # atomic: a concurrent dlopen sees the old or the new file, never half of one
    _stamp(KERNEL_LIB, digest)
    return KERNEL_LIB


def build_host_extension(force: bool = False) -> Path:
    """g++-compile ``csrc/reducer/*.cpp`` + ``csrc/runtime/*.cpp`` into ``_C``."""
    import torch
    from torch.utils import cpp_extension as ce

    srcs = sorted((CSRC / "reducer").glob("*.cpp")) + sorted((CSRC / "runtime").glob("*.cpp"))
    hdrs = sorted((CSRC / "reducer").glob("*.h")) + sorted((CSRC / "runtime").glob("*.h"))
    flags = ["-O2", "-std=c++17", "-fPIC", "-DTORCH_EXTENSION_NAME=_C", "-DTORCH_API_INCLUDE_EXTENSION_H",
             f"-D_GLIBCXX_USE_CXX11_ABI={int(torch._C._GLIBCXX_USE_CXX11_ABI)}", "-Wno-attributes"]
    cxx = _host_cxx()
    flags.append(f"-DB200DDP_CXX_ID=\"{cxx}\"")
    digest = _digest(srcs + hdrs, " ".join(flags) + torch.__version__)
    if not force and not _stale(HOST_EXT, digest):
        return HOST_EXT
    with _BuildLock("host"):
        if not force and not _stale(HOST_EXT, digest):
            return HOST_EXT
        return _build_host_locked(srcs, hdrs, flags, digest, cxx, force)


def _build_host_locked(srcs: List[Path], hdrs: List[Path], flags: List[str], digest: str, cxx: str, force: bool) -> Path:
    import torch
    from torch.utils import cpp_extension as ce

    inc = [f"-I{p}" for p in ce.include_paths()] + [f"-I{sysconfig.get_paths()['include']}",
                                                     f"-I{CSRC / 'reducer'}", f"-I{CSRC / 'runtime'}",
                                                     "-I/usr/local/cuda/include"]
    objdir = ROOT / "build" / "host"
    objdir.mkdir(parents=True, exist_ok=True)
    objs, procs = [], []
    for s in srcs:
        o = objdir / (s.stem + ".o")
        sd = _digest([s] + hdrs, " ".join(flags) + torch.__version__)
        objs.append(o)
        if not force and not _stale(o, sd):
            continue
        cmd = [cxx, *flags, *inc, "-c", str(s), "-o", str(o)]
        procs.append((s, o, sd, cmd, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    for s, o, sd, cmd, p in procs:
        out, _ = p.communicate()
        (objdir / (s.stem + ".log")).write_text(" ".join(cmd) + "\n" + out)
        if p.returncode != 0:
            sys.stderr.write(out[-8000:])
            raise RuntimeError(f"g++ failed for {s.name}")
        _stamp(o, sd)
    torch_lib = Path(torch.__file__).parent / "lib"
    tmp = HOST_EXT.with_name(HOST_EXT.name + f".tmp{os.getpid()}")
    link = [cxx, "-shared", "-o", str(tmp), *map(str, objs), f"-L{torch_lib}", "-lc10", "-ltorch_cpu",
            "-ltorch", "-ltorch_python", "-ldl", f"-Wl,-rpath,{torch_lib}"]
    _run(link, objdir / "link.log")
    os.replace(tmp, HOST_EXT)
    _stamp(HOST_EXT, digest)
    return HOST_EXT


REF_DIR = ROOT / "baseline" / "_ref"


def reference_sources() -> List[Path]:
    """Places the unmodified reference script may be read from (first hit wins)."""
    cands = []
    if os.environ.get("B200DDP_REFERENCE"):
        cands.append(Path(os.environ["B200DDP_REFERENCE"]))
    cands.append(Path("/root/reference/dpp.py"))
    return [c / "dpp.py" if c.is_dir() else c for c in cands]


def install_reference(verbose: bool = False) -> Optional[Path]:
    """Put the UNMODIFIED reference (``dpp.py``) under ``baseline/_ref`` so ``bench.py --impl reference``
    finds it on the GPU box (the directory is git-ignored but travels with the snapshot).

    The reference is a bare script: ``pip install --target baseline/_ref /root/reference`` refuses it
    ("not installable": no setup.py / pyproject.toml).  It is therefore installed through pip from a
    scratch copy carrying a minimal ``setup.py`` (``py_modules=["dpp"]``), and, if pip is unavailable,
    copied.  Either way the result is compared byte for byte with the source.
    """
    dst = REF_DIR / "dpp.py"
    src = next((c for c in reference_sources() if c.is_file()), None)
    if src is None:
        return dst if dst.is_file() else None
    if dst.is_file() and dst.read_bytes() == src.read_bytes():
        return dst
    REF_DIR.mkdir(parents=True, exist_ok=True)
    import tempfile

    with tempfile.TemporaryDirectory(prefix="b200ddp_ref_") as tmp:
        shutil.copyfile(src, Path(tmp) / "dpp.py")
        (Path(tmp) / "setup.py").write_text(
            "from setuptools import setup\nsetup(name='dpp-reference', version='0', py_modules=['dpp'])\n")
        cmd = [sys.executable, "-m", "pip", "install", "--no-index", "--no-build-isolation", "--no-deps", "--quiet",
               "--find-links", "/opt/wheelhouse", "--upgrade", "--target", str(REF_DIR), tmp]
        proc = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
        if verbose or os.environ.get("B200DDP_BUILD_VERBOSE"):
            print(f"[build] pip install reference -> rc {proc.returncode}", file=sys.stderr)
    if not (dst.is_file() and dst.read_bytes() == src.read_bytes()):
        shutil.copyfile(src, dst)                      # pip unavailable / refused: plain copy
    if dst.read_bytes() != src.read_bytes():
        raise RuntimeError(f"{dst} differs from {src}")
    return dst


def build_all(force: bool = False) -> None:
    build_host_extension(force)
    if kernel_sources():
        build_kernels(force)
    try:
        install_reference()
    except Exception as e:  # the reference arm is a benchmark baseline, never a build blocker
        print(f"[build] reference install skipped: {e}", file=sys.stderr)


if __name__ == "__main__":
    os.environ.setdefault("B200DDP_BUILD_VERBOSE", "1")
    build_all(force="--force" in sys.argv)
    print("built:", HOST_EXT.name, KERNEL_LIB if KERNEL_LIB.exists() else "(no kernels yet)")
