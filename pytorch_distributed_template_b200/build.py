"""In-tree build of libpdt_b200.so (sm_100a only).

``python -m pytorch_distributed_template_b200.build`` compiles every
``csrc/*.cu`` with ``nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo``
and links ``lib/libpdt_b200.so`` next to the package, so the built library
travels with the source tree.  nvcc cross-compiles without a GPU.  Staleness is
decided by CONTENT hashes (of each source + the headers for an object, of all of
them for the library, recorded in ``*.id`` files and inside the library as
``pdt_build_id()``), never by mtimes.
"""
import concurrent.futures
import glob
import hashlib
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
INCLUDE = os.path.join(os.path.dirname(HERE), "include")
LIBDIR = os.path.join(HERE, "lib")
OBJDIR = os.path.join(HERE, "build")
LIB = os.path.join(LIBDIR, "libpdt_b200.so")

NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "--expt-relaxed-constexpr", "-Xcompiler", "-fPIC", "-I", INCLUDE, "-I", CSRC]


def nvcc_path():
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found; set NVCC=/path/to/nvcc")


def _digest(paths, extra=""):
    h = hashlib.sha256(extra.encode())
    for p in sorted(paths):
        h.update(os.path.basename(p).encode())
        with open(p, "rb") as f:
            h.update(f.read())
    return h.hexdigest()[:20]


def _sources():
    return sorted(glob.glob(os.path.join(CSRC, "*.cu")))


def _headers():
    return sorted(glob.glob(os.path.join(CSRC, "*.cuh")) + glob.glob(os.path.join(INCLUDE, "*.h")))


def source_id():
    """Hash of everything the library is built from (contents, not mtimes: a checkout, a copy to another box
    and eight ranks starting at once all see the same value)."""
    return _digest(_sources() + _headers(), " ".join(NVCC_FLAGS[:8]))


def built_id():
    """Id recorded next to the library by the build that produced it, or None."""
    try:
        with open(LIB + ".id") as f:
            return f.read().strip() if os.path.exists(LIB) else None
    except OSError:
        return None


def _read(path):
    try:
        with open(path) as f:
            return f.read().strip()
    except OSError:
        return None


def _compile(nvcc, src, obj, verbose, defines=()):
    cmd = [nvcc] + NVCC_FLAGS + list(defines) + ["-c", src, "-o", obj]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("nvcc failed for {}:\n{}\n{}".format(src, r.stdout, r.stderr))
    if verbose and (r.stdout or r.stderr):
        print(r.stdout, r.stderr)
    return obj


def build(force=False, verbose=False):
    nvcc = nvcc_path()
    os.makedirs(LIBDIR, exist_ok=True)
    os.makedirs(OBJDIR, exist_ok=True)
    sources, headers = _sources(), _headers()
    want = source_id()
    if not force and built_id() == want:
        return LIB
    jobs, objs = [], []
    for src in sources:
        name = os.path.basename(src)[:-3]
        obj = os.path.join(OBJDIR, name + ".o")
        objs.append(obj)
        defines = ['-DPDT_BUILD_ID="{}"'.format(want)] if name == "buildid" else []
        oid = _digest([src] + headers, want if name == "buildid" else "")
        if force or not os.path.exists(obj) or _read(obj + ".id") != oid:
            jobs.append((src, obj, defines, oid))
    if jobs:
        with concurrent.futures.ThreadPoolExecutor(max_workers=min(8, len(jobs))) as ex:
            list(ex.map(lambda j: _compile(nvcc, j[0], j[1], verbose, j[2]), jobs))
        for _, obj, _, oid in jobs:
            with open(obj + ".id", "w") as f:
                f.write(oid)
    tmp = LIB + ".tmp.{}".format(os.getpid())
    cmd = [nvcc, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", tmp] + objs
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("link failed:\n{}\n{}".format(r.stdout, r.stderr))
    os.replace(tmp, LIB)          # readers see the old or the new library, never a half-written one
    with open(LIB + ".id", "w") as f:
        f.write(want)
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
