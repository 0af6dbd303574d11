"""In-tree build of the CUDA library (nvcc, sm_100a only).  `python -m gpflow_b200.build [--force]`."""
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIBDIR = os.path.join(HERE, "lib")
LIB = os.path.join(LIBDIR, "libgpflow_b200.so")
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17", "--extended-lambda",
              "-Xcompiler", "-fPIC",
              "-I", os.path.join(os.path.dirname(HERE), "include")]


def _sources():
    return sorted(os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith(".cu"))


def _deps():
    d = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cu", ".cuh", ".h"))]
    d.append(os.path.join(os.path.dirname(HERE), "include", "gpflow_b200.h"))
    return d


def source_hash():
    """sha256 over the CUDA sources and the public header: ties a built library to the committed sources."""
    import hashlib
    h = hashlib.sha256()
    for p in sorted(_deps()):
        h.update(os.path.basename(p).encode())
        h.update(open(p, "rb").read())
    return h.hexdigest()


HASH_FILE = os.path.join(LIBDIR, "libgpflow_b200.sources.sha256")


def built_hash():
    try:
        return open(HASH_FILE).read().strip()
    except OSError:
        return None


def up_to_date():
    if not os.path.exists(LIB) or built_hash() != source_hash():
        return False
    t = os.path.getmtime(LIB)
    return all(os.path.getmtime(p) <= t for p in _deps())


def build(force=False, verbose=False):
    # objects of an experiment build (GPFLOW_NVCC_EXTRA) must not be re-linked into a default library, and vice versa
    if "+extra" in (built_hash() or "") or os.environ.get("GPFLOW_NVCC_EXTRA", "").strip():
        force = True
    if not force and up_to_date():
        return LIB
    nvcc = os.environ.get("NVCC", "nvcc")
    os.makedirs(LIBDIR, exist_ok=True)
    objdir = os.path.join(HERE, "build")
    os.makedirs(objdir, exist_ok=True)
    srcs = _sources()
    objs = [os.path.join(objdir, os.path.basename(s)[:-3] + ".o") for s in srcs]
    newest_hdr = max(os.path.getmtime(p) for p in _deps() if not p.endswith(".cu"))

    def compile_one(so):
        s, o = so
        if not force and os.path.exists(o) and os.path.getmtime(o) >= max(os.path.getmtime(s), newest_hdr):
            return
        extra = os.environ.get("GPFLOW_NVCC_EXTRA", "").split()   # experiments only (e.g. -DGPF_SF2_CONST=1); use with --force
        cmd = [nvcc] + NVCC_FLAGS + extra + (["-Xptxas", "-v"] if verbose else []) + ["-c", s, "-o", o]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if verbose:
            sys.stderr.write(r.stderr)
        if r.returncode != 0:
            raise RuntimeError("nvcc failed for %s:\n%s" % (s, r.stderr[-4000:]))

    with ThreadPoolExecutor(max_workers=min(8, len(srcs))) as ex:
        list(ex.map(compile_one, zip(srcs, objs)))
    # -cudart shared: the runtime stays a separate libcudart.so (torch ships one); a static cudart would put the
    # names of CUDA runtime entry points this library never calls into the shipped binary
    r = subprocess.run([nvcc, "-shared", "-cudart", "shared", "-o", LIB] + objs + ["-gencode", "arch=compute_100a,code=sm_100a", "-ldl"],
                       capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("link failed:\n" + r.stderr[-4000:])
    extra = os.environ.get("GPFLOW_NVCC_EXTRA", "").strip()
    with open(HASH_FILE, "w") as f:
        # an experiment build (extra -D flags) is marked: up_to_date() is then false and the next plain build() restores
        # the default library, smoke() refuses a marked one
        f.write(source_hash() + (" +extra:" + extra if extra else "") + "\n")
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
