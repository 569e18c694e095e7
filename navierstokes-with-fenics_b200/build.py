"""Build recipe of libnsb200.so (hand-written sm_100a CUDA + the C-ABI of include/nsb200.h).

In-tree build with an explicit nvcc command so that the built library travels with the
repository snapshot to the GPU box.  ``python -m`` is not usable on a hyphenated directory:
call ``build()`` through ``__graft_entry__.build()`` or run this file directly.
"""
import hashlib
import os
import shutil
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libnsb200.so")
STAMP = os.path.join(HERE, ".libnsb200.stamp")
SOURCES = ["setup_host.cu", "context.cu", "assemble.cu", "assemble_jac.cu", "assemble_res.cu", "linalg.cu", "krylov.cu", "comm.cu", "direct.cu", "functionals.cu", "project.cu", "cg_persistent.cu"]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
              "--use_fast_math=false", "-Xcompiler", "-fPIC,-fopenmp,-O3", "-shared", "-lgomp", "-ldl"]


def _nvcc():
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found: libnsb200.so cannot be built (there is no CPU fallback)")


def _file_digest(paths, extra=""):
    h = hashlib.sha256()
    for p in paths:
        h.update(os.path.basename(p).encode())
        with open(p, "rb") as f:
            h.update(f.read())
    h.update(extra.encode())
    return h.hexdigest()


def _headers():
    hs = [os.path.join(CSRC, n) for n in sorted(os.listdir(CSRC)) if n.endswith((".cuh", ".h"))]
    return hs + [os.path.join(HERE, "..", "include", "nsb200.h")]


def _digest():
    srcs = [os.path.join(CSRC, n) for n in sorted(os.listdir(CSRC)) if os.path.isfile(os.path.join(CSRC, n))]
    return _file_digest(srcs + [os.path.join(HERE, "..", "include", "nsb200.h")], " ".join(NVCC_FLAGS + SOURCES))


def _compile_one(nvcc, src, obj, verbose):
    flags = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
             "-Xcompiler", "-fPIC,-fopenmp,-O3", "-c"]
    cmd = [nvcc] + flags + (["-Xptxas", "-v"] if verbose else []) + ["-o", obj, src]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError("nvcc failed:\n%s\n%s" % (" ".join(cmd), res.stderr[-8000:]))
    return res.stderr


def build(force=False, verbose=False):
    """Compile every CUDA source for sm_100a into libnsb200.so (no-op when up to date).  Each source is
    compiled to its own object (in parallel, cached under build/ by content digest), then linked."""
    from concurrent.futures import ThreadPoolExecutor
    dig = _digest()
    if not force and os.path.exists(LIB) and os.path.exists(STAMP):
        with open(STAMP) as f:
            if f.read().strip() == dig:
                return LIB
    nvcc = _nvcc()
    objdir = os.path.join(HERE, "build")
    os.makedirs(objdir, exist_ok=True)
    hdr = _headers()
    jobs, objs = [], []
    for s in SOURCES:
        src = os.path.join(CSRC, s)
        d = _file_digest([src] + hdr, "v2" + ("-v" if verbose else ""))[:16]
        obj = os.path.join(objdir, "%s.%s.o" % (s[:-3], d))
        objs.append(obj)
        if force or verbose or not os.path.exists(obj):
            for old in os.listdir(objdir):
                if old.startswith(s[:-3] + ".") and old.endswith(".o"):
                    os.remove(os.path.join(objdir, old))
            jobs.append((src, obj))
    with ThreadPoolExecutor(max_workers=min(8, max(1, len(jobs)))) as ex:
        outs = list(ex.map(lambda j: _compile_one(nvcc, j[0], j[1], verbose), jobs))
    if verbose:
        print("\n".join(outs))
    cmd = [nvcc, "-gencode", "arch=compute_100a,code=sm_100a", "-shared", "-Xcompiler", "-fPIC,-fopenmp", "-o", LIB] + objs + \
          ["-lgomp", "-ldl"]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError("nvcc link failed:\n%s\n%s" % (" ".join(cmd), res.stderr[-8000:]))
    with open(STAMP, "w") as f:
        f.write(dig)
    return LIB


if __name__ == "__main__":
    import sys
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
