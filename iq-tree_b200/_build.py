"""In-tree builds of the native libraries.

* ``build_cuda()``   -> iq-tree_b200/lib/liblktree_cuda.so : THE PRODUCT: the sm_100a kernels behind the C ABI of
  include/lkcuda.h and include/lkpars.h, nothing else.  nvcc cross-compiles it without a GPU.
* ``build_harness()`` -> harness/_build/liblktree_harness.so : TEST / BENCH DRIVER.  harness/*.cpp is a stand-alone C++ mirror of the
  reference's PhyloTree call sequences (flat C view: harness/lktree.h) so that tests and bench.py can drive the engine from
  Python exactly as the reference would; linked against the product library.  The real binding is integration/.
* ``build_oracle()`` -> oracle/_build/liblktree_oracle.so  : TEST INFRASTRUCTURE.  The same driver linked
  against the scalar C restatement oracle/lk_oracle.c instead of the CUDA engine; only tests/,
  __graft_entry__.smoke() and bench.py's cpu_baseline leg load it.
* ``build_reference()`` -> oracle/_ref/ref_harness (+ ref_harness_cuda, ref_harness_oracle) : the UNMODIFIED reference compiled from
  /root/reference by oracle/Makefile.ref, and the same sources with integration/ applied, linked against the product library
  (only where /root/reference exists; the GPU box uses the prebuilt binaries).
"""
import os
import shutil
import subprocess
import sys

PKG = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(PKG)
CSRC = os.path.join(PKG, "csrc")
LIBDIR = os.path.join(PKG, "lib")
CUDA_LIB = os.path.join(LIBDIR, "liblktree_cuda.so")
HARNESS_DIR = os.path.join(ROOT, "harness")
HARNESS_LIB = os.path.join(HARNESS_DIR, "_build", "liblktree_harness.so")
ORACLE_DIR = os.path.join(ROOT, "oracle")
ORACLE_LIB = os.path.join(ORACLE_DIR, "_build", "liblktree_oracle.so")
REF_HARNESS = os.path.join(ORACLE_DIR, "_ref", "ref_harness")
REF_HARNESS_CUDA = os.path.join(ORACLE_DIR, "_ref", "ref_harness_cuda")
REFERENCE = "/root/reference"

CUDA_SOURCES = ["lkcuda_engine.cu", "lkcuda_partial.cu", "lkcuda_partial_mma.cu", "lkcuda_partial_mmak.cu", "lkcuda_reduce.cu", "lkcuda_aux.cu", "lkcuda_sitemodel.cu", "lkcuda_pars.cu", "lkcuda_noalloc.cu"]
HOST_SOURCES = ["lk_phylotree.cpp", "lk_parsimony.cpp", "lk_rate_em.cpp", "lk_capi.cpp"]
ORACLE_SOURCES = ["lk_oracle.c", "lkpars_oracle.c"]
HEADERS = [os.path.join(CSRC, "lkcuda_common.cuh"), os.path.join(CSRC, "lkcuda_traverse.cuh"),
           os.path.join(ROOT, "include", "lkcuda.h"), os.path.join(ROOT, "include", "lkpars.h")]
HARNESS_HEADERS = HEADERS + [os.path.join(HARNESS_DIR, "lk_phylotree.h"), os.path.join(HARNESS_DIR, "lktree.h")]


def _newer(target, sources):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(s) > t for s in sources if os.path.exists(s))


def _run(cmd, **kw):
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, **kw)
    if r.returncode != 0:
        sys.stderr.write(r.stdout)
        raise RuntimeError("build failed: " + " ".join(cmd))
    return r.stdout


def nvcc_path():
    for c in (shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if c and os.path.exists(c):
            return c
    return None


def build_cuda(force=False, verbose=False):
    """Compile every CUDA source for sm_100a into one shared library (in-tree, travels to the GPU box)."""
    srcs = [os.path.join(CSRC, s) for s in CUDA_SOURCES]
    if not force and not _newer(CUDA_LIB, srcs + HEADERS):
        return CUDA_LIB
    nvcc = nvcc_path()
    if nvcc is None:
        if os.path.exists(CUDA_LIB):
            return CUDA_LIB       # a GPU box without nvcc uses the prebuilt library
        raise RuntimeError("nvcc not found and %s is not built" % CUDA_LIB)
    os.makedirs(LIBDIR, exist_ok=True)
    # one object per source (only the changed ones are recompiled, all of them in parallel), then one link
    objdir = os.path.join(LIBDIR, "obj")
    os.makedirs(objdir, exist_ok=True)
    flags = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "-Xcompiler", "-fPIC",
             "-Xptxas", "-v" if verbose else "-O3"]
    procs, objs = [], []
    for s in srcs:
        o = os.path.join(objdir, os.path.basename(s).replace(".cu", ".o"))
        objs.append(o)
        if force or _newer(o, [s] + HEADERS):
            procs.append((s, subprocess.Popen([nvcc] + flags + ["-c", s, "-o", o], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    failed = []
    for s, pr in procs:
        out = pr.communicate()[0]
        if pr.returncode != 0:
            sys.stderr.write(out)
            failed.append(s)
        elif verbose:
            print(out)
    if failed:
        raise RuntimeError("build failed: " + " ".join(failed))
    _run([nvcc, "-gencode", "arch=compute_100a,code=sm_100a", "-shared", "-o", CUDA_LIB] + objs)
    return CUDA_LIB


def build_tools(force=False):
    """tools/fp64_peak: the FP64 / HBM peak probe bench.py runs for the roofline denominators of the protein / codon configs"""
    src, exe = os.path.join(ROOT, "tools", "fp64_peak.cu"), os.path.join(ROOT, "tools", "fp64_peak")
    if not force and not _newer(exe, [src]):
        return exe
    nvcc = nvcc_path()
    if nvcc is None:
        return exe if os.path.exists(exe) else None
    _run([nvcc, "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-o", exe, src])
    return exe


def build_write_peak(force=False):
    """tools/write_peak: ceiling of a write-only HBM stream issued with the DNA traversal kernel's store (bench.py: roofline.write_only_peak)"""
    src, exe = os.path.join(ROOT, "tools", "write_peak.cu"), os.path.join(ROOT, "tools", "write_peak")
    if not force and not _newer(exe, [src]):
        return exe
    nvcc = nvcc_path()
    if nvcc is None:
        return exe if os.path.exists(exe) else None
    _run([nvcc, "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-std=c++17", "-o", exe, src])
    return exe


def build_harness(force=False):
    """TEST / BENCH DRIVER: the C++ mirror of the reference's PhyloTree call sequences, linked against the product library."""
    cuda = build_cuda()
    srcs = [os.path.join(HARNESS_DIR, s) for s in HOST_SOURCES]
    if not force and not _newer(HARNESS_LIB, srcs + HARNESS_HEADERS + [cuda]):
        return HARNESS_LIB
    if shutil.which("g++") is None:
        if os.path.exists(HARNESS_LIB):
            return HARNESS_LIB
        raise RuntimeError("g++ not found and %s is not built" % HARNESS_LIB)
    bdir = os.path.dirname(HARNESS_LIB)
    os.makedirs(bdir, exist_ok=True)
    objs = []
    for s in HOST_SOURCES:
        o = os.path.join(bdir, s.replace(".cpp", ".o"))
        _run(["g++", "-O2", "-fPIC", "-std=c++17", "-c", os.path.join(HARNESS_DIR, s), "-o", o])
        objs.append(o)
    _run(["g++", "-shared", "-o", HARNESS_LIB] + objs + ["-L" + LIBDIR, "-llktree_cuda", "-Wl,-rpath,$ORIGIN/../../iq-tree_b200/lib", "-lm"])
    return HARNESS_LIB


def build_oracle(force=False):
    """TEST INFRASTRUCTURE: harness mirror + scalar C oracle (no CUDA)."""
    srcs = [os.path.join(HARNESS_DIR, s) for s in HOST_SOURCES] + [os.path.join(ORACLE_DIR, s) for s in ORACLE_SOURCES]
    if not force and not _newer(ORACLE_LIB, srcs + HARNESS_HEADERS):
        return ORACLE_LIB
    if shutil.which("gcc") is None or shutil.which("g++") is None:
        if os.path.exists(ORACLE_LIB):
            return ORACLE_LIB
        raise RuntimeError("gcc/g++ not found and %s is not built" % ORACLE_LIB)
    bdir = os.path.dirname(ORACLE_LIB)
    os.makedirs(bdir, exist_ok=True)
    objs = []
    for s in ORACLE_SOURCES:
        o = os.path.join(bdir, s.replace(".c", ".o"))
        _run(["gcc", "-O2", "-fPIC", "-ffp-contract=off", "-c", os.path.join(ORACLE_DIR, s), "-o", o])
        objs.append(o)
    for s in HOST_SOURCES:
        o = os.path.join(bdir, s.replace(".cpp", ".o"))
        _run(["g++", "-O2", "-fPIC", "-std=c++17", "-c", os.path.join(HARNESS_DIR, s), "-o", o])
        objs.append(o)
    _run(["g++", "-shared", "-o", ORACLE_LIB] + objs + ["-lm"])
    return ORACLE_LIB


def build_reference(jobs=8):
    """TEST INFRASTRUCTURE: compile the unmodified reference into oracle/_ref/ref_harness (needs /root/reference)."""
    if not os.path.isdir(REFERENCE):
        return REF_HARNESS if os.path.exists(REF_HARNESS) else None
    _run(["make", "-f", os.path.join("oracle", "Makefile.ref"), "-j%d" % jobs], cwd=ROOT)
    # the same sources with the reference-side binding (integration/) applied, linked against the product and the oracle libraries
    build_cuda(); build_oracle()
    _run(["make", "-f", os.path.join("oracle", "Makefile.ref"), "-j%d" % jobs, "cuda"], cwd=ROOT)
    # the second CPU baseline of SURVEY 8(d): the reference built as IQTREE_FLAGS="omp fma"
    _run(["make", "-f", os.path.join("oracle", "Makefile.ref"), "-j%d" % jobs, "fma"], cwd=ROOT)
    return REF_HARNESS
