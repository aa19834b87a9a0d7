"""ctypes loader for libmddctc.so (the sm_100a CUDA kernels behind include/mddctc.h).

The library is built in-tree by ``build()`` (``nvcc -gencode arch=compute_100a,code=sm_100a``) so the ``.so`` travels
with the repository snapshot.  There is NO fallback: if the library is missing or fails to load, every product entry
point raises ``MddCtcLibraryError``.
"""
from __future__ import annotations

import ctypes as C
import glob
import os
import shutil
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(_HERE, "csrc")
SO_PATH = os.environ.get("MDD_SO_PATH") or os.path.join(_HERE, "libmddctc.so")   # override: experiment builds only
INCLUDE = os.path.join(os.path.dirname(_HERE), "include")

NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17", "-Xcompiler", "-fPIC",
              "--expt-relaxed-constexpr", "-Xptxas", "-v"]


class MddCtcLibraryError(RuntimeError):
    pass


def _nvcc():
    for cand in (shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise MddCtcLibraryError("nvcc not found: cannot build libmddctc.so")


def sources():
    return sorted(glob.glob(os.path.join(CSRC, "*.cu")))


def needs_build() -> bool:
    if not os.path.exists(SO_PATH):
        return True
    t = os.path.getmtime(SO_PATH)
    deps = sources() + glob.glob(os.path.join(CSRC, "*.cuh")) + glob.glob(os.path.join(INCLUDE, "*.h"))
    return any(os.path.getmtime(d) > t for d in deps)


def build(force: bool = False, verbose: bool = False) -> str:
    """Compile every .cu under csrc/ for sm_100a into libmddctc.so (parallel object builds, then one link)."""
    if not force and not needs_build():
        return SO_PATH
    nvcc = _nvcc()
    objdir = os.path.join(_HERE, "build")
    os.makedirs(objdir, exist_ok=True)
    procs = []
    objs = []
    hdr_t = max(os.path.getmtime(d) for d in glob.glob(os.path.join(CSRC, "*.cuh")) + glob.glob(os.path.join(INCLUDE, "*.h")))
    for src in sources():
        obj = os.path.join(objdir, os.path.basename(src)[:-3] + ".o")
        objs.append(obj)
        inc = [os.path.join(CSRC, l.split('"')[1]) for l in open(src) if l.startswith('#include "') and l.split('"')[1].endswith(".cu")]
        dep_t = max([os.path.getmtime(src), hdr_t] + [os.path.getmtime(f) for f in inc])    # a .cu may include another
        if not force and os.path.exists(obj) and os.path.getmtime(obj) > dep_t:
            continue
        log = open(obj + ".log", "w")
        extra = os.environ.get("MDD_NVCC_EXTRA", "").split()          # e.g. -DMDD_PAIR_TIMING for a debug build
        procs.append((src, log, subprocess.Popen([nvcc] + NVCC_FLAGS + extra + ["-c", src, "-o", obj], stdout=log,
                                                 stderr=subprocess.STDOUT)))
    failed = []
    for src, log, pr in procs:
        rc = pr.wait()
        log.close()
        if rc != 0:
            failed.append(src)
    if failed:
        msgs = "\n".join(open(os.path.join(objdir, os.path.basename(s)[:-3] + ".o.log")).read()[-4000:] for s in failed)
        raise MddCtcLibraryError("nvcc failed for %s:\n%s" % (failed, msgs))
    subprocess.check_call([nvcc, "-shared", "-o", SO_PATH] + objs + ["-lcudart"])
    if verbose:
        for o in objs:
            if os.path.exists(o + ".log"):
                print(open(o + ".log").read())
    return SO_PATH


_lib = None

_i32p, _i64p, _f32p, _f64p, _u8p = (C.POINTER(t) for t in (C.c_int32, C.c_int64, C.c_float, C.c_double, C.c_uint8))


class PostOut(C.Structure):
    _fields_ = [("a_dec", C.c_void_p), ("a_can", C.c_void_p), ("a_ops", C.c_void_p), ("ins", C.c_void_p),
                ("sub", C.c_void_p), ("del_", C.c_void_p), ("counts", C.c_void_p)]


EXPORTS = ["mdd_version", "mdd_error_string", "mdd_launch_count", "mdd_launch_count_reset",
           "mdd_ctc_loss_workspace_bytes", "mdd_ctc_loss", "mdd_ctc_loss_static", "mdd_ctc_loss_logits", "mdd_row_lse", "mdd_copy_h2d_2d", "mdd_ctc_scale_grad", "mdd_greedy_decode", "mdd_collapse", "mdd_post_align", "mdd_corpus_counts",
           "mdd_edit_align_workspace_bytes", "mdd_edit_align", "mdd_beam_search_workspace_bytes", "mdd_beam_search",
           "mdd_exp_transpose", "mdd_prefix_init", "mdd_prefix_probs", "mdd_prefix_score", "mdd_prefix_advance", "mdd_prefix_select"]


def lib():
    """Returns the loaded library; raises MddCtcLibraryError when it is absent (no CPU fallback exists)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(SO_PATH):
        raise MddCtcLibraryError(
            "libmddctc.so is missing (%s). Build it with `python -c 'import __graft_entry__ as g; g.build()'`; "
            "there is no CPU fallback." % SO_PATH)
    try:
        L = C.CDLL(SO_PATH)
    except OSError as e:
        raise MddCtcLibraryError("cannot load %s: %s" % (SO_PATH, e)) from e
    vp, i, i64, f, d, sz = C.c_void_p, C.c_int, C.c_int64, C.c_float, C.c_double, C.c_size_t
    L.mdd_version.restype = i
    L.mdd_error_string.restype = C.c_char_p
    L.mdd_error_string.argtypes = [i]
    L.mdd_launch_count.restype = i64
    L.mdd_launch_count_reset.restype = None
    L.mdd_ctc_loss_workspace_bytes.argtypes = [i, i, i, i, i, C.POINTER(sz)]
    L.mdd_ctc_loss.argtypes = [vp, i, i, i, i64, i64, vp, i64, vp, vp, i, i, i, vp, f, vp, vp, vp, vp, i, vp, sz, vp]
    L.mdd_ctc_loss_static.argtypes = [vp, i, i, i, i64, i64, vp, i64, vp, vp, i, i, i, vp, f, vp, vp, vp, vp, i, vp]
    L.mdd_ctc_loss_logits.argtypes = [vp, i, i, i, i64, i64, vp, i64, vp, vp, i, i, i, vp, f, vp, vp, vp, vp, vp, vp, i,
                                      vp, sz, vp]
    L.mdd_row_lse.argtypes = [vp, i, i, i, i64, i64, vp, vp, vp]
    L.mdd_copy_h2d_2d.argtypes = [vp, sz, vp, sz, sz, sz, vp]
    L.mdd_ctc_scale_grad.argtypes = [vp, i, i, i, vp, f, vp]
    L.mdd_greedy_decode.argtypes = [vp, i, i, i, vp, i, vp, vp, vp, i, vp, vp]
    L.mdd_collapse.argtypes = [vp, i, i, vp, i, vp, vp, i, vp, vp]
    L.mdd_post_align.argtypes = [vp, i, vp, vp, i, vp, i, i, i, vp, i, vp, C.POINTER(PostOut), i, vp]
    L.mdd_corpus_counts.argtypes = [vp, i, vp, vp, i, vp, i, i, i, i, vp, vp, i, vp, i, vp, vp, i, vp, vp, vp, vp, vp, vp, vp]
    L.mdd_edit_align_workspace_bytes.argtypes = [i, i, i, C.POINTER(sz)]
    L.mdd_edit_align.argtypes = [vp, i, vp, vp, i, vp, i, i, i, vp, vp, i, vp, vp, i, vp, vp, C.POINTER(PostOut), vp,
                                 sz, vp]
    L.mdd_beam_search_workspace_bytes.argtypes = [i, i, i, i, C.POINTER(sz)]
    L.mdd_beam_search.argtypes = [vp, i, i, i, vp, i, i, vp, d, vp, i, vp, vp, vp, vp, sz, vp]
    L.mdd_exp_transpose.argtypes = [vp, i, i, i, vp, vp]
    L.mdd_prefix_init.argtypes = [vp, i, i, i, vp, i, i, vp, vp]
    L.mdd_prefix_probs.argtypes = [vp, i, i, i, vp, vp]
    L.mdd_prefix_score.argtypes = [vp, i, i, i, vp, i, i, vp, vp, vp, vp, vp, vp, vp]
    L.mdd_prefix_advance.argtypes = [vp, i, i, i, vp, i, i, vp, vp, vp, vp, vp, vp, vp, vp]
    L.mdd_prefix_select.argtypes = [i, i, i, i, i, d, i, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp]
    for name in EXPORTS:
        fn = getattr(L, name)            # AttributeError here = the .so is stale / an export is missing
        if name not in ("mdd_error_string", "mdd_launch_count", "mdd_launch_count_reset") and fn.argtypes is None \
                and name != "mdd_version":
            raise MddCtcLibraryError("no ctypes signature declared for %s" % name)
    _lib = L
    return L


def check(rc: int, what: str):
    if rc != 0:
        msg = lib().mdd_error_string(rc).decode()
        raise RuntimeError("%s failed: %s (code %d)" % (what, msg, rc))
