"""Loads the REAL reference code from /root/reference (build container only).  TEST INFRASTRUCTURE ONLY.

The reference tree does not exist on the GPU box: nothing in ``tests -m gpu``, ``smoke()`` or ``bench.py`` may call
this at run time.  It is used by ``oracle/gen_golden.py`` (writes tests/golden/) and by the container-only pin
tests, which skip when the tree is absent.

* ``utils/ctcDecoder.py``, ``utils/BeamSearch.py``, ``utils/NgramLM.py`` import as-is (torch, numpy, math, re).
* ``infer.py`` cannot be imported (argparse at import time, infer.py:58; missing third-party deps), so its pure
  functions are AST-extracted and exec'd: align_canonical_decoded (:162-209), stastics (:405-433),
  print_aligned_string (:155-160).
"""
from __future__ import annotations

import ast
import importlib
import math
import os
import sys

REF_ROOT = os.environ.get("MDD_REF_ROOT", "/root/reference")


def available() -> bool:
    return os.path.isdir(os.path.join(REF_ROOT, "egs", "attention_aug", "utils"))


def eg_dir(eg="attention_aug"):
    return os.path.join(REF_ROOT, "egs", eg)


_cache = {}


def decoder_module(eg="attention_aug"):
    """Returns the reference's utils.ctcDecoder module for one recipe (its BeamDecoder imports utils.* by name)."""
    if eg in _cache:
        return _cache[eg]
    sys.dont_write_bytecode = True
    root = eg_dir(eg)
    for m in [k for k in sys.modules if k == "utils" or k.startswith("utils.")]:
        del sys.modules[m]
    sys.path.insert(0, root)
    try:
        mod = importlib.import_module("utils.ctcDecoder")
        importlib.import_module("utils.BeamSearch")
        importlib.import_module("utils.NgramLM")
    finally:
        sys.path.remove(root)
    _cache[eg] = mod
    return mod


def infer_functions():
    """dict of the pure functions of egs/attention_aug/infer.py."""
    if "infer" in _cache:
        return _cache["infer"]
    src = open(os.path.join(eg_dir(), "infer.py")).read()
    tree = ast.parse(src)
    want = {"align_canonical_decoded", "stastics", "print_aligned_string"}
    body = [n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name in want]
    ns = {"math": math}
    exec(compile(ast.Module(body=body, type_ignores=[]), "infer.py", "exec"), ns)
    _cache["infer"] = {k: ns[k] for k in want}
    return _cache["infer"]


def units_vocab(eg="attention_aug"):
    """int2char as data_loader.Vocab builds it (data_loader.py:13-52): 0 'blank', 1 'UNK', then data/units order."""
    int2char = {0: "blank", 1: "UNK"}
    with open(os.path.join(eg_dir(eg), "data", "units")) as f:
        for line in f:
            parts = line.strip().split(" ")
            w = " ".join(parts[1:]) if len(parts) > 1 else parts[0]
            for word in w.split(" "):
                if word and word not in int2char.values():
                    int2char[len(int2char)] = word
    return int2char


def arpa_path(eg="attention_aug"):
    return os.path.join(eg_dir(eg), "data", "lm_phone_bg.arpa")
