"""CPU: the reference-side binding shipped under julia/ (Julia itself is absent from this image, so the files are checked
structurally): every symbol shopt_b200.jl ccalls is declared in include/shopt_b200.h, exported by the built library and
called with as many arguments as the ctypes signature table holds; the patch lands on the loops it names."""
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SHIM = os.path.join(ROOT, "julia", "shopt_b200.jl")
REF = "/root/reference/shopt/shopt.jl"


def split_top_level(s):
    out, depth, cur = [], 0, ""
    for ch in s:
        if ch in "({[":
            depth += 1
        elif ch in ")}]":
            depth -= 1
        if ch == "," and depth == 0:
            out.append(cur)
            cur = ""
        else:
            cur += ch
    if cur.strip():
        out.append(cur)
    return [x.strip() for x in out if x.strip()]


def ccalls(text):
    """(symbol, number of argument types) for every ccall((:sym, LIBSHOPT), ret, (types...), args...)"""
    res = []
    for m in re.finditer(r"ccall\(\(:(\w+), LIBSHOPT\),\s*(\w+),\s*\(", text):
        i, depth = m.end(), 1
        while depth:
            depth += {"(": 1, ")": -1}.get(text[i], 0)
            i += 1
        types = text[m.end():i - 1]
        res.append((m.group(1), len(split_top_level(types))))
    return res


def test_shim_symbols_and_arity():
    from shopt_b200 import _lib
    text = open(SHIM).read()
    header = open(os.path.join(ROOT, "include", "shopt_b200.h")).read()
    calls = ccalls(text)
    assert len(calls) >= 18
    L = _lib.lib()
    for sym, nargs in calls:
        assert re.search(r"\b%s\s*\(" % sym, header), sym + " is not declared in include/shopt_b200.h"
        assert hasattr(L, sym), sym + " is not exported by libshopt_b200.so"
        assert sym in _lib.SIGNATURES, sym
        assert nargs == len(_lib.SIGNATURES[sym][1]), "%s: %d argument types in the shim, %d in the C ABI" % (
            sym, nargs, len(_lib.SIGNATURES[sym][1]))


def test_shim_structs_mirror_the_header():
    """StageParams / StageIO fields appear in the header's struct order."""
    from shopt_b200 import _lib
    text = open(SHIM).read()
    body = text[text.index("mutable struct StageParams"):text.index("const MODES")]
    jl_fields = re.findall(r"^\s+(\w+)::C", body, re.M)
    assert jl_fields == [f[0] for f in _lib.Params._fields_]
    io = text[text.index("mutable struct StageIO"):text.index('"""\n    stage_run')]
    jl_io = re.findall(r"(\w+)::(?:Ptr|Cint|Int64|Int32|NTuple)", io)
    assert jl_io == [f[0] for f in _lib.StageIO._fields_]


@pytest.mark.skipif(not os.path.exists(REF), reason="the reference checkout is not on this machine")
def test_patch_lands_on_the_loops():
    import importlib.util
    spec = importlib.util.spec_from_file_location("patch_shopt_jl", os.path.join(ROOT, "julia", "patch_shopt_jl.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    src = open(REF, encoding="utf-8").read()
    lines = src.split("\n")
    # the first line of every replaced range is what the docstring says it is
    starts = {79: 'include("reader.jl")', 131: "@time begin", 194: 'if mode == "chisq"', 281: 'if mode == "PCA"',
              316: 'if mode == "smooth"', 359: "@time begin", 534: "@time begin"}
    for (a, b) in mod.BLOCKS:
        assert lines[a - 1].strip() == starts[a]
        assert lines[b - 1].strip() in ("end", 'include("reader.jl")')
        # the replaced range is balanced on its own: as many block openers as `end`s
        blk = "\n".join(lines[a - 1:b])
        if a != 79:
            opens = len(re.findall(r"^\s*(?:@time begin|@threads for|for |if |try\b|function )", blk, re.M))
            assert opens == len(re.findall(r"^\s*end\s*$", blk, re.M)), (a, b)
            new = mod.BLOCKS[(a, b)]
            opens = len(re.findall(r"^\s*(?:@time begin|for |if |try\b|function )", new, re.M))
            assert opens == len(re.findall(r"^\s*end\s*$", new, re.M)), (a, b, "replacement")
    out = mod.apply(src)
    assert out.count("ShoptB200.") >= 7 and "optimize(cost," not in out and "optimize(costD," not in out
    # the one call left sits in the reference's `#= Future Work ... =#` block comment after the loop
    assert src.count("polynomial_optimizer(degree, training_u_coords") == 2
    assert out.count("polynomial_optimizer(degree, training_u_coords") == 1
    # untouched: the splices, the field fits, the writer
    for keep in ("splice!(starCatalog, i)", "optimize(interpCostS", "writeFitsData()"):
        assert keep in out
