"""The Haskell glue (accelerate-blas_b200/haskell/B200/*.hs) cannot be compiled here (no GHC in the image), so nothing
but this test keeps it in step with include/accblas_b200.h: every `foreign import ccall` is compared with the C
prototype of the same name -- arity, and per position the class of the type (pointer / CInt / Int64 / CSize) -- and
the saturated call sites in Hooks.hs are counted against the import they use.  Editing a prototype in the header
without the glue (or the other way round) fails here.

What the reference's hooks pass, for comparison: LLVM/PTX/Level3.hs:88-106 (gemm: handle, opB, opA, n, m, k, alpha, B,
ldb, A, lda, beta, C, ldc -- cuBLAS's column-major argument order) and LLVM/PTX/Level2.hs:105-124 (gemv).  The C ABI
takes the row-major operands as they are, so the glue passes (stream, dtype, opA, opB, rowsA, colsA, rowsB, colsB,
alpha, A, B, C); the order checked here is the header's.
"""
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "accblas_b200.h")
HS = os.path.join(ROOT, "accelerate-blas_b200", "haskell", "B200")

# entry points that block on a collective or a stream: a GHC `unsafe` call would stall every Haskell thread of the
# capability for that long, so these must be imported `safe`
BLOCKING = {"accblas_mg_init", "accblas_mg_finalize", "accblas_mg_gemm", "accblas_mg_gemm_plan",
            "accblas_mg_gemm_plan_destroy"}


def c_prototypes():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", " ", src, flags=re.S)
    src = re.sub(r"^\s*#.*$", " ", src, flags=re.M)
    protos = {}
    for ret, name, params in re.findall(r"([A-Za-z_][\w\s\*]*?)\b(accblas_\w+)\s*\(([^()]*)\)\s*;", src):
        if "typedef" in ret:
            continue
        plist = [p.strip() for p in params.split(",")] if params.strip() not in ("", "void") else []
        protos[name] = (c_class(ret.strip(), is_return=True), [c_class(p) for p in plist])
    return protos


def c_class(decl, is_return=False):
    decl = decl.replace("const", " ").strip()
    if "*" in decl:
        return "ptr"
    ty = decl.split()[0] if is_return or len(decl.split()) == 1 else " ".join(decl.split()[:-1])
    ty = ty.strip()
    if ty in ("accblas_stream_t", "accblas_mg_comm_t", "accblas_mg_plan_t"):
        return "ptr"
    return {"int": "int", "int64_t": "i64", "size_t": "size", "uint64_t": "u64", "void": "void"}[ty]


def hs_class(ty):
    ty = ty.strip()
    if ty != "()" and ty.startswith("(") and ty.endswith(")"):
        ty = ty[1:-1].strip()
    if ty.startswith("Ptr") or ty in ("Comm", "Plan", "CString"):
        return "ptr"
    return {"CInt": "int", "Status": "int", "Int64": "i64", "CSize": "size", "Word64": "u64", "CULLong": "u64",
            "()": "void"}[ty]


def hs_imports():
    src = open(os.path.join(HS, "FFI.hs")).read()
    src = re.sub(r"--.*$", "", src, flags=re.M)
    out = {}
    pat = re.compile(r'foreign import ccall\s+(safe|unsafe)\s+"(accblas_\w+)"\s+(\w+)\s*::(.*?)(?=\n\s*\n|\nforeign import|\Z)', re.S)
    for safety, cname, hsname, sig in pat.findall(src):
        parts = [p.strip() for p in sig.replace("\n", " ").split("->")]
        assert parts[-1].startswith("IO "), (cname, parts[-1])
        out[cname] = (safety, hsname, hs_class(parts[-1][3:]), [hs_class(p) for p in parts[:-1]])
    return out


def test_every_foreign_import_matches_its_prototype():
    protos, imports = c_prototypes(), hs_imports()
    assert len(imports) >= 25, "the parser lost the foreign imports"
    for cname, (safety, hsname, ret, args) in imports.items():
        assert cname in protos, f"{hsname} imports {cname}, which include/accblas_b200.h does not declare"
        cret, cargs = protos[cname]
        assert hsname == "c_" + cname
        assert len(args) == len(cargs), f"{cname}: header has {len(cargs)} parameters, FFI.hs {len(args)}"
        for i, (h, c) in enumerate(zip(args, cargs)):
            assert h == c, f"{cname}: parameter {i + 1} is {c} in the header, {h} in FFI.hs"
        assert ret == cret, f"{cname}: returns {cret} in the header, IO {ret} in FFI.hs"
        if cname in BLOCKING:
            assert safety == "safe", f"{cname} blocks on the other ranks: import it `safe`"


def test_the_hot_path_is_imported_and_exported():
    imports = hs_imports()
    for need in ("accblas_gemm", "accblas_gemv", "accblas_dot", "accblas_gemm_alpha_dev", "accblas_gemv_alpha_dev",
                 "accblas_dot_static", "accblas_sdot", "accblas_asum", "accblas_amax", "accblas_amin"):
        assert need in imports, need
    src = open(os.path.join(HS, "FFI.hs")).read()
    export_list = src[src.index("module "):src.index(") where")]
    for _, hsname, _, _ in imports.values():
        used_inside = len(re.findall(r"\b%s\b" % hsname, src)) > 1          # e.g. checkStatus uses c_accblas_status_string
        assert re.search(r"\b%s\b" % hsname, export_list) or used_inside, f"{hsname} is imported but neither exported nor used"


def call_sites(path):
    """(name, argument count) of every saturated application `c_accblas_x a b (c d) ...` that starts a line."""
    lines = [re.sub(r"--.*$", "", ln).rstrip() for ln in open(path).read().split("\n")]
    out = []
    for i, ln in enumerate(lines):
        m = re.match(r"^(\s*)(c_accblas_\w+)\s+(.*)$", ln)
        if not m:
            continue
        indent, text = len(m.group(1)), m.group(3)
        j = i + 1
        while j < len(lines) and (not lines[j].strip() or len(lines[j]) - len(lines[j].lstrip()) > indent):
            text += " " + lines[j].strip()
            j += 1
        depth, count, in_tok = 0, 0, False
        for ch in text:
            if ch == "(":
                if depth == 0 and not in_tok:
                    count += 1
                depth += 1
            elif ch == ")":
                depth -= 1
                in_tok = False
            elif ch.isspace():
                if depth == 0:
                    in_tok = False
            elif depth == 0 and not in_tok:
                in_tok = True
                count += 1
        assert depth == 0, (path, i + 1)
        out.append((m.group(2)[2:], count, i + 1))
    return out


def test_call_sites_are_saturated():
    protos = c_prototypes()
    seen = set()
    for f in ("Hooks.hs", "Level1.hs"):
        for cname, count, line in call_sites(os.path.join(HS, f)):
            assert count == len(protos[cname][1]), f"{f}:{line}: {cname} applied to {count} arguments, header has {len(protos[cname][1])}"
            seen.add(cname)
    # the three hooks of the hot path and the device-alpha product (LLVM/PTX/Level3.hs:88-106, Level2.hs:105-124, Level1.hs:73-92)
    assert {"accblas_gemm", "accblas_gemv", "accblas_dot", "accblas_gemm_alpha_dev"} <= seen


def test_the_check_notices_a_changed_prototype(tmp_path, monkeypatch):
    """The point of the file: an edited header prototype must turn the comparison red."""
    import sys
    mod = sys.modules[__name__]
    edited = open(HEADER).read().replace("int accblas_dot(accblas_stream_t stream, int dtype, int conj_x, int64_t n,",
                                         "int accblas_dot(accblas_stream_t stream, int dtype, int64_t n,")
    assert edited != open(HEADER).read()
    p = tmp_path / "accblas_b200.h"
    p.write_text(edited)
    monkeypatch.setattr(mod, "HEADER", str(p))
    try:
        test_every_foreign_import_matches_its_prototype()
    except AssertionError as e:
        assert "accblas_dot" in str(e)
    else:
        raise AssertionError("a dropped parameter went unnoticed")
