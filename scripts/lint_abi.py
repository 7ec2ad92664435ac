#!/usr/bin/env python
"""Every ``_lib.call("tdl_x", ...)`` in the Python tree against the prototype of ``tdl_x`` in include/tdl_kernels.h:
the symbol must be declared (and exported by the built library when it is present) and the argument count must match.
ctypes does not check either.

  python scripts/lint_abi.py
"""
from __future__ import annotations

import ast
import pathlib
import re
import subprocess
import sys

ROOT = pathlib.Path(__file__).resolve().parent.parent


def prototypes() -> dict:
    text = (ROOT / "include" / "tdl_kernels.h").read_text()
    text = re.sub(r"/\*.*?\*/", " ", text, flags=re.S)
    text = re.sub(r"//[^\n]*", " ", text)
    out = {}
    for m in re.finditer(r"\b(tdl_\w+)\s*\(([^;{]*?)\)\s*;", text, flags=re.S):
        args = m.group(2).strip()
        out[m.group(1)] = [] if args in ("", "void") else [kind_of_c(a) for a in args.split(",")]
    return out


def kind_of_c(decl: str) -> str:
    d = " ".join(decl.split())
    if "*" in d or "cudaStream_t" in d:
        return "ptr"
    if re.search(r"\b(int64_t|uint64_t|size_t|long long)\b", d):
        return "i64"
    if re.search(r"\bdouble\b", d):
        return "f64"
    if re.search(r"\bfloat\b", d):
        return "f32"
    if re.search(r"\b(int|int32_t|uint32_t|unsigned)\b", d):
        return "i32"
    return "?"


def kind_of_py(node: ast.AST) -> str:
    """ptr / i64 / i32 / f32 / f64 for the ctypes wrappers of ops/_lib.py, '?' for anything else."""
    if isinstance(node, ast.IfExp):
        a, b = kind_of_py(node.body), kind_of_py(node.orelse)
        return a if a == b else "?"
    if isinstance(node, ast.Call):
        f = node.func
        name = f.id if isinstance(f, ast.Name) else (f.attr if isinstance(f, ast.Attribute) else "")
        if name in ("ptr", "stream_ptr", "c_void_p", "_p", "sptr"):
            return "ptr"
        if name in ("i64", "c_int64"):
            return "i64"
        if name in ("i32", "c_int"):
            return "i32"
        if name in ("f32", "c_float"):
            return "f32"
        if name in ("f64", "c_double"):
            return "f64"
    return "?"


def calls():
    for f in sorted(list((ROOT / "trustworthy_dl").rglob("*.py")) + list((ROOT / "bench").rglob("*.py")) +
                    list((ROOT / "tests").rglob("*.py")) + [ROOT / "bench.py", ROOT / "__graft_entry__.py"]):
        for node in ast.walk(ast.parse(f.read_text())):
            if isinstance(node, ast.Call) and isinstance(node.func, ast.Attribute) and node.func.attr == "call" \
                    and node.args and isinstance(node.args[0], ast.Constant) and isinstance(node.args[0].value, str) \
                    and node.args[0].value.startswith("tdl_"):
                starred = any(isinstance(a, ast.Starred) for a in node.args[1:])
                yield f.relative_to(ROOT), node.lineno, node.args[0].value, \
                    None if starred else [kind_of_py(a) for a in node.args[1:]]


def main() -> int:
    protos = prototypes()
    exported = None
    so = ROOT / "trustworthy_dl" / "_C" / "libtdl_kernels.so"
    if so.exists():
        nm = subprocess.run(["nm", "-D", "--defined-only", str(so)], capture_output=True, text=True).stdout
        exported = {ln.split()[-1] for ln in nm.splitlines() if ln.strip()}
    problems, unknown, seen = [], [], set()
    for f, line, name, n in calls():
        seen.add(name)
        if name not in protos:
            problems.append(f"{f}:{line}: {name} is not declared in include/tdl_kernels.h")
        elif n is not None and len(n) != len(protos[name]):
            problems.append(f"{f}:{line}: {name} called with {len(n)} arguments, declared with {len(protos[name])}")
        elif n is not None:
            for i, (a, b) in enumerate(zip(n, protos[name])):
                if a == "?":
                    unknown.append(f"{f}:{line}: {name} argument {i + 1} ({b}) is not a recognised ctypes wrapper")
                elif b != "?" and a != b:
                    problems.append(f"{f}:{line}: {name} argument {i + 1} passed as {a}, declared {b}")
        if exported is not None and name not in exported:
            problems.append(f"{f}:{line}: {name} is not exported by {so.name}")
    if exported is not None:
        for name in sorted(protos):
            if name not in exported:
                problems.append(f"include/tdl_kernels.h: {name} is declared but not exported by {so.name}")
    if "-v" in sys.argv:
        print("\n".join(unknown))
    print("\n".join(problems) if problems else
          f"{len(seen)} entry points called from Python, {len(protos)} declared"
          + (f", {len([e for e in exported if e.startswith('tdl_')])} exported" if exported is not None else "") + ": consistent")
    return 1 if problems else 0


if __name__ == "__main__":
    raise SystemExit(main())
