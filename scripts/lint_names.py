#!/usr/bin/env python
"""Undefined-name check for code paths the CPU tests cannot reach (GPU-only branches): every name a scope reads as a
global must be bound at module level or be a builtin.  No third-party linter is installed in this image.

  python scripts/lint_names.py [paths...]      # exit status 1 if anything is reported
"""
from __future__ import annotations

import builtins
import pathlib
import symtable
import sys


def check(path: pathlib.Path) -> list:
    src = path.read_text()
    try:
        top = symtable.symtable(src, str(path), "exec")
    except SyntaxError as e:
        return [f"{path}:{e.lineno}: syntax error: {e.msg}"]
    bound = {s.get_name() for s in top.get_symbols() if s.is_assigned() or s.is_imported() or s.is_namespace()}
    star = "import *" in src
    out = []

    def walk(tab):
        for s in tab.get_symbols():
            n = s.get_name()
            if s.is_referenced() and s.is_global() and not s.is_assigned() and n not in bound and not hasattr(builtins, n) \
                    and n not in ("__file__", "__name__", "__doc__", "__spec__", "__path__", "__class__") and not star:
                out.append(f"{path}:{tab.get_lineno()}: '{n}' read in {tab.get_type()} '{tab.get_name()}' is never bound")
        for c in tab.get_children():
            walk(c)

    for c in top.get_children():
        walk(c)
    # module-level reads
    for s in top.get_symbols():
        n = s.get_name()
        if s.is_referenced() and not (s.is_assigned() or s.is_imported() or s.is_namespace()) and not hasattr(builtins, n) \
                and n not in ("__file__", "__name__", "__doc__", "__spec__", "__path__") and not star:
            out.append(f"{path}: module reads '{n}' which is never bound")
    return out


def main() -> int:
    roots = [pathlib.Path(p) for p in (sys.argv[1:] or ["trustworthy_dl", "bench", "bench.py", "__graft_entry__.py", "baseline",
                                                         "experiments", "tests", "scripts"])]
    files = []
    for r in roots:
        files += [r] if r.is_file() else sorted(p for p in r.rglob("*.py") if "_ref" not in p.parts)
    problems = [m for f in files for m in check(f)]
    print("\n".join(problems) if problems else f"{len(files)} files: no unbound names")
    return 1 if problems else 0


if __name__ == "__main__":
    raise SystemExit(main())
