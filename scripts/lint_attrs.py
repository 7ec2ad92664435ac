#!/usr/bin/env python
"""``self.<attr>`` reads that no method of the class (or of a base class in the same package) ever binds.  Complements
lint_names.py for GPU-only branches that the CPU tests cannot execute.

  python scripts/lint_attrs.py [paths...]
"""
from __future__ import annotations

import ast
import pathlib
import sys


def class_info(tree):
    info = {}
    for node in ast.walk(tree):
        if not isinstance(node, ast.ClassDef):
            continue
        bound, reads = set(), []
        for item in node.body:
            if isinstance(item, (ast.FunctionDef, ast.AsyncFunctionDef, ast.ClassDef)):
                bound.add(item.name)
            elif isinstance(item, (ast.Assign, ast.AnnAssign, ast.AugAssign)):
                for t in (item.targets if isinstance(item, ast.Assign) else [item.target]):
                    for n in ast.walk(t):
                        if isinstance(n, ast.Name):
                            bound.add(n.id)
        for sub in ast.walk(node):
            if isinstance(sub, ast.Attribute) and isinstance(sub.value, ast.Name) and sub.value.id == "self":
                if isinstance(sub.ctx, (ast.Store, ast.Del)):
                    bound.add(sub.attr)
                else:
                    reads.append((sub.attr, sub.lineno))
            # setattr(self, "name", ...) / for name in (...): setattr(self, name, ...)
            if isinstance(sub, ast.Call) and isinstance(sub.func, ast.Name) and sub.func.id == "setattr":
                bound.add("*dynamic*")
        bases = [b.id if isinstance(b, ast.Name) else (b.attr if isinstance(b, ast.Attribute) else None) for b in node.bases]
        info[node.name] = (bound, reads, bases, any(isinstance(d, ast.Name) and d.id == "dataclass" or
                                                    isinstance(d, ast.Call) and getattr(d.func, "id", "") == "dataclass"
                                                    for d in node.decorator_list))
    return info


def main() -> int:
    roots = [pathlib.Path(p) for p in (sys.argv[1:] or ["trustworthy_dl", "bench", "bench.py", "baseline"])]
    files = []
    for r in roots:
        files += [r] if r.is_file() else sorted(p for p in r.rglob("*.py") if "_ref" not in p.parts)
    per_file = {f: class_info(ast.parse(f.read_text())) for f in files}
    every = {}
    for f, info in per_file.items():
        for name, v in info.items():
            every.setdefault(name, []).append(v)
    problems = []
    for f, info in per_file.items():
        for name, (bound, reads, bases, _) in info.items():
            known, todo, external = set(bound), list(bases), False
            while todo:
                b = todo.pop()
                if b in every:
                    for bb, _, b2, _ in every[b]:
                        known |= bb
                        todo += b2
                elif b not in (None, "object"):
                    external = True           # torch.nn.Module, Enum, ...: attributes we cannot see
            if external or "*dynamic*" in known or "__getattr__" in known:
                continue
            for attr, line in reads:
                if attr not in known and not (attr.startswith("__") and attr.endswith("__")):
                    problems.append(f"{f}:{line}: {name}.{attr} is read but never bound")
    print("\n".join(sorted(set(problems))) if problems else f"{len(files)} files: every self.<attr> read has a binding")
    return 1 if problems else 0


if __name__ == "__main__":
    raise SystemExit(main())
