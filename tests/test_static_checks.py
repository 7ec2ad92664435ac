"""Static consistency of the paths the CPU suite cannot execute (GPU-only branches): unbound names, ``self.<attr>`` reads
without a binding, and every ``_lib.call("tdl_x", ...)`` against the prototype in include/tdl_kernels.h (ctypes checks
neither the symbol's arity nor the argument kinds)."""
import pathlib
import subprocess
import sys

import pytest

ROOT = pathlib.Path(__file__).resolve().parent.parent


@pytest.mark.parametrize("script", ["lint_names.py", "lint_attrs.py", "lint_abi.py"])
def test_static_check(script):
    r = subprocess.run([sys.executable, str(ROOT / "scripts" / script)], cwd=ROOT, capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
