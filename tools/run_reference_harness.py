#!/usr/bin/env python
"""Runs the reference's OWN, unmodified run_test.py / tester.py / lab*_processor.py against
this repository's binaries (run from the repository root, same arguments as run_test.py).

The reference harness imports matplotlib at module top (not installed here) and resolves its
processors relative to its own directory, so this launcher (a) installs a stub `matplotlib`
module, (b) puts the reference tree first on sys.path, (c) runs its run_test.py as __main__.
Note: the reference harness writes converted images next to the inputs in ./lab2/data etc.
and crashes on lab1 (SURVEY §0 defect 1) — that is its behaviour, not ours.
"""
import os
import runpy
import sys
import types
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
CANDIDATES = [ROOT / "baseline" / "_ref" / "src", Path("/root/reference")]


def main():
    ref = next((p for p in CANDIDATES if (p / "run_test.py").exists()), None)
    if ref is None:
        raise SystemExit("reference tree not found (run baseline/install_reference.sh)")
    try:
        import matplotlib  # noqa: F401
    except ImportError:
        mpl, plt = types.ModuleType("matplotlib"), types.ModuleType("matplotlib.pyplot")

        class _Anything:
            def __getattr__(self, _):
                return self

            def __call__(self, *a, **k):
                return self

            def __iter__(self):
                return iter(())

        any_ = _Anything()
        plt.subplots = lambda *a, **k: (any_, any_)
        for name in ("text", "tight_layout", "savefig", "close"):
            setattr(plt, name, lambda *a, **k: None)
        mpl.pyplot = plt
        sys.modules["matplotlib"], sys.modules["matplotlib.pyplot"] = mpl, plt
    sys.path.insert(0, str(ref))
    for shadow in ("tester", "arg_parsing", "utils", "lab1", "lab2", "lab3"):
        sys.modules.pop(shadow, None)
    os.chdir(ROOT)
    sys.argv = [str(ref / "run_test.py")] + sys.argv[1:]
    runpy.run_path(str(ref / "run_test.py"), run_name="__main__")


if __name__ == "__main__":
    main()
