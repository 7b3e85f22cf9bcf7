"""Loader of the staged reference (``oracle/_ref/``, made by ``oracle/make_ref.py``).

``load()`` returns the reference's own ``speechpy`` modules (processing, feature, functions)
imported under the private package name ``spb_ref_speechpy`` -- so they never collide with the
drop-in's ``speechpy`` -- or ``None`` when ``oracle/_ref`` has not been staged.  The only change to
the environment is ``np.lib.pad = np.pad`` (numpy >= 2 dropped the alias the reference calls at
processing.py:223,301,314).

Test infrastructure only (tests/, bench.py's CPU legs).
"""
from __future__ import annotations

import importlib
import os
import sys
import types

HERE = os.path.dirname(os.path.abspath(__file__))
REF_DIR = os.path.join(HERE, "_ref")
PKG = "spb_ref_speechpy"


def available() -> bool:
    return all(os.path.exists(os.path.join(REF_DIR, "speechpy", f + ".py")) for f in ("processing", "feature", "functions"))


def load():
    if not available():
        return None
    if PKG in sys.modules:
        return sys.modules[PKG]
    import numpy as np
    if not hasattr(np.lib, "pad"):
        np.lib.pad = np.pad
    prev = sys.dont_write_bytecode
    sys.dont_write_bytecode = True
    try:
        pkg = types.ModuleType(PKG)
        pkg.__path__ = [os.path.join(REF_DIR, "speechpy")]
        sys.modules[PKG] = pkg
        for sub in ("processing", "functions", "feature"):
            setattr(pkg, sub, importlib.import_module(f"{PKG}.{sub}"))
    except Exception:
        sys.modules.pop(PKG, None)
        raise
    finally:
        sys.dont_write_bytecode = prev
    return pkg
