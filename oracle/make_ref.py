#!/usr/bin/env python
"""Recipe for ``oracle/_ref/``: the reference's OWN speechpy path, staged so that it travels to the
GPU box (``/root/reference`` does not exist there).

    python oracle/make_ref.py            # in the build container, where /root/reference exists

The reference's hot path is three pure-Python files with nothing to compile, so the "build" is a
verbatim copy, read from where they lie under ``/root/reference``:

    speech_feature_extraction/speechpy/{processing,feature,functions}.py  -> oracle/_ref/speechpy/
    speech_feature_extraction/tests/test_speechpy.py                      -> oracle/_ref/tests/

``oracle/_ref/`` is git-ignored (reference sources never enter the history) but not gpurun-ignored.
A manifest with the sha256 of every staged file is written beside them.  Nothing is edited: the one
shim the files need under numpy >= 2 (``np.lib.pad = np.pad``, processing.py:223,301,314) is applied
by the loader ``oracle/ref_loader.py`` at import time.

Test infrastructure only: used by ``tests/`` (the reference's test file run against the drop-in,
reference == oracle port) and by ``bench.py``'s CPU legs (``cpu_baseline.kind`` = "reference").
"""
from __future__ import annotations

import hashlib
import json
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "_ref")
REF = os.environ.get("SPB_REFERENCE", "/root/reference")
FILES = [
    ("speech_feature_extraction/speechpy/processing.py", "speechpy/processing.py"),
    ("speech_feature_extraction/speechpy/feature.py", "speechpy/feature.py"),
    ("speech_feature_extraction/speechpy/functions.py", "speechpy/functions.py"),
    ("speech_feature_extraction/tests/test_speechpy.py", "tests/test_speechpy.py"),
]


def make(verbose: bool = False) -> str | None:
    """Stage the files; returns the directory, or None when the reference is not present."""
    if not os.path.isdir(REF):
        return None
    manifest = {"source": REF, "files": {}}
    for src, dst in FILES:
        s, d = os.path.join(REF, src), os.path.join(OUT, dst)
        os.makedirs(os.path.dirname(d), exist_ok=True)
        if os.path.exists(d):
            os.chmod(d, 0o644)
        shutil.copyfile(s, d)
        os.chmod(d, 0o644)
        with open(d, "rb") as fh:
            manifest["files"][dst] = {"from": src, "sha256": hashlib.sha256(fh.read()).hexdigest()}
        if verbose:
            print(f"{s} -> {d}")
    with open(os.path.join(OUT, "MANIFEST.json"), "w") as fh:
        json.dump(manifest, fh, indent=1)
    return OUT


if __name__ == "__main__":
    out = make(verbose=True)
    if out is None:
        print(f"{REF} is not present: nothing staged", file=sys.stderr)
        sys.exit(1)
    print(out)
