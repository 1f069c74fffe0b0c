"""Import the reference's own hot-path modules from ``/root/reference``.

TEST INFRASTRUCTURE ONLY, and usable ONLY in the build container:
``/root/reference`` does not exist on the GPU box, so nothing that runs there
(``-m gpu`` tests, ``smoke()``, ``bench.py``) may call into this module.  It is
used by ``oracle/make_golden.py`` (fixture generation) and by the CPU tests
that pin ``oracle/engine.py`` / ``oracle/graphs.py`` against the reference
when it is present (they skip otherwise).

The reference targets python2.7 / torch 1.2 and imports three packages that
are absent here; they are replaced by stubs before import:

* ``pywrapfst``  -> ``oracle.minifst`` (restated OpenFst slice)
* ``torchtext``  -> empty stub (only ``egs/vocab.py`` touches it, off-path)
* ``kaldi_io``   -> empty stub (feature IO, off-path)
"""
from __future__ import annotations

import os
import sys
import types

REFERENCE_ROOT = os.environ.get("CTCG_REFERENCE_ROOT", "/root/reference")

_cache = {}


def available():
    return os.path.isfile(os.path.join(REFERENCE_ROOT, "att_speech", "fst_utils.py"))


def _install_stubs():
    from . import minifst
    sys.modules.setdefault("pywrapfst", minifst)
    if "torchtext" not in sys.modules:
        tt = types.ModuleType("torchtext")
        tt.vocab = types.ModuleType("torchtext.vocab")
        sys.modules["torchtext"] = tt
        sys.modules["torchtext.vocab"] = tt.vocab
    sys.modules.setdefault("kaldi_io", types.ModuleType("kaldi_io"))


def load():
    """Return ``(fst_utils, ctc_losses)`` modules of the reference, unmodified."""
    if "mods" in _cache:
        return _cache["mods"]
    if not available():
        raise RuntimeError("reference tree not present at %s" % REFERENCE_ROOT)
    _install_stubs()
    if REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, REFERENCE_ROOT)
    import numpy as np
    if not hasattr(np, "int"):  # py2-era alias used by egs/wsj/data.py
        np.int = int
    import importlib
    fst_utils = importlib.import_module("att_speech.fst_utils")
    ctc_losses = importlib.import_module("att_speech.modules.ctc_losses")
    _cache["mods"] = (fst_utils, ctc_losses)
    return _cache["mods"]
