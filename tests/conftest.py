import glob
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def golden_names():
    return sorted(os.path.splitext(os.path.basename(p))[0] for p in glob.glob(os.path.join(GOLDEN, "*.npz")))


def load_golden(name):
    d = dict(np.load(os.path.join(GOLDEN, name + ".npz")))
    d["name"] = name
    d["S"] = int(d["S"])
    d["order"] = int(d["order"])
    nbd = int(d["normalize_by_dim"])
    d["normalize_by_dim"] = None if nbd < 0 else nbd
    d["use_den"] = bool(int(d["use_den"]))
    d["args"] = {str(kv).split("=")[0]: bool(int(str(kv).split("=")[1])) for kv in d["args"].tolist()}
    d["num_mats"] = [d["num%d" % i] for i in range(8)]
    d["den_mats"] = [d["den%d" % i] for i in range(8)]
    tl = d["text_lens"]
    syms, seqs, o = d["symbols"], [], 0
    for n in tl:
        seqs.append(syms[o:o + int(n)].tolist())
        o += int(n)
    d["label_seqs"] = seqs
    return d


@pytest.fixture(scope="session")
def lib_built():
    import __graft_entry__ as ge
    ge.build()
    import pytorch_asr_b200 as P
    return P
