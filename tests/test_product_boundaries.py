"""Guards of the tier's rules that can be checked without a GPU: the product never touches ``oracle/`` (test
infrastructure; only tests, ``__graft_entry__.smoke`` and ``bench.py``'s CPU-baseline leg may), and there is no
CPU fallback -- every compute entry point refuses CPU tensors instead of computing something else."""
import ast
import os
import re

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "pytorch-asr_b200")


def _imports(path):
    tree = ast.parse(open(path).read())
    for node in ast.walk(tree):
        if isinstance(node, ast.Import):
            for a in node.names:
                yield a.name
        elif isinstance(node, ast.ImportFrom):
            yield ("." * node.level) + (node.module or "")


def test_product_package_never_imports_the_oracle():
    offenders = []
    for base, _, files in os.walk(PKG):
        for f in files:
            if f.endswith(".py"):
                for name in _imports(os.path.join(base, f)):
                    if name.split(".")[0] == "oracle" or "/reference" in name:
                        offenders.append((f, name))
            if f.endswith((".cu", ".cpp", ".cuh", ".h")):
                text = open(os.path.join(base, f)).read()
                if re.search(r"#include\s+[<\"][^>\"]*oracle", text):
                    offenders.append((f, "#include oracle"))
    assert offenders == []
    # bench.py: only inside cpu_path_time (the pass that cpu_baseline and --impl reference time); __graft_entry__: only inside smoke()
    for fname, allowed in (("bench.py", {"cpu_path_time"}), ("__graft_entry__.py", {"smoke"})):
        tree = ast.parse(open(os.path.join(ROOT, fname)).read())
        for node in ast.walk(tree):
            if isinstance(node, (ast.FunctionDef,)):
                uses = [n for n in ast.walk(node) if isinstance(n, ast.ImportFrom) and (n.module or "").split(".")[0] == "oracle"]
                if uses:
                    assert node.name in allowed, (fname, node.name)
        top = [n for n in tree.body if isinstance(n, (ast.Import, ast.ImportFrom)) and
               "oracle" in (getattr(n, "module", None) or "") + "".join(a.name for a in n.names)]
        assert top == [], fname


def test_no_cpu_fallback(lib_built):
    P = lib_built
    gen = P.CTCGraphGen(context_order=2, graph_build_args={}, num_symbols=5)
    num, den = gen.get_training_graphs([[1, 2], [3]]), gen.get_decoding_graph()
    x = torch.zeros(6, 2, 25)
    with pytest.raises(RuntimeError):
        P.fst_loss(x, [6, 4], num, den)
    with pytest.raises(RuntimeError):
        P._lib.path_sum(x, [6, 4], den)
    with pytest.raises(RuntimeError):
        P._lib.viterbi(x, [6, 4], den)
    with pytest.raises(RuntimeError):
        P.path_reduction(x, torch.tensor([6, 4]), gen.get_decoding_matrices())
    with pytest.raises(RuntimeError):
        P.postprocess_logprobs(x, block_normalize=True)
    if not torch.cuda.is_available():
        with pytest.raises(RuntimeError):
            P.upload_logits(x, [6, 4])
