"""Repository rules that can be checked without a GPU: the product package never imports the CPU oracle, and it has
no CPU path -- without a CUDA device its entry points raise instead of computing something else."""
import ast
import os

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "jumbo_mae_tpu_b200")


def _imports(path):
    tree = ast.parse(open(path).read())
    for node in ast.walk(tree):
        if isinstance(node, ast.Import):
            for a in node.names:
                yield a.name
        elif isinstance(node, ast.ImportFrom):
            yield ("." * node.level) + (node.module or "")


def test_product_package_never_imports_the_oracle():
    for name in sorted(os.listdir(PKG)):
        if name.endswith(".py"):
            for mod in _imports(os.path.join(PKG, name)):
                assert not mod.lstrip(".").startswith("oracle"), f"{name} imports {mod}"
    for name in sorted(os.listdir(os.path.join(PKG, "csrc"))):
        src = open(os.path.join(PKG, "csrc", name)).read()
        assert "oracle/" not in src.replace("the oracle", ""), name


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the behaviour WITHOUT a CUDA device")
def test_entry_points_fail_loudly_without_a_gpu():
    import argparse
    from jumbo_mae_tpu_b200 import _lib
    from jumbo_mae_tpu_b200.finetuning import create_train_state as ft_create
    from jumbo_mae_tpu_b200.pretraining import create_train_state as pt_create
    with pytest.raises(_lib.JmError, match="no CUDA device"):
        _lib.require_device()
    for create in (pt_create, ft_create):
        with pytest.raises(_lib.JmError):
            create(argparse.Namespace())
