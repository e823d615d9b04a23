"""The oracle is test infrastructure: nothing under legged_tracking_b200/ (the product) may import, load or execute it, and the
product must fail loudly -- not fall back -- when the CUDA library is missing."""
import ast
import glob
import os

import pytest

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(REPO, "legged_tracking_b200")


def test_product_never_imports_the_oracle_or_the_tests():
    for path in glob.glob(os.path.join(PKG, "**", "*.py"), recursive=True):
        tree = ast.parse(open(path).read(), path)
        for node in ast.walk(tree):
            names = []
            if isinstance(node, ast.Import):
                names = [a.name for a in node.names]
            elif isinstance(node, ast.ImportFrom):
                names = [node.module or ""]
            for n in names:
                assert not n.split(".")[0] in ("oracle", "tests"), f"{path} imports {n}"
    for path in glob.glob(os.path.join(PKG, "csrc", "*")):
        for line in open(path):
            if line.lstrip().startswith("#include"):
                assert "oracle" not in line and "tests/" not in line, f"{path}: {line.strip()}"


def test_missing_library_is_an_error_not_a_fallback(tmp_path):
    from legged_tracking_b200 import _abi
    with pytest.raises(_abi.LtError, match="no CPU fallback"):
        saved = _abi._lib
        _abi._lib = None
        try:
            _abi.load(str(tmp_path / "absent.so"))
        finally:
            _abi._lib = saved
