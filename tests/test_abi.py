"""CPU tests of the boundary: libb2bfs.so loads, exports every symbol include/b2bfs.h declares, and
fails loudly (no CPU fallback) when there is no CUDA device."""
import ctypes
import os
import re

import pytest
import torch


def test_library_exports_every_declared_symbol(b2):
    header = open(b2.HEADER_PATH).read()
    header = re.sub(r"/\*.*?\*/", "", header, flags=re.S)
    declared = set(re.findall(r"\b(b2bfs_[a-z0-9_]+)\s*\(", header))
    assert len(declared) >= 20
    L = ctypes.CDLL(b2.LIB_PATH)
    missing = [s for s in sorted(declared) if not hasattr(L, s)]
    assert not missing, missing
    assert b"sm_100a" in b2.lib().b2bfs_version()


def test_stats_struct_layout_matches_header(b2):
    assert ctypes.sizeof(b2.Stats) == 4 * 4 + 5 * 8 + 2 * 8


@pytest.mark.skipif(torch.cuda.is_available(), reason="only meaningful on a box without a GPU")
def test_no_gpu_means_loud_failure_not_fallback(b2):
    with pytest.raises(b2.B2BfsError) as e:
        b2.Bfs(0)
    assert e.value.code == -5 and "no CPU path" in str(e.value)


def test_cli_binary_exists_and_prints_usage(b2):
    import subprocess

    r = subprocess.run([b2.CLI_PATH], capture_output=True, text=True)
    assert r.returncode == 2 and "usage:" in r.stderr


def test_product_never_imports_the_oracle(b2):
    root = os.path.dirname(b2.HEADER_PATH) + "/.."
    pkg = os.path.join(root, "distributed-cuda-bfs_b200")
    for dirpath, _, files in os.walk(pkg):
        if any(part in dirpath for part in ("/build", "/lib", "/bin", "__pycache__")):
            continue
        for f in files:
            text = open(os.path.join(dirpath, f), errors="ignore").read()
            assert "import oracle" not in text and "oracle/" not in text and "liboracle" not in text, f
