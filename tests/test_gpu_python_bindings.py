"""f4: the reference's Python bindings, built unmodified from /root/reference/python by `make -C oracle py`, pick
the plugin up through NVIMGCODEC_EXTENSIONS_PATH; tools/py_bindings_check.py mirrors the decode part of the reference's
test/python/test_python.py and compares every image with the CPU backend of the same Python API, bit for bit."""
import os
import subprocess
import sys

import pytest

import util

PY = os.path.join(util.ROOT, "oracle", "_ref", "py", "nvidia", "nvimgcodec")
pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not os.path.isdir(PY), reason="oracle/_ref/py not built (make -C oracle py)")]


@pytest.mark.timeout(600)
def test_reference_python_layer_picks_the_plugin_up():
    r = subprocess.run([sys.executable, os.path.join(util.ROOT, "tools", "py_bindings_check.py")], capture_output=True, text=True, timeout=550)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-5000:]
    assert "python bindings OK" in r.stdout
