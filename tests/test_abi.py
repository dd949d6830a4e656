"""CPU tests of the drop-in boundary: the C-ABI library loads, exports every symbol include/pwc_b200.h declares,
the ctypes table covers them, the product never imports the oracle, and there is no CPU fallback."""
import ctypes
import os
import re
import subprocess
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    names = set()
    for fn in os.listdir(os.path.join(ROOT, "include")):
        txt = open(os.path.join(ROOT, "include", fn)).read()
        txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
        names |= set(re.findall(r"\b(pwc_[a-z0-9_]+)\s*\(", txt))
    return names


def test_library_exports_every_declared_symbol():
    from opticalflow_b200 import _lib
    lib = _lib.load()
    decl = _declared()
    assert len(decl) >= 10
    raw = ctypes.CDLL(_lib.LIB_PATH)
    for name in decl:
        assert hasattr(raw, name), f"{name} declared in include/ but not exported"
    assert decl == set(_lib.SIGNATURES), (decl ^ set(_lib.SIGNATURES))
    assert lib.pwc_b200_arch() == 100
    assert b"sm_100a" in lib.pwc_b200_version()


def test_sass_is_sm100a_only():
    from opticalflow_b200 import _lib
    _lib.load()
    cuobjdump = "/usr/local/cuda/bin/cuobjdump"
    if not os.path.exists(cuobjdump):
        pytest.skip("cuobjdump not present")
    out = subprocess.run([cuobjdump, "-lelf", _lib.LIB_PATH], capture_output=True, text=True).stdout
    archs = set(re.findall(r"sm_(\d+a?)", out))
    assert archs == {"100a"}, archs


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "opticalflow_b200")
    for dp, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h")):
                txt = open(os.path.join(dp, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", txt, flags=re.M), f"{f} imports oracle"
                # the reference tree is cited in comments / docstrings only: nothing may open, import or list it at run time
                for ln in txt.splitlines():
                    if "/root/reference" in ln:
                        assert not re.search(r"\b(open|listdir|sys\.path|import_module|exists|isdir|isfile|glob)\b", ln), \
                            f"{f}: touches /root/reference at run time: {ln.strip()}"


def test_no_cpu_fallback():
    from opticalflow_b200.models.PWCNet import PWCDCNet
    from opticalflow_b200.models.correlation_package.correlation import Correlation
    m = PWCDCNet().eval()
    with pytest.raises(RuntimeError):
        m(torch.zeros(1, 6, 64, 64))
    c = Correlation(pad_size=4, kernel_size=1, max_displacement=4, stride1=1, stride2=1, corr_multiply=1)
    with pytest.raises(RuntimeError):
        c(torch.zeros(1, 8, 8, 8), torch.zeros(1, 8, 8, 8))


def test_state_dict_layout_and_init_match_reference_golden(golden_dir):
    import hashlib
    import numpy as np
    from opticalflow_b200.models.PWCNet import PWCDCNet
    g = np.load(os.path.join(golden_dir, "rand_64x128.npz"))
    torch.manual_seed(0)
    sd = PWCDCNet().state_dict()
    h = hashlib.sha256()
    for k in sd:
        h.update(k.encode()); h.update(sd[k].contiguous().numpy().tobytes())
    assert h.hexdigest() == str(g["sd_sha256"])     # same keys, order, shapes and random init as the reference
    assert len(sd) == 128 and sum(v.numel() for v in sd.values()) == 9374340


def test_concat_buffer_channel_maps():
    from opticalflow_b200 import engine
    for lvl in (6, 5, 4, 3, 2):
        wd = engine.concat_width(lvl)
        assert wd % 32 == 0
        for start in engine.DEC_START + [0]:
            cm = engine.chan_map(lvl, start)
            assert len(set(cm)) == len(cm) and 0 <= min(cm) and max(cm) < wd - start
        n_in = 81 + (0 if lvl == 6 else engine.PYR_CH[lvl] + 4)
        assert len(engine.chan_map(lvl, 0)) == 448 + n_in       # PWCNet.py:83,93,103,113,123
