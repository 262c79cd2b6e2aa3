"""CPU tests: the C-ABI library builds for sm_100a, loads, exports every symbol the header declares,
and fails loudly (no CPU fallback) when there is no CUDA device."""
import ctypes as C
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols(name="tamcmc_gpu.h"):
    txt = open(os.path.join(ROOT, "include", name)).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(tg_[a-z0-9_]+)\s*\(", txt)))


def test_library_builds_and_exports_header_symbols(tb):
    path = tb.build_library()
    assert os.path.exists(path)
    lib = C.CDLL(path)
    syms = header_symbols()
    assert len(syms) >= 18
    for s in syms:
        assert hasattr(lib, s), "header declares %s but the library does not export it" % s
    from tamcmc_b200.gpu import EXPORTS, EXPORTS_OUT, EXPORTS_FE
    assert sorted(EXPORTS) == syms, "python binding and header disagree"
    # the host-code headers of the same library: outputs / restore formats and the input front end
    for hdr, exp in (("tamcmc_outputs.h", EXPORTS_OUT), ("tamcmc_frontend.h", EXPORTS_FE)):
        hs = header_symbols(hdr)
        for sname in hs:
            assert hasattr(lib, sname), "%s declares %s but the library does not export it" % (hdr, sname)
        assert sorted(exp) == hs, "python binding and %s disagree" % hdr


def test_library_is_sm100a_only(tb):
    import subprocess
    out = subprocess.run(["cuobjdump", "--list-elf", tb.build_library()], stdout=subprocess.PIPE, text=True).stdout
    archs = set(re.findall(r"sm_\d+a?", out))
    assert archs == {"sm_100a"}, archs


def test_no_cuda_device_fails_loudly(tb, W):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    prob = W.s_local(Nx=600)
    with pytest.raises(tb.TgError) as e:
        tb.GpuModelDef(prob, tb.MalaConfig(Nchains=2))
    assert "no CUDA device" in str(e.value) or "-4" in str(e.value)


def test_product_package_does_not_import_the_oracle():
    pkg = os.path.join(ROOT, "tamcmc-c-_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                txt = open(os.path.join(dirpath, f)).read()
                assert "import oracle" not in txt and "from oracle" not in txt and "liboracle" not in txt, os.path.join(dirpath, f)
