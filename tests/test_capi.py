"""The C-ABI boundary: libmrts_b200.so loads without a GPU and exports every function that
include/mrts_b200.h declares; argument validation is host-side and needs no device."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from microrts_py_b200 import _capi
from tests.common import MAP_8, MAP_16, templates_for

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "mrts_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(mrts_[a-z_0-9]+)\s*\(", src)))


def test_header_and_binding_agree():
    assert sorted(_capi.SYMBOLS) == _declared()


def test_library_exports_every_declared_symbol():
    _capi.build_library()
    lib = C.CDLL(_capi.LIB_PATH)
    for name in _declared():
        assert getattr(lib, name) is not None, name
    lib.mrts_version.restype = C.c_char_p
    assert b"sm_100a" in lib.mrts_version()


def test_library_is_sm100a_only():
    import subprocess

    out = subprocess.run(["cuobjdump", "-lelf", _capi.LIB_PATH], capture_output=True, text=True).stdout
    assert "sm_100a" in out and "sm_90" not in out


@pytest.mark.parametrize(
    "kw,msg",
    [
        (dict(nsp=1), "even"),
        (dict(nsp=0, nbot=0), "no environments"),
        (dict(max_steps=0), "max_steps"),
        (dict(elem=2), "elem_bytes"),
        (dict(mixed=True), "same size"),
    ],
)
def test_create_rejects_bad_arguments_without_a_device(kw, msg):
    """mrts_create validates before touching CUDA (JNIGridnetVecClient would throw in its constructor)."""
    L = _capi.lib()
    nsp, nbot = kw.get("nsp", 2), kw.get("nbot", 0)
    n = max(nsp + nbot, 1)
    paths = [MAP_16] * n
    if kw.get("mixed"):
        paths[-1] = MAP_8
    t = templates_for(paths)
    cfg, keep = _capi.make_config(nsp, nbot, 100, False, t[: nsp + nbot] if nsp + nbot else [], [0] * nbot, elem_bytes=kw.get("elem", 4)) if nsp + nbot else (None, None)
    if cfg is None:
        cfg, keep = _capi.make_config(2, 0, 100, False, templates_for([MAP_16] * 2), [])
        cfg.num_selfplay_envs = 0
    if "max_steps" in kw:
        cfg.max_steps = kw["max_steps"]
    h = C.c_void_p()
    rc = L.mrts_create(C.byref(cfg), C.byref(h))
    assert rc == -1 and not h.value
    assert msg in L.mrts_last_error().decode()


def test_no_cpu_fallback_in_product():
    """The product package must not reference the oracle or the host debug build."""
    pkg = os.path.join(ROOT, "microrts_py_b200")
    for d, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".h", ".cu", ".cpp")):
                s = open(os.path.join(d, f)).read()
                assert "oracle_py" not in s and "libmrts_oracle" not in s and "import oracle" not in s, f
                assert "from tests" not in s and "libmrts_hostsim" not in s, f


def test_env_class_needs_cuda():
    import torch

    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    from microrts_py_b200.envs import MicroRTSGridModeVecEnv

    with pytest.raises(RuntimeError, match="no CPU fallback"):
        MicroRTSGridModeVecEnv(2, 0, map_paths=[MAP_16])
