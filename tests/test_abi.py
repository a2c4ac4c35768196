"""CPU-side checks of the boundary: the shared library loads, exports every symbol that
include/honk_b200.h declares, and fails loudly (no CPU fallback) without a GPU."""
import ctypes
import os
import re

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "honk_b200.h")).read()
    return sorted(set(re.findall(r"HONK_API[^;(]*?\b(honk_\w+)\s*\(", src)))


def test_header_declares_expected_surface():
    syms = declared_symbols()
    assert len(syms) >= 18
    for s in ("honk_frontend_mfcc", "honk_model_create", "honk_model_forward", "honk_model_forward_pcm",
              "honk_last_error", "honk_label_from_logits"):
        assert s in syms


def test_library_exports_every_declared_symbol(built_lib):
    lib = ctypes.CDLL(built_lib)
    for s in declared_symbols():
        assert hasattr(lib, s), "missing export " + s
    assert lib.honk_abi_version() == 2


def test_python_binding_covers_header(built_lib):
    from honk_b200 import _native
    assert sorted(_native.EXPORTS) == declared_symbols()
    _native.lib()


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU failure mode")
def test_no_cpu_fallback(built_lib):
    from honk_b200 import _native, AudioPreprocessor, find_config, find_model
    assert _native.lib().honk_device_check(0) == -3
    assert b"no CPU fallback" in _native.lib().honk_last_error()
    with pytest.raises(RuntimeError):
        AudioPreprocessor()
    m = find_model("res8-narrow")(find_config("res8-narrow")).eval()
    with pytest.raises(RuntimeError):
        m(torch.zeros(1, 101, 40))


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "honk_b200")
    for fn in os.listdir(pkg):
        if fn.endswith(".py"):
            assert "oracle" not in open(os.path.join(pkg, fn)).read().replace("no oracle", ""), fn


def test_api_surface_matches_reference_names():
    import honk_b200 as hb
    assert [c.value for c in hb.ConfigType] == [
        "cnn-trad-pool2", "cnn-one-stride1", "cnn-one-fpool3", "cnn-one-fstride4", "cnn-one-fstride8",
        "cnn-tpool2", "cnn-tpool3", "cnn-tstride2", "cnn-tstride4", "cnn-tstride8",
        "res15", "res26", "res8", "res15-narrow", "res8-narrow", "res26-narrow"]
    assert hb.find_model(hb.ConfigType.RES15) is hb.SpeechResModel
    assert hb.find_model("cnn-tpool2") is hb.SpeechModel
    assert hb.find_config("res15") is hb.find_config(hb.ConfigType.RES15)      # shared mutable dict
    with pytest.raises(KeyError):
        hb.find_config("nope")
    from honk_b200 import synth
    for ct in hb.ConfigType:
        cfg = dict(hb.find_config(ct))
        sd = hb.find_model(ct)(cfg).state_dict()
        shapes = synth.state_dict_shapes(ct.value, cfg)
        assert list(sd.keys()) == list(shapes.keys())
        assert all(tuple(sd[k].shape) == tuple(shapes[k]) for k in sd)


def test_every_package_module_imports():
    import importlib
    pkg = os.path.join(ROOT, "honk_b200")
    for fn in sorted(os.listdir(pkg)):
        if fn.endswith(".py") and fn != "build.py":
            importlib.import_module("honk_b200." + fn[:-3] if fn != "__init__.py" else "honk_b200")
