#!/usr/bin/env python
"""Generate the committed golden vectors.  Run ONLY in the build container.

It imports the reference's own ``utils/model.py`` from ``/root/reference`` (read-only,
not present on the GPU box) with stub modules for its four uninstallable imports
(chainmap, librosa, pcen, pyaudio - SURVEY.md §8c), loads seeded weights into the
reference classes with a strict ``load_state_dict`` (which also proves the key names
and shapes of ``honk_b200.synth.state_dict_shapes``), runs the reference forward on
CPU in float32 and float64, and freezes:

* ``cnn_golden.npz``    logits of all 16 architectures on 6 synthetic clips, the
                        calibrated BN statistics, and each config dict (JSON);
* ``mfcc_golden.npz``   oracle MFCCs for the edge-case clips and 8 random clips
                        (front-end parity is unpinned by the reference - see
                        oracle/honk_oracle.py - these freeze the restatement).

Usage:  python tests/golden/make_golden.py
"""
import collections
import json
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from honk_b200 import synth  # noqa: E402
from oracle import honk_oracle as orc  # noqa: E402


def import_reference_model():
    ref = "/root/reference"
    if not os.path.isdir(ref):
        raise SystemExit("reference tree not present; goldens can only be regenerated in the build container")
    chainmap = types.ModuleType("chainmap")
    chainmap.ChainMap = collections.ChainMap
    librosa = types.ModuleType("librosa")
    librosa.filters = types.SimpleNamespace(dct=orc.dct_filters)
    librosa.core = types.SimpleNamespace()
    librosa.feature = types.SimpleNamespace()
    pcen = types.ModuleType("pcen")
    pcen.StreamingPCENTransform = lambda **kw: None
    pyaudio = types.ModuleType("pyaudio")
    pyaudio.paInt16 = 8
    for name, m in (("chainmap", chainmap), ("librosa", librosa), ("pcen", pcen), ("pyaudio", pyaudio)):
        sys.modules.setdefault(name, m)
    sys.path.insert(0, ref)
    import utils.model as refmod  # noqa
    return refmod


def main():
    refmod = import_reference_model()
    torch.manual_seed(0)
    torch.set_num_threads(8)
    pre = orc.AudioPreprocessorOracle()

    # ---- front-end goldens -------------------------------------------------
    mf = {}
    specials = synth.special_clips()
    for k, pcm in specials.items():
        mf["special/" + k] = pre.compute_mfccs(orc.pcm16_to_float(pcm))[:, :, 0]
    rnd = synth.make_clips(8, start=0, seed=1234)
    mf["random_pcm_start0_seed1234"] = np.stack(
        [pre.compute_mfccs(orc.pcm16_to_float(c))[:, :, 0] for c in rnd])
    # float32 input path (model.py:295-303 hands float32/float64 arrays)
    mf["random_f32in"] = np.stack(
        [pre.compute_mfccs((c / 32768.0).astype(np.float32))[:, :, 0] for c in rnd[:2]])
    np.savez_compressed(os.path.join(HERE, "mfcc_golden.npz"), **mf)
    print("mfcc_golden:", {k: v.shape for k, v in mf.items()})

    # ---- CNN goldens -------------------------------------------------------
    clips = synth.make_clips(6, start=100, seed=1234)
    x = pre.compute_mfccs_batch([orc.pcm16_to_float(c) for c in clips]).astype(np.float32)
    calib = pre.compute_mfccs_batch([orc.pcm16_to_float(c) for c in synth.make_clips(16, start=200)])
    out = {"x": x}
    for ct in refmod.ConfigType:
        name = ct.value
        cfg = dict(refmod.find_config(ct))
        if not name.startswith("res"):
            cfg["n_labels"] = 12                      # server uses 12 labels (server.py:119-121)
        sd = synth.make_state_dict(name, cfg, seed=1, calib_mfcc=calib.astype(np.float32))
        model = refmod.find_model(ct)(cfg)
        tsd = {k: torch.from_numpy(np.asarray(v)) for k, v in sd.items()}
        model.load_state_dict(tsd, strict=True)
        model.eval()
        with torch.no_grad():
            y32 = model(torch.from_numpy(x)).numpy()
            y64 = model.double()(torch.from_numpy(x).double()).numpy()
            o64 = orc.model_forward(name, sd, cfg, x, torch.float64).numpy()
        print("%-18s ref32-ref64 %.2e  oracle64-ref64 %.2e  |logit| %.3f" %
              (name, np.abs(y32 - y64).max(), np.abs(o64 - y64).max(), np.abs(y64).mean()))
        out[name + "/logits_f32"] = y32
        out[name + "/logits_f64"] = y64
        out[name + "/config"] = np.frombuffer(json.dumps(cfg).encode(), dtype=np.uint8)
        for k, v in sd.items():
            if "running_" in k:
                out[name + "/sd/" + k] = v
    np.savez_compressed(os.path.join(HERE, "cnn_golden.npz"), **out)
    print("wrote", os.path.join(HERE, "cnn_golden.npz"))


if __name__ == "__main__":
    main()
