"""The two-convolution cnn-* configs (model.py:385-407) run conv1 with the MaxPool fused into its epilogue and conv2 with
several clips sharing one pass of the weight stream (csrc/gemm_tc.cu).  bias + ReLU + f16 rounding are monotonic, so pooling
before them must not change a bit, and neither may the number of clips per pass: the logits of the default path are compared
BIT FOR BIT with the same library run with those two choices switched off (separate pooling pass, one clip per pass).

The switches are read once per process, so each variant runs in its own interpreter."""
import os
import subprocess
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

SCRIPT = r"""
import sys
import numpy as np
import torch
sys.path.insert(0, sys.argv[1])
import honk_b200 as hb
from honk_b200 import synth
name, out = sys.argv[2], sys.argv[3]
pre = hb.AudioPreprocessor()
clips = synth.make_clips(37, start=900)
mf = pre.compute_mfccs_batch(torch.from_numpy(clips).cuda())
cfg = dict(hb.find_config(name))
cfg["n_labels"] = 12
sd = synth.make_state_dict(name, cfg, seed=5, calib_mfcc=mf[:16].cpu().numpy())
m = hb.find_model(name)(cfg)
m.load_state_dict({k: torch.from_numpy(np.asarray(v)) for k, v in sd.items()})
m = m.cuda().eval()
np.save(out, m(mf).cpu().numpy())
"""


def run_variant(name, out, extra_env):
    env = {k: v for k, v in os.environ.items() if not k.startswith("HONK_")}
    env.update(extra_env)
    r = subprocess.run([sys.executable, "-c", SCRIPT, ROOT, name, out], env=env, capture_output=True, text=True, timeout=240)
    assert r.returncode == 0, r.stderr[-2000:]
    return np.load(out)


@pytest.mark.parametrize("name", ["cnn-tpool2", "cnn-tstride4"])
def test_fused_pool_and_shared_passes_change_no_bit(name, tmp_path):
    fused = run_variant(name, str(tmp_path / "fused.npy"), {})
    plain = run_variant(name, str(tmp_path / "plain.npy"), {"HONK_C1_FUSE": "0", "HONK_CR_G": "1"})
    assert fused.shape == (37, 12) and np.isfinite(fused).all()
    assert np.array_equal(fused, plain), float(np.abs(fused - plain).max())
