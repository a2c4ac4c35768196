"""Seeded synthetic 1-second 16 kHz clips.

There is no Speech Commands data on the box, so tests and ``bench.py`` use these.
Each clip has its own RNG substream (``seed + index``) so that any rank can
regenerate exactly its shard.  Clips are int16 PCM like the wire format the
reference's label service receives (service.py:98); ``/ 32768`` gives the float
form ``SpeechDataset.load_audio`` would produce (model.py:295-303).
"""
from __future__ import annotations

import numpy as np

SR = 16000
N_SAMPLES = 16000
N_CLASSES = 12


def _burst(rng, t, f0, f1, centre, width):
    """Hann-enveloped tone/chirp burst."""
    env = np.zeros_like(t)
    a, b = centre - width / 2, centre + width / 2
    m = (t >= a) & (t < b)
    env[m] = 0.5 - 0.5 * np.cos(2 * np.pi * (t[m] - a) / width)
    k = (f1 - f0) / max(width, 1e-3)
    phase = 2 * np.pi * (f0 * (t - a) + 0.5 * k * (t - a) ** 2) + rng.uniform(0, 2 * np.pi)
    return env * np.sin(phase)


def make_clip(index, seed=1234, label=None):
    """One int16 clip.  ``label`` (0..11) selects a class-structured tone pattern."""
    rng = np.random.default_rng(seed + int(index))
    t = np.arange(N_SAMPLES) / SR
    kind = rng.uniform()
    if label == 0 or (label is None and kind < 0.01):
        return np.zeros(N_SAMPLES, dtype=np.int16)          # silence clip (model.py:292)
    x = np.zeros(N_SAMPLES)
    if label is None:
        for _ in range(rng.integers(1, 4)):
            f0 = np.exp(rng.uniform(np.log(50), np.log(7000)))
            f1 = f0 * np.exp(rng.uniform(-0.5, 0.5))
            x += _burst(rng, t, f0, f1, rng.uniform(0.1, 0.9), rng.uniform(0.1, 0.6))
    else:
        # class pattern: two bursts whose frequencies and order depend on the label
        base = 200.0 * (1.35 ** label)
        jit = np.exp(rng.normal(0, 0.03))
        x += _burst(rng, t, base * jit, base * jit * (1.0 + 0.1 * (label % 3)),
                    0.3 + 0.03 * rng.normal(), 0.3)
        x += 0.7 * _burst(rng, t, base * 2.2 * jit, base * 2.2 * jit, 0.65 + 0.03 * rng.normal(), 0.25)
    x /= max(np.max(np.abs(x)), 1e-9)
    noise_db = rng.uniform(-60, -20)
    x = x + 10 ** (noise_db / 20) * rng.standard_normal(N_SAMPLES)
    peak_db = rng.uniform(-30, 0)
    x = x / max(np.max(np.abs(x)), 1e-9) * 10 ** (peak_db / 20)
    pcm = np.clip(np.round(x * 32767.0), -32768, 32767).astype(np.int16)
    if label is None and 0.01 <= kind < 0.02:
        pcm[rng.integers(2000, 15000):] = 0               # short clip zero-padded (model.py:297)
    return pcm


def make_clips(n, start=0, seed=1234, labelled=False):
    """``[n, 16000]`` int16 clips (+ labels when ``labelled``)."""
    out = np.empty((n, N_SAMPLES), dtype=np.int16)
    labels = np.empty(n, dtype=np.int64)
    for i in range(n):
        lab = None
        if labelled:
            lab = int(np.random.default_rng(seed + 7919 * (start + i)).integers(0, N_CLASSES))
            labels[i] = lab
        out[i] = make_clip(start + i, seed=seed, label=lab)
    return (out, labels) if labelled else out


def special_clips():
    """Edge-case clips used by the parity tests: name -> int16 ``[16000]``."""
    t = np.arange(N_SAMPLES) / SR
    out = {}
    out["zeros"] = np.zeros(N_SAMPLES, dtype=np.int16)
    out["square_fullscale"] = np.where(np.sin(2 * np.pi * 440 * t) >= 0, 32767, -32768).astype(np.int16)
    out["impulse"] = np.zeros(N_SAMPLES, dtype=np.int16)
    out["impulse"][8000] = 32767
    for b in (3, 30, 60, 119, 130):                         # bin-centred sinusoids (bin = 100/3 Hz)
        f = b * SR / 480.0
        out["sine_bin%d" % b] = np.round(0.5 * 32767 * np.sin(2 * np.pi * f * t)).astype(np.int16)
    short = make_clip(5, seed=99)
    short[6000:] = 0
    out["short_padded"] = short
    out["int16_extremes"] = np.tile(np.array([-32768, 32767, 0, 1, -1], dtype=np.int16), 3200)
    q = np.zeros(N_SAMPLES, dtype=np.int16)
    q[::7] = 1                                              # 1-LSB signal: tiny but non-zero energies
    out["one_lsb"] = q
    return out


# --------------------------------------------------------------------------
# Seeded weights with the reference's state_dict key names (model.py:85-102, 126-184)
# --------------------------------------------------------------------------

def _conv_out(n, k, s):
    return (n - k) // s + 1


def state_dict_shapes(name, config):
    """Ordered ``{key: shape}`` of the reference model's state_dict for ``config``."""
    shapes = {}
    if name.startswith("res"):
        C, L = config["n_feature_maps"], config["n_labels"]
        shapes["conv0.weight"] = (C, 1, 3, 3)
        for i in range(1, config["n_layers"] + 1):
            shapes["bn%d.running_mean" % i] = (C,)
            shapes["bn%d.running_var" % i] = (C,)
            shapes["bn%d.num_batches_tracked" % i] = ()
            shapes["conv%d.weight" % i] = (C, C, 3, 3)
        shapes["output.weight"] = (L, C)
        shapes["output.bias"] = (L,)
        return shapes
    H, W = config["height"], config["width"]
    C1 = config["n_feature_maps1"]
    kh, kw = config["conv1_size"]
    sh, sw = config["conv1_stride"]
    shapes["conv1.weight"] = (C1, 1, kh, kw)
    shapes["conv1.bias"] = (C1,)
    H, W = _conv_out(H, kh, sh), _conv_out(W, kw, sw)
    ph, pw = config["conv1_pool"]
    H, W = H // ph, W // pw
    C = C1
    if "conv2_size" in config:
        C2 = config["n_feature_maps2"]
        kh, kw = config["conv2_size"]
        sh, sw = config["conv2_stride"]
        shapes["conv2.weight"] = (C2, C, kh, kw)
        shapes["conv2.bias"] = (C2,)
        H, W = _conv_out(H, kh, sh), _conv_out(W, kw, sw)
        ph, pw = config["conv2_pool"]
        H, W = H // ph, W // pw
        C = C2
    flat = C * H * W
    last = flat
    tf_variant = bool(config.get("tf_variant"))
    if not tf_variant:
        shapes["lin.weight"] = (32, flat)
        shapes["lin.bias"] = (32,)
    if "dnn1_size" in config:
        d1 = config["dnn1_size"]
        shapes["dnn1.weight"] = (d1, flat if tf_variant else 32)
        shapes["dnn1.bias"] = (d1,)
        last = d1
        if "dnn2_size" in config:
            d2 = config["dnn2_size"]
            shapes["dnn2.weight"] = (d2, d1)
            shapes["dnn2.bias"] = (d2,)
            last = d2
    shapes["output.weight"] = (config["n_labels"], last)
    shapes["output.bias"] = (config["n_labels"],)
    return shapes


def make_state_dict(name, config, seed=0, calib_mfcc=None):
    """Seeded float32 weights (numpy RNG, not torch's) under the reference key names.

    Conv/linear weights are He-scaled.  For ``res*`` models the BatchNorm running
    statistics are calibrated on ``calib_mfcc`` ([B,101,40] float32) the way
    training would set them, then perturbed (mean shifted by N(0, 0.3 sigma), var
    scaled by U(0.7, 1.4)) so that BN is exercised with non-trivial, non-matching
    statistics.  Without ``calib_mfcc`` the stats are N(0, 0.5) / U(0.5, 2).
    """
    import torch
    import torch.nn.functional as F
    rng = np.random.default_rng(10_000 + seed)
    sd = {}
    for key, shp in state_dict_shapes(name, config).items():
        if key.endswith("num_batches_tracked"):
            sd[key] = np.asarray(100, dtype=np.int64)
        elif key.endswith("running_mean"):
            sd[key] = rng.normal(0, 0.5, shp).astype(np.float32)
        elif key.endswith("running_var"):
            sd[key] = rng.uniform(0.5, 2.0, shp).astype(np.float32)
        elif key.endswith("bias"):
            sd[key] = rng.normal(0, 0.1, shp).astype(np.float32)
        else:
            fan_in = int(np.prod(shp[1:]))
            sd[key] = (rng.standard_normal(shp) * np.sqrt(2.0 / fan_in)).astype(np.float32)
    if name.startswith("res") and calib_mfcc is not None:
        dil = bool(config["use_dilation"])
        with torch.no_grad():
            x = torch.as_tensor(np.asarray(calib_mfcc), dtype=torch.float32).unsqueeze(1)
            old = None
            for i in range(config["n_layers"] + 1):
                d = 1 if (i == 0 or not dil) else int(2 ** ((i - 1) // 3))
                y = F.relu(F.conv2d(x, torch.from_numpy(sd["conv%d.weight" % i]), None, 1, d, d))
                if i == 0:
                    if "res_pool" in config:
                        y = F.avg_pool2d(y, tuple(config["res_pool"]))
                    old = y
                if i > 0 and i % 2 == 0:
                    x = y + old
                    old = x
                else:
                    x = y
                if i > 0:
                    mu = x.mean(dim=(0, 2, 3))
                    var = x.var(dim=(0, 2, 3), unbiased=True)
                    sig = var.sqrt()
                    mu = mu + torch.from_numpy(rng.normal(0, 0.3, mu.shape).astype(np.float32)) * sig
                    var = var * torch.from_numpy(rng.uniform(0.7, 1.4, var.shape).astype(np.float32))
                    var = var.clamp_min(1e-6)
                    sd["bn%d.running_mean" % i] = mu.numpy().copy()
                    sd["bn%d.running_var" % i] = var.numpy().copy()
                    x = (x - mu[None, :, None, None]) / torch.sqrt(var[None, :, None, None] + 1e-5)
    return sd
