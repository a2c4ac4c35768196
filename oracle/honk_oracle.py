"""CPU oracle for the honk_b200 hot path.  TEST INFRASTRUCTURE ONLY.

This file restates, on the CPU, the arithmetic of the reference's batched
keyword-spotting inference path so that the CUDA kernels can be checked against
it.  Only ``tests/``, ``__graft_entry__.smoke()`` and the ``cpu_baseline`` /
``--impl reference`` legs of ``bench.py`` may import it; the product package
``honk_b200`` never does (it raises when its CUDA library is missing).

Parity status
-------------
* CNN half (``res_forward`` / ``cnn_forward``): PINNED.  The reference's own
  ``utils/model.py`` classes run in the build container (with stub modules for
  the four missing imports) and ``tests/golden/make_golden.py`` froze their
  logits for every architecture; ``tests/test_oracle.py`` checks this file
  against those vectors.
* Front end (``compute_mfccs``): parity UNPINNED by the reference.  The
  arithmetic lives in librosa 0.6.2-0.6.3 (requirements.txt:3 of the
  reference), which is not vendored and cannot be installed offline (its
  ``filters.dct`` was removed in 0.7, so no newer release can stand in), and the
  reference ships no tests or golden MFCCs.  The restatement below follows the
  published librosa-0.6.3 algorithm at the reference's call sites
  (utils/manage_audio.py:19-42) and is cross-checked against an independent
  witness (torchaudio's Slaney mel filterbank and a direct O(N^2) DFT) in
  ``tests/test_oracle.py``.

Every function cites the reference lines it follows.
"""
from __future__ import annotations

import math

import numpy as np

try:  # scipy.fftpack is what librosa 0.6 calls; numpy.fft agrees to 1 ulp-ish
    from scipy import fftpack as _fftpack

    def _fft(x, axis=0):
        return _fftpack.fft(x, axis=axis)
except Exception:  # pragma: no cover
    def _fft(x, axis=0):
        return np.fft.fft(x, axis=axis)


# --------------------------------------------------------------------------
# Front end: librosa-0.6.3 melspectrogram -> log -> DCT  (manage_audio.py:30-42)
# --------------------------------------------------------------------------

def hz_to_mel(freqs):
    """librosa 0.6 ``hz_to_mel(htk=False)`` (Slaney / Auditory-toolbox scale)."""
    freqs = np.asanyarray(freqs, dtype=np.float64)
    f_sp = 200.0 / 3
    mels = freqs / f_sp
    min_log_hz = 1000.0
    min_log_mel = min_log_hz / f_sp
    logstep = np.log(6.4) / 27.0
    if freqs.ndim:
        big = freqs >= min_log_hz
        mels[big] = min_log_mel + np.log(freqs[big] / min_log_hz) / logstep
    elif freqs >= min_log_hz:
        mels = min_log_mel + np.log(freqs / min_log_hz) / logstep
    return mels


def mel_to_hz(mels):
    """librosa 0.6 ``mel_to_hz(htk=False)``."""
    mels = np.asanyarray(mels, dtype=np.float64)
    f_sp = 200.0 / 3
    freqs = f_sp * mels
    min_log_hz = 1000.0
    min_log_mel = min_log_hz / f_sp
    logstep = np.log(6.4) / 27.0
    if mels.ndim:
        big = mels >= min_log_mel
        freqs[big] = min_log_hz * np.exp(logstep * (mels[big] - min_log_mel))
    elif mels >= min_log_mel:
        freqs = min_log_hz * np.exp(logstep * (mels - min_log_mel))
    return freqs


def mel_filterbank(sr=16000, n_fft=480, n_mels=40, fmin=20.0, fmax=4000.0):
    """librosa 0.6 ``filters.mel(sr, n_fft, n_mels, fmin, fmax, htk=False, norm=1)``.

    Called through ``librosa.feature.melspectrogram`` at manage_audio.py:31-38.
    Returns float64 ``[n_mels, 1 + n_fft//2]`` (0.6.x has no dtype argument).
    """
    if fmax is None:
        fmax = float(sr) / 2
    n_bins = 1 + n_fft // 2
    weights = np.zeros((n_mels, n_bins), dtype=np.float64)
    fftfreqs = np.linspace(0, float(sr) / 2, n_bins, endpoint=True)
    mel_pts = np.linspace(hz_to_mel(fmin), hz_to_mel(fmax), n_mels + 2)
    mel_f = mel_to_hz(mel_pts)
    fdiff = np.diff(mel_f)
    ramps = np.subtract.outer(mel_f, fftfreqs)
    for i in range(n_mels):
        lower = -ramps[i] / fdiff[i]
        upper = ramps[i + 2] / fdiff[i + 1]
        weights[i] = np.maximum(0, np.minimum(lower, upper))
    enorm = 2.0 / (mel_f[2:n_mels + 2] - mel_f[:n_mels])
    weights *= enorm[:, np.newaxis]
    return weights


def dct_filters(n_filters=40, n_input=40):
    """librosa 0.6 ``filters.dct`` (orthonormal DCT-II), manage_audio.py:22."""
    basis = np.empty((n_filters, n_input), dtype=np.float64)
    basis[0, :] = 1.0 / np.sqrt(n_input)
    samples = np.arange(1, 2 * n_input, 2) * np.pi / (2.0 * n_input)
    for i in range(1, n_filters):
        basis[i, :] = np.cos(i * samples) * np.sqrt(2.0 / n_input)
    return basis


def hann_periodic(n):
    """``scipy.signal.get_window('hann', n, fftbins=True)``."""
    k = np.arange(n, dtype=np.float64)
    return 0.5 - 0.5 * np.cos(2.0 * np.pi * k / n)


def stft_power(y, n_fft=480, hop_length=160):
    """librosa 0.6 ``_spectrogram(y, n_fft, hop_length, power=2)``.

    centre=True reflect padding, periodic Hann, float64 FFT stored as complex64,
    ``np.abs`` (float32 hypot) then ``**2`` in float32.  Returns float32
    ``[1 + n_fft//2, n_frames]``.
    """
    y = np.asarray(y)
    if y.ndim != 1:
        raise ValueError("expected a 1-D clip")
    window = hann_periodic(n_fft).reshape(-1, 1)
    ypad = np.pad(y, n_fft // 2, mode="reflect")
    n_frames = 1 + (len(ypad) - n_fft) // hop_length
    idx = np.arange(n_fft)[:, None] + hop_length * np.arange(n_frames)[None, :]
    frames = ypad[idx]                       # [n_fft, n_frames], y's dtype
    spec = _fft(window * frames, axis=0)[: 1 + n_fft // 2]
    spec = spec.astype(np.complex64)         # librosa stft dtype=np.complex64
    return np.abs(spec) ** 2.0               # float32


def log_mel(y, sr=16000, n_fft=480, hop_length=160, n_mels=40, fmin=20.0, fmax=4000.0,
            mel_basis=None):
    """melspectrogram + ``data[data > 0] = log(data[data > 0])`` (manage_audio.py:31-39).

    Returns float64 ``[n_mels, n_frames]``.
    """
    if mel_basis is None:
        mel_basis = mel_filterbank(sr, n_fft, n_mels, fmin, fmax)
    data = np.dot(mel_basis, stft_power(y, n_fft, hop_length))
    pos = data > 0
    data[pos] = np.log(data[pos])
    return data


class AudioPreprocessorOracle:
    """CPU restatement of ``AudioPreprocessor`` (manage_audio.py:18-42)."""

    def __init__(self, sr=16000, n_dct_filters=40, n_mels=40, f_max=4000, f_min=20, n_fft=480,
                 hop_ms=10):
        self.n_mels = n_mels
        self.dct_filters = dct_filters(n_dct_filters, n_mels)
        self.sr = sr
        self.f_max = f_max if f_max is not None else sr // 2
        self.f_min = f_min
        self.n_fft = n_fft
        self.hop_length = sr // 1000 * hop_ms
        self._mel = mel_filterbank(sr, n_fft, n_mels, f_min, self.f_max)

    def compute_mfccs(self, data):
        """One clip -> float32 ``(n_frames, n_dct, 1)`` Fortran-ordered (manage_audio.py:30-42)."""
        lm = log_mel(data, self.sr, self.n_fft, self.hop_length, self.n_mels, self.f_min,
                     self.f_max, mel_basis=self._mel)
        cols = [np.matmul(self.dct_filters, x) for x in np.split(lm, lm.shape[1], axis=1)]
        return np.array(cols, order="F").astype(np.float32)

    def compute_mfccs_batch(self, clips):
        """``[B, n_samples]`` -> float32 ``[B, n_frames, n_dct]`` (collate_fn shape, model.py:258)."""
        return np.stack([self.compute_mfccs(c).reshape(-1, self.dct_filters.shape[0]) for c in clips])


_POOL_PRE = None


def _pool_mfcc(clip):
    """Worker of ``MfccPool``: one int16 clip -> float32 ``[n_frames, n_dct]`` exactly as ``collate_fn`` calls it."""
    global _POOL_PRE
    if _POOL_PRE is None:
        _POOL_PRE = AudioPreprocessorOracle()
    return _POOL_PRE.compute_mfccs(pcm16_to_float(clip)).reshape(-1, _POOL_PRE.dct_filters.shape[0])


class MfccPool:
    """Per-clip ``compute_mfccs`` (manage_audio.py:30-42, called one clip at a time like model.py:253-265) spread over a
    PROCESS pool: the per-clip loop holds the GIL (101 small matmuls + np.split), so threads do not scale it
    (SURVEY.md 8d asks for the all-cores number).  Workers are forked (they only run numpy / scipy, never CUDA, so a
    live CUDA context in the parent is not touched); ``HONK_ORACLE_MP=spawn`` selects spawned workers instead."""

    def __init__(self, workers=None):
        import multiprocessing as mp
        import os
        if workers is None:
            try:
                workers = len(os.sched_getaffinity(0))
            except AttributeError:
                workers = os.cpu_count() or 1
        self.workers = max(1, int(workers))
        self._pool = mp.get_context(os.environ.get("HONK_ORACLE_MP", "fork")).Pool(self.workers)

    def compute(self, clips):
        clips = [np.ascontiguousarray(c) for c in clips]
        chunk = max(1, min(64, len(clips) // (4 * self.workers) or 1))
        return np.stack(self._pool.map(_pool_mfcc, clips, chunksize=chunk))

    def close(self):
        self._pool.close()
        self._pool.join()

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()


def pcm16_to_float(pcm):
    """``np.frombuffer(wav, int16) / 32768.`` (service.py:98) -> float64."""
    return np.asarray(pcm, dtype=np.int16) / 32768.0


# --------------------------------------------------------------------------
# CNN half: functional restatement of SpeechResModel / SpeechModel forward
# --------------------------------------------------------------------------

def _t(x, dtype):
    import torch
    return torch.as_tensor(np.asarray(x) if not hasattr(x, "detach") else x.detach().cpu()).to(dtype)


def res_forward(state, config, x, dtype=None):
    """``SpeechResModel.forward`` in eval mode (utils/model.py:104-121).

    ``state`` uses the reference's state_dict key names (model.py:85-102).
    ``x``: ``[B, T, F]``.  Returns ``[B, n_labels]`` logits (torch tensor).
    """
    import torch
    import torch.nn.functional as F
    dtype = dtype or torch.float32
    x = _t(x, dtype).unsqueeze(1)
    n_layers = config["n_layers"]
    dil = bool(config["use_dilation"])
    old_x = None
    for i in range(n_layers + 1):
        w = _t(state["conv%d.weight" % i], dtype)
        if i == 0:
            d = 1
        else:
            d = int(2 ** ((i - 1) // 3)) if dil else 1
        y = F.relu(F.conv2d(x, w, None, 1, d, d))
        if i == 0:
            if "res_pool" in config:
                y = F.avg_pool2d(y, tuple(config["res_pool"]))
            old_x = y
        if i > 0 and i % 2 == 0:
            x = y + old_x
            old_x = x
        else:
            x = y
        if i > 0:
            mean = _t(state["bn%d.running_mean" % i], dtype)
            var = _t(state["bn%d.running_var" % i], dtype)
            x = F.batch_norm(x, mean, var, None, None, False, 0.0, 1e-5)
    x = x.reshape(x.size(0), x.size(1), -1).mean(2)
    return F.linear(x, _t(state["output.weight"], dtype), _t(state["output.bias"], dtype))


def res_train_logits(params, stats, config, x, momentum=0.1, eps=1e-5):
    """``SpeechResModel.forward`` in TRAINING mode (utils/model.py:104-121 after ``model.train()``, train.py:128):
    ``nn.BatchNorm2d(n_maps, affine=False)`` (model.py:98) normalises with the batch mean / biased variance and moves
    ``stats`` (running_mean / running_var torch tensors, updated IN PLACE) by ``momentum`` towards the batch mean /
    unbiased variance.  ``params``: torch tensors (leaf, requires_grad) under the reference's key names."""
    import torch.nn.functional as F
    x = x.unsqueeze(1)
    dil = bool(config["use_dilation"])
    old_x = None
    for i in range(config["n_layers"] + 1):
        d = 1 if (i == 0 or not dil) else int(2 ** ((i - 1) // 3))
        y = F.relu(F.conv2d(x, params["conv%d.weight" % i], None, 1, d, d))
        if i == 0:
            if "res_pool" in config:
                y = F.avg_pool2d(y, tuple(config["res_pool"]))
            old_x = y
        if i > 0 and i % 2 == 0:
            x = y + old_x
            old_x = x
        else:
            x = y
        if i > 0:
            x = F.batch_norm(x, stats["bn%d.running_mean" % i], stats["bn%d.running_var" % i], None, None, True, momentum, eps)
    x = x.reshape(x.size(0), x.size(1), -1).mean(2)
    return F.linear(x, params["output.weight"], params["output.bias"])


def cnn_train_logits(params, config, x, masks=None):
    """``SpeechModel.forward`` in TRAINING mode (utils/model.py:186-205): ``nn.Dropout(dropout_prob)`` after relu(conv1),
    relu(conv2), dnn1 and dnn2, each BEFORE the max-pool where there is one.  ``masks``: keep masks (1 = keep) for those four
    sites in that order (entries of absent layers ignored); the kept values are scaled by 1 / (1 - p) like nn.Dropout does.
    ``masks=None`` means p = 0 (no dropout)."""
    import torch.nn.functional as F
    p = float(config.get("dropout_prob", 0.5)) if masks is not None else 0.0

    def drop(t, site):
        if masks is None or p == 0.0:
            return t
        m = _t(masks[site], t.dtype).reshape(t.shape)
        return t * m * (1.0 / (1.0 - p))

    x = F.relu(F.conv2d(x.unsqueeze(1), params["conv1.weight"], params["conv1.bias"], tuple(config["conv1_stride"])))
    x = F.max_pool2d(drop(x, 0), tuple(config["conv1_pool"]))
    if "conv2.weight" in params:
        x = F.relu(F.conv2d(x, params["conv2.weight"], params["conv2.bias"], tuple(config["conv2_stride"])))
        x = F.max_pool2d(drop(x, 1), tuple(config["conv2_pool"]))
    x = x.reshape(x.size(0), -1)
    if "lin.weight" in params:
        x = F.linear(x, params["lin.weight"], params["lin.bias"])
    if "dnn1.weight" in params:
        x = F.linear(x, params["dnn1.weight"], params["dnn1.bias"])
        if not config.get("tf_variant"):
            x = F.relu(x)
        x = drop(x, 2)
    if "dnn2.weight" in params:
        x = drop(F.linear(x, params["dnn2.weight"], params["dnn2.bias"]), 3)
    return F.linear(x, params["output.weight"], params["output.bias"])


def sgd_train_steps(state, config, batches, lr=0.001, momentum=0.9, weight_decay=0.00001, nesterov=False, dtype=None):
    """The reference's inner training loop (utils/train.py:127-139) for a res* model, restated without nn.Module /
    torch.optim: forward in training mode, mean CrossEntropyLoss (train.py:104), autograd for the gradients, then the
    update rule of ``torch.optim.SGD`` (dampening 0): g += wd * p; buf = g on the first step, else momentum * buf + g;
    g = g + momentum * buf (Nesterov) or buf; p -= lr * g.

    ``state``: numpy state_dict; ``batches``: list of ``(mfcc [B, T, F], labels [B])`` - for a cnn-* model (a state_dict
    with ``conv1.bias``) optionally ``(mfcc, labels, masks)`` with the Dropout keep masks of ``cnn_train_logits``.
    Returns a dict with the final ``state`` (numpy, num_batches_tracked advanced), per-step ``losses``, ``scores`` and
    ``grads`` (list of dicts)."""
    import torch
    import torch.nn.functional as F
    dtype = dtype or torch.float32
    params, stats, tracked = {}, {}, {}
    for k, v in state.items():
        if k.endswith("num_batches_tracked"):
            tracked[k] = int(np.asarray(v))
        elif "running_" in k:
            stats[k] = _t(v, dtype).clone()
        else:
            params[k] = _t(v, dtype).clone().requires_grad_()
    bufs = {}
    losses, scores, grads = [], [], []
    for batch in batches:
        x, y = batch[0], batch[1]
        for p in params.values():
            p.grad = None
        if "conv1.bias" in params:                       # SpeechModel (the res* convolutions have no bias)
            logits = cnn_train_logits(params, config, _t(x, dtype), batch[2] if len(batch) > 2 else None)
        else:
            logits = res_train_logits(params, stats, config, _t(x, dtype))
        loss = F.cross_entropy(logits, torch.as_tensor(np.asarray(y), dtype=torch.int64))
        loss.backward()
        losses.append(float(loss.detach()))
        scores.append(logits.detach().numpy().copy())
        grads.append({k: p.grad.detach().numpy().copy() for k, p in params.items()})
        with torch.no_grad():
            for k, p in params.items():
                g = p.grad
                if weight_decay != 0:
                    g = g + weight_decay * p
                if momentum != 0:
                    bufs[k] = g.clone() if k not in bufs else bufs[k] * momentum + g
                    g = g + momentum * bufs[k] if nesterov else bufs[k]
                p -= lr * g
        for k in tracked:
            tracked[k] += 1
    out = {k: v.detach().numpy().copy() for k, v in params.items()}
    out.update({k: v.numpy().copy() for k, v in stats.items()})
    out.update({k: np.asarray(v, dtype=np.int64) for k, v in tracked.items()})
    return {"state": out, "losses": losses, "scores": scores, "grads": grads,
            "momentum": {k: v.numpy().copy() for k, v in bufs.items()}}


def cnn_forward(state, config, x, dtype=None):
    """``SpeechModel.forward`` in eval mode (utils/model.py:186-205); dropout is identity."""
    import torch
    import torch.nn.functional as F
    dtype = dtype or torch.float32
    tf_variant = bool(config.get("tf_variant"))
    x = _t(x, dtype).unsqueeze(1)
    x = F.relu(F.conv2d(x, _t(state["conv1.weight"], dtype), _t(state["conv1.bias"], dtype),
                        tuple(config["conv1_stride"])))
    x = F.max_pool2d(x, tuple(config["conv1_pool"]))
    if "conv2.weight" in state:
        x = F.relu(F.conv2d(x, _t(state["conv2.weight"], dtype), _t(state["conv2.bias"], dtype),
                            tuple(config["conv2_stride"])))
        x = F.max_pool2d(x, tuple(config["conv2_pool"]))
    x = x.reshape(x.size(0), -1)
    if "lin.weight" in state:
        x = F.linear(x, _t(state["lin.weight"], dtype), _t(state["lin.bias"], dtype))
    if "dnn1.weight" in state:
        x = F.linear(x, _t(state["dnn1.weight"], dtype), _t(state["dnn1.bias"], dtype))
        if not tf_variant:
            x = F.relu(x)
    if "dnn2.weight" in state:
        x = F.linear(x, _t(state["dnn2.weight"], dtype), _t(state["dnn2.bias"], dtype))
    return F.linear(x, _t(state["output.weight"], dtype), _t(state["output.bias"], dtype))


def model_forward(name, state, config, x, dtype=None):
    """Dispatch like ``find_model`` (utils/model.py:51-57)."""
    if name.startswith("res"):
        return res_forward(state, config, x, dtype)
    return cnn_forward(state, config, x, dtype)


def softmax_label(logits_row):
    """``TorchLabelService.label`` tail (service.py:103-104): softmax -> (argmax, max prob)."""
    z = np.asarray(logits_row, dtype=np.float64)
    e = np.exp(z - z.max())
    p = e / e.sum()
    return int(np.argmax(p)), float(np.max(p))


# --------------------------------------------------------------------------
# Evaluation data path: PINNED by tests/golden/eval_golden.npz, which
# tests/golden/make_eval_golden.py produced by running the reference's own
# SpeechDataset._timeshift_audio / load_audio, print_eval and nn.CrossEntropyLoss.
# --------------------------------------------------------------------------

def timeshift_audio(data, shift):
    """``SpeechDataset._timeshift_audio`` with the drawn ``shift`` given (model.py:267-273)."""
    a = -min(0, shift)
    b = max(0, shift)
    data = np.pad(data, (a, b), "constant")
    return data[:len(data) - a] if a else data[b:]


def load_audio(data, in_len=16000, shift=0, noise_gain=None, bg_noise=None):
    """The arithmetic of ``SpeechDataset.load_audio`` (model.py:296-303) with the random draws given:
    zero-pad to ``in_len``, time shift, and, when ``noise_gain`` is not None,
    ``clip(noise_gain * bg_noise + data, -1, 1)`` (``bg_noise`` already cut to ``in_len`` samples)."""
    data = np.asarray(data, dtype=np.float32)
    data = np.pad(data, (0, max(0, in_len - len(data))), "constant")
    data = timeshift_audio(data, shift)
    if noise_gain is not None:
        bg = np.zeros(in_len, dtype=np.float32) if bg_noise is None else np.asarray(bg_noise, dtype=np.float32)
        data = np.clip(np.float32(noise_gain) * bg + data, -1, 1)
    return data[:in_len]


def eval_metrics(scores, labels):
    """``print_eval``'s accuracy and ``nn.CrossEntropyLoss`` (train.py:41-46, 66-83) -> (accuracy, mean loss)."""
    z = np.asarray(scores, dtype=np.float64)
    y = np.asarray(labels, dtype=np.int64)
    acc = float((z.argmax(1) == y).sum()) / len(y)
    m = z.max(1)
    lse = np.log(np.exp(z - m[:, None]).sum(1)) + m
    return acc, float((lse - z[np.arange(len(y)), y]).mean())


def listen(label_fn, wav_bytes, stride_ms=500, min_keyword_prob=0.85, keyword="command", method="all_label"):
    """``ListenEndpoint.POST`` (server.py:99-112) over ``stride`` (service.py:106-110); ``label_fn(bytes)`` is
    ``label_service.label``."""
    step, window = int(2 * 16000 * stride_ms / 1000), 2 * 16000
    labels = {}
    i = 0
    while i + window <= len(wav_bytes):
        label, prob = label_fn(wav_bytes[i:i + window])
        labels[label] = labels.get(label, 0.0) + float(prob)
        if label == keyword and prob >= min_keyword_prob and method == "command_tagging":
            return dict(contains_command=True)
        i += step
    return dict(contains_command=False) if method == "command_tagging" else labels


# --------------------------------------------------------------------------
# Algorithmic work used by bench.py for roofline arithmetic
# --------------------------------------------------------------------------

def conv_flops_per_clip(name, config):
    """2*MAC of the unpadded reference layers for one clip (matches SURVEY.md §8d)."""
    H, W = 101, 40
    if name.startswith("res"):
        C = config["n_feature_maps"]
        fl = 2 * 9 * C * H * W
        if "res_pool" in config:
            ph, pw = config["res_pool"]
            H, W = H // ph, W // pw
        fl += config["n_layers"] * 2 * 9 * C * C * H * W
        fl += 2 * C * config["n_labels"]
        return fl
    fl = 0
    kh, kw = config["conv1_size"]
    sh, sw = config["conv1_stride"]
    C1 = config["n_feature_maps1"]
    H, W = (H - kh) // sh + 1, (W - kw) // sw + 1
    fl += 2 * kh * kw * C1 * H * W
    ph, pw = config["conv1_pool"]
    H, W = H // ph, W // pw
    C = C1
    if "conv2_size" in config:
        kh, kw = config["conv2_size"]
        sh, sw = config["conv2_stride"]
        C2 = config["n_feature_maps2"]
        H, W = (H - kh) // sh + 1, (W - kw) // sw + 1
        fl += 2 * kh * kw * C * C2 * H * W
        ph, pw = config["conv2_pool"]
        H, W = H // ph, W // pw
        C = C2
    last = C * H * W
    if not config.get("tf_variant"):
        fl += 2 * last * 32
        last = 32
    if "dnn1_size" in config:
        fl += 2 * last * config["dnn1_size"]
        last = config["dnn1_size"]
        if "dnn2_size" in config:
            fl += 2 * last * config["dnn2_size"]
            last = config["dnn2_size"]
    fl += 2 * last * config["n_labels"]
    return fl
