"""CPU tests of the oracle itself: it must reproduce the committed golden vectors (the CNN
vectors were produced by the reference's own classes, tests/golden/make_golden.py) and agree
with independent witnesses for the front end (whose parity the reference leaves unpinned)."""
import json
import os

import numpy as np
import pytest
import torch

from honk_b200 import synth
from oracle import honk_oracle as orc

ARCHS = ["cnn-trad-pool2", "cnn-one-stride1", "cnn-one-fpool3", "cnn-one-fstride4", "cnn-one-fstride8",
         "cnn-tpool2", "cnn-tpool3", "cnn-tstride2", "cnn-tstride4", "cnn-tstride8",
         "res15", "res26", "res8", "res15-narrow", "res8-narrow", "res26-narrow"]


def golden_model(golden_dir, name):
    g = np.load(os.path.join(golden_dir, "cnn_golden.npz"))
    cfg = json.loads(bytes(g[name + "/config"]).decode())
    sd = synth.make_state_dict(name, cfg, seed=1)
    for k in g.files:
        if k.startswith(name + "/sd/"):
            sd[k[len(name) + 4:]] = g[k]
    return g, cfg, sd


@pytest.mark.parametrize("name", ARCHS)
def test_cnn_oracle_matches_reference_golden(golden_dir, name):
    g, cfg, sd = golden_model(golden_dir, name)
    y64 = orc.model_forward(name, sd, cfg, g["x"], torch.float64).numpy()
    ref64 = g[name + "/logits_f64"]
    scale = max(1.0, np.abs(ref64).max())
    assert np.abs(y64 - ref64).max() <= 1e-9 * scale
    y32 = orc.model_forward(name, sd, cfg, g["x"], torch.float32).numpy()
    assert np.abs(y32 - ref64).max() <= 2e-4 * scale


def test_mfcc_golden_frozen(golden_dir):
    g = np.load(os.path.join(golden_dir, "mfcc_golden.npz"))
    pre = orc.AudioPreprocessorOracle()
    for k, pcm in synth.special_clips().items():
        got = pre.compute_mfccs(orc.pcm16_to_float(pcm))
        assert got.shape == (101, 40, 1) and got.dtype == np.float32 and got.flags["F_CONTIGUOUS"]
        np.testing.assert_allclose(got[:, :, 0], g["special/" + k], rtol=0, atol=2e-5)
    rnd = synth.make_clips(8, start=0, seed=1234)
    got = pre.compute_mfccs_batch([orc.pcm16_to_float(c) for c in rnd])
    np.testing.assert_allclose(got, g["random_pcm_start0_seed1234"], rtol=0, atol=2e-5)


def test_zero_clip_stays_zero():
    # data[data > 0] = log(...) leaves exact zeros alone (manage_audio.py:39)
    out = orc.AudioPreprocessorOracle().compute_mfccs(np.zeros(16000))
    assert np.all(out == 0)


def test_mel_basis_against_torchaudio():
    ta = pytest.importorskip("torchaudio")
    fb = ta.functional.melscale_fbanks(241, 20.0, 4000.0, 40, 16000, norm="slaney", mel_scale="slaney").T.numpy()
    W = orc.mel_filterbank()
    assert np.abs(fb - W).max() <= 1e-5 * W.max()
    nz = np.nonzero(W.sum(0))[0]
    assert nz.min() == 1 and nz.max() == 119 and (W >= 0).all()


def test_dct_orthonormal():
    D = orc.dct_filters(40, 40)
    assert np.abs(D @ D.T - np.eye(40)).max() < 1e-13


def test_stft_against_direct_dft():
    rng = np.random.default_rng(3)
    y = rng.standard_normal(16000) * 0.1
    P = orc.stft_power(y)
    assert P.shape == (241, 101) and P.dtype == np.float32
    ypad = np.pad(y, 240, mode="reflect")
    n = np.arange(480)
    win = 0.5 - 0.5 * np.cos(2 * np.pi * n / 480)
    for t in (0, 50, 100):
        fr = ypad[160 * t:160 * t + 480] * win
        k = np.arange(241)[:, None]
        X = (fr[None, :] * np.exp(-2j * np.pi * k * n[None, :] / 480)).sum(1)
        np.testing.assert_allclose(P[:, t], np.abs(X) ** 2, rtol=2e-6, atol=1e-12)


def test_whole_chain_against_independent_implementation():
    # Front-end parity is UNPINNED by the reference (librosa 0.6.x is neither vendored nor installable, and the reference
    # holds no golden MFCCs).  This is a whole-chain WITNESS, not a pin: transformers.audio_utils is an independent numpy
    # implementation written to match librosa (reflect-padded centred STFT, periodic Hann, power spectrum, Slaney mel
    # filters with area norm); log where > 0 and scipy's orthonormal DCT-II complete manage_audio.py:30-42.  Its spectrum
    # stays float64 where librosa 0.6 rounds to complex64, hence the 2e-5 bound (measured 3.9e-6).
    tr = pytest.importorskip("transformers.audio_utils")
    fftpack = pytest.importorskip("scipy.fftpack")
    mel = tr.mel_filter_bank(241, 40, 20.0, 4000.0, 16000, norm="slaney", mel_scale="slaney")
    assert np.abs(mel.T - orc.mel_filterbank()).max() <= 1e-15
    win = tr.window_function(480, "hann", periodic=True)
    pre = orc.AudioPreprocessorOracle()
    special = synth.special_clips()
    clips = [special[k] for k in ("square_fullscale", "impulse", "short_padded", "one_lsb")] + list(synth.make_clips(64, start=20000, seed=7))
    worst = 0.0
    for pcm in clips:
        y = orc.pcm16_to_float(pcm)
        S = tr.spectrogram(y, win, frame_length=480, hop_length=160, fft_length=480, power=2.0, center=True,
                           pad_mode="reflect", mel_filters=mel, mel_floor=1e-300)
        assert S.shape == (40, 101)
        S[S <= 1e-300] = 0.0                  # undo the witness's mel floor: exact zeros stay zero (manage_audio.py:39)
        pos = S > 0
        S[pos] = np.log(S[pos])
        # bands that only hold rounding noise of an exactly-zero spectrum differ by construction (float64 vs complex64)
        live = S > S.max(axis=0, keepdims=True) - 30.0
        got = pre.compute_mfccs(y)[:, :, 0]
        if live.all():
            worst = max(worst, float(np.abs(fftpack.dct(S, type=2, norm="ortho", axis=0).T - got).max()))
        else:
            lm_got = (got.astype(np.float64) @ orc.dct_filters(40, 40)).T
            worst = max(worst, float(np.abs(lm_got - S)[live].max()))
    assert worst <= 2e-5, worst
    assert np.all(pre.compute_mfccs(orc.pcm16_to_float(special["zeros"])) == 0)


def test_listen_oracle_follows_reference_loop():
    # oracle.listen restates ListenEndpoint.POST (server.py:99-112) over stride() (service.py:106-110): byte strides,
    # window = 2 * 16000 bytes, per-label sums of the winning probability, early exit for command_tagging
    calls = []

    def label_fn(b):
        calls.append(len(b))
        k = len(calls)
        return ("command" if k == 3 else "other", 0.9 if k == 3 else 0.5)
    wav = bytes(2 * 16000 * 3)
    out = orc.listen(label_fn, wav, stride_ms=500)
    assert len(calls) == 5 and set(calls) == {32000} and out == {"other": 2.0, "command": 0.9}
    calls.clear()
    assert orc.listen(label_fn, wav, stride_ms=500, method="command_tagging") == dict(contains_command=True) and len(calls) == 3
    calls.clear()
    assert orc.listen(label_fn, wav, stride_ms=500, min_keyword_prob=0.95, method="command_tagging") == dict(contains_command=False)
    assert orc.listen(label_fn, wav[:31999]) == {}


def test_sine_known_answer():
    # a bin-centred sinusoid of amplitude A gives |X_b| = A*N/4 under a periodic Hann window
    A, b = 0.5, 30
    t = np.arange(16000) / 16000.0
    y = A * np.sin(2 * np.pi * b * 16000 / 480.0 * t)
    P = orc.stft_power(y)
    np.testing.assert_allclose(P[b, 50], (A * 480 / 4) ** 2, rtol=1e-5)


def test_flops_table():
    from honk_b200 import model as mod
    exp = {"res15": 1917.63e6, "res26": 878.07e6, "res8": 74.35e6, "res8-narrow": 14.05e6,
           "res15-narrow": 342.66e6, "res26-narrow": 157.33e6, "cnn-trad-pool2": 191.95e6}
    for name, fl in exp.items():
        got = orc.conv_flops_per_clip(name, mod.find_config(name))
        assert abs(got - fl) / fl < 2e-3, (name, got)


def test_eval_path_oracle_matches_reference_golden(golden_dir):
    # the restated load_audio / _timeshift_audio / print_eval arithmetic against outputs of the
    # reference's own functions (tests/golden/make_eval_golden.py), bit for bit
    g = np.load(os.path.join(golden_dir, "eval_golden.npz"))
    bg = g["bg_noise"]
    for k, (n, shift, noisy, gain, off) in enumerate(g["load_audio/params"]):
        x = g["load_audio/%d/x" % k]
        y = orc.load_audio(x, 16000, int(shift), gain if noisy else None, bg[int(off):int(off) + 16000])
        ref = g["load_audio/%d/y" % k]
        assert y.dtype == ref.dtype == np.float32 and np.array_equal(y, ref), k
    gain, off = g["load_audio/silence/params"]
    y = orc.load_audio(np.zeros(0, np.float32), 16000, 0, gain, bg[int(off):int(off) + 16000])
    assert np.array_equal(y, g["load_audio/silence/y"])
    for sh in (-1600, -1, 0, 1, 1600):
        assert np.array_equal(orc.timeshift_audio(g["timeshift/x"], sh), g["timeshift/%d" % sh])
    acc, loss = orc.eval_metrics(g["eval/scores"], g["eval/labels"])
    assert abs(acc - float(g["eval/accuracy"])) < 1e-6 and abs(loss - float(g["eval/loss"])) < 1e-5
