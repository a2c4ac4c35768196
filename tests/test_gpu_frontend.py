"""GPU parity of the front end (through the C ABI) against the CPU oracle and golden vectors.

Tolerance: the DFT is exact (integer limbs on the tensor cores for int16 PCM) or FP64 (float inputs) and
mel / log / DCT are FP64 like the reference, so the bound is set by the float32 spectrum / power the reference
itself uses and the final float32 cast: |d MFCC| <= 2e-4 absolute (MFCC magnitudes reach ~150); measured 8e-6."""
import os

import numpy as np
import pytest
import torch

from honk_b200 import synth

pytestmark = pytest.mark.gpu
ATOL = 2e-4
# Bands whose mel energy is more than e^-46 (1e-20) below the frame's strongest band hold only
# FP64 rounding noise of the DFT in the reference itself (e.g. a bin-centred sinusoid has exact
# spectral nulls), so their logarithm is algorithm-dependent noise; they are excluded.
LOG_FLOOR = 46.0


def assert_logmel_close(got_mfcc, ref_mfcc, name):
    from oracle import honk_oracle
    D = honk_oracle.dct_filters(40, 40)
    lm_got, lm_ref = got_mfcc.astype(np.float64) @ D, ref_mfcc.astype(np.float64) @ D   # D orthonormal
    live = lm_ref > (lm_ref.max(axis=1, keepdims=True) - LOG_FLOOR)
    assert live.mean() > 0.02, name
    err = np.abs(lm_got - lm_ref)[live].max()
    assert err <= 5 * ATOL, (name, err)


@pytest.fixture(scope="module")
def pre():
    from honk_b200 import AudioPreprocessor
    return AudioPreprocessor()


@pytest.fixture(scope="module")
def orc():
    from oracle import honk_oracle
    return honk_oracle


def test_tables_match_oracle(pre, orc):
    np.testing.assert_allclose(pre.dct_filters, orc.dct_filters(40, 40), rtol=0, atol=1e-15)
    np.testing.assert_allclose(pre.mel_basis(), orc.mel_filterbank(), rtol=0, atol=1e-15)
    assert (pre.n_mels, pre.sr, pre.f_max, pre.f_min, pre.n_fft, pre.hop_length) == (40, 16000, 4000, 20, 480, 160)


def test_golden_special_clips(pre, golden_dir):
    g = np.load(os.path.join(golden_dir, "mfcc_golden.npz"))
    for name, pcm in synth.special_clips().items():
        out = pre.compute_mfccs(pcm / 32768.0)                  # float64 like service.py:98
        assert out.shape == (101, 40, 1) and out.dtype == np.float32 and out.flags["F_CONTIGUOUS"]
        if name.startswith("sine_bin") or name == "int16_extremes":      # line spectra with exact nulls
            assert_logmel_close(out[:, :, 0], g["special/" + name], name)
        else:
            err = np.abs(out[:, :, 0] - g["special/" + name]).max()
            assert err <= ATOL, (name, err)
    assert np.all(pre.compute_mfccs(np.zeros(16000)) == 0)      # the data > 0 guard (manage_audio.py:39)


def test_golden_random_int16_batch(pre, golden_dir):
    g = np.load(os.path.join(golden_dir, "mfcc_golden.npz"))
    pcm = torch.from_numpy(synth.make_clips(8, start=0, seed=1234)).cuda()
    out = pre.compute_mfccs_batch(pcm)
    assert out.shape == (8, 101, 40) and out.is_cuda
    assert np.abs(out.cpu().numpy() - g["random_pcm_start0_seed1234"]).max() <= ATOL
    f32 = (pcm[:2].to(torch.float32) / 32768.0)
    assert np.abs(pre.compute_mfccs_batch(f32).cpu().numpy() - g["random_f32in"]).max() <= ATOL


def test_against_oracle_fresh_clips(pre, orc):
    clips = synth.make_clips(48, start=5000, seed=77)
    ref = orc.AudioPreprocessorOracle().compute_mfccs_batch([orc.pcm16_to_float(c) for c in clips])
    got = pre.compute_mfccs_batch(torch.from_numpy(clips).cuda()).cpu().numpy()
    err = np.abs(got - ref)
    assert err.max() <= ATOL, err.max()
    assert not np.isnan(got).any()


def test_dtypes_agree_and_strided_rows(pre):
    pcm = torch.from_numpy(synth.make_clips(5, start=900)).cuda()
    a = pre.compute_mfccs_batch(pcm)
    b = pre.compute_mfccs_batch(pcm.to(torch.float64) / 32768.0)
    # int16 / 32768 is exact in FP64, so both inputs describe the same signal; int16 takes the integer-limb tensor-core DFT
    # (exact spectrum, one float32 rounding), float64 the FP64 DFT: they differ by float32 rounding noise of the spectrum
    assert pre.native_path == "tcgen05-i8-limbs"
    assert (a - b).abs().max().item() <= 2e-5
    wide = torch.zeros((5, 20000), dtype=torch.int16, device="cuda")
    wide[:, :16000] = pcm
    assert torch.equal(pre.compute_mfccs_batch(wide[:, :16000]), a)   # row stride != n_samples


def test_large_batch_properties(pre):
    # size-independent properties at a BASELINE-size batch: batch invariance and scale law
    base = torch.from_numpy(synth.make_clips(64, start=300)).cuda()
    big = base.repeat(64, 1)                                    # 4096 clips
    out = pre.compute_mfccs_batch(big)
    assert torch.equal(out[:64], out[-64:]) and torch.equal(out[:64], pre.compute_mfccs_batch(base))
    # doubling the amplitude adds 2*ln(2)*sqrt(40) to c0 only (log-mel shifts by a constant)
    x = (base[:8].to(torch.float64) / 32768.0) * 0.25
    m1, m2 = pre.compute_mfccs_batch(x), pre.compute_mfccs_batch(2 * x)
    d = (m2 - m1).cpu().numpy()
    live = (base[:8].abs().amax(1) > 0).cpu().numpy()
    assert np.abs(d[live][:, :, 0] - 2 * np.log(2) * np.sqrt(40)).max() < 1e-3
    assert np.abs(d[live][:, :, 1:]).max() < 1e-3


def test_other_parameters(pre, orc):
    from honk_b200 import AudioPreprocessor
    p2 = AudioPreprocessor(n_mels=32, n_dct_filters=20, f_max=7600, f_min=60)
    ref = orc.AudioPreprocessorOracle(n_mels=32, n_dct_filters=20, f_max=7600, f_min=60)
    clip = synth.make_clip(11) / 32768.0
    got = p2.compute_mfccs(clip)
    assert got.shape == (101, 20, 1)
    assert np.abs(got - ref.compute_mfccs(clip)).max() <= ATOL


def test_errors(pre):
    with pytest.raises(ValueError):
        pre.compute_mfccs(np.zeros((2, 16000)))
    with pytest.raises(ValueError):
        pre.compute_mfccs_batch(torch.zeros(1, 100, device="cuda"))      # shorter than the reflect pad
    assert pre.compute_mfccs_batch(torch.zeros(0, 16000, device="cuda")).shape == (0, 101, 40)
    with pytest.raises(NotImplementedError):
        pre.compute_pcen(np.zeros(16000))


def test_int16_tensor_core_path_shapes_and_edges(pre, orc):
    # integer-limb tcgen05 DFT (fe_i8.cu): ragged tile pairs (clip counts around multiples of 128 rows), clips shorter than
    # one hop group, lengths that are no multiple of 16 or 160 (per-sample builder path), overlapping rows, odd alignment
    from honk_b200 import AudioPreprocessor
    cases = [(1, 16000), (2, 16000), (3, 15999), (5, 8000), (7, 1600), (2, 320), (1, 241), (130, 16000), (4, 16016), (3, 12345)]
    for n, ns in cases:
        clips = synth.make_clips(n, start=9000 + n)[:, :ns].copy() if ns <= 16000 else np.concatenate(
            [synth.make_clips(n, start=9000 + n), synth.make_clips(n, start=9100 + n)[:, :ns - 16000]], axis=1)
        ref = np.stack([orc.AudioPreprocessorOracle().compute_mfccs(orc.pcm16_to_float(c))[:, :, 0] for c in clips])
        got = pre.compute_mfccs_batch(torch.from_numpy(clips).cuda()).cpu().numpy()
        assert got.shape == ref.shape, (n, ns)
        assert np.abs(got - ref).max() <= ATOL, (n, ns, np.abs(got - ref).max())
    base = torch.from_numpy(np.concatenate(list(synth.make_clips(3, start=40)))).cuda()
    for hop, off in ((4000, 0), (164, 0), (8000, 1), (1, 3)):          # overlapping rows; buffers that are not 16-byte aligned
        buf = base[off:]
        n = min(9, 1 + (buf.numel() - 16000) // hop)
        win = buf.as_strided((n, 16000), (hop, 1))
        got = pre.compute_mfccs_batch(win).cpu().numpy()
        ref = np.stack([orc.AudioPreprocessorOracle().compute_mfccs(orc.pcm16_to_float(w))[:, :, 0] for w in win.cpu().numpy()])
        assert np.abs(got - ref).max() <= ATOL, (hop, off)
    # the FP64 DMMA kernel (HONK_FE_I8=0) stays available and agrees
    os.environ["HONK_FE_I8"] = "0"
    try:
        p64 = AudioPreprocessor()
    finally:
        os.environ.pop("HONK_FE_I8")
    assert p64.native_path == "fp64-dmma"
    x = torch.from_numpy(synth.make_clips(33, start=77)).cuda()
    assert (p64.compute_mfccs_batch(x) - pre.compute_mfccs_batch(x)).abs().max().item() <= 2e-5
