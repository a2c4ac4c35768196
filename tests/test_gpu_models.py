"""GPU parity of every architecture (through the C ABI) against the committed golden logits
(produced by the reference's own classes) and the CPU oracle.

Tolerances, relative to max|logit| of the batch:
  cuda-core-f32 path : 2e-4   (float32 everywhere, different summation order)
  tcgen05-f16 path   : 2e-3   (f16 operands and f16 activation storage, f32 accumulate, 7-25 stacked layers; measured
                               1.2e-4 ... 4e-4, so the bound is 5-10x the measurement - a dropped tap or a wrong BatchNorm
                               fold shows up at 1e-2 and above)"""
import json
import os

import numpy as np
import pytest
import torch

from honk_b200 import synth

pytestmark = pytest.mark.gpu

ARCHS = ["cnn-trad-pool2", "cnn-one-stride1", "cnn-one-fpool3", "cnn-one-fstride4", "cnn-one-fstride8",
         "cnn-tpool2", "cnn-tpool3", "cnn-tstride2", "cnn-tstride4", "cnn-tstride8",
         "res15", "res26", "res8", "res15-narrow", "res8-narrow", "res26-narrow"]
TOL = {"cuda-core-f32": 2e-4, "tcgen05-f16": 2e-3}


def load_golden(golden_dir, name):
    g = np.load(os.path.join(golden_dir, "cnn_golden.npz"))
    cfg = json.loads(bytes(g[name + "/config"]).decode())
    sd = synth.make_state_dict(name, cfg, seed=1)
    for k in g.files:
        if k.startswith(name + "/sd/"):
            sd[k[len(name) + 4:]] = g[k]
    return g, cfg, sd


def build(name, cfg, sd, precision=0):
    import honk_b200 as hb
    m = hb.find_model(name)(cfg)
    m.load_state_dict({k: torch.from_numpy(np.asarray(v)) for k, v in sd.items()}, strict=True)
    m.precision = precision
    return m.cuda().eval()


@pytest.mark.parametrize("precision", [1, 0])
@pytest.mark.parametrize("name", ARCHS)
def test_golden_logits(golden_dir, name, precision):
    g, cfg, sd = load_golden(golden_dir, name)
    m = build(name, cfg, sd, precision)
    x = torch.from_numpy(g["x"]).cuda()
    y = m(x).cpu().numpy()
    ref = g[name + "/logits_f64"]
    assert y.shape == ref.shape and y.dtype == np.float32
    path = m.native_path
    if precision == 1:
        assert path == "cuda-core-f32"
    err = np.abs(y - ref).max() / np.abs(ref).max()
    assert err <= TOL[path], (name, path, err)
    if path == "cuda-core-f32":
        assert (y.argmax(1) == ref.argmax(1)).all()


@pytest.mark.parametrize("name", ["res8-narrow", "res8", "cnn-trad-pool2", "res15-narrow", "res15"])
def test_against_oracle_batch64(name):
    # BASELINE config 1 shape: batch 64 synthetic clips through front end + CNN
    import honk_b200 as hb
    from oracle import honk_oracle as orc
    cfg = dict(hb.find_config(name))
    cfg["n_labels"] = 12
    clips = synth.make_clips(64, start=4000)
    mf = orc.AudioPreprocessorOracle().compute_mfccs_batch([orc.pcm16_to_float(c) for c in clips])
    sd = synth.make_state_dict(name, cfg, seed=3, calib_mfcc=mf[:16])
    ref = orc.model_forward(name, sd, cfg, mf, torch.float64).numpy()
    pre = hb.AudioPreprocessor()
    for precision in (1, 0):
        m = build(name, cfg, sd, precision)
        y = m.forward_pcm(torch.from_numpy(clips).cuda(), pre).cpu().numpy()
        err = np.abs(y - ref).max() / np.abs(ref).max()
        assert err <= TOL[m.native_path] * 1.5, (name, m.native_path, err)
        y2 = m(pre.compute_mfccs_batch(torch.from_numpy(clips).cuda())).cpu().numpy()
        # fused PCM entry == two-step entry (atomics reorder sums).  For the un-pooled res* models on the tcgen05 path the
        # fused entry never materialises MFCCs: the front end writes conv0's operand image itself (fe_forward_x0)
        assert np.allclose(y, y2, rtol=0, atol=1e-5 * np.abs(ref).max())


@pytest.mark.parametrize("name,n", [("res15-narrow", 1500), ("res8-narrow", 70)])
def test_host_pipeline_matches_device_entry(name, n, monkeypatch):
    # honk_model_forward_pcm_host: chunked H2D under the kernels == the device-buffer entry, row for row
    import honk_b200 as hb
    monkeypatch.setenv("HONK_HOST_CHUNKS", "3")
    cfg = dict(hb.find_config(name))
    cfg["n_labels"] = 12
    base = synth.make_clips(100, start=9000)
    clips = np.ascontiguousarray(np.tile(base, (n // 100 + 1, 1))[:n])
    clips[1::2] = clips[1::2][:, ::-1]                 # make neighbouring rows differ
    pre = hb.AudioPreprocessor()
    mf = pre.compute_mfccs_batch(torch.from_numpy(clips[:16]).cuda()).cpu().numpy()
    sd = synth.make_state_dict(name, cfg, seed=5, calib_mfcc=mf)
    m = build(name, cfg, sd, 0)
    ref = m.forward_pcm(torch.from_numpy(clips).cuda(), pre).cpu().numpy()
    scale = np.abs(ref).max()
    host = torch.from_numpy(clips).pin_memory()
    y = m.forward_pcm_host(host, pre).numpy()
    assert y.shape == ref.shape and np.allclose(y, ref, rtol=0, atol=1e-5 * scale)
    # strided rows (clip_stride > n_samples) and pageable memory take the 2-D copy path
    wide = np.zeros((n, clips.shape[1] + 37), dtype=clips.dtype)
    wide[:, :clips.shape[1]] = clips
    y2 = m.forward_pcm_host(torch.from_numpy(wide)[:, :clips.shape[1]], pre).numpy()
    assert np.allclose(y2, ref, rtol=0, atol=1e-5 * scale)
    with pytest.raises(RuntimeError):
        m.forward_pcm_host(host.cuda(), pre)


def test_top1_agreement_trained_res15_16384(golden_dir):
    # top-1 identical on >= 99.9 % of >= 16384 clips, with briefly trained weights so that the
    # logit margins are realistic (tests/golden/make_trained.py -> res15_trained.npz).  Three arms on the same clips:
    # the f16-operand tcgen05 path, the all-FP32 CUDA-core path, and the CPU oracle in float32 fed by the float64 oracle
    # front end - all three on all 16384 clips.
    import honk_b200 as hb
    from oracle import honk_oracle as orc
    cfg = dict(hb.find_config("res15"))
    cfg["n_labels"] = 12
    sd = dict(np.load(os.path.join(golden_dir, "res15_trained.npz")))
    lab, _ = synth.make_clips(2048, start=300000, labelled=True)
    unl = synth.make_clips(2048, start=400000)
    base = np.concatenate([lab, unl])                                    # 4096 unique clips
    clips = np.concatenate([base, (-(base.astype(np.int32) // 2)).astype(np.int16), np.ascontiguousarray(base[:, ::-1]),
                            np.roll(base, 1234, axis=1)])                 # + half-gain inverted, reversed, shifted
    assert clips.shape == (16384, 16000)
    pre = hb.AudioPreprocessor()
    x = torch.from_numpy(clips).cuda()
    y16 = build("res15", cfg, sd, 0).forward_pcm(x, pre).cpu().numpy()
    y32 = build("res15", cfg, sd, 1).forward_pcm(x, pre).cpu().numpy()
    scale = np.abs(y32).max()
    err = np.abs(y16 - y32).max() / scale
    agree = float((y16.argmax(1) == y32.argmax(1)).mean())
    assert err <= TOL["tcgen05-f16"], err
    assert agree >= 0.999, (agree, err)
    # the CPU oracle on ALL 16384 clips (float64 front end per clip over a process pool, reference-shaped float32 forward
    # in batches of 64 on every core: about a minute on the GPU box's host)
    with orc.MfccPool() as pool:
        mf = pool.compute(clips)
    torch.set_num_threads(max(1, len(os.sched_getaffinity(0))))
    with torch.no_grad():
        ref = np.concatenate([orc.model_forward("res15", sd, cfg, mf[i:i + 64], torch.float32).numpy() for i in range(0, len(mf), 64)])
    for y in (y16, y32):
        assert float((y.argmax(1) == ref.argmax(1)).mean()) >= 0.999
    assert np.abs(y32 - ref).max() / np.abs(ref).max() <= 1e-3
    assert np.abs(y16 - ref).max() / np.abs(ref).max() <= TOL["tcgen05-f16"]


@pytest.mark.parametrize("n", [1, 3, 149, 300])
def test_cnn_trad_pool2_tensor_path_batch_sizes(golden_dir, n):
    # persistent one-CTA-per-SM kernel: fewer clips than SMs, a ragged last round, several rounds per CTA
    g, cfg, sd = load_golden(golden_dir, "cnn-trad-pool2")
    rng = np.random.default_rng(n)
    x = torch.from_numpy((rng.standard_normal((n, 101, 40)) * np.abs(g["x"]).std()).astype(np.float32)).cuda()
    m16, m32 = build("cnn-trad-pool2", cfg, sd, 0), build("cnn-trad-pool2", cfg, sd, 1)
    y16, y32 = m16(x).cpu().numpy(), m32(x).cpu().numpy()
    assert m16.native_path == "tcgen05-f16" and m32.native_path == "cuda-core-f32"
    assert np.abs(y16 - y32).max() / np.abs(y32).max() <= 5e-3
    assert np.array_equal(m16(x).cpu().numpy(), y16)          # deterministic (no atomics on this path)


def test_strict_state_dict_and_reload(golden_dir):
    import honk_b200 as hb
    g, cfg, sd = load_golden(golden_dir, "res8-narrow")
    m = build("res8-narrow", cfg, sd, 1)
    x = torch.from_numpy(g["x"]).cuda()
    y0 = m(x)
    bad = {k: torch.from_numpy(np.asarray(v)) for k, v in sd.items()}
    bad.pop("conv3.weight")
    with pytest.raises(RuntimeError):
        hb.SpeechResModel(cfg).load_state_dict(bad)
    # weights changed in place -> the native handle is rebuilt
    with torch.no_grad():
        m.output.bias.add_(1.0)
    assert torch.allclose(m(x), y0 + 1.0, atol=1e-5)
    # non-contiguous / Fortran-ordered input (manage_audio.py:41)
    xt = x.transpose(1, 2).contiguous().transpose(1, 2)
    assert torch.equal(m(xt), m(x))
    assert m(x[:0]).shape == (0, 12)
    m.train()
    with pytest.raises(RuntimeError):
        m(x)


def test_save_load_roundtrip(tmp_path, golden_dir):
    import honk_b200 as hb
    g, cfg, sd = load_golden(golden_dir, "cnn-trad-pool2")
    m = build("cnn-trad-pool2", cfg, sd, 1)
    f = str(tmp_path / "m.pt")
    m.save(f)
    m2 = hb.SpeechModel(cfg)
    m2.load(f)
    m2.precision = 1                  # same arithmetic path as m: the round trip must be bit-exact
    m2.cuda().eval()
    x = torch.from_numpy(g["x"]).cuda()
    assert torch.equal(m(x), m2(x))


def test_label_kernel_matches_softmax():
    from honk_b200 import ops
    z = torch.randn(1000, 12, device="cuda") * 3
    idx, p = ops.label_from_logits(z)
    sm = torch.softmax(z.double(), 1)
    assert torch.equal(idx.long(), sm.argmax(1))
    assert torch.allclose(p.double(), sm.max(1).values, atol=1e-6)


@pytest.mark.parametrize("maps", [45, 19])
@pytest.mark.parametrize("n_layers,pool", [(1, None), (2, None), (3, None), (4, None), (7, None), (10, None),
                                           (13, None), (2, (4, 3)), (5, (2, 2))])
def test_res_depth_sweep(n_layers, pool, maps):
    # isolates each dilation (1,2,4,8,16) and the residual/BN epilogue of the tcgen05 path
    import honk_b200 as hb
    from oracle import honk_oracle as orc
    cfg = dict(n_labels=12, use_dilation=True, n_layers=n_layers, n_feature_maps=maps)
    if pool:
        cfg["res_pool"] = pool
    name = "res-sweep"
    clips = synth.make_clips(5, start=40 + n_layers)
    mf = orc.AudioPreprocessorOracle().compute_mfccs_batch([orc.pcm16_to_float(c) for c in clips])
    sd = synth.make_state_dict(name, cfg, seed=n_layers, calib_mfcc=mf)
    ref = orc.res_forward(sd, cfg, mf, torch.float64).numpy()
    m = hb.SpeechResModel(cfg)
    m.load_state_dict({k: torch.from_numpy(np.asarray(v)) for k, v in sd.items()})
    m.cuda().eval()
    y = m(torch.from_numpy(mf).cuda()).cpu().numpy()
    err = np.abs(y - ref).max() / np.abs(ref).max()
    assert err <= TOL[m.native_path], (n_layers, pool, maps, m.native_path, err)


@pytest.mark.parametrize("G,K", [("1", "3"), ("2", "2"), ("4", "4")])
@pytest.mark.parametrize("name,n", [("res8", 1), ("res8", 7), ("res8", 2500), ("res26", 3), ("res26", 1000),
                                    ("res8-narrow", 1301), ("res26-narrow", 66)])
def test_pooled_stacked_groups(name, n, G, K, monkeypatch):
    # pooled res8 / res26 on the persistent kernel: S clips stacked along time per 128-row tile (res8: 4, res26: 2).
    # Ragged last groups, several waves of groups per team (slot reuse) and every team size, row for row against
    # the FP32 CUDA-core path (itself pinned to the golden logits above).
    import honk_b200 as hb
    monkeypatch.setenv("HONK_TC2_G", G)
    monkeypatch.setenv("HONK_TC2_K", K)
    cfg = dict(hb.find_config(name))
    cfg["n_labels"] = 12
    base = synth.make_clips(64, start=7000)
    pre = hb.AudioPreprocessor()
    mf64 = pre.compute_mfccs_batch(torch.from_numpy(base).cuda())
    sd = synth.make_state_dict(name, cfg, seed=11, calib_mfcc=mf64[:16].cpu().numpy())
    reps = n // 64 + 1
    mf = mf64.repeat(reps, 1, 1)[:n].contiguous()
    mf[1::3] = mf[1::3].flip(1) * 0.5                     # neighbouring clips of a group differ
    m32 = build(name, cfg, sd, 1)
    m16 = build(name, cfg, sd, 0)
    ref = m32(mf).cpu().numpy()
    y = m16(mf).cpu().numpy()
    assert m16.native_path == "tcgen05-f16" and m32.native_path == "cuda-core-f32"
    err = np.abs(y - ref).max() / np.abs(ref).max()
    assert err <= TOL["tcgen05-f16"], (name, n, err)
    # the same clip gives the same logits wherever it sits in a group and whichever team ran it
    if n > 64 * 3:
        assert np.abs(y[0] - y[192]).max() <= 1e-4 * np.abs(ref).max()


@pytest.mark.parametrize("name,n", [("res26", 1), ("res26", 5), ("res26", 74), ("res26", 75), ("res8", 2), ("res8", 148),
                                    ("res8", 149), ("res8-narrow", 9), ("res15", 1), ("res15", 37), ("res15", 38)])
def test_request_sized_batches_latency_mode(name, n):
    # small batches switch the persistent kernel to one group per team (K = 1) and, for the pooled models, to teams
    # of four CTAs per stacked group; sizes on both sides of each switch-over, against the FP32 path
    import honk_b200 as hb
    cfg = dict(hb.find_config(name))
    cfg["n_labels"] = 12
    base = synth.make_clips(max(16, min(n, 64)), start=8100)
    pre = hb.AudioPreprocessor()
    mf64 = pre.compute_mfccs_batch(torch.from_numpy(base).cuda())
    sd = synth.make_state_dict(name, cfg, seed=12, calib_mfcc=mf64[:16].cpu().numpy())
    mf = mf64.repeat(n // len(base) + 1, 1, 1)[:n].contiguous()
    mf[1::2] = mf[1::2].flip(2) * 0.7
    m32, m16 = build(name, cfg, sd, 1), build(name, cfg, sd, 0)
    ref, y = m32(mf).cpu().numpy(), m16(mf).cpu().numpy()
    assert m16.native_path == "tcgen05-f16"
    assert np.abs(y - ref).max() / np.abs(ref).max() <= TOL["tcgen05-f16"], (name, n)


def test_f16_activation_range_and_saturation_counter(capfd, monkeypatch):
    # The tcgen05 path stores activations in f16 and the residual stream is un-normalised (BatchNorm is not applied to it,
    # model.py:112-118), so it grows over the 6 residual adds of res15.  f16 storage saturates at +-65504 instead of
    # overflowing (cvt.rn.satfinite); the debug build of the kernel counts accumulator values beyond that range.
    # Supported dynamic range: |activation| < 6.5e4.  Scaling conv0 and conv2..n by 2^k (conv1 reads the raw conv0 output and
    # inherits the factor; every later conv reads a BatchNorm output, which is scale-free once the running statistics are
    # scaled too) multiplies every stored activation by 2^k and leaves the logits unchanged.  So the logits must not move
    # below the limit - and the counter must fire, not stay silent, above it.
    import honk_b200 as hb
    from oracle import honk_oracle as orc
    monkeypatch.setenv("HONK_TC2_DEBUG", "1")
    cfg = dict(hb.find_config("res15"))
    clips = synth.make_clips(6, start=77)
    mf = orc.AudioPreprocessorOracle().compute_mfccs_batch([orc.pcm16_to_float(c) for c in clips])
    sd = synth.make_state_dict("res15", cfg, seed=9, calib_mfcc=mf)

    def scaled(k):
        out = dict(sd)
        out["conv0.weight"] = sd["conv0.weight"] * np.float32(k)
        for i in range(1, cfg["n_layers"] + 1):
            if i >= 2:
                out["conv%d.weight" % i] = sd["conv%d.weight" % i] * np.float32(k)
            out["bn%d.running_mean" % i] = sd["bn%d.running_mean" % i] * np.float32(k)
            out["bn%d.running_var" % i] = (sd["bn%d.running_var" % i] + np.float32(1e-5)) * np.float32(k * k) - np.float32(1e-5)
        return out

    def run(k):
        capfd.readouterr()
        y = build("res15", cfg, scaled(k), 0)(torch.from_numpy(mf).cuda()).cpu().numpy()
        err = capfd.readouterr().err
        return y, float(err.rsplit("f16 saturations", 1)[1].split()[0])
    ref = orc.res_forward(sd, cfg, mf, torch.float64).numpy()
    peak = max(float(np.abs(v).max()) for v in _res_stream(orc, sd, cfg, mf))
    y1, sat1 = run(1.0)
    assert sat1 == 0 and np.abs(y1 - ref).max() / np.abs(ref).max() <= TOL["tcgen05-f16"]
    k_ok = 2.0 ** np.floor(np.log2(3.0e4 / peak))              # residual stream peaks at <= 3e4: inside the f16 range
    y2, sat2 = run(k_ok)
    assert sat2 == 0 and np.abs(y2 - ref).max() / np.abs(ref).max() <= 2 * TOL["tcgen05-f16"], (k_ok, peak)
    k_bad = 2.0 ** np.ceil(np.log2(2.6e5 / peak))               # peaks at >= 2.6e5: must saturate, and the counter must say so
    _, sat3 = run(k_bad)
    assert sat3 > 0


def _res_stream(orc, sd, cfg, mf):
    """Raw (un-normalised) activations x_i of SpeechResModel.forward (model.py:104-121) as the tcgen05 path stores them."""
    import torch.nn.functional as F
    x = torch.as_tensor(mf, dtype=torch.float64).unsqueeze(1)
    old = None
    for i in range(cfg["n_layers"] + 1):
        d = 1 if i == 0 else int(2 ** ((i - 1) // 3))
        y = F.relu(F.conv2d(x, torch.as_tensor(sd["conv%d.weight" % i], dtype=torch.float64), None, 1, d, d))
        if i == 0:
            old = y
        if i > 0 and i % 2 == 0:
            x = y + old
            old = x
        else:
            x = y
        yield x.numpy()
        if i > 0:
            mu = torch.as_tensor(sd["bn%d.running_mean" % i], dtype=torch.float64)[None, :, None, None]
            var = torch.as_tensor(sd["bn%d.running_var" % i], dtype=torch.float64)[None, :, None, None]
            x = (x - mu) / torch.sqrt(var + 1e-5)


CNN_GEMM = ["cnn-one-stride1", "cnn-one-fpool3", "cnn-one-fstride4", "cnn-one-fstride8", "cnn-tpool2", "cnn-tpool3", "cnn-tstride2",
            "cnn-tstride4", "cnn-tstride8"]


@pytest.mark.parametrize("name", CNN_GEMM)
def test_cnn_family_on_tensor_cores(golden_dir, name, monkeypatch):
    # the nine cnn-* configs other than cnn-trad-pool2 (model.py:385-407) run every layer as a tcgen05 GEMM
    # (gemm_tc.cu).  Golden logits of the reference's own classes, ragged batch sizes (one clip, a partial M tile, more clips
    # than one pass of the im2col workspace holds) against the FP32 CUDA-core path, and the oracle at 12 labels.
    from oracle import honk_oracle as orc
    g, cfg, sd = load_golden(golden_dir, name)
    m16, m32 = build(name, cfg, sd, 0), build(name, cfg, sd, 1)
    x = torch.from_numpy(g["x"]).cuda()
    y = m16(x).cpu().numpy()
    m32(x)
    assert m16.native_path == "tcgen05-f16" and m32.native_path == "cuda-core-f32"
    ref = g[name + "/logits_f64"]
    assert np.abs(y - ref).max() / np.abs(ref).max() <= TOL["tcgen05-f16"]
    rng = np.random.default_rng(7)
    monkeypatch.setenv("HONK_GEMM_CHUNK", "256")                      # several passes without a multi-GB batch
    m16b = build(name, cfg, sd, 0)
    for n in (1, 3, 130, 700):
        xb = torch.from_numpy((rng.standard_normal((n, 101, 40)) * np.abs(g["x"]).std()).astype(np.float32)).cuda()
        y16, y32 = m16b(xb).cpu().numpy(), m32(xb).cpu().numpy()
        assert y16.shape == (n, cfg["n_labels"])
        assert np.abs(y16 - y32).max() / np.abs(y32).max() <= TOL["tcgen05-f16"], (name, n)
        assert np.array_equal(m16b(xb).cpu().numpy(), y16)            # no atomics on this path: deterministic
    cfg12 = dict(cfg)
    cfg12["n_labels"] = 12
    sd12 = synth.make_state_dict(name, cfg12, seed=17)
    clips = synth.make_clips(40, start=6100)
    mf = orc.AudioPreprocessorOracle().compute_mfccs_batch([orc.pcm16_to_float(c) for c in clips])
    ref = orc.cnn_forward(sd12, cfg12, mf, torch.float64).numpy()
    import honk_b200 as hb
    yp = build(name, cfg12, sd12, 0).forward_pcm(torch.from_numpy(clips).cuda(), hb.AudioPreprocessor()).cpu().numpy()
    assert np.abs(yp - ref).max() / np.abs(ref).max() <= TOL["tcgen05-f16"]


@pytest.mark.parametrize("name,n", [("res15", 8192), ("res8", 16384), ("cnn-trad-pool2", 4096), ("cnn-tpool2", 4096)])
def test_full_size_batch_properties(name, n):
    """Size-independent properties at BASELINE.json's batch sizes (where the oracle would take minutes): a clip's logits do
    not depend on its position in the batch or on its neighbours (permutation equivariance; sub-batch == slice of the
    batch), duplicated clips give identical rows, and the fused PCM entry equals front end + model.  Sums that go through
    atomics (the res* global mean, then the folded last BatchNorm's mean - mu cancellation) differ in the last bits from run
    to run - 2.3e-5 of the logit scale measured - hence 1e-4 of the scale for res* and 1e-5 for cnn-*, not equality."""
    import honk_b200 as hb
    cfg = dict(hb.find_config(name))
    cfg["n_labels"] = 12
    pre = hb.AudioPreprocessor()
    base = synth.make_clips(1024, start=500000)
    mf = pre.compute_mfccs_batch(torch.from_numpy(base[:16]).cuda()).cpu().numpy()
    m = build(name, cfg, synth.make_state_dict(name, cfg, seed=9, calib_mfcc=mf), 0)
    clips = np.tile(base, (n // 1024, 1))                                # every clip appears n / 1024 times
    rng = np.random.default_rng(5)
    perm = rng.permutation(n)
    x = torch.from_numpy(clips).cuda()
    y = m.forward_pcm(x, pre)
    tol = (1e-4 if name.startswith("res") else 1e-5) * float(y.abs().max())
    assert torch.isfinite(y).all()
    yp = m.forward_pcm(x[torch.from_numpy(perm).cuda()], pre)
    assert float((yp - y[torch.from_numpy(perm).cuda()]).abs().max()) <= tol             # permutation equivariance
    assert float((y.view(n // 1024, 1024, -1) - y[:1024]).abs().max()) <= tol             # duplicates agree
    sub = m.forward_pcm(x[1000:1000 + 777], pre)                                         # a ragged sub-batch
    assert float((sub - y[1000:1000 + 777]).abs().max()) <= tol
    two_step = m(pre.compute_mfccs_batch(x[:2048]))
    assert float((two_step - y[:2048]).abs().max()) <= tol
