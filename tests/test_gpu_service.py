"""TorchLabelService-shaped calls (service.py:72-104, server.py:99-112) through the drop-in."""
import numpy as np
import pytest
import torch

from honk_b200 import synth

pytestmark = pytest.mark.gpu
LABELS = ["_silence_", "_unknown_"] + ["w%d" % i for i in range(10)]


def test_label_matches_oracle():
    import honk_b200 as hb
    from honk_b200.service import TorchLabelService
    from oracle import honk_oracle as orc
    cfg = dict(hb.find_config("cnn-trad-pool2"))
    cfg["n_labels"] = 12
    sd = synth.make_state_dict("cnn-trad-pool2", cfg, seed=5)
    tsd = {k: torch.from_numpy(np.asarray(v)) for k, v in sd.items()}
    svc = TorchLabelService(tsd, labels=LABELS)
    assert hb.find_config("cnn-trad-pool2")["n_labels"] == 12        # shared dict mutated like service.py:82
    pre = orc.AudioPreprocessorOracle()
    clips = synth.make_clips(6, start=700)
    for c in clips:
        label, prob = svc.label(c.tobytes())
        mf = pre.compute_mfccs(orc.pcm16_to_float(c)).squeeze(2)[None]
        z = orc.cnn_forward(sd, cfg, mf, torch.float64).numpy()[0]
        idx, p = orc.softmax_label(z)
        assert label == LABELS[idx] and abs(prob - p) < 1e-3
    hb.find_config("cnn-trad-pool2")["n_labels"] = 4


def test_listen_windows_and_model_type():
    import honk_b200 as hb
    from honk_b200.service import TorchLabelService, stride
    cfg = dict(hb.find_config("res8-narrow"))
    sd = synth.make_state_dict("res8-narrow", cfg, seed=6)
    tsd = {k: torch.from_numpy(np.asarray(v)) for k, v in sd.items()}
    svc = TorchLabelService(tsd, labels=LABELS, model_type=hb.ConfigType.RES8_NARROW)
    long = np.concatenate([synth.make_clip(i, seed=5) for i in range(3)]).tobytes()     # 3 s
    wins = list(stride(long, 16000, 32000))
    assert len(wins) == 5
    agg = svc.listen(long)
    per = [svc.label(w) for w in wins]
    ref = {}
    for lab, p in per:
        ref[lab] = ref.get(lab, 0.0) + p
    assert set(agg) == set(ref) and all(abs(agg[k] - ref[k]) < 1e-5 for k in ref)
    # the in-place device path (honk_listen: overlapping rows of one buffer) == the host-window path, for every
    # stride, both methods, thresholds on either side of the winning probabilities, and the degenerate lengths
    for ms in (500, 250, 100, 37):
        step = int(2 * 16000 * ms / 1000)
        host = svc._listen_host(long, step, 0.85, "command", "all_label")
        dev = svc.listen(long, stride_ms=ms)
        assert set(dev) == set(host) and all(abs(dev[k] - host[k]) < 1e-5 for k in host), ms
    top = max(per, key=lambda lp: lp[1])
    for thr in (top[1] - 1e-3, top[1] + 1e-3):
        for kw in (top[0], "not-a-label"):
            host = svc._listen_host(long, 16000, thr, kw, "command_tagging")
            assert svc.listen(long, min_keyword_prob=thr, keyword=kw, method="command_tagging") == host
    one, one_host = svc.listen(long[:32000]), svc._listen_host(long[:32000], 16000, 0.85, "x", "all_label")
    assert svc.listen(long[:31998]) == {} and set(one) == set(one_host) and len(one) == 1
    assert all(abs(one[k] - one_host[k]) < 1e-5 for k in one)      # (the global-mean sums meet by atomics: ~1e-7 run to run)
    assert svc.listen(long[:100], method="command_tagging") == dict(contains_command=False)
    assert svc.listen(long[:32001]).keys() == svc.listen(long[:32000]).keys()
    with pytest.raises(RuntimeError):
        TorchLabelService(tsd, no_cuda=True, labels=LABELS, model_type=hb.ConfigType.RES8_NARROW)


def _oracle_label_fn(orc, name, sd, cfg, labels, log=None):
    """``TorchLabelService.label`` restated on the CPU (service.py:89-104): int16 / 32768 -> float64 MFCC -> model in
    float64 -> softmax -> (label, prob).  ``log`` collects every window's full probability vector."""
    pre = orc.AudioPreprocessorOracle()

    def label(wav_bytes):
        x = np.frombuffer(wav_bytes, dtype=np.int16) / 32768.0
        mf = pre.compute_mfccs(x).squeeze(2)[None]
        z = orc.model_forward(name, sd, cfg, mf, torch.float64).numpy()[0]
        idx, p = orc.softmax_label(z)
        if log is not None:
            e = np.exp(z - z.max())
            log.append(e / e.sum())
        return labels[idx], p
    return label


def _check_all_label(got, probs, labels, tol, amb):
    """``got`` (label -> summed winning probability) against the oracle's per-window probability vectors.  Windows whose
    runner-up is within ``amb`` of the winner may legitimately go either way on the f16 path (seeded random weights put
    some windows within 1e-3 of a tie); every other window must be attributed exactly as server.py:105-108 does."""
    lo = {l: 0.0 for l in labels}
    hi = {l: 0.0 for l in labels}
    for p in probs:
        top = float(p.max())
        near = [i for i in range(len(p)) if p[i] >= top - amb]
        if len(near) == 1:
            lo[labels[near[0]]] += top
            hi[labels[near[0]]] += top
        else:
            for i in near:
                hi[labels[i]] += top
    for l in labels:
        g = got.get(l, 0.0)
        assert lo[l] - tol * max(1.0, lo[l]) <= g <= hi[l] + tol * max(1.0, hi[l]), (l, g, lo[l], hi[l])
        if lo[l] > tol:
            assert l in got


@pytest.mark.parametrize("precision", [1, 0])
@pytest.mark.parametrize("name", ["res8-narrow", "cnn-trad-pool2"])
def test_listen_matches_oracle_listen(name, precision):
    # the device path (honk_listen: windows read in place, listen_reduce_kernel, early exit) against the
    # reference's loop restated in oracle.listen (server.py:99-112 over service.py:106-110) with an ORACLE label function
    # (float64 MFCC + float64 model + softmax), not against the product's own host path.  precision 1 = all-FP32 CNN
    # (tight bound), 0 = the f16-operand tcgen05 CNN (probabilities to ~1e-3, near-tie windows may flip).
    import honk_b200 as hb
    from honk_b200.service import TorchLabelService
    from oracle import honk_oracle as orc
    cfg = dict(hb.find_config(name))
    cfg["n_labels"] = 12
    clips = [synth.make_clip(900 + i, seed=5) for i in range(4)]
    long = np.concatenate(clips)[: 16000 * 3 + 4321].tobytes()          # 3.27 s: the tail is shorter than one hop
    mf = orc.AudioPreprocessorOracle().compute_mfccs_batch([orc.pcm16_to_float(c) for c in clips])
    sd = synth.make_state_dict(name, cfg, seed=21, calib_mfcc=mf)
    tsd = {k: torch.from_numpy(np.asarray(v)) for k, v in sd.items()}
    keep = hb.find_config(name)["n_labels"]
    svc = TorchLabelService(tsd, labels=LABELS, model_type=hb.ConfigType(name))
    svc.model.precision = precision
    tol, amb = (5e-3, 1.5e-2) if precision == 0 else (2e-5, 2e-4)      # f16-operand CNN: probabilities to ~3e-3
    # 10.25 ms: a hop (164 samples) that is no multiple of 8; 500.03125 ms: an ODD byte step (16001), which the reference's
    # byte-wise stride() turns into windows at odd byte offsets (host-window path of the drop-in)
    for ms in (500, 250, 100, 37, 10.25, 500.03125, 1000, 1):
        buf = long[: 2 * 16000 + 2 * 16 * 20] if ms == 1 else long       # 1 ms hop = 16 samples: keep it to 21 windows
        probs = []
        ref = orc.listen(_oracle_label_fn(orc, name, sd, cfg, LABELS, probs), buf, stride_ms=ms)
        got = svc.listen(buf, stride_ms=ms)
        assert len(probs) == 1 + (len(buf) - 32000) // int(2 * 16000 * ms / 1000)
        _check_all_label(got, probs, LABELS, tol, amb)
        if all(np.sort(p)[-1] - np.sort(p)[-2] > amb for p in probs):    # no near-tie window: the dicts agree key for key
            assert set(got) == set(ref) and all(abs(got[k] - ref[k]) <= tol * max(1.0, ref[k]) for k in ref), (ms, got, ref)
    # command_tagging (server.py:109-110): the loop ends at the first window whose label is the keyword with prob >= threshold
    probs = []
    label_fn = _oracle_label_fn(orc, name, sd, cfg, LABELS, probs)
    orc.listen(label_fn, long)
    clear = [(int(p.argmax()), float(p.max())) for p in probs if np.sort(p)[-1] - np.sort(p)[-2] > amb]
    assert clear
    kw_idx, kw_p = max(clear, key=lambda t: t[1])
    for thr in (kw_p - 2 * amb, kw_p + 2 * amb, 0.0, 1.1):
        if any(abs(float(p.max()) - thr) < amb for p in probs):
            continue                                                      # a window sits on the threshold: either answer is right
        for kw in (LABELS[kw_idx], "not-a-label"):
            if kw != "not-a-label" and any(p[kw_idx] >= p.max() - amb and p.argmax() != kw_idx for p in probs):
                continue
            ref = orc.listen(label_fn, long, min_keyword_prob=thr, keyword=kw, method="command_tagging")
            got = svc.listen(long, min_keyword_prob=thr, keyword=kw, method="command_tagging")
            assert got == ref, (thr, kw)
    # degenerate lengths: shorter than one window, exactly one window, one byte more, an odd byte count
    for n in (100, 31998, 32000, 32001, 48001):
        probs = []
        ref = orc.listen(_oracle_label_fn(orc, name, sd, cfg, LABELS, probs), long[:n])
        got = svc.listen(long[:n])
        assert len(probs) == (0 if n < 32000 else 1 + (n - 32000) // 16000) and (ref == {}) == (got == {})
        _check_all_label(got, probs, LABELS, tol, amb)
        assert svc.listen(long[:n], method="command_tagging", keyword="not-a-label") == dict(contains_command=False)
    hb.find_config(name)["n_labels"] = keep


def test_concurrent_threads_and_streams():
    # server.py:144-146: CherryPy calls label() from a thread pool with one shared model.  Four threads, each on its own
    # CUDA stream, hammer label_batch (request-sized: persistent kernel in latency mode), forward_pcm (device entry) and
    # label() of the same service; every result must equal the single-threaded one.
    import threading
    import honk_b200 as hb
    from honk_b200.service import TorchLabelService
    cfg = dict(hb.find_config("res8"))
    cfg["n_labels"] = 12
    base = synth.make_clips(96, start=5100)
    pre0 = hb.AudioPreprocessor()
    mf = pre0.compute_mfccs_batch(torch.from_numpy(base[:16]).cuda()).cpu().numpy()
    sd = synth.make_state_dict("res8", cfg, seed=31, calib_mfcc=mf)
    tsd = {k: torch.from_numpy(np.asarray(v)) for k, v in sd.items()}
    svc = TorchLabelService(tsd, labels=LABELS, model_type=hb.ConfigType.RES8)
    reqs = [c.tobytes() for c in base]
    want_small = svc.label_batch(reqs[:5])
    want_big = svc.model.forward_pcm(torch.from_numpy(base).cuda(), svc.audio_processor).cpu().numpy()
    want_one = [svc.label(r) for r in reqs[:3]]
    errors = []

    def worker(k):
        try:
            st = torch.cuda.Stream()
            with torch.cuda.stream(st):
                x = torch.from_numpy(base).cuda()
                for it in range(12):
                    got = svc.label_batch(reqs[:5])
                    assert [g[0] for g in got] == [w[0] for w in want_small]
                    assert np.allclose([g[1] for g in got], [w[1] for w in want_small], atol=1e-5)
                    y = svc.model.forward_pcm(x, svc.audio_processor)
                    st.synchronize()
                    assert np.allclose(y.cpu().numpy(), want_big, rtol=0, atol=1e-4 * np.abs(want_big).max())
                    lab = svc.label(reqs[(k + it) % 3])
                    assert lab[0] == want_one[(k + it) % 3][0] and abs(lab[1] - want_one[(k + it) % 3][1]) < 1e-5
        except BaseException as e:      # noqa: BLE001 - reported to the main thread
            errors.append((k, repr(e)))

    threads = [threading.Thread(target=worker, args=(k,)) for k in range(4)]
    for t in threads:
        t.start()
    for t in threads:
        t.join(300)
    assert not errors, errors
    # a model whose handle was never built: the first requests arrive concurrently (_ensure_native race)
    m = hb.find_model("res8-narrow")(dict(hb.find_config("res8-narrow")))
    m.cuda().eval()
    x = pre0.compute_mfccs_batch(torch.from_numpy(base[:4]).cuda())
    outs, handles = [None] * 6, [None] * 6

    def first(k):
        with torch.cuda.stream(torch.cuda.Stream()):
            outs[k] = m(x).cpu().numpy()
            handles[k] = m._native
    ts = [threading.Thread(target=first, args=(k,)) for k in range(6)]
    for t in ts:
        t.start()
    for t in ts:
        t.join(120)
    assert len(set(handles)) == 1 and handles[0] is not None
    # (the global-mean sums of the tcgen05 path meet by atomics: ~1e-7 relative run to run, not bit-equal)
    assert all(np.allclose(o, outs[0], rtol=0, atol=1e-5 * np.abs(outs[0]).max()) for o in outs)


def test_label_nan_and_shape_errors():
    # one NaN logit makes every softmax entry NaN: np.argmax names entry 0, max() is NaN (service.py:103-104); the drop-in
    # returns that instead of an out-of-range index.  Wrong input shapes are refused, never read out of bounds.
    import honk_b200 as hb
    from honk_b200 import ops
    z = torch.randn(5, 12, device="cuda")
    z[1, 7] = float("nan")
    z[3, :] = float("nan")
    idx, p = ops.label_from_logits(z)
    idx, p = idx.cpu().numpy(), p.cpu().numpy()
    assert idx[1] == 0 and idx[3] == 0 and np.isnan(p[1]) and np.isnan(p[3])
    sm = torch.softmax(z.double(), 1).cpu().numpy()
    for r in (0, 2, 4):
        assert idx[r] == sm[r].argmax() and abs(p[r] - sm[r].max()) < 1e-6
    zi = torch.full((2, 4), float("-inf"), device="cuda")
    zi[1, 2] = 0.0
    idx, p = ops.label_from_logits(zi)
    assert int(idx[1]) == 2 and 0 <= int(idx[0]) < 4
    m = hb.find_model("res8-narrow")(dict(hb.find_config("res8-narrow"))).cuda().eval()
    good = torch.zeros(3, 101, 40, device="cuda")
    assert m(good).shape == (3, 12)
    for bad in (torch.zeros(101, 40, device="cuda"), torch.zeros(3, 40, 101, device="cuda"), torch.zeros(3, 99, 40, device="cuda"),
                torch.zeros(3, 101, 20, device="cuda"), torch.zeros(3, 1, 101, 40, device="cuda")):
        with pytest.raises(ValueError):
            m(bad)
    with pytest.raises(ValueError):
        ops.model_forward(torch.zeros(3, 50, 40, device="cuda"), m.warm(), 12)       # the C ABI itself checks the clip shape
    assert m(good).shape == (3, 12)                                                   # the context survived


@pytest.mark.parametrize("n", [1, 3, 700])
def test_label_batch_single_call_matches_two_step(n):
    # honk_model_label_pcm_host (one native call: copy, front end, model, softmax/argmax, copy back) == forward_pcm +
    # label_from_logits on device buffers; n = 700 takes the chunked two-stream path
    import honk_b200 as hb
    from honk_b200 import ops
    from honk_b200.service import TorchLabelService
    cfg = dict(hb.find_config("res8-narrow"))
    sd = synth.make_state_dict("res8-narrow", cfg, seed=8)
    tsd = {k: torch.from_numpy(np.asarray(v)) for k, v in sd.items()}
    svc = TorchLabelService(tsd, labels=LABELS, model_type=hb.ConfigType.RES8_NARROW)
    base = synth.make_clips(50, start=1200)
    clips = [base[i % 50][: 16000 - (i % 7) * 900].tobytes() for i in range(n)]      # some short: zero-padded
    got = svc.label_batch(clips)
    logits = svc.model.forward_pcm(svc._pcm_tensor(clips), svc.audio_processor)
    idx, prob = ops.label_from_logits(logits)
    idx, prob = idx.cpu().numpy(), prob.cpu().numpy()
    assert [g[0] for g in got] == [LABELS[int(i)] for i in idx]
    assert np.allclose([g[1] for g in got], prob, atol=1e-5)
    assert svc.label_batch([]) == []
