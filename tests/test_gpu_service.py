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
