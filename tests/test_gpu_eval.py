"""Evaluation data path on the device: load_audio arithmetic, print_eval / CrossEntropyLoss,
evaluate() over PCM or MFCC batches - against the reference's own outputs (tests/golden/eval_golden.npz) and the
CPU oracle."""
import os

import numpy as np
import pytest
import torch

from honk_b200 import synth

pytestmark = pytest.mark.gpu


def test_load_audio_bit_exact_vs_reference_golden(golden_dir):
    from honk_b200 import ops
    g = np.load(os.path.join(golden_dir, "eval_golden.npz"))
    bg = torch.from_numpy(g["bg_noise"]).cuda()
    params = g["load_audio/params"]
    B = len(params) + 1
    pcm = np.zeros((B, 16000), np.float32)
    lens, shifts, gains, offs, refs = [], [], [], [], []
    for k, (n, shift, noisy, gain, off) in enumerate(params):
        x = g["load_audio/%d/x" % k]
        pcm[k, :len(x)] = x
        pcm[k, len(x):] = 7.0                          # garbage past the clip's length must never be read
        lens.append(len(x)); shifts.append(int(shift)); gains.append(np.float32(gain) if noisy else -1.0); offs.append(int(off))
        refs.append(g["load_audio/%d/y" % k])
    gain, off = g["load_audio/silence/params"]         # silence example: no samples, noise always mixed
    pcm[B - 1] = 3.0
    lens.append(0); shifts.append(0); gains.append(np.float32(gain)); offs.append(int(off))
    refs.append(g["load_audio/silence/y"])
    y = ops.load_audio(torch.from_numpy(pcm).cuda(), lens, shifts, gains, offs, bg).cpu().numpy()
    for k in range(B):
        assert np.array_equal(y[k], refs[k]), k
    # strided rows, defaults (no shift / noise / lengths) = plain copy of the first input_length samples
    wide = torch.from_numpy(np.arange(3 * 20000, dtype=np.float32).reshape(3, 20000) / 1e5).cuda()
    z = ops.load_audio(wide[:, :17000])
    assert torch.equal(z, wide[:, :16000])
    for sh in (-1600, -1, 0, 1, 1600):
        x = torch.from_numpy(g["timeshift/x"])[None].cuda()
        assert np.array_equal(ops.load_audio(x, shifts=[sh]).cpu().numpy()[0], g["timeshift/%d" % sh])
    with pytest.raises(ValueError):
        ops.load_audio(wide, None, None, [0.1, 0.1, 0.1], [0, 0, 40000], bg)


def test_eval_metrics_vs_reference_golden(golden_dir, capsys):
    from honk_b200 import ops, train
    g = np.load(os.path.join(golden_dir, "eval_golden.npz"))
    scores, labels = torch.from_numpy(g["eval/scores"]).cuda(), torch.from_numpy(g["eval/labels"]).cuda()
    pred, correct, loss_sum = ops.eval_metrics(scores, labels)
    assert np.array_equal(pred.cpu().numpy(), g["eval/pred"])           # ties: first maximum, like torch.max
    n = len(g["eval/labels"])
    assert abs(correct.item() / n - float(g["eval/accuracy"])) < 1e-6
    assert abs(loss_sum.item() / n - float(g["eval/loss"])) < 1e-5
    acc = train.print_eval("test", scores, labels)
    assert abs(acc - float(g["eval/accuracy"])) < 1e-6
    assert capsys.readouterr().out.startswith("test accuracy: ")
    big = torch.randn(100003, 12, device="cuda")
    yb = torch.randint(0, 12, (100003,), device="cuda")
    _, c, ls = ops.eval_metrics(big, yb)
    assert c.item() == int((big.argmax(1) == yb).sum())
    assert abs(ls.item() / 100003 - torch.nn.functional.cross_entropy(big.double(), yb).item()) < 1e-9
    # a target outside [0, n_labels) is an error in torch (device assert): not skipped silently here either
    bad = yb[:64].clone()
    bad[5] = 12
    _, _, ls = ops.eval_metrics(big[:64], bad)
    assert np.isnan(ls.item())
    with pytest.raises(IndexError):
        train.eval_metrics(big[:64], bad)
    # NaN scores: torch.max names the first NaN entry (print_eval's argmax, train.py:42)
    zn = big[:4].clone()
    zn[2, 5] = float("nan")
    pred, _, _ = ops.eval_metrics(zn, yb[:4])
    assert int(pred[2]) == 5 and int(pred[0]) == int(zn[0].argmax())


def test_evaluate_pcm_and_mfcc_loaders():
    # utils/train.py:evaluate with a loader of raw PCM batches (fused front end) == a loader of MFCC batches ==
    # the CPU oracle's accuracy
    import honk_b200 as hb
    from honk_b200 import train
    from oracle import honk_oracle as orc
    name = "res8-narrow"
    cfg = dict(hb.find_config(name))
    cfg["n_labels"] = 12
    clips = synth.make_clips(96, start=300)
    mf = orc.AudioPreprocessorOracle().compute_mfccs_batch([orc.pcm16_to_float(c) for c in clips])
    sd = synth.make_state_dict(name, cfg, seed=4, calib_mfcc=mf[:16])
    ref = orc.model_forward(name, sd, cfg, mf, torch.float64).numpy()
    labels = ref.argmax(1)
    labels[::5] = (labels[::5] + 1) % 12                      # 20 of 96 wrong on purpose
    want, _ = orc.eval_metrics(ref, labels)
    m = hb.find_model(name)(cfg)
    m.load_state_dict({k: torch.from_numpy(np.asarray(v)) for k, v in sd.items()})
    conf = dict(no_cuda=False, gpu_no=0)
    pcm_loader = [(torch.from_numpy(clips[i:i + 40]), torch.from_numpy(labels[i:i + 40])) for i in range(0, 96, 40)]
    mf_loader = [(torch.from_numpy(mf[i:i + 32].astype(np.float32)), torch.from_numpy(labels[i:i + 32])) for i in range(0, 96, 32)]
    a1 = train.evaluate(conf, m, pcm_loader)
    a2 = train.evaluate(conf, m, mf_loader)
    assert abs(a1 - want) <= 1.5 / 96 and abs(a2 - want) <= 1.5 / 96 and abs(a1 - a2) < 1e-9
    with pytest.raises(RuntimeError):
        train.evaluate(dict(no_cuda=True), m, pcm_loader)
