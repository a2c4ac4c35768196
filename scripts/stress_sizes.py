#!/usr/bin/env python
"""Odd batch sizes through the persistent kernel, tcgen05 path against the FP32 path (hang / race hunt).
python scripts/stress_sizes.py"""
import sys
import numpy as np
import torch
sys.path.insert(0, ".")
import honk_b200 as hb                      # noqa: E402
from honk_b200 import synth                 # noqa: E402

pre = hb.AudioPreprocessor()
base = synth.make_clips(128, start=500)
mf0 = pre.compute_mfccs_batch(torch.from_numpy(base).cuda())
worst = 0.0
for name, sizes in (("res15", (73, 74, 75, 147, 149, 1111, 2957)), ("res26", (295, 297, 593, 3001)), ("res8", (591, 1185, 4099)),
                    ("res15-narrow", (223, 1500)), ("cnn-trad-pool2", (147, 149, 1000)),
                    ("cnn-tpool2", (1, 2, 147, 149, 1000, 4100)), ("cnn-tstride2", (3, 148, 297, 5000)), ("cnn-tstride4", (1, 151, 4097)),
                    ("cnn-tpool3", (5, 150, 1025)), ("cnn-tstride8", (149, 1000)), ("cnn-one-stride1", (149, 1000))):
    cfg = dict(hb.find_config(name)); cfg["n_labels"] = 12
    sd = synth.make_state_dict(name, cfg, seed=21, calib_mfcc=mf0[:16].cpu().numpy())
    ms = []
    for prec in (0, 1):
        m = hb.find_model(name)(cfg)
        m.load_state_dict({k: torch.from_numpy(np.asarray(v)) for k, v in sd.items()})
        m.precision = prec
        ms.append(m.cuda().eval())
    for n in sizes:
        mf = mf0.repeat(n // 128 + 1, 1, 1)[:n].contiguous()
        mf[1::2] = mf[1::2].flip(1) * 0.8
        ref = ms[1](mf).cpu().numpy()
        for rep in range(3):
            y = ms[0](mf).cpu().numpy()
            err = float(np.abs(y - ref).max() / np.abs(ref).max())
            worst = max(worst, err)
            assert err < 3e-3, (name, n, rep, err)
        print(name, n, "ok", "%.2e" % err, flush=True)
print("worst relative error %.2e" % worst)
