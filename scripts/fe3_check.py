"""Integer-limb tcgen05 front end (fe_i8.cu) against the FP64 DMMA front end and the CPU oracle: error histogram of the
log-mel energies and MFCCs, and kernel timing.  Writes JSON to stdout."""
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import honk_b200 as hb                 # noqa: E402
from honk_b200 import synth, _native as N   # noqa: E402
from oracle import honk_oracle as orc  # noqa: E402

n_time = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
os.environ["HONK_FE_I8"] = "1"
pre_i8 = hb.AudioPreprocessor()
os.environ["HONK_FE_I8"] = "0"
pre_f64 = hb.AudioPreprocessor()
os.environ.pop("HONK_FE_I8")
out = {"paths": [pre_i8.native_path, pre_f64.native_path]}
D = orc.dct_filters(40, 40)


def stats(got, ref, floor):
    lm_g, lm_r = got.astype(np.float64) @ D, ref.astype(np.float64) @ D
    live = lm_r > lm_r.max(axis=-1, keepdims=True) - floor
    e = np.abs(lm_g - lm_r)
    return {"mfcc_max_abs": float(np.abs(got - ref).max()), "logmel_live_max": float(e[live].max()), "live_frac": float(live.mean()),
            "logmel_live_p99": float(np.percentile(e[live], 99))}


sp = synth.special_clips()
names = list(sp)
spec = np.stack([sp[k] for k in names])
ref = orc.AudioPreprocessorOracle().compute_mfccs_batch([orc.pcm16_to_float(c) for c in spec])
x = torch.from_numpy(spec).cuda()
gi, gf = pre_i8.compute_mfccs_batch(x).cpu().numpy(), pre_f64.compute_mfccs_batch(x).cpu().numpy()
out["special"] = {k: {"i8_vs_oracle": stats(gi[i], ref[i], 46.0), "f64_vs_oracle": stats(gf[i], ref[i], 46.0)} for i, k in enumerate(names)}
clips = synth.make_clips(256, start=12000, seed=3)
ref = orc.AudioPreprocessorOracle().compute_mfccs_batch([orc.pcm16_to_float(c) for c in clips])
x = torch.from_numpy(clips).cuda()
gi, gf = pre_i8.compute_mfccs_batch(x).cpu().numpy(), pre_f64.compute_mfccs_batch(x).cpu().numpy()
out["random256"] = {"i8_vs_oracle": stats(gi, ref, 46.0), "f64_vs_oracle": stats(gf, ref, 46.0), "i8_vs_f64_max_abs": float(np.abs(gi - gf).max()),
                    "bit_equal_frac": float((gi == gf).mean())}
hist, edges = np.histogram(np.log10(np.abs(gi - ref).ravel() + 1e-12), bins=np.arange(-12, 0.5, 1.0))
out["random256"]["i8_abs_err_log10_hist"] = {"%d" % e: int(h) for e, h in zip(edges[:-1], hist)}
# odd sizes: ragged tile pairs, clip counts below one tile, short clips
for n, ns in ((1, 16000), (3, 16000), (5, 8000), (130, 16000), (257, 16000), (7, 1600), (2, 320)):
    c = synth.make_clips(n, start=500 + n)[:, :ns].copy()
    xx = torch.from_numpy(c).cuda()
    a, b = pre_i8.compute_mfccs_batch(xx), pre_f64.compute_mfccs_batch(xx)
    out.setdefault("sizes", {})["%dx%d" % (n, ns)] = float((a - b).abs().max())
# strided / overlapping rows (listen windows)
buf = torch.from_numpy(np.concatenate(list(synth.make_clips(3, start=40)))).cuda()
win = buf.as_strided((9, 16000), (4000, 1))
out["overlapping_rows"] = float((pre_i8.compute_mfccs_batch(win) - pre_f64.compute_mfccs_batch(win)).abs().max())
# timing
big = torch.from_numpy(np.tile(synth.make_clips(512, start=0), (n_time // 512, 1))).cuda()
for name, pre in (("i8", pre_i8), ("f64", pre_f64)):
    for _ in range(3):
        pre.compute_mfccs_batch(big)
    torch.cuda.synchronize()
    N.profile_reset(); N.profile_enable(True)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        pre.compute_mfccs_batch(big)
    e1.record()
    torch.cuda.synchronize()
    N.profile_enable(False)
    ms = e0.elapsed_time(e1) / 10
    out["time_" + name] = {"clips": int(big.shape[0]), "ms": ms, "clips_per_s": big.shape[0] / ms * 1e3,
                           "spans_ms": {k: N.profile_get(k)[0] / 10 for k in ("frontend", "fe_dft", "fe_post") if N.profile_get(k)[1]}}
print(json.dumps(out, indent=1))
