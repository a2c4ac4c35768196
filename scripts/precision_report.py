"""Prints tcgen05-f16 vs FP64-oracle logit errors and top-1 agreement (run on the GPU box)."""
import json, os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import honk_b200 as hb
from honk_b200 import synth
from oracle import honk_oracle as orc

n = int(sys.argv[1]) if len(sys.argv) > 1 else 256
clips = synth.make_clips(n, start=20000)
pre_o = orc.AudioPreprocessorOracle()
mf = pre_o.compute_mfccs_batch([orc.pcm16_to_float(c) for c in clips])
x = torch.from_numpy(mf).cuda()
for name in ["res8-narrow", "res8", "res26-narrow", "res26", "res15-narrow", "res15"]:
    cfg = dict(hb.find_config(name))
    sd = synth.make_state_dict(name, cfg, seed=4, calib_mfcc=mf[:32])
    ref = orc.model_forward(name, sd, cfg, mf, torch.float64).numpy()
    out = {}
    for prec in (1, 0):
        m = hb.find_model(name)(cfg)
        m.load_state_dict({k: torch.from_numpy(np.asarray(v)) for k, v in sd.items()})
        m.precision = prec
        m.cuda().eval()
        y = m(x).cpu().numpy()
        out[m.native_path] = y
    srt = np.sort(ref, 1)
    margin = srt[:, -1] - srt[:, -2]
    line = "%-13s |logit|max %.3f std %.3f margin p1 %.4f p50 %.3f" % (name, np.abs(ref).max(), ref.std(), np.percentile(margin, 1), np.median(margin))
    for path, y in out.items():
        e = np.abs(y - ref)
        line += " | %s: max %.2e rms %.2e rel %.2e top1 %.4f" % (path, e.max(), np.sqrt((e ** 2).mean()), e.max() / np.abs(ref).max(), (y.argmax(1) == ref.argmax(1)).mean())
    print(line, flush=True)
