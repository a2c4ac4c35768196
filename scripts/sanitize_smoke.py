import sys, numpy as np, torch
sys.path.insert(0, ".")
import honk_b200 as hb
from honk_b200 import synth
pre = hb.AudioPreprocessor()
clips = synth.make_clips(9, start=0)
for name in ("res8-narrow", "res26", "res15", "cnn-trad-pool2"):
    cfg = dict(hb.find_config(name)); cfg["n_labels"] = 12
    sd = synth.make_state_dict(name, cfg, seed=2)
    m = hb.find_model(name)(cfg)
    m.load_state_dict({k: torch.from_numpy(np.asarray(v)) for k, v in sd.items()})
    m.cuda().eval()
    y = m.forward_pcm(torch.from_numpy(clips).cuda(), pre)
    torch.cuda.synchronize()
    print(name, m.native_path, float(y.abs().max()))
