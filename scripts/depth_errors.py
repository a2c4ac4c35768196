import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import honk_b200 as hb
from honk_b200 import synth
from oracle import honk_oracle as orc
clips = synth.make_clips(8, start=77)
mf = orc.AudioPreprocessorOracle().compute_mfccs_batch([orc.pcm16_to_float(c) for c in clips])
for maps in (45, 19):
    for nl in range(1, 14):
        cfg = dict(n_labels=12, use_dilation=True, n_layers=nl, n_feature_maps=maps)
        sd = synth.make_state_dict("res-x", cfg, seed=nl, calib_mfcc=mf)
        ref = orc.res_forward(sd, cfg, mf, torch.float64).numpy()
        m = hb.SpeechResModel(cfg); m.load_state_dict({k: torch.from_numpy(np.asarray(v)) for k, v in sd.items()}); m.cuda().eval()
        y = m(torch.from_numpy(mf).cuda()).cpu().numpy()
        print("maps %d layers %2d d=%2d rel err %.2e" % (maps, nl, 2 ** ((nl - 1) // 3), np.abs(y - ref).max() / np.abs(ref).max()), flush=True)
