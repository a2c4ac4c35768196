import sys, torch, numpy as np
sys.path.insert(0, "/root/repo")
import honk_b200 as hb
from honk_b200 import synth
pre = hb.AudioPreprocessor()
big = torch.from_numpy(np.tile(synth.make_clips(256, start=0), (16, 1))).cuda()
for _ in range(2):
    pre.compute_mfccs_batch(big)
torch.cuda.synchronize()
