#!/usr/bin/env python
"""Front end alone: MFCC of 8192 device-resident int16 clips, CUDA-event timing (development aid)."""
import sys
import numpy as np
import torch
sys.path.insert(0, ".")
import honk_b200 as hb                      # noqa: E402
from honk_b200 import synth                 # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
base = synth.make_clips(256, start=7000)
pcm = torch.from_numpy(np.ascontiguousarray(np.tile(base, (B // 256 + 1, 1))[:B])).cuda()
pre = hb.AudioPreprocessor()
for _ in range(3):
    pre.compute_mfccs_batch(pcm)
torch.cuda.synchronize()
ts = []
for _ in range(5):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        pre.compute_mfccs_batch(pcm)
    e1.record()
    torch.cuda.synchronize()
    ts.append(e0.elapsed_time(e1) / 5)
print("front end %d clips: min %.3f ms, median %.3f ms  (%.2f M clips/s)" % (B, min(ts), sorted(ts)[2], B / min(ts) / 1e3))
