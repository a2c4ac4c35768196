#!/usr/bin/env python
"""Development aid: wall time of the host-buffer call against the device spans inside it."""
import sys, time
import numpy as np
import torch
sys.path.insert(0, ".")
import honk_b200 as hb                      # noqa: E402
from honk_b200 import synth, _native as N   # noqa: E402

B = 8192
base = synth.make_clips(256, start=7000)
pre = hb.AudioPreprocessor()
cfg = dict(hb.find_config("res15")); cfg["n_labels"] = 12
mf = pre.compute_mfccs_batch(torch.from_numpy(base[:16]).cuda()).cpu().numpy()
sd = synth.make_state_dict("res15", cfg, seed=2, calib_mfcc=mf)
m = hb.find_model("res15")(cfg)
m.load_state_dict({k: torch.from_numpy(np.asarray(v)) for k, v in sd.items()})
m.cuda().eval()
pcm = torch.from_numpy(np.ascontiguousarray(np.tile(base, (B // 256, 1)))).pin_memory()
out = torch.empty((B, 12), dtype=torch.float32).pin_memory()
dev = pcm.cuda()
with torch.no_grad():
    for _ in range(3):
        m.forward_pcm_host(pcm, pre, out=out)
        m.forward_pcm(dev, pre)
    torch.cuda.synchronize()
    for mode in ("host", "device"):
        N.profile_reset(); N.profile_enable(True)
        t0 = time.perf_counter()
        for _ in range(5):
            if mode == "host":
                m.forward_pcm_host(pcm, pre, out=out)
            else:
                m.forward_pcm(dev, pre)
        torch.cuda.synchronize()
        wall = (time.perf_counter() - t0) / 5 * 1e3
        N.profile_enable(False)
        tot = 0.0
        parts = []
        for sp in ("frontend", "res_pack_x0", "res_conv", "res_tail"):
            t, n = N.profile_get(sp)
            tot += t / 5
            parts.append("%s %.2f (%d)" % (sp, t / 5, n // 5))
        print("%s: wall %.2f ms, spans %.2f ms: %s" % (mode, wall, tot, ", ".join(parts)))
    # pure copy time for reference
    buf = torch.empty_like(dev)
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(5):
        buf.copy_(pcm, non_blocking=True)
    torch.cuda.synchronize()
    print("H2D of %d MB alone: %.2f ms" % (pcm.numel() * 2 // 2**20, (time.perf_counter() - t0) / 5 * 1e3))
