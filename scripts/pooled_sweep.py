#!/usr/bin/env python
"""Team size G / groups in flight K sweep of the persistent kernel on the pooled models (res8, res26).

    python scripts/pooled_sweep.py > gpurun_out/pooled_sweep.json

HONK_TC2_G / HONK_TC2_K are read when the model handle is created.  CUDA-event timing of
forward_pcm (front end + conv0/pool + persistent kernel + tail), 3 warm-up + 10 timed iterations."""
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, ".")
import honk_b200 as hb                                  # noqa: E402
from honk_b200 import synth                             # noqa: E402


def run(name, B, G, K, pre, base, sd_cache):
    os.environ["HONK_TC2_G"] = str(G)
    os.environ["HONK_TC2_K"] = str(K)
    cfg = dict(hb.find_config(name))
    cfg["n_labels"] = 12
    if name not in sd_cache:
        mf = pre.compute_mfccs_batch(torch.from_numpy(base[:16]).cuda()).cpu().numpy()
        sd_cache[name] = synth.make_state_dict(name, cfg, seed=2, calib_mfcc=mf)
    m = hb.find_model(name)(cfg)
    m.load_state_dict({k: torch.from_numpy(np.asarray(v)) for k, v in sd_cache[name].items()})
    m.cuda().eval()
    pcm = torch.from_numpy(np.ascontiguousarray(np.tile(base, (B // len(base) + 1, 1))[:B])).cuda()
    mf = pre.compute_mfccs_batch(pcm)
    out = {"config": name, "batch": B, "G": G, "K": K, "path": m.native_path}
    for what, fn in (("pcm", lambda: m.forward_pcm(pcm, pre)), ("mfcc", lambda: m(mf))):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10):
            fn()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 10
        out["ms_" + what] = ms
        out["clips_per_s_" + what] = B / ms * 1e3
    return out


def main():
    base = synth.make_clips(1024, start=0)
    pre = hb.AudioPreprocessor()
    cache = {}
    combos = [(1, 2), (1, 3), (2, 2), (2, 3), (2, 4), (4, 2), (4, 4), (4, 6), (4, 8)]
    if len(sys.argv) > 1:
        combos = [tuple(int(v) for v in a.split(",")) for a in sys.argv[1:]]
    for name, B in (("res26", 8192), ("res8", 16384)):
        for G, K in combos:
            print(json.dumps(run(name, B, G, K, pre, base, cache)), flush=True)


if __name__ == "__main__":
    main()
