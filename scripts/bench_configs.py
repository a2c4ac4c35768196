#!/usr/bin/env python
"""The other BASELINE.json / SURVEY.md 8(d) configurations on one GPU (bench.py times the headline one).

    python scripts/bench_configs.py > gpurun_out/r02_configs.json

  (2) cnn-trad-pool2, B = 4096      MFCC + CNN clips/s, device-resident int16 PCM
  (3) res8, B = 16384               same (the R = 1 point of the scaling line)
  (1') res8-narrow, B = 64          the reference's own CPU-runnable case, GPU side
  (5) res26 through TorchLabelService.label_batch, B = 1..32: p50 / p99 latency of the whole call
      (host int16 bytes in, Python (label, prob) tuples out: H2D, front end, CNN, softmax/argmax, D2H, sync)

One JSON object per line.  CUDA-event timing, 3 warm-up + 10 timed iterations for throughput;
200 calls per batch size for latency (wall clock around the synchronous call, like a client sees it).
"""
import json
import sys
import time

import numpy as np
import torch

sys.path.insert(0, ".")
import honk_b200 as hb                                  # noqa: E402
from honk_b200 import synth, service                    # noqa: E402
from oracle import honk_oracle as orc                   # noqa: E402  (flop counts only)


def weights(name, cfg, pre, clips):
    mf = pre.compute_mfccs_batch(torch.from_numpy(clips[:16]).cuda()).cpu().numpy()
    return synth.make_state_dict(name, cfg, seed=2, calib_mfcc=mf)


def throughput(name, B, pre, base):
    cfg = dict(hb.find_config(name))
    cfg["n_labels"] = 12
    sd = weights(name, cfg, pre, base)
    m = hb.find_model(name)(cfg)
    m.load_state_dict({k: torch.from_numpy(np.asarray(v)) for k, v in sd.items()})
    m.cuda().eval()
    pcm = torch.from_numpy(np.ascontiguousarray(np.tile(base, (B // len(base) + 1, 1))[:B])).cuda()
    with torch.no_grad():
        for _ in range(3):
            m.forward_pcm(pcm, pre)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10):
            m.forward_pcm(pcm, pre)
        e1.record()
        torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    flops = orc.conv_flops_per_clip(name, cfg)
    return {"config": name, "batch": B, "path": m.native_path, "ms_per_step": ms, "clips_per_s": B / ms * 1e3,
            "algorithmic_tflops": flops * B / ms / 1e9, "inputs": "device-resident int16 PCM, %d unique clips tiled" % len(base)}


def label_latency(pre, base):
    name = "res26"
    cfg = dict(hb.find_config(name))
    cfg["n_labels"] = 12
    sd = weights(name, cfg, pre, base)
    labels = ["l%d" % i for i in range(12)]
    svc = service.TorchLabelService({k: torch.from_numpy(np.asarray(v)) for k, v in sd.items()}, labels=labels,
                                    model_type=name)
    out = []
    for B in (1, 2, 4, 8, 16, 32):
        req = [base[i % len(base)].tobytes() for i in range(B)]
        for _ in range(20):
            svc.label_batch(req)
        ts = []
        for _ in range(200):
            t0 = time.perf_counter()
            svc.label_batch(req)
            ts.append((time.perf_counter() - t0) * 1e3)
        ts.sort()
        out.append({"config": "res26 TorchLabelService.label_batch", "batch": B, "path": svc.model.native_path,
                    "p50_ms": ts[len(ts) // 2], "p99_ms": ts[int(len(ts) * 0.99) - 1], "calls": len(ts)})
    return out


def main():
    base = synth.make_clips(256, start=7000)
    pre = hb.AudioPreprocessor()
    for name, B in (("res8-narrow", 64), ("cnn-trad-pool2", 4096), ("res8", 16384), ("res26", 8192), ("res15-narrow", 8192)):
        print(json.dumps(throughput(name, B, pre, base)), flush=True)
    for row in label_latency(pre, base):
        print(json.dumps(row), flush=True)


if __name__ == "__main__":
    main()
