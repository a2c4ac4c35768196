#!/usr/bin/env python
"""Where the time of one TorchLabelService.label_batch call goes (res26, B = 1..32): host staging, device time of
the fused forward (CUDA events), label + device->host.  python scripts/latency_probe.py > gpurun_out/latency.json"""
import json
import sys
import time

import numpy as np
import torch

sys.path.insert(0, ".")
import honk_b200 as hb                                  # noqa: E402
from honk_b200 import synth, service, ops               # noqa: E402


def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "res26"
    base = synth.make_clips(64, start=0)
    pre = hb.AudioPreprocessor()
    cfg = dict(hb.find_config(name))
    cfg["n_labels"] = 12
    mf = pre.compute_mfccs_batch(torch.from_numpy(base[:16]).cuda()).cpu().numpy()
    sd = synth.make_state_dict(name, cfg, seed=2, calib_mfcc=mf)
    svc = service.TorchLabelService({k: torch.from_numpy(np.asarray(v)) for k, v in sd.items()},
                                    labels=["l%d" % i for i in range(12)], model_type=name)
    for B in (1, 8, 32):
        req = [base[i % 64].tobytes() for i in range(B)]
        for _ in range(30):
            svc.label_batch(req)
        t_stage, t_dev, t_fwd_host, t_tail, t_all = [], [], [], [], []
        for _ in range(200):
            t0 = time.perf_counter()
            pcm = svc._pcm_tensor(req)
            t1 = time.perf_counter()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            logits = svc.model.forward_pcm(pcm, svc.audio_processor)
            e1.record()
            t2 = time.perf_counter()
            idx, prob = ops.labels_to_host(logits)
            t3 = time.perf_counter()
            t_stage.append(t1 - t0); t_fwd_host.append(t2 - t1); t_tail.append(t3 - t2); t_dev.append(e0.elapsed_time(e1))
            t0 = time.perf_counter()
            svc.label_batch(req)
            t_all.append(time.perf_counter() - t0)
        from honk_b200 import _native as N
        N.profile_reset()
        N.profile_enable(True)
        for _ in range(50):
            svc.label_batch(req)
        N.profile_enable(False)
        spans = {}
        for sp in ("frontend", "res_conv0_pool", "res_pack_x0", "res_conv", "res_tail", "cnn_conv", "cnn_fc"):
            t, n = N.profile_get(sp)
            if n:
                spans[sp] = t / n
        med = lambda v: float(np.median(v))   # noqa: E731
        print(json.dumps({"config": name, "batch": B, "host_staging_ms": med(t_stage) * 1e3, "forward_host_enqueue_ms": med(t_fwd_host) * 1e3,
                          "forward_device_ms": med(t_dev), "label_and_d2h_ms": med(t_tail) * 1e3, "label_batch_ms": med(t_all) * 1e3,
                          "p99_label_batch_ms": float(np.percentile(t_all, 99)) * 1e3, "device_spans_ms": spans}), flush=True)


if __name__ == "__main__":
    main()
