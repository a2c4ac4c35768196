#!/usr/bin/env python
"""Training-step throughput (reference utils/train.py:127-139) on one GPU, with the CPU reference
(the oracle's restatement of the same step: torch autograd + the SGD rule, all host threads) timed beside it.

    python scripts/bench_train.py [config:batch ...] > gpurun_out/r04_train.json

One JSON object per line: clips/s of honk_trainer_step (CUDA events, 3 warm-up + 10 timed steps, device-resident MFCCs
and labels), the forward / backward / SGD spans, launches per step, and the CPU figure on a bounded number of steps.
"""
import json
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, ".")
import honk_b200 as hb                                  # noqa: E402
from honk_b200 import _native as N, synth, train        # noqa: E402
from oracle import honk_oracle as orc                   # noqa: E402  (CPU baseline only)

DEFAULT = ["res15:64", "res8:64", "res26:64", "res15-narrow:64", "res8-narrow:64", "res26-narrow:64", "cnn-trad-pool2:64",
           "cnn-one-stride1:64", "cnn-tstride2:64", "cnn-tpool2:64", "res15:256", "cnn-trad-pool2:1024"]


def span(name):
    ms, n = N.C.c_double(), N.C.c_int64()
    N.lib().honk_profile_get(name.encode(), N.C.byref(ms), N.C.byref(n))
    return ms.value, n.value


def main():
    cases = sys.argv[1:] or DEFAULT
    pre = hb.AudioPreprocessor()
    for case in cases:
        name, B = case.split(":")
        B = int(B)
        cfg = dict(hb.find_config(name))
        cfg["n_labels"] = 12
        clips, labels = synth.make_clips(B, start=900000, labelled=True)
        mfcc = pre.compute_mfccs_batch(torch.from_numpy(clips).cuda()).float()
        y = torch.from_numpy(labels.astype(np.int64)).cuda()
        state = synth.make_state_dict(name, cfg, seed=3, calib_mfcc=mfcc[:8].cpu().numpy())
        if name.startswith("cnn"):
            state = {k: (v * np.float32(0.3) if k.endswith("weight") else v) for k, v in state.items()}
        model = hb.find_model(name)(cfg)
        model.load_state_dict({k: torch.from_numpy(np.asarray(v)) for k, v in state.items()})
        model.cuda()
        tr = train.SgdTrainer(model, lr=0.001, momentum=0.9, weight_decay=1e-5)
        for _ in range(3):
            tr.step(mfcc, y, check_labels=False)
        torch.cuda.synchronize()
        N.lib().honk_profile_enable(1)
        N.lib().honk_profile_reset()
        l0 = N.launch_count()
        steps = 10
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            tr.step(mfcc, y, check_labels=False)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / steps
        launches = (N.launch_count() - l0) / steps
        spans = {k: span(k)[0] / steps for k in ("train_fwd", "train_bwd", "train_sgd")}
        N.lib().honk_profile_enable(0)
        # CPU reference: the same step with torch autograd on the host cores
        torch.set_num_threads(os.cpu_count() or 1)
        xb, yb = mfcc.cpu().numpy(), labels.astype(np.int64)
        cpu_steps = 1 if orc.conv_flops_per_clip(name, cfg) * B > 2e10 else 3
        rng = np.random.default_rng(0)
        masks = None
        if name.startswith("cnn"):                                              # the CPU step pays for Dropout too
            shapes = [(B, model.conv1.out_channels, (101 - cfg["conv1_size"][0]) // cfg["conv1_stride"][0] + 1,
                       (40 - cfg["conv1_size"][1]) // cfg["conv1_stride"][1] + 1)]
            masks = [rng.integers(0, 2, shapes[0]).astype(np.float32), None, None, None]
            if "conv2_size" in cfg:
                h1, w1 = shapes[0][2] // cfg["conv1_pool"][0], shapes[0][3] // cfg["conv1_pool"][1]
                masks[1] = rng.integers(0, 2, (B, cfg["n_feature_maps2"], h1 - cfg["conv2_size"][0] + 1, w1 - cfg["conv2_size"][1] + 1)).astype(np.float32)
            if "dnn1_size" in cfg:
                masks[2] = rng.integers(0, 2, (B, cfg["dnn1_size"])).astype(np.float32)
            if "dnn2_size" in cfg:
                masks[3] = rng.integers(0, 2, (B, cfg["dnn2_size"])).astype(np.float32)
        batch = (xb, yb, masks) if masks else (xb, yb)
        orc.sgd_train_steps(state, cfg, [tuple(b[:8] if isinstance(b, np.ndarray) else ([m[:8] if m is not None else None for m in b]) for b in batch)])
        t0 = time.perf_counter()
        orc.sgd_train_steps(state, cfg, [batch] * cpu_steps, lr=0.001, momentum=0.9, weight_decay=1e-5)
        cpu_ms = (time.perf_counter() - t0) * 1e3 / cpu_steps
        flops = 3 * orc.conv_flops_per_clip(name, cfg) * B                      # forward + data gradient + weight gradient
        print(json.dumps({"config": name, "batch": B, "ms_per_step": ms, "clips_per_s": B / ms * 1e3, "spans_ms": spans,
                          "launches_per_step": launches, "algorithmic_tflops": flops / ms / 1e9, "dtype": "f32 (CUDA cores)",
                          "workspace_mb": int(N.lib().honk_trainer_workspace_bytes(tr.handle, B)) / 2 ** 20,
                          "cpu_baseline": {"ms_per_step": cpu_ms, "clips_per_s": B / cpu_ms * 1e3, "cores": os.cpu_count(),
                                           "kind": "port", "sample": "%d step(s) of the same batch" % cpu_steps},
                          "speedup_vs_cpu": cpu_ms / ms}), flush=True)
        del tr, model


if __name__ == "__main__":
    main()
