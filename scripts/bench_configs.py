#!/usr/bin/env python
"""The other BASELINE.json / SURVEY.md 8(d) configurations on one GPU (bench.py times the headline one).

    python scripts/bench_configs.py [config ... | latency] > gpurun_out/r04_configs.json

  (2) cnn-trad-pool2, B = 4096      MFCC + CNN clips/s, device-resident int16 PCM
  (3) res8, B = 16384               same (the R = 1 point of the scaling line)
  (1') res8-narrow, B = 64          the reference's own CPU-runnable case, GPU side
  (5) res26 through TorchLabelService.label_batch, B = 1..32: p50 / p99 latency of the whole call
      (host int16 bytes in, Python (label, prob) tuples out: H2D, front end, CNN, softmax/argmax, D2H, sync)

One JSON object per line.  CUDA-event timing, 3 warm-up + 10 timed iterations for throughput;
200 calls per batch size for latency (wall clock around the synchronous call, like a client sees it).
"""
import json
import os
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


def peaks():
    try:
        with open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")) as f:
            d = json.load(f)
        return float(d.get("bf16_tflops_sustained", d["bf16_tflops"])), "measured"
    except (OSError, KeyError, ValueError):
        return 1400.0, "fallback"


def cpu_reference(name, cfg, sd, clips, seconds=4.0):
    """clips/s of the reference's CPU path (oracle port: per-clip float64 MFCC over a process pool + reference-shaped
    float32 forward on all cores) on a bounded sample of the same clips."""
    cores = len(os.sched_getaffinity(0))
    torch.set_num_threads(cores)
    pool = orc.MfccPool(cores)

    def run(x):
        mf = pool.compute(x)
        with torch.no_grad():
            return torch.cat([orc.model_forward(name, sd, cfg, mf[i:i + 64], torch.float32) for i in range(0, len(mf), 64)]).numpy()
    t0 = time.perf_counter()
    run(clips[:64])
    t = time.perf_counter() - t0
    n = int(max(64, min(len(clips), seconds / max(t, 1e-6) * 64))) // 64 * 64
    t0 = time.perf_counter()
    ref = run(clips[:n])
    dt = time.perf_counter() - t0
    pool.close()
    return {"value": n / dt, "unit": "clips/s", "cores": cores, "kind": "port", "sample": "%d clips" % n}, ref, n


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
        from honk_b200 import _native as N
        N.profile_reset()
        N.profile_enable(True)
        y = m.forward_pcm(pcm, pre)
        torch.cuda.synchronize()
        N.profile_enable(False)
        spans = {k: N.profile_get(k)[0] for k in ("frontend", "res_conv0_pool", "res_conv", "cnn_pack_x8", "cnn_conv", "cnn_fc", "f32_model") +
                 tuple("gt_%s%d" % (k, i) for k in ("pack", "gemm") for i in range(6))
                 if N.profile_get(k)[1]}
    ms = e0.elapsed_time(e1) / 10
    flops = orc.conv_flops_per_clip(name, cfg)
    cpu, ref, n = cpu_reference(name, cfg, sd, base[:min(len(base), B)])
    got = y[:n].cpu().numpy()
    tf, src = peaks()
    return {"config": name, "batch": B, "path": m.native_path, "frontend_path": pre.native_path, "ms_per_step": ms, "clips_per_s": B / ms * 1e3,
            "algorithmic_tflops": flops * B / ms / 1e9, "frac_of_tensor_peak_whole_step": flops * B / ms / 1e9 / tf, "peak_source": src,
            "spans_ms": spans, "dominant_span": max(spans, key=spans.get) if spans else None,
            "cpu_baseline": cpu, "speedup_vs_cpu": B / ms * 1e3 / cpu["value"],
            "max_logit_err_rel_vs_cpu": float(np.abs(got - ref).max() / np.abs(ref).max()),
            "top1_agreement_vs_cpu": float((got.argmax(1) == ref.argmax(1)).mean()),
            "inputs": "device-resident int16 PCM, %d unique clips tiled" % len(base)}


def label_latency(pre, base):
    name = "res26"
    cfg = dict(hb.find_config(name))
    cfg["n_labels"] = 12
    sd = weights(name, cfg, pre, base)
    labels = ["l%d" % i for i in range(12)]
    svc = service.TorchLabelService({k: torch.from_numpy(np.asarray(v)) for k, v in sd.items()}, labels=labels,
                                    model_type=name)
    # the reference's own loop on the host (service.py:89-104): one request at a time, float64 MFCC -> float32 model -> softmax
    opre = orc.AudioPreprocessorOracle()
    torch.set_num_threads(len(os.sched_getaffinity(0)))

    def cpu_label(wav):
        x = np.frombuffer(wav, dtype=np.int16) / 32768.0
        mf = opre.compute_mfccs(x).squeeze(2)[None]
        with torch.no_grad():
            z = orc.model_forward(name, sd, cfg, mf, torch.float32).numpy()[0]
        return orc.softmax_label(z)
    cts = []
    for i in range(40):
        t0 = time.perf_counter()
        cpu_label(base[i % len(base)].tobytes())
        cts.append((time.perf_counter() - t0) * 1e3)
    cts = sorted(cts[5:])
    cpu_ms = {"p50_ms": cts[len(cts) // 2], "p99_ms": cts[-1], "calls": len(cts), "cores": len(os.sched_getaffinity(0)),
              "what": "reference-shaped label() on the host CPU, one request per call (oracle port)"}
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
                    "p50_ms": ts[len(ts) // 2], "p99_ms": ts[int(len(ts) * 0.99) - 1], "calls": len(ts),
                    "clips_per_s_at_p50": B / ts[len(ts) // 2] * 1e3, "cpu_reference_label": cpu_ms,
                    "cpu_reference_ms_for_this_batch": B * cpu_ms["p50_ms"]})
    return out


def main():
    base = synth.make_clips(256, start=7000)
    pre = hb.AudioPreprocessor()
    only = sys.argv[1:]
    runs = [("res8-narrow", 64), ("cnn-trad-pool2", 4096), ("res8", 16384), ("res26", 8192), ("res15-narrow", 8192), ("res26-narrow", 8192)]
    runs += [(n, 4096) for n in ("cnn-one-stride1", "cnn-one-fpool3", "cnn-one-fstride4", "cnn-one-fstride8", "cnn-tpool2", "cnn-tpool3",
                                 "cnn-tstride2", "cnn-tstride4", "cnn-tstride8")]
    for name, B in runs:
        if only and name not in only:
            continue
        print(json.dumps(throughput(name, B, pre, base)), flush=True)
    if only and "latency" not in only:
        return
    for row in label_latency(pre, base):
        print(json.dumps(row), flush=True)


if __name__ == "__main__":
    main()
