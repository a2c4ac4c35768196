#!/usr/bin/env python
"""bench.py - headline benchmark of the honk_b200 hot path (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--model res15] [--clips-per-gpu B]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...
    python bench.py --impl reference ...       # the reference's CPU path on the host cores

One "step" = one pass of the hot path (PCM -> MFCC front end -> CNN -> logits) over one batch
of synthetic 1-s 16 kHz clips.  Workload: BASELINE.json configs[3], res15 with 65536 clips
sharded over 8 GPUs = 8192 clips per GPU, kept fixed per GPU as N grows (weak scaling).
Rank 0 prints ONE JSON line.  See DESIGN.md "Measurement" for the accounting.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

SPANS = ("frontend", "fe_dft", "fe_post", "res_zero", "res_pack_x0", "res_conv0_pool", "res_conv0", "res_conv", "res_tail", "cnn_pack_x8", "cnn_conv",
         "cnn_fc", "f32_model")
FRONTEND_KERNEL = "fe3_dft_kernel + fe3_post_kernel (DFT as exact int8-limb tcgen05 MMAs with int32 accumulators, cta_group::2; then FP64 mel + log + DCT)"
FE_DTYPE = "exact int8-limb DFT (s32 accumulate) + f64 mel/log/DCT"
METRIC = "clips_per_s_mfcc_cnn"
UNIT = "clips/s"
N_UNIQUE = 2048          # distinct synthetic clips generated per rank, tiled to the batch


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="honk_b200", choices=["honk_b200", "reference"])
    ap.add_argument("--model", default="res15")
    ap.add_argument("--clips-per-gpu", type=int, default=8192)
    ap.add_argument("--global-clips", type=int, default=0,
                    help="strong scaling: this many clips in total, split evenly over the ranks (BASELINE configs[2]: "
                         "--model res8 --global-clips 16384); overrides --clips-per-gpu")
    ap.add_argument("--precision", type=int, default=0, help="0 = tcgen05 f16 operands, 1 = all-FP32 CUDA cores")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-seconds", type=float, default=15.0, help="target CPU time of the cpu_baseline sample")
    args = ap.parse_args()
    args.scaling = "weak"
    if args.global_clips > 0:
        world = int(os.environ.get("WORLD_SIZE", "1"))
        args.clips_per_gpu = max(1, args.global_clips // world)
        args.scaling = "strong"
    return args


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return dict(hbm_gbs=float(d["hbm_gbs"]), tf_burst=float(d["bf16_tflops"]),
                    tf_sustained=float(d.get("bf16_tflops_sustained", d["bf16_tflops"])), source="measured")
    return dict(hbm_gbs=6650.0, tf_burst=1590.0, tf_sustained=1400.0, source="fallback")


# ----------------------------------------------------------------------------------------------
# clocks sampler (NVML), runs during the timed region
# ----------------------------------------------------------------------------------------------
class ClockSampler(threading.Thread):
    REASONS = {0x1: "gpu_idle", 0x2: "applications_clocks_setting", 0x4: "sw_power_cap", 0x8: "hw_slowdown",
               0x10: "sync_boost", 0x20: "sw_thermal_slowdown", 0x40: "hw_thermal_slowdown",
               0x80: "hw_power_brake_slowdown", 0x100: "display_clock_setting"}

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index = index
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop_evt = threading.Event()
        self.ok = False
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = int(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
            self.ok = True
        except Exception:
            self.ok = False

    def run(self):
        if not self.ok:
            return
        nv = self.nv
        while not self._stop_evt.is_set():
            try:
                self.samples.append(int(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)))
                try:
                    r = int(nv.nvmlDeviceGetCurrentClocksEventReasons(self.h))
                except Exception:
                    r = int(nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h))
                for bit, name in self.REASONS.items():
                    if r & bit and name != "gpu_idle":
                        self.reasons.add(name)
            except Exception:
                pass
            self._stop_evt.wait(0.05)

    def finish(self):
        self._stop_evt.set()
        if self.is_alive():
            self.join(2.0)
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": ["nvml_unavailable"]}
        return {"sm_mhz": statistics.median(self.samples), "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(self.samples)}


# ----------------------------------------------------------------------------------------------
# synthetic data / weights
# ----------------------------------------------------------------------------------------------
def make_pcm(n, start):
    import numpy as np
    from honk_b200 import synth
    uniq = synth.make_clips(min(n, N_UNIQUE), start=start)
    reps = (n + len(uniq) - 1) // len(uniq)
    return np.ascontiguousarray(np.tile(uniq, (reps, 1))[:n])


TRAINED = {"res15": os.path.join(os.path.dirname(os.path.abspath(__file__)), "tests", "golden", "res15_trained.npz")}


def weights_kind(name):
    return ("briefly trained fixture tests/golden/res15_trained.npz (realistic logit margins, SURVEY 8d)"
            if name in TRAINED and os.path.exists(TRAINED[name]) else "seeded random init with calibrated BatchNorm statistics")


def make_weights(name, cfg, calib_mfcc):
    """Seeded random weights of the architecture; for the headline model the briefly trained fixture, so that the
    top-1 agreement reported next to the CPU baseline is measured at realistic logit margins (random weights put
    every clip within ~1e-3 of a tie).  Weight values do not change the work done."""
    import numpy as np
    from honk_b200 import synth
    if name in TRAINED and os.path.exists(TRAINED[name]) and cfg.get("n_labels") == 12:
        return dict(np.load(TRAINED[name]))
    return synth.make_state_dict(name, cfg, seed=0, calib_mfcc=calib_mfcc)


# ----------------------------------------------------------------------------------------------
# CPU path of the reference (oracle port): per-clip MFCC like collate_fn + torch CPU forward
# ----------------------------------------------------------------------------------------------
def cpu_model_name():
    try:
        with open("/proc/cpuinfo") as f:
            for line in f:
                if line.lower().startswith("model name"):
                    return line.split(":", 1)[1].strip()
    except OSError:
        pass
    import platform
    return platform.processor() or "unknown"


class CpuPath:
    """The reference's CPU implementation of the path, restated in oracle/ (kind "port": the
    reference's librosa 0.6 dependency is not installable and /root/reference does not exist on
    the GPU box).  MFCC one clip at a time like SpeechDataset.collate_fn (model.py:253-265) -
    spread over a PROCESS pool, one worker per core (the per-clip loop holds the GIL) -
    then the model forward in float32, batch 64, all torch threads."""

    def __init__(self, name, cfg, sd):
        import torch
        from oracle import honk_oracle as orc
        self.orc, self.torch = orc, torch
        self.name, self.cfg, self.sd = name, cfg, sd
        self.cores = os.cpu_count() or 1
        try:
            self.cores = len(os.sched_getaffinity(0))
        except Exception:
            pass
        torch.set_num_threads(self.cores)
        self.pool = orc.MfccPool(self.cores)
        self.cpu_model = cpu_model_name()

    def run(self, pcm):
        import numpy as np
        torch = self.torch
        mf = self.pool.compute(pcm)
        outs = []
        with torch.no_grad():
            for i in range(0, len(mf), 64):
                outs.append(self.orc.model_forward(self.name, self.sd, self.cfg, mf[i:i + 64], torch.float32))
        return torch.cat(outs, 0).numpy()

    def timed(self, pcm):
        t0 = time.perf_counter()
        out = self.run(pcm)
        return time.perf_counter() - t0, out

    def close(self):
        self.pool.close()


def cpu_sample_size(cpu, pcm, target_s, lo=64, hi=4096):
    """Picks a sample size whose CPU time is about target_s, from a 64-clip calibration run."""
    t, _ = cpu.timed(pcm[:lo])
    n = int(max(lo, min(hi, target_s / max(t, 1e-6) * lo)))
    return (n // 64) * 64, t


# ----------------------------------------------------------------------------------------------
def run_reference(args):
    """--impl reference: the reference's own CPU path (oracle port) on the host cores."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    import honk_b200.model as hm        # config table only (no CUDA call)
    name = args.model
    cfg = dict(hm.find_config(name))
    cfg["n_labels"] = 12
    pcm = make_pcm(N_UNIQUE, 0)
    from oracle import honk_oracle as orc
    mf = orc.AudioPreprocessorOracle().compute_mfccs_batch([orc.pcm16_to_float(c) for c in pcm[:32]])
    sd = make_weights(name, cfg, mf)
    cpu = CpuPath(name, cfg, sd)
    total = max(1, args.steps + args.warmup)
    per_step_s = min(20.0, 150.0 / total)
    n, _ = cpu_sample_size(cpu, pcm, per_step_s, hi=N_UNIQUE)
    sample = pcm[:n]
    for _ in range(args.warmup):
        cpu.run(sample)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        cpu.run(sample)
    dt = (time.perf_counter() - t0) / max(1, args.steps)
    value = n / dt
    sample_desc = "%d clips per step (bounded sample of the %d-clip/GPU workload), float64 MFCC per clip over a %d-process pool + %s fp32 forward batch 64, torch %d threads" % (
        n, args.clips_per_gpu, cpu.cores, name, cpu.cores)
    cpu.close()
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f64 front end / f32 CNN", "data": "synthetic",
        "config": workload_config(args, args.gpus),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cpu.cores, "cpu_model": cpu.cpu_model, "kind": "port", "sample": sample_desc},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)
    return 0


def workload_config(args, world):
    return {
        "workload": "%s MFCC+CNN inference, %d synthetic 1-s 16 kHz int16 clips per GPU "
                    "(%s)" % (args.model, args.clips_per_gpu,
                              "BASELINE configs[3]: res15 batch 65536 over 8 GPUs = 8192/GPU" if args.model == "res15" else
                              "BASELINE configs[2]: res8 batch 16384 at 1/2/4/8 GPUs" if args.model == "res8" else
                              "not the headline configuration"),
        "clips_per_gpu": args.clips_per_gpu, "global_clips": args.clips_per_gpu * world,
        "n_samples": 16000, "n_labels": 12, "parallelism": "clip shards, %d independent rank(s), one all_gather of logits" % world,
        "l2": "inputs %.0f MB per GPU per step > 126 MB L2 (no flush needed)" % (args.clips_per_gpu * 32000 / 1e6),
        "unique_clips_per_rank": min(N_UNIQUE, args.clips_per_gpu),
        "weights": weights_kind(args.model),
    }


def recorded_traffic(kernel, clips):
    """DRAM read+write bytes per launch of `kernel` from the committed ncu capture (profiles/traffic.json),
    scaled to this run's clips per launch; None when no capture is recorded."""
    try:
        with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "profiles", "traffic.json")) as f:
            t = json.load(f)[kernel]
        return (t["dram_bytes_read"] + t["dram_bytes_write"]) * clips / 8192.0
    except (OSError, KeyError, ValueError):
        return None


def main():
    args = parse_args()
    if args.impl == "reference":
        return run_reference(args)

    import numpy as np
    import torch
    import torch.distributed as dist
    import honk_b200 as hb
    from honk_b200 import _native as N
    from honk_b200 import parallel

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; honk_b200 has no CPU fallback (use --impl reference for the CPU path)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        # NCCL writes its version banner to stdout when the communicator is created: keep fd 1 for the JSON line only
        sys.stdout.flush()
        saved_stdout = os.dup(1)
        os.dup2(2, 1)
        dist.init_process_group("nccl", device_id=dev)
    N.lib()                                            # fails loudly if the extension is missing

    name = args.model
    cfg = dict(hb.find_config(name))
    cfg["n_labels"] = 12
    B = args.clips_per_gpu
    pcm_host = torch.from_numpy(make_pcm(B, rank * B)).pin_memory()
    pcm = pcm_host.to(dev)
    pre = hb.AudioPreprocessor()
    calib = pre.compute_mfccs_batch(pcm[:32]).cpu().numpy()
    sd = make_weights(name, cfg, calib)
    model = hb.find_model(name)(cfg)
    model.load_state_dict({k: torch.from_numpy(np.asarray(v)) for k, v in sd.items()})
    model.precision = args.precision
    model.cuda().eval()

    gatherer = parallel.LogitGatherer(B * world, cfg["n_labels"], dev) if world > 1 else None

    def step_device():
        # front end + CNN on this rank's shard; the all_gather of the logits runs on a side stream into a preallocated
        # [B * world, L] buffer and is consumed one step later, so it never re-synchronises the ranks mid-loop
        logits = model.forward_pcm(pcm, pre)
        return gatherer.submit(logits) if gatherer is not None else logits

    host_out = torch.empty((B, cfg["n_labels"]), dtype=torch.float32).pin_memory()

    def step_e2e():
        # the public host-buffer call: pinned PCM in, logits out (chunked H2D under the kernels, one D2H)
        # (each rank keeps its own shard's logits on its host: the shards are independent)
        return model.forward_pcm_host(pcm_host, pre, out=host_out)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    with torch.no_grad():
        for _ in range(max(args.warmup, 0)):
            out = step_device()
        if gatherer is not None:
            gatherer.flush()
            gatherer.seconds_in_submit = 0.0       # the first call created the NCCL communicator: not part of a step
        barrier()
        # ---------------- timed region (device-resident inputs) ----------------
        N.profile_reset()
        N.profile_enable(True)
        sampler = ClockSampler(local)
        sampler.start()
        launches0 = N.launch_count()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        e0.record()
        for _ in range(args.steps):
            out = step_device()
        if gatherer is not None:
            out = gatherer.flush()                 # the last step's gather completes inside the timed region
        e1.record()
        barrier()
        launches = N.launch_count() - launches0
        clocks = sampler.finish()
        N.profile_enable(False)
        ms = e0.elapsed_time(e1) / max(1, args.steps)
        spans = {}
        for sp in SPANS:
            t, n = N.profile_get(sp)
            if n:
                spans[sp] = (t, n)
        N.profile_reset()
        # ---------------- end-to-end (pinned host PCM -> logits on host) ----------------
        for _ in range(2):
            step_e2e()
        barrier()
        f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        f0.record()
        for _ in range(args.steps):
            host_logits = step_e2e()
        f1.record()
        barrier()
        e2e_wall = (time.perf_counter() - t0) / max(1, args.steps) * 1e3
        e2e_ms = max(f0.elapsed_time(f1) / max(1, args.steps), e2e_wall)

    conv_span = next((k for k in ("res_conv", "cnn_conv", "f32_model") if k in spans), None)
    per_rank = None
    ms_local = ms
    if world > 1:
        # per-rank breakdown: whose step sets the max, and is it the conv kernel (power-cap spread) or something else?
        mine = torch.tensor([ms, e2e_ms, spans.get(conv_span, (0.0, 0))[0] / max(1, args.steps), spans.get("frontend", (0.0, 0))[0] / max(1, args.steps),
                             (clocks.get("sm_mhz") or 0.0), gatherer.seconds_in_submit * 1e3 / max(1, args.steps)],
                            device=dev, dtype=torch.float64)
        allr = torch.empty((world, mine.numel()), device=dev, dtype=torch.float64)
        dist.all_gather_into_tensor(allr, mine)
        allr = allr.cpu().numpy()
        per_rank = {"ms_per_step": [round(float(v), 4) for v in allr[:, 0]], "e2e_ms_per_step": [round(float(v), 4) for v in allr[:, 1]],
                    (conv_span or "conv") + "_ms": [round(float(v), 4) for v in allr[:, 2]], "frontend_ms": [round(float(v), 4) for v in allr[:, 3]],
                    "sm_mhz_median": [float(v) for v in allr[:, 4]], "gather_submit_host_ms": [round(float(v), 4) for v in allr[:, 5]]}
        ms, e2e_ms = float(allr[:, 0].max()), float(allr[:, 1].max())
        lt = torch.tensor([launches], device=dev, dtype=torch.int64)
        dist.all_reduce(lt)
        launches = int(lt[0])

    if world > 1:
        sys.stdout.flush()
        os.dup2(saved_stdout, 1)
        os.close(saved_stdout)
    if rank == 0:
        from oracle import honk_oracle as orc          # checker / CPU baseline only
        peaks = measured_peaks()
        total_clips = B * world
        value = total_clips / (ms * 1e-3)
        step_ms_local = ms_local * args.steps           # rank 0's own timed region: shares are of THIS rank's step

        def tensor_roof(kernel, flops, t_ms, n_launch, traffic_key=None):
            achieved = flops / (t_ms * 1e-3) / 1e12
            tr = recorded_traffic(traffic_key, B) if traffic_key else None
            return {"bound": "tensor", "kernel": kernel, "achieved": achieved, "peak": peaks["tf_sustained"], "unit": "TFLOP/s",
                    "frac": achieved / peaks["tf_sustained"], "frac_burst": achieved / peaks["tf_burst"],
                    "peak_source": peaks["source"] + " bf16 sustained (frac) / burst (frac_burst), MEASURED_PEAKS.json",
                    "traffic": tr, "traffic_source": "profiles/traffic.json: ncu --set full capture of this kernel on this workload, scaled by clips per launch; NOT measured in this run" if tr else None,
                    "avg_launch_ms": t_ms / n_launch, "launches_per_step": n_launch / args.steps, "flops_per_launch": flops / n_launch,
                    "share_of_step": t_ms / step_ms_local}

        # ---- roofline candidates (rank 0's spans); the reported one is the kernel with the largest share of the step ----
        n_layers = cfg.get("n_layers", 0)
        cands = []
        if "res_conv" in spans and name.startswith("res"):
            C = cfg["n_feature_maps"]
            H, W = 101, 40
            if "res_pool" in cfg:
                H, W = H // cfg["res_pool"][0], W // cfg["res_pool"][1]
            flops_layer_clip = 2 * 9 * C * C * H * W                     # unpadded 2*MAC of one C->C 3x3 layer
            t_ms, n_spans = spans["res_conv"]
            persistent = "res_conv0" not in spans          # one launch runs conv0 + all C->C layers (res_tc2.cu)
            kname = (("tc2_res_kernel (persistent: %d 3x3 C->C layers on %d-clip time-stacked tiles, tcgen05 ring-accumulate)"
                      % (n_layers, 113 // (H + 1)) if "res_pool" in cfg else
                      "tc2_res_kernel (persistent: conv0 + %d dilated 3x3 C->C layers, tcgen05 ring-accumulate)" % n_layers)
                     if persistent else "tc_conv_kernel (dilated 3x3 C->C implicit GEMM, tcgen05)")
            cands.append(tensor_roof(kname, flops_layer_clip * n_layers * B * args.steps, t_ms, n_spans if persistent else n_spans * n_layers,
                                     ("tc2_res_kernel" if name == "res15" else "tc2_res_kernel:" + name) if persistent else None))
        elif "cnn_conv" in spans:
            t_ms, n_spans = spans["cnn_conv"]
            cands.append(tensor_roof("cnn_tc_kernel (persistent: conv1 + maxpool + conv2 [+ lin/dnn] per clip, tcgen05 implicit GEMM)",
                                     orc.conv_flops_per_clip(name, cfg) * B * args.steps, t_ms, n_spans))
        elif "f32_model" in spans:
            t_ms, n_spans = spans["f32_model"]
            cands.append(tensor_roof("conv_gemm_f32_kernel (CUDA-core FP32 path)", orc.conv_flops_per_clip(name, cfg) * B * args.steps, t_ms, n_spans))
        fe = None
        if "frontend" in spans:
            t_ms, n_spans = spans["frontend"]
            bytes_clip = 32000 + 101 * 40 * 4                              # int16 PCM in + float32 MFCC out
            gbs = bytes_clip * B * args.steps / (t_ms * 1e-3) / 1e9
            fe = {"bound": "hbm", "kernel": FRONTEND_KERNEL, "achieved": gbs, "peak": peaks["hbm_gbs"], "unit": "GB/s",
                  "frac": gbs / peaks["hbm_gbs"], "peak_source": peaks["source"] + " copy bandwidth, MEASURED_PEAKS.json",
                  "traffic": None, "clips_per_s": B * args.steps / (t_ms * 1e-3), "avg_launch_ms": t_ms / n_spans,
                  "bytes_per_launch": bytes_clip * B * args.steps / n_spans, "share_of_step": t_ms / step_ms_local,
                  "note": "algorithmic HBM bytes (48 160 B/clip) over the front end's span; the kernels are arithmetic-bound by "
                          "construction (dense DFT), see DESIGN.md section 4"}
            cands.append(fe)
        roof = max(cands, key=lambda r: r["share_of_step"]) if cands else None
        others = [r for r in cands if r is not roof]
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": args.scaling,
            "vs_baseline": None, "dtype": FE_DTYPE + " front end; f16 operands / f32 accumulate CNN" if args.precision == 0 else FE_DTYPE + " front end; f32 CNN",
            "data": "synthetic", "config": workload_config(args, world),
            "path": model.native_path, "gpu_launches": launches, "clocks": clocks,
            "e2e": {"value": total_clips / (e2e_ms * 1e-3), "unit": UNIT, "ms_per_step": e2e_ms,
                    "h2d_bytes_per_step": int(pcm_host.numel() * 2), "d2h_bytes_per_step": int(host_logits.numel() * 4)},
            "roofline": roof, "other_kernels": others,
            "spans_ms_per_step": {k: v[0] / args.steps for k, v in spans.items()},
            "algorithmic_flops_per_clip": orc.conv_flops_per_clip(name, cfg),
        }
        if per_rank is not None:
            line["per_rank"] = per_rank
        if world == 1 and not args.no_cpu_baseline:
            pcm_np = pcm_host.numpy()[:N_UNIQUE]
            cpu = CpuPath(name, cfg, sd)
            n, _ = cpu_sample_size(cpu, pcm_np, args.cpu_seconds, hi=N_UNIQUE)
            dt, ref = cpu.timed(pcm_np[:n])
            cpu.close()
            got = host_logits[:n].numpy()
            agree = float((got.argmax(1) == ref.argmax(1)).mean())
            line["cpu_baseline"] = {"value": n / dt, "unit": UNIT, "cores": cpu.cores, "cpu_model": cpu.cpu_model, "kind": "port",
                                    "sample": "%d clips of the same batch: per-clip float64 MFCC over a %d-process pool + %s fp32 forward, batch 64, torch %d threads"
                                              % (n, cpu.cores, name, cpu.cores),
                                    "top1_agreement_vs_gpu": agree,
                                    "max_logit_err_rel": float(np.abs(got - ref).max() / np.abs(ref).max())}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
