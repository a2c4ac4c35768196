"""torch custom ops (``torch.ops.honk_b200.*``) over the C ABI.

PyTorch is plumbing here: it allocates the output/workspace tensors and supplies the
current CUDA stream; the arithmetic happens in libhonk_b200.so.  Handles travel as
Python ints (the opaque pointers of include/honk_b200.h).
"""
from __future__ import annotations

import torch

from . import _native as N


def _prep_pcm(pcm):
    N.require_cuda(pcm, "pcm")
    if pcm.dim() != 2:
        raise ValueError("honk_b200: pcm must be [n_clips, n_samples]")
    if pcm.stride(1) != 1 or (pcm.shape[0] > 1 and pcm.stride(0) < pcm.shape[1]):
        pcm = pcm.contiguous()
    return pcm


def _row_stride(pcm):
    # a size-1 batch dimension may carry any stride (numpy's a[None, :] gives 0)
    return pcm.stride(0) if pcm.shape[0] > 1 else pcm.shape[1]


@torch.library.custom_op("honk_b200::frontend_mfcc", mutates_args=(), device_types="cuda")
def frontend_mfcc(pcm: torch.Tensor, frontend: int, n_frames: int, n_dct: int) -> torch.Tensor:
    pcm = _prep_pcm(pcm)
    out = torch.empty((pcm.shape[0], n_frames, n_dct), dtype=torch.float32, device=pcm.device)
    L = N.lib()
    code = N.pcm_dtype_code(pcm)
    ws_bytes = int(L.honk_frontend_workspace_bytes(frontend, pcm.shape[0], pcm.shape[1], code))
    ws = torch.empty((max(ws_bytes, 1),), dtype=torch.uint8, device=pcm.device)
    with torch.cuda.device(pcm.device):
        N.check(L.honk_frontend_mfcc(frontend, pcm.data_ptr(), code, pcm.shape[0], pcm.shape[1], _row_stride(pcm),
                                     out.data_ptr(), ws.data_ptr(), ws_bytes, N.stream_ptr(pcm.device)))
    return out


@frontend_mfcc.register_fake
def _(pcm, frontend, n_frames, n_dct):
    return pcm.new_empty((pcm.shape[0], n_frames, n_dct), dtype=torch.float32)


@torch.library.custom_op("honk_b200::model_forward", mutates_args=(), device_types="cuda")
def model_forward(mfcc: torch.Tensor, model: int, n_labels: int) -> torch.Tensor:
    N.require_cuda(mfcc, "model input")
    if mfcc.dim() != 3:
        raise ValueError("honk_b200: model input must be [n_clips, frames, coefficients], got %s" % (tuple(mfcc.shape),))
    x = mfcc.to(torch.float32).contiguous()          # inputs may arrive Fortran-ordered (manage_audio.py:41)
    B = x.shape[0]
    out = torch.empty((B, n_labels), dtype=torch.float32, device=x.device)
    L = N.lib()
    ws_bytes = int(L.honk_model_workspace_bytes(model, B))
    ws = torch.empty((max(ws_bytes, 1),), dtype=torch.uint8, device=x.device)
    with torch.cuda.device(x.device):
        # the clip shape travels across the C ABI: a wrong shape is HONK_ERR_INVALID (ValueError), never an out-of-bounds read
        N.check(L.honk_model_forward(model, x.data_ptr(), B, x.shape[1], x.shape[2], out.data_ptr(), ws.data_ptr(), ws_bytes,
                                     N.stream_ptr(x.device)))
    return out


@model_forward.register_fake
def _(mfcc, model, n_labels):
    return mfcc.new_empty((mfcc.shape[0], n_labels), dtype=torch.float32)


@torch.library.custom_op("honk_b200::model_forward_pcm", mutates_args=(), device_types="cuda")
def model_forward_pcm(pcm: torch.Tensor, model: int, frontend: int, n_labels: int) -> torch.Tensor:
    pcm = _prep_pcm(pcm)
    B = pcm.shape[0]
    out = torch.empty((B, n_labels), dtype=torch.float32, device=pcm.device)
    L = N.lib()
    ws_bytes = int(L.honk_model_workspace_bytes_pcm(model, frontend, B, pcm.shape[1]))
    ws = torch.empty((max(ws_bytes, 1),), dtype=torch.uint8, device=pcm.device)
    with torch.cuda.device(pcm.device):
        N.check(L.honk_model_forward_pcm(model, frontend, pcm.data_ptr(), N.pcm_dtype_code(pcm), B, pcm.shape[1],
                                         _row_stride(pcm), out.data_ptr(), ws.data_ptr(), ws_bytes,
                                         N.stream_ptr(pcm.device)))
    return out


@model_forward_pcm.register_fake
def _(pcm, model, frontend, n_labels):
    return pcm.new_empty((pcm.shape[0], n_labels), dtype=torch.float32)


def model_forward_pcm_host(pcm: torch.Tensor, model: int, frontend: int, n_labels: int, device, out=None) -> torch.Tensor:
    """Host PCM ``[B, n_samples]`` (pinned memory overlaps the copies) -> host logits ``[B, n_labels]``.

    One C-ABI call (honk_model_forward_pcm_host): chunked double-buffered H2D copies on a second stream
    under the front-end and conv kernels, then one D2H copy of the logits.  Not a custom op: it takes
    and returns host tensors and synchronises before returning.
    """
    if pcm.is_cuda:
        raise RuntimeError("honk_b200: model_forward_pcm_host takes a host tensor")
    if pcm.dim() == 1:
        pcm = pcm[None]
    if pcm.stride(-1) != 1:
        pcm = pcm.contiguous()
    B = pcm.shape[0]
    if out is None:
        out = torch.empty((B, n_labels), dtype=torch.float32, pin_memory=True)
    if B == 0:
        return out
    L = N.lib()
    code = N.pcm_dtype_code(pcm)
    ws_bytes = int(L.honk_model_workspace_bytes_pcm_host(model, frontend, B, pcm.shape[1], code))
    ws = torch.empty((max(ws_bytes, 1),), dtype=torch.uint8, device=device)
    with torch.cuda.device(device):
        N.check(L.honk_model_forward_pcm_host(model, frontend, pcm.data_ptr(), code, B, pcm.shape[1], _row_stride(pcm),
                                              out.data_ptr(), ws.data_ptr(), ws_bytes, N.stream_ptr(device)))
    return out


@torch.library.custom_op("honk_b200::label_from_logits", mutates_args=(), device_types="cuda")
def label_from_logits(logits: torch.Tensor) -> tuple[torch.Tensor, torch.Tensor]:
    N.require_cuda(logits, "logits")
    z = logits.to(torch.float32).contiguous()
    labels = torch.empty((z.shape[0],), dtype=torch.int32, device=z.device)
    probs = torch.empty((z.shape[0],), dtype=torch.float32, device=z.device)
    with torch.cuda.device(z.device):
        N.check(N.lib().honk_label_from_logits(z.data_ptr(), z.shape[0], z.shape[1], labels.data_ptr(),
                                               probs.data_ptr(), N.stream_ptr(z.device)))
    return labels, probs


def labels_to_host(logits):
    """``(labels int32 [B], probs float32 [B])`` as numpy arrays through ONE device->host copy (the serving path's
    tail: both outputs of honk_label_from_logits share a buffer)."""
    N.require_cuda(logits, "logits")
    z = logits.to(torch.float32).contiguous()
    B = z.shape[0]
    buf = torch.empty((2, B), dtype=torch.int32, device=z.device)
    with torch.cuda.device(z.device):
        N.check(N.lib().honk_label_from_logits(z.data_ptr(), B, z.shape[1], buf[0].data_ptr(), buf[1].data_ptr(),
                                               N.stream_ptr(z.device)))
    h = buf.cpu().numpy()
    return h[0], h[1].view("float32")


@label_from_logits.register_fake
def _(logits):
    return (logits.new_empty((logits.shape[0],), dtype=torch.int32),
            logits.new_empty((logits.shape[0],), dtype=torch.float32))


@torch.library.custom_op("honk_b200::listen", mutates_args=(), device_types="cuda")
def listen(pcm: torch.Tensor, model: int, frontend: int, n_labels: int, window: int, hop: int, keyword: int,
           min_keyword_prob: float) -> tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
    """One device buffer of samples -> per-label probability sums (float64), per-label window counts and the
    first window that ends the ``command_tagging`` loop (-1: none); server.py:99-112 in one C-ABI call."""
    N.require_cuda(pcm, "pcm")
    if pcm.dim() != 1:
        raise ValueError("honk_b200: listen takes one 1-D buffer of samples")
    pcm = pcm.contiguous()
    sums = torch.empty((n_labels,), dtype=torch.float64, device=pcm.device)
    counts = torch.empty((n_labels,), dtype=torch.int32, device=pcm.device)
    first = torch.empty((1,), dtype=torch.int32, device=pcm.device)
    L = N.lib()
    ws_bytes = int(L.honk_listen_workspace_bytes(model, frontend, pcm.shape[0], window, hop))
    ws = torch.empty((max(ws_bytes, 1),), dtype=torch.uint8, device=pcm.device)
    with torch.cuda.device(pcm.device):
        N.check(L.honk_listen(model, frontend, pcm.data_ptr(), N.pcm_dtype_code(pcm), pcm.shape[0], window, hop, keyword,
                              min_keyword_prob, sums.data_ptr(), counts.data_ptr(), first.data_ptr(), ws.data_ptr(),
                              ws_bytes, N.stream_ptr(pcm.device)))
    return sums, counts, first


@listen.register_fake
def _(pcm, model, frontend, n_labels, window, hop, keyword, min_keyword_prob):
    return (pcm.new_empty((n_labels,), dtype=torch.float64), pcm.new_empty((n_labels,), dtype=torch.int32),
            pcm.new_empty((1,), dtype=torch.int32))


@torch.library.custom_op("honk_b200::eval_metrics", mutates_args=(), device_types="cuda")
def eval_metrics(logits: torch.Tensor, labels: torch.Tensor) -> tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
    """``(pred int32 [B], correct int64 [1], loss_sum float64 [1])``: print_eval's argmax == label count and
    CrossEntropyLoss's sum (utils/train.py:41-46, 66-83) in one kernel."""
    N.require_cuda(logits, "logits")
    N.require_cuda(labels, "labels")
    z = logits.to(torch.float32).contiguous()
    y = labels.to(torch.int64).contiguous()
    if y.dim() != 1 or z.dim() != 2 or y.shape[0] != z.shape[0]:
        raise ValueError("honk_b200: eval_metrics takes logits [B, L] and labels [B]")
    pred = torch.empty((z.shape[0],), dtype=torch.int32, device=z.device)
    correct = torch.empty((1,), dtype=torch.int64, device=z.device)
    loss = torch.empty((1,), dtype=torch.float64, device=z.device)
    with torch.cuda.device(z.device):
        N.check(N.lib().honk_eval_metrics(z.data_ptr(), y.data_ptr(), z.shape[0], z.shape[1], pred.data_ptr(),
                                          correct.data_ptr(), loss.data_ptr(), N.stream_ptr(z.device)))
    return pred, correct, loss


@eval_metrics.register_fake
def _(logits, labels):
    return (logits.new_empty((logits.shape[0],), dtype=torch.int32), logits.new_empty((1,), dtype=torch.int64),
            logits.new_empty((1,), dtype=torch.float64))


def load_audio(pcm, lengths=None, shifts=None, noise_gain=None, noise_offset=None, bg_noise=None, input_length=16000):
    """Batched arithmetic of ``SpeechDataset.load_audio`` (utils/model.py:281-303) on the device; the caller
    makes the random draws.  ``pcm`` float32 ``[B, >=1]`` CUDA; returns float32 ``[B, input_length]``."""
    N.require_cuda(pcm, "pcm")
    if pcm.dim() != 2 or pcm.dtype != torch.float32:
        raise ValueError("honk_b200: load_audio takes float32 pcm [B, n]")
    if pcm.stride(1) != 1:
        pcm = pcm.contiguous()
    dev = pcm.device
    B = pcm.shape[0]

    def arg(t, dtype, what):
        if t is None:
            return None
        t = torch.as_tensor(t, dtype=dtype, device=dev).contiguous()
        if t.shape != (B,):
            raise ValueError("honk_b200: %s must have one entry per clip" % what)
        return t
    if lengths is None:
        lengths = torch.full((B,), pcm.shape[1], dtype=torch.int32, device=dev)
    lengths = arg(lengths, torch.int32, "lengths").clamp(max=pcm.shape[1])
    shifts = arg(shifts, torch.int32, "shifts")
    noise_gain = arg(noise_gain, torch.float32, "noise_gain")
    noise_offset = arg(noise_offset, torch.int64, "noise_offset")
    if noise_gain is not None and bg_noise is not None:
        N.require_cuda(bg_noise, "bg_noise")
        bg_noise = bg_noise.to(torch.float32).contiguous()
        if noise_offset is None:
            raise ValueError("honk_b200: noise_offset is required with bg_noise")
        if int(noise_offset.min()) < 0 or int(noise_offset.max()) + input_length > bg_noise.numel():
            raise ValueError("honk_b200: noise window outside bg_noise")
    else:
        bg_noise = None
    out = torch.empty((B, input_length), dtype=torch.float32, device=dev)
    ptr = lambda t: t.data_ptr() if t is not None else None      # noqa: E731
    with torch.cuda.device(dev):
        N.check(N.lib().honk_load_audio(pcm.data_ptr(), B, pcm.stride(0) if B > 1 else pcm.shape[1], ptr(lengths), ptr(shifts),
                                        ptr(noise_gain), ptr(noise_offset), ptr(bg_noise), input_length, out.data_ptr(),
                                        N.stream_ptr(dev)))
    return out


_ws_cache = __import__("threading").local()


def model_label_pcm_host(rows, model: int, frontend: int, device):
    """Host int16/float32/float64 numpy ``[B, n_samples]`` -> ``(labels int32 [B], probs float32 [B])`` numpy arrays
    through ONE C-ABI call (honk_model_label_pcm_host: copy in, front end, model, softmax/argmax, copy out, one
    synchronisation).  The serving path: no custom-op dispatch, the device scratch is cached per thread."""
    import numpy as np
    if rows.ndim != 2 or not rows.flags.c_contiguous:
        raise ValueError("honk_b200: label_pcm_host takes a C-contiguous [B, n_samples] array")
    code = {np.dtype(np.int16): 0, np.dtype(np.float32): 1, np.dtype(np.float64): 2}.get(rows.dtype)
    if code is None:
        raise ValueError("honk_b200: PCM must be int16, float32 or float64")
    B, n = rows.shape
    labels = np.empty((B,), dtype=np.int32)
    probs = np.empty((B,), dtype=np.float32)
    if B == 0:
        return labels, probs
    L = N.lib()
    need = int(L.honk_model_workspace_bytes_pcm_host(model, frontend, B, n, code))
    ws = getattr(_ws_cache, "ws", None)
    if ws is None or ws.device != device or ws.numel() < need:
        ws = torch.empty((max(need, 1 << 20),), dtype=torch.uint8, device=device)
        _ws_cache.ws = ws
    with torch.cuda.device(device):
        N.check(L.honk_model_label_pcm_host(model, frontend, rows.ctypes.data, code, B, n, n, labels.ctypes.data,
                                            probs.ctypes.data, ws.data_ptr(), ws.numel(), N.stream_ptr(device)))
    return labels, probs
