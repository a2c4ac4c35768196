"""Drop-in for the reference's ``utils.manage_audio.AudioPreprocessor`` (manage_audio.py:18-42).

Same constructor keywords, same attributes, and ``compute_mfccs(data)`` returns the same
float32 ``(n_frames, n_dct_filters, 1)`` Fortran-ordered array for one 1-D clip.  The
arithmetic runs in libhonk_b200.so on the GPU (FP64, see csrc/frontend.cu); there is no
librosa and no CPU fallback.  ``compute_mfccs_batch`` is the batched extension that
replaces the per-clip loop of ``SpeechDataset.collate_fn`` (model.py:253-265).
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import _native as N
from . import ops


class AudioPreprocessor(object):
    def __init__(self, sr=16000, n_dct_filters=40, n_mels=40, f_max=4000, f_min=20, n_fft=480, hop_ms=10,
                 device=None):
        super().__init__()
        self.n_mels = n_mels
        self.sr = sr
        self.f_max = f_max if f_max is not None else sr // 2
        self.f_min = f_min
        self.n_fft = n_fft
        self.hop_length = sr // 1000 * hop_ms
        self.n_dct_filters = n_dct_filters
        if not torch.cuda.is_available():
            raise RuntimeError("honk_b200: AudioPreprocessor needs a CUDA device (no CPU fallback)")
        self.device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        cfg = N.FrontendConfig(sr, n_fft, self.hop_length, n_mels, n_dct_filters, float(f_min), float(self.f_max))
        h = C.c_void_p()
        N.check(N.lib().honk_frontend_create(C.byref(cfg), self.device.index or 0, C.byref(h)))
        self._handle = h.value
        dct = np.empty((n_dct_filters, n_mels), dtype=np.float64)
        N.check(N.lib().honk_frontend_get_dct(self._handle, dct.ctypes.data))
        self.dct_filters = dct          # same attribute as manage_audio.py:22

    def __del__(self):
        h = getattr(self, "_handle", None)
        if h:
            try:
                N.lib().honk_frontend_destroy(h)
            except Exception:
                pass
            self._handle = None

    @property
    def handle(self):
        return self._handle

    @property
    def native_path(self):
        """How int16 PCM is transformed: "tcgen05-i8-limbs" (exact integer DFT on the tensor cores), "fp64-dmma" or "fp64-dfma"."""
        return N.lib().honk_frontend_path(self._handle).decode()

    def mel_basis(self):
        out = np.empty((self.n_mels, 1 + self.n_fft // 2), dtype=np.float64)
        N.check(N.lib().honk_frontend_get_mel(self._handle, out.ctypes.data))
        return out

    def num_frames(self, n_samples):
        return int(N.lib().honk_frontend_num_frames(self._handle, int(n_samples)))

    def compute_mfccs_batch(self, clips):
        """``[B, n_samples]`` int16 / float32 / float64 (tensor or array) -> CUDA float32 ``[B, n_frames, n_dct]``."""
        if not isinstance(clips, torch.Tensor):
            clips = torch.from_numpy(np.ascontiguousarray(clips))
        if clips.dim() == 1:
            clips = clips.unsqueeze(0)
        if clips.dtype not in (torch.int16, torch.float32, torch.float64):
            clips = clips.to(torch.float64)
        if not clips.is_cuda:
            clips = clips.to(self.device, non_blocking=True)
        return ops.frontend_mfcc(clips, self._handle, self.num_frames(clips.shape[1]), self.n_dct_filters)

    def compute_mfccs(self, data):
        """One 1-D clip -> numpy float32 ``(n_frames, n_dct_filters, 1)``, Fortran-ordered (manage_audio.py:30-42)."""
        data = np.asarray(data)
        if data.ndim != 1:
            raise ValueError("compute_mfccs expects a 1-D clip; use compute_mfccs_batch for [B, n_samples]")
        out = self.compute_mfccs_batch(data[None, :])[0].cpu().numpy()      # [n_frames, n_dct]
        return np.asfortranarray(out[:, :, None])

    def compute_pcen(self, data):
        raise NotImplementedError("honk_b200: the PCEN front end (manage_audio.py:44-47) is out of scope; "
                                  "it depends on the un-vendored pytorch-pcen package")
