"""Drop-in for ``service.TorchLabelService`` (service.py:28-51, 72-110 of the reference).

``label(wav_bytes) -> (label, prob)`` keeps the reference semantics (int16 PCM / 32768 ->
MFCC -> model -> softmax -> argmax) but the whole chain, including softmax/argmax, runs on
the GPU through one fused C-ABI call; ``label_batch`` is the batched extension used for the
sliding windows of the ``/listen`` endpoint (server.py:99-112).  Thread-safe: the model
handle is immutable, it is built in ``reload()`` (not inside the first request), every thread keeps its own pinned
staging buffer and device workspace, and every call runs on the calling thread's current stream.
"""
from __future__ import annotations

import os
import threading
import wave

import numpy as np
import torch

from . import model as mod
from . import ops
from .manage_audio import AudioPreprocessor


def stride(array, stride_size, window_size):
    """Sliding windows (service.py:106-110)."""
    i = 0
    while i + window_size <= len(array):
        yield array[i:i + window_size]
        i += stride_size


class LabelService(object):
    def evaluate(self, speech_dirs, indices=[]):
        """Folder accuracy (service.py:29-48); one batched call per folder instead of per file."""
        dir_labels = {}
        if indices:
            real_labels = [self.labels[i] for i in indices]
        else:
            real_labels = [os.path.basename(os.path.normpath(d)) for d in speech_dirs]
        for i, label in enumerate(real_labels):
            if label not in self.labels:
                real_labels[i] = "_unknown_"
            dir_labels[speech_dirs[i]] = real_labels[i]
        accuracy = []
        for folder in speech_dirs:
            clips = []
            for filename in sorted(os.listdir(folder)):
                with wave.open(os.path.join(folder, filename)) as f:
                    clips.append(f.readframes(16000))
            for label, _ in self.label_batch(clips):
                accuracy.append(int(label == dir_labels[folder]))
        return sum(accuracy) / len(accuracy)

    def label(self, wav_data):
        raise NotImplementedError


class TorchLabelService(LabelService):
    def __init__(self, model_filename, no_cuda=False, labels=["_silence_", "_unknown_", "command", "random"],
                 model_type=mod.ConfigType.CNN_TRAD_POOL2, n_samples=16000):
        if no_cuda:
            raise RuntimeError("honk_b200: no_cuda=True is not supported (there is no CPU fallback)")
        self.labels = labels
        self.model_filename = model_filename
        self.no_cuda = no_cuda
        self.model_type = model_type      # extension: the reference hard-codes cnn-trad-pool2 (service.py:81)
        self.n_samples = n_samples
        self.audio_processor = AudioPreprocessor()
        self._tls = threading.local()
        self.reload()

    def reload(self):
        config = mod.find_config(self.model_type)
        config["n_labels"] = len(self.labels)          # mutates the shared dict like service.py:82
        self.model = mod.find_model(self.model_type)(config)
        self.model.cuda()
        if isinstance(self.model_filename, dict):
            self.model.load_state_dict(self.model_filename)
        else:
            self.model.load(self.model_filename)
        self.model.eval()
        self.model.warm()            # build the native handle here, not inside the first (possibly concurrent) request

    def _pcm_rows(self, clips):
        """Request bodies -> int16 rows.  Request-sized batches are assembled in a page-locked buffer kept per thread
        (CherryPy calls label() from a thread pool, server.py:144-146), so the copy to the GPU needs no staging."""
        B = len(clips)
        rows = None
        if B <= 1024:
            pinned = getattr(self._tls, "pinned", None)
            if pinned is None or pinned.shape[0] < B or pinned.shape[1] != self.n_samples:
                pinned = torch.empty((max(32, B), self.n_samples), dtype=torch.int16).pin_memory()
                self._tls.pinned = pinned
            rows = pinned.numpy()[:B]
        else:
            rows = np.empty((B, self.n_samples), dtype=np.int16)
        for i, c in enumerate(clips):
            a = np.frombuffer(c, dtype=np.int16) if isinstance(c, (bytes, bytearray, memoryview)) else np.asarray(c, dtype=np.int16)
            n = min(len(a), self.n_samples)
            rows[i, :n] = a[:n]
            rows[i, n:] = 0                            # short clips are zero-padded (model.py:297)
        return rows

    def _pcm_tensor(self, clips):
        return torch.from_numpy(self._pcm_rows(clips)).to(self.audio_processor.device)

    def label_batch(self, clips):
        """List of int16 PCM byte strings -> list of ``(label, prob)``; one native call (copy in, front end, model,
        softmax/argmax, copy out, one synchronisation)."""
        if len(clips) == 0:
            return []
        idx, prob = self.model.label_pcm_host(self._pcm_rows(clips), self.audio_processor)
        return [(self.labels[int(i)], float(p)) for i, p in zip(idx, prob)]

    def label(self, wav_data):
        """Labels audio data as one of the specified trained labels (service.py:89-104)."""
        return self.label_batch([wav_data])[0]

    def listen(self, wav_data, stride_ms=500, min_keyword_prob=0.85, keyword="command", method="all_label"):
        """Batched form of ``ListenEndpoint.POST`` (server.py:99-112): the decompressed int16 buffer is copied to
        the GPU once, its 1-s windows are read in place as overlapping rows (no window copies), and softmax,
        argmax, the per-label sums and the ``command_tagging`` early-exit test run on the device (honk_listen);
        one small device->host read returns the result."""
        step = int(2 * 16000 * stride_ms / 1000)           # bytes, like the reference's stride() over the byte string
        if step <= 0 or step % 2 or self.n_samples != 16000:
            return self._listen_host(wav_data, step, min_keyword_prob, keyword, method)
        a = np.frombuffer(wav_data, dtype=np.int16, count=len(wav_data) // 2)
        if len(a) < self.n_samples:
            return dict(contains_command=False) if method == "command_tagging" else {}
        kw = self.labels.index(keyword) if keyword in self.labels else -1
        with torch.no_grad():
            pcm = torch.from_numpy(a.copy()).pin_memory().to(self.audio_processor.device, non_blocking=True)
            sums, counts, first = self.model.listen(pcm, self.audio_processor, self.n_samples, step // 2, kw, min_keyword_prob)
            sums, counts, first = sums.cpu().numpy(), counts.cpu().numpy(), int(first.cpu()[0])
        if method == "command_tagging":
            return dict(contains_command=first >= 0)
        return {self.labels[i]: float(sums[i]) for i in range(len(self.labels)) if counts[i] > 0}

    def _listen_host(self, wav_data, step, min_keyword_prob, keyword, method):
        """Window extraction on the host for strides the in-place path does not cover (odd byte steps)."""
        windows = list(stride(wav_data, max(step, 1), 2 * 16000))
        labels = {}
        for label, prob in self.label_batch(windows):
            labels[label] = labels.get(label, 0.0) + float(prob)
            if label == keyword and prob >= min_keyword_prob and method == "command_tagging":
                return dict(contains_command=True)
        return dict(contains_command=False) if method == "command_tagging" else labels
