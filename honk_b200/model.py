"""Drop-in for the inference half of the reference's ``utils/model.py``.

Keeps ``ConfigType`` / ``find_model`` / ``find_config`` (model.py:33-62, 381-414) and the
``SpeechResModel`` / ``SpeechModel`` classes (model.py:82-205) as ``nn.Module``s whose
parameters and buffers carry the reference's state_dict key names and shapes, so honk-models
``.pt`` files load with a strict ``load_state_dict``.  ``forward`` never runs a torch conv:
the weights are handed to libhonk_b200.so (BN folded / repacked there) and the CUDA kernels
compute the logits.  Inference only (eval mode); there is no CPU fallback.
"""
from __future__ import annotations

import ctypes as C
import threading
from enum import Enum

import numpy as np
import torch
import torch.nn as nn

from . import _native as N
from . import ops


class ConfigType(Enum):
    CNN_TRAD_POOL2 = "cnn-trad-pool2"
    CNN_ONE_STRIDE1 = "cnn-one-stride1"
    CNN_ONE_FPOOL3 = "cnn-one-fpool3"
    CNN_ONE_FSTRIDE4 = "cnn-one-fstride4"
    CNN_ONE_FSTRIDE8 = "cnn-one-fstride8"
    CNN_TPOOL2 = "cnn-tpool2"
    CNN_TPOOL3 = "cnn-tpool3"
    CNN_TSTRIDE2 = "cnn-tstride2"
    CNN_TSTRIDE4 = "cnn-tstride4"
    CNN_TSTRIDE8 = "cnn-tstride8"
    RES15 = "res15"
    RES26 = "res26"
    RES8 = "res8"
    RES15_NARROW = "res15-narrow"
    RES8_NARROW = "res8-narrow"
    RES26_NARROW = "res26-narrow"


def _cnn(maps1, conv1, pool1, stride1=(1, 1), maps2=None, conv2=None, dnn=True, tf_variant=False):
    """One Google ``cnn-*`` entry of the reference's ``_configs`` (model.py:382-407)."""
    cfg = dict(dropout_prob=0.5, height=101, width=40, n_labels=4, n_feature_maps1=maps1,
               conv1_size=conv1, conv1_pool=pool1, conv1_stride=stride1)
    if maps2 is not None:
        cfg.update(n_feature_maps2=maps2, conv2_size=conv2, conv2_stride=(1, 1), conv2_pool=(1, 1))
    if dnn:
        cfg.update(dnn1_size=128, dnn2_size=128)
    if tf_variant:
        cfg["tf_variant"] = True
    return cfg


def _res(n_layers, maps, dilation, pool=None):
    """One ``res*`` entry (model.py:408-413)."""
    cfg = dict(n_labels=12, n_layers=n_layers, n_feature_maps=maps, use_dilation=dilation)
    if pool is not None:
        cfg["res_pool"] = pool
    return cfg


# Same keys and values as the reference table; ``find_config`` hands out these shared,
# mutable dicts exactly like the reference does (callers mutate n_labels, service.py:82).
_configs = {
    "cnn-trad-pool2": _cnn(64, (20, 8), (2, 2), maps2=64, conv2=(10, 4), dnn=False, tf_variant=True),
    "cnn-one-stride1": _cnn(186, (101, 8), (1, 1), tf_variant=True),
    "cnn-tstride2": _cnn(78, (16, 8), (1, 3), (2, 1), maps2=78, conv2=(9, 4)),
    "cnn-tstride4": _cnn(100, (16, 8), (1, 3), (4, 1), maps2=78, conv2=(5, 4)),
    "cnn-tstride8": _cnn(126, (16, 8), (1, 3), (8, 1), maps2=78, conv2=(5, 4)),
    "cnn-tpool2": _cnn(94, (21, 8), (2, 3), maps2=94, conv2=(6, 4)),
    "cnn-tpool3": _cnn(94, (15, 8), (3, 3), maps2=94, conv2=(6, 4)),
    "cnn-one-fpool3": _cnn(54, (101, 8), (1, 3)),
    "cnn-one-fstride4": _cnn(186, (101, 8), (1, 1), (1, 4)),
    "cnn-one-fstride8": _cnn(336, (101, 8), (1, 1), (1, 8)),
    "res15": _res(13, 45, True),
    "res8": _res(6, 45, False, (4, 3)),
    "res26": _res(24, 45, False, (2, 2)),
    "res15-narrow": _res(13, 19, True),
    "res8-narrow": _res(6, 19, False, (4, 3)),
    "res26-narrow": _res(24, 19, False, (2, 2)),
}


def find_model(conf):
    if isinstance(conf, ConfigType):
        conf = conf.value
    return SpeechResModel if conf.startswith("res") else SpeechModel


def find_config(conf):
    if isinstance(conf, ConfigType):
        conf = conf.value
    return _configs[conf]


def _pair(v):
    return (int(v[0]), int(v[1]))


def native_tensors(sd):
    """``honk_tensor[]`` over host copies of a state_dict; returns ``(host tensors to keep alive, ctypes array)``."""
    keep, arr = [], (N.Tensor * len(sd))()
    for i, (name, t) in enumerate(sd.items()):
        h = t.detach().to("cpu")
        h = h.to(torch.float32).contiguous() if h.is_floating_point() else h.to(torch.int64).contiguous()
        keep.append(h)
        arr[i].name = name.encode()
        arr[i].h_data = h.data_ptr()
        arr[i].ndim = h.dim()
        for d in range(h.dim()):
            arr[i].dims[d] = h.shape[d]
    return keep, arr


class SerializableModule(nn.Module):
    """``save`` / ``load`` with the reference's bare-state_dict format (model.py:72-80)."""

    def __init__(self):
        super().__init__()
        self._native = None           # honk_model* as int
        self._native_key = None       # (device index, parameter versions) the handle was built from
        self._native_lock = threading.Lock()   # label() is entered from CherryPy's thread pool (server.py:144-146)
        self.precision = 0            # 0: tcgen05 f16 operands where available; 1: all-FP32 CUDA cores

    def save(self, filename):
        torch.save(self.state_dict(), filename)

    def load(self, filename):
        self.load_state_dict(torch.load(filename, map_location=lambda storage, loc: storage))

    # -- native handle management ---------------------------------------------------------
    def _native_config(self):
        raise NotImplementedError

    def _release(self):
        """Destroys the handle.  Callers hold ``_native_lock`` (or are the finaliser)."""
        if getattr(self, "_native", None):
            try:
                N.lib().honk_model_destroy(self._native)
            except Exception:
                pass
        self._native = None
        self._native_key = None

    def __getstate__(self):                      # locks do not pickle / deepcopy; the handle belongs to this object
        d = self.__dict__.copy()
        d["_native"], d["_native_key"], d["_native_lock"] = None, None, None
        return d

    def __setstate__(self, d):
        super().__setstate__(d)
        self._native_lock = threading.Lock()

    def warm(self):
        """Builds the native handle now (weights validated, BN folded, repacked) instead of inside the first request."""
        return self._ensure_native(self._device())

    def __del__(self):
        try:
            self._release()
        except Exception:       # interpreter shutdown
            pass

    def _state_key(self, device):
        ts = list(self.state_dict(keep_vars=True).values())          # one pass: (version, storage) of every entry
        return (device.index, self.precision, tuple(t._version for t in ts), tuple(t.data_ptr() for t in ts))

    def _ensure_native(self, device):
        """The native handle for the current weights, built on first use and rebuilt when a parameter changed.
        Serialised by a lock: two first requests from a thread pool must not both build (and one leak) a handle, nor
        may one destroy the handle the other is about to use.  A handle that is in use by a concurrent forward is only
        replaced when the weights really changed, which inference callers do not do mid-request."""
        key = self._state_key(device)
        if self._native is not None and key == self._native_key:
            return self._native
        with self._native_lock:
            if self._native is not None and key == self._native_key:
                return self._native
            return self._build_native(device, key)

    def _build_native(self, device, key):
        self._release()
        cfg = self._native_config()
        cfg.precision = int(self.precision)
        sd = self.state_dict()
        keep, arr = native_tensors(sd)
        handle = C.c_void_p()
        N.check(N.lib().honk_model_create(C.byref(cfg), arr, len(sd), device.index or 0, C.byref(handle)))
        self._native = handle.value
        self._native_key = key
        return self._native

    @property
    def native_path(self):
        """"tcgen05-f16" or "cuda-core-f32" once the handle exists (after the first forward)."""
        return N.lib().honk_model_path(self._native).decode() if self._native else None

    def _device(self):
        p = next(self.parameters())
        if not p.is_cuda:
            raise RuntimeError("honk_b200: the model must be on a CUDA device - call .cuda() "
                               "(no_cuda / CPU inference is not supported, there is no CPU fallback)")
        return p.device

    def forward(self, x):
        if self.training:
            raise RuntimeError("honk_b200 implements the inference path only; call model.eval()")
        dev = self._device()
        if not x.is_cuda:
            raise RuntimeError("honk_b200: model input must be a CUDA tensor")
        if x.dim() != 3 or tuple(x.shape[1:]) != (self.height, self.width):
            # the reference raises a size mismatch for the cnn-* models (model.py:195-205); the kernels here are built
            # for the configured height x width for every family, so any other shape is refused instead of read
            raise ValueError("honk_b200: model input must be [n_clips, %d, %d] (frames x coefficients), got %s"
                             % (self.height, self.width, tuple(x.shape)))
        return ops.model_forward(x, self._ensure_native(dev), self.n_labels)

    def forward_pcm(self, pcm, audio_processor):
        """Fused extension: ``[B, n_samples]`` PCM on the GPU -> logits (front end + CNN)."""
        if self.training:
            raise RuntimeError("honk_b200 implements the inference path only; call model.eval()")
        dev = self._device()
        return ops.model_forward_pcm(pcm, self._ensure_native(dev), audio_processor.handle, self.n_labels)

    def listen(self, pcm, audio_processor, window=16000, hop=8000, keyword=-1, min_keyword_prob=0.85):
        """One 1-D device buffer of samples -> ``(sums float64 [L], counts int32 [L], first_hit int32 [1])`` over
        all sliding windows (server.py:99-112): the windows are read in place as overlapping rows."""
        if self.training:
            raise RuntimeError("honk_b200 implements the inference path only; call model.eval()")
        dev = self._device()
        return ops.listen(pcm, self._ensure_native(dev), audio_processor.handle, self.n_labels, int(window), int(hop),
                          int(keyword), float(min_keyword_prob))

    def label_pcm_host(self, rows, audio_processor):
        """Host numpy PCM ``[B, n_samples]`` -> numpy ``(labels, probs)``: ``TorchLabelService.label`` for a batch in
        one native call (service.py:88-104)."""
        if self.training:
            raise RuntimeError("honk_b200 implements the inference path only; call model.eval()")
        dev = self._device()
        return ops.model_label_pcm_host(rows, self._ensure_native(dev), audio_processor.handle, dev)

    def forward_pcm_host(self, pcm, audio_processor, out=None):
        """Host ``[B, n_samples]`` PCM -> host logits, copies overlapped with the kernels chunk by chunk."""
        if self.training:
            raise RuntimeError("honk_b200 implements the inference path only; call model.eval()")
        dev = self._device()
        return ops.model_forward_pcm_host(pcm, self._ensure_native(dev), audio_processor.handle, self.n_labels, dev, out)


class SpeechResModel(SerializableModule):
    """res8 / res15 / res26 [-narrow] (model.py:82-121).  The torch sub-modules only HOLD the
    weights under the reference key names; the arithmetic is in csrc/ (res_tc.cu, model.cu)."""

    def __init__(self, config):
        super().__init__()
        self.n_labels = n_labels = int(config["n_labels"])
        self.n_maps = n_maps = int(config["n_feature_maps"])
        self.n_layers = n_layers = int(config["n_layers"])
        self.use_dilation = bool(config["use_dilation"])
        self.res_pool = _pair(config["res_pool"]) if "res_pool" in config else None
        self.height = int(config.get("height", 101))
        self.width = int(config.get("width", 40))
        self.conv0 = nn.Conv2d(1, n_maps, (3, 3), padding=(1, 1), bias=False)
        for i in range(n_layers):
            d = int(2 ** (i // 3)) if self.use_dilation else 1
            self.add_module("bn{}".format(i + 1), nn.BatchNorm2d(n_maps, affine=False))
            self.add_module("conv{}".format(i + 1),
                            nn.Conv2d(n_maps, n_maps, (3, 3), padding=d, dilation=d, bias=False))
        self.output = nn.Linear(n_maps, n_labels)

    def _native_config(self):
        cfg = N.ModelConfig()
        cfg.family = N.FAMILY_RES
        cfg.n_labels = self.n_labels
        cfg.height, cfg.width = self.height, self.width
        cfg.n_feature_maps = self.n_maps
        cfg.n_layers = self.n_layers
        cfg.use_dilation = int(self.use_dilation)
        if self.res_pool is not None:
            cfg.res_pool[0], cfg.res_pool[1] = self.res_pool
        return cfg


class SpeechModel(SerializableModule):
    """The Google ``cnn-*`` family (model.py:123-205)."""

    def __init__(self, config):
        super().__init__()
        self.n_labels = n_labels = int(config["n_labels"])
        self.dropout_prob = float(config.get("dropout_prob", 0.5))      # nn.Dropout (model.py:183): used by the training step only
        n_featmaps1 = int(config["n_feature_maps1"])
        self.conv1_size = _pair(config["conv1_size"])
        self.conv1_pool = _pair(config["conv1_pool"])
        self.conv1_stride = _pair(config["conv1_stride"])
        self.width = int(config["width"])
        self.height = int(config["height"])
        self.tf_variant = bool(config.get("tf_variant"))
        self.conv1 = nn.Conv2d(1, n_featmaps1, self.conv1_size, stride=self.conv1_stride)
        h = ((self.height - self.conv1_size[0]) // self.conv1_stride[0] + 1) // self.conv1_pool[0]
        w = ((self.width - self.conv1_size[1]) // self.conv1_stride[1] + 1) // self.conv1_pool[1]
        maps = n_featmaps1
        self.has_conv2 = "conv2_size" in config
        if self.has_conv2:
            self.conv2_size = _pair(config["conv2_size"])
            self.conv2_pool = _pair(config["conv2_pool"])
            self.conv2_stride = _pair(config["conv2_stride"])
            maps = int(config["n_feature_maps2"])
            self.conv2 = nn.Conv2d(n_featmaps1, maps, self.conv2_size, stride=self.conv2_stride)
            h = ((h - self.conv2_size[0]) // self.conv2_stride[0] + 1) // self.conv2_pool[0]
            w = ((w - self.conv2_size[1]) // self.conv2_stride[1] + 1) // self.conv2_pool[1]
        if h < 1 or w < 1:
            raise ValueError("feature map vanished for this config")
        conv_net_size = maps * h * w
        last_size = conv_net_size
        if not self.tf_variant:
            self.lin = nn.Linear(conv_net_size, 32)
        self.dnn1_size = self.dnn2_size = 0
        if "dnn1_size" in config:
            self.dnn1_size = last_size = int(config["dnn1_size"])
            self.dnn1 = nn.Linear(conv_net_size if self.tf_variant else 32, self.dnn1_size)
            if "dnn2_size" in config:
                self.dnn2_size = last_size = int(config["dnn2_size"])
                self.dnn2 = nn.Linear(self.dnn1_size, self.dnn2_size)
        self.output = nn.Linear(last_size, n_labels)
        if self.tf_variant:                      # TF-style init (model.py:137-183); weights are
            for mod in self.children():          # normally overwritten by load()
                nn.init.trunc_normal_(mod.weight, std=0.01, a=-0.02, b=0.02)
                nn.init.zeros_(mod.bias)

    def _native_config(self):
        cfg = N.ModelConfig()
        cfg.family = N.FAMILY_CNN
        cfg.n_labels = self.n_labels
        cfg.height, cfg.width = self.height, self.width
        cfg.n_feature_maps1 = self.conv1.out_channels
        for k in range(2):
            cfg.conv1_size[k] = self.conv1_size[k]
            cfg.conv1_pool[k] = self.conv1_pool[k]
            cfg.conv1_stride[k] = self.conv1_stride[k]
        cfg.has_conv2 = int(self.has_conv2)
        if self.has_conv2:
            cfg.n_feature_maps2 = self.conv2.out_channels
            for k in range(2):
                cfg.conv2_size[k] = self.conv2_size[k]
                cfg.conv2_pool[k] = self.conv2_pool[k]
                cfg.conv2_stride[k] = self.conv2_stride[k]
        cfg.tf_variant = int(self.tf_variant)
        cfg.dnn1_size = self.dnn1_size
        cfg.dnn2_size = self.dnn2_size
        return cfg
