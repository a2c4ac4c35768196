"""The inference half of ``utils/train.py``: ``print_eval`` (train.py:41-46) and ``evaluate`` (train.py:56-85).

``evaluate`` keeps the reference's loop (one ``model(model_in)`` per loader batch, accuracy weighted by the batch
size) but argmax, the label comparison and the CrossEntropyLoss run in one kernel on the device
(``honk_eval_metrics``), and a loader may hand raw PCM ``[B, 16000]`` instead of MFCCs: the batched front end then
replaces ``SpeechDataset.collate_fn``'s per-clip ``compute_mfccs`` + ``torch.cat`` (model.py:253-265).

The training half (train.py:87-163): ``SgdTrainer.step`` replaces the body of the reference's inner
loop (model.train(); scores = model(model_in); loss = criterion(scores, labels); loss.backward(); optimizer.step();
train.py:127-139) with one native call, ``train`` keeps the loop around it (learning-rate schedule that rebuilds the
optimiser, dev-set evaluation every ``dev_every`` epochs, best model saved).  Dataset walking (SpeechDataset.splits) stays
out of scope: the loaders are handed in.  Both families train; the cnn-* family's Dropout masks are drawn by the library
(reproducible from ``dropout_seed``, not torch's stream) unless the caller hands them in (``step(..., dropout_masks=...)``).
"""
from __future__ import annotations

import copy
import ctypes as C

import numpy as np
import torch

from . import _native as N
from . import ops
from .manage_audio import AudioPreprocessor


def eval_metrics(scores, labels):
    """``(accuracy, mean CrossEntropyLoss)`` of one batch as Python floats (train.py:41-46, 66-83)."""
    _, correct, loss_sum = ops.eval_metrics(scores, labels)
    n = max(int(labels.shape[0]), 1)
    loss = float(loss_sum.item()) / n
    if loss != loss:                 # the kernel poisons the sum for a target outside [0, n_labels) (or a NaN score)
        lab = labels.detach().cpu()
        if lab.numel() and (int(lab.min()) < 0 or int(lab.max()) >= scores.shape[1]):
            raise IndexError("Target %d is out of bounds for %d labels" % (int(lab.max() if lab.max() >= scores.shape[1] else lab.min()),
                                                                          scores.shape[1]))
    return float(correct.item()) / n, loss


def print_eval(name, scores, labels, loss=None, end="\n"):
    """Same line and return value as the reference; ``loss`` may be omitted (it is then computed here)."""
    accuracy, ce = eval_metrics(scores, labels)
    loss = ce if loss is None else (loss.item() if hasattr(loss, "item") else float(loss))
    print("{} accuracy: {:>5}, loss: {:<25}".format(name, accuracy, loss), end=end)
    return accuracy


def evaluate(config, model=None, test_loader=None, audio_processor=None):
    """``utils/train.py:evaluate``: returns the final test accuracy (the reference only prints it).

    ``test_loader`` yields ``(model_in, labels)``; ``model_in`` is either MFCCs ``[B, 101, 40]`` (what
    ``collate_fn`` produces) or PCM ``[B, n_samples]`` int16/float32/float64, on the host or the GPU."""
    if test_loader is None:
        raise ValueError("honk_b200.train.evaluate needs a test_loader (dataset walking is out of scope)")
    if config.get("no_cuda", False):
        raise RuntimeError("honk_b200: no_cuda=True is not supported (there is no CPU fallback)")
    if model is None:
        model = config["model_class"](config)
        model.load(config["input_file"])
    torch.cuda.set_device(config.get("gpu_no", 0) or 0)
    model.cuda()
    model.eval()
    results = []
    total = 0
    with torch.no_grad():
        for model_in, labels in test_loader:
            model_in = torch.as_tensor(model_in).cuda()
            labels = torch.as_tensor(labels).cuda()
            if model_in.dim() == 2:                       # raw PCM: fused front end + CNN
                if audio_processor is None:
                    audio_processor = AudioPreprocessor()
                scores = model.forward_pcm(model_in, audio_processor)
            else:
                scores = model(model_in)
            results.append(print_eval("test", scores, labels) * model_in.size(0))
            total += model_in.size(0)
    final = sum(results) / max(total, 1)
    print("final test accuracy: {}".format(final))
    return final


class SgdTrainer:
    """The reference's ``torch.optim.SGD(model.parameters(), lr, momentum, weight_decay, nesterov)`` plus the
    forward / loss / backward it is stepped with (train.py:100, 127-139), for a ``SpeechResModel`` or ``SpeechModel`` on the GPU.

    The trainer owns device copies of the parameters, gradients, momentum buffers and BatchNorm running statistics
    (``honk_trainer``); ``sync()`` writes the current weights and statistics back into the module (done automatically by
    ``train`` before every evaluation / save), after which ``model.eval()`` inference uses them."""

    def __init__(self, model, lr=0.001, momentum=0.9, weight_decay=0.00001, nesterov=False, dropout_seed=0):
        if not hasattr(model, "_native_config"):
            raise TypeError("SgdTrainer needs a honk_b200 model")
        self.model = model
        self.device = model._device()
        self.opt = N.Sgd(float(lr), float(momentum), float(weight_decay), int(bool(nesterov)))
        cfg = model._native_config()
        sd = model.state_dict()
        keep, arr = _native_tensors(sd)
        handle = C.c_void_p()
        with torch.cuda.device(self.device):
            N.check(N.lib().honk_trainer_create(C.byref(cfg), arr, len(sd), self.device.index or 0, C.byref(handle)))
        self.handle = handle.value
        self._ws = None
        self._synced_steps = 0
        if hasattr(model, "dropout_prob"):                      # SpeechModel: nn.Dropout(config["dropout_prob"])
            N.check(N.lib().honk_trainer_set_dropout(self.handle, float(model.dropout_prob), int(dropout_seed)))

    def __del__(self):
        h, self.handle = getattr(self, "handle", None), None
        if h:
            try:
                N.lib().honk_trainer_destroy(h)
            except Exception:
                pass

    @property
    def lr(self):
        return self.opt.lr

    def new_optimizer(self, lr):
        """train.py:141-143: a NEW SGD with the next learning rate - the momentum buffers start over."""
        self.opt.lr = float(lr)
        N.check(N.lib().honk_trainer_reset_optimizer(self.handle))

    @property
    def steps(self):
        return int(N.lib().honk_trainer_steps(self.handle))

    def step(self, model_in, labels, check_labels=True, dropout_masks=None):
        """One optimisation step; returns ``(scores [B, n_labels], loss [])`` on the device: the training-mode scores
        and the batch's mean CrossEntropyLoss before the update, the two things the reference prints (train.py:144).
        ``dropout_masks`` (cnn-* only): up to four keep masks (nonzero = keep) for the Dropout after conv1, conv2, dnn1,
        dnn2 (``None`` entries are drawn by the library); they replace the library's own draws for this step."""
        m = self.model
        x = torch.as_tensor(model_in)
        y = torch.as_tensor(labels)
        if x.dim() != 3 or tuple(x.shape[1:]) != (m.height, m.width):
            raise ValueError("honk_b200: model input must be [n_clips, %d, %d], got %s" % (m.height, m.width, tuple(x.shape)))
        if y.dim() != 1 or y.shape[0] != x.shape[0]:
            raise ValueError("honk_b200: labels must be [n_clips]")
        x = x.to(self.device, torch.float32).contiguous()
        y = y.to(self.device, torch.int64).contiguous()
        if check_labels and y.numel():
            lo, hi = int(y.min()), int(y.max())
            if lo < 0 or hi >= m.n_labels:                      # what nn.CrossEntropyLoss raises before any update
                raise IndexError("Target %d is out of bounds." % (hi if hi >= m.n_labels else lo))
        B = int(x.shape[0])
        need = int(N.lib().honk_trainer_workspace_bytes(self.handle, B))
        if self._ws is None or self._ws.numel() < need:
            self._ws = None
            self._ws = torch.empty(need, dtype=torch.uint8, device=self.device)
        scores = torch.empty((B, m.n_labels), dtype=torch.float32, device=self.device)
        loss = torch.empty((), dtype=torch.float32, device=self.device)
        keep = []
        if dropout_masks is not None:
            ptrs = (C.c_void_p * 4)()
            for i, mk in enumerate(list(dropout_masks)[:4]):
                if mk is None:
                    continue
                mk = (torch.as_tensor(mk).to(self.device) != 0).to(torch.uint8).contiguous()
                if mk.shape[0] != B:
                    raise ValueError("honk_b200: dropout mask %d is for %d clips, the batch has %d" % (i, mk.shape[0], B))
                keep.append(mk)
                ptrs[i] = mk.data_ptr()
            N.check(N.lib().honk_trainer_set_dropout_masks(self.handle, ptrs, 4))
        with torch.cuda.device(self.device):
            N.check(N.lib().honk_trainer_step(self.handle, x.data_ptr(), y.data_ptr(), B, m.height, m.width, C.byref(self.opt),
                                              scores.data_ptr(), loss.data_ptr(), self._ws.data_ptr(), self._ws.numel(),
                                              N.stream_ptr(self.device)))
        return scores, loss

    def read(self, name):
        """One tensor of the trainer as a numpy array: a state_dict key, ``grad:<parameter>`` or ``momentum:<parameter>``."""
        key = name.split(":", 1)[-1]
        shape = tuple(self.model.state_dict()[key].shape)
        out = np.empty(shape, dtype=np.float32)
        N.check(N.lib().honk_trainer_read(self.handle, name.encode(), out.ctypes.data, out.size))
        return out

    def sync(self):
        """Writes the trained parameters, running statistics and ``num_batches_tracked`` back into the module."""
        sd = self.model.state_dict()
        advanced = self.steps - self._synced_steps
        with torch.no_grad():
            for k, t in sd.items():
                if k.endswith("num_batches_tracked"):
                    t.add_(advanced)
                else:
                    t.copy_(torch.from_numpy(self.read(k)))
        self._synced_steps = self.steps
        return self.model


def _native_tensors(sd):
    from .model import native_tensors
    return native_tensors(sd)


def train(config, train_loader, dev_loader=None, test_loader=None, model=None):
    """``utils/train.py:train`` (87-163) around the native step.  ``config`` carries the reference's keys (lr list,
    schedule, n_epochs, dev_every, momentum, weight_decay, use_nesterov, output_file, input_file, model_class); the
    loaders yield ``(model_in [B, 101, 40], labels [B])``.  Returns the best model (by dev accuracy), like the
    reference keeps it for its final ``evaluate``."""
    if config.get("no_cuda", False):
        raise RuntimeError("honk_b200: no_cuda=True is not supported (there is no CPU fallback)")
    if model is None:
        model = config["model_class"](config)
        if config.get("input_file"):
            model.load(config["input_file"])
    torch.cuda.set_device(config.get("gpu_no", 0) or 0)
    model.cuda()
    lrs = list(config["lr"])
    schedule_steps = list(config.get("schedule", [])) + [np.inf]
    trainer = SgdTrainer(model, lr=lrs[0], nesterov=config.get("use_nesterov", False),
                         weight_decay=config.get("weight_decay", 0.0), momentum=config.get("momentum", 0.0))
    sched_idx, step_no, max_acc, best_model = 0, 0, 0, None
    dev_every = int(config.get("dev_every", 10))
    for epoch_idx in range(int(config["n_epochs"])):
        for model_in, labels in train_loader:
            labels = torch.as_tensor(labels).cuda()
            scores, loss = trainer.step(model_in, labels)
            step_no += 1
            if step_no > schedule_steps[sched_idx]:
                sched_idx += 1
                print("changing learning rate to {}".format(lrs[sched_idx]))
                trainer.new_optimizer(lrs[sched_idx])
            print_eval("train step #{}".format(step_no), scores, labels, loss)
        if dev_loader is not None and epoch_idx % dev_every == dev_every - 1:
            trainer.sync()
            model.eval()
            accs = []
            with torch.no_grad():
                for model_in, labels in dev_loader:
                    model_in, labels = torch.as_tensor(model_in).cuda(), torch.as_tensor(labels).cuda()
                    accs.append(print_eval("dev", model(model_in), labels))
            avg_acc = float(np.mean(accs))
            print("final dev accuracy: {}".format(avg_acc))
            if avg_acc > max_acc:
                print("saving best model...")
                max_acc = avg_acc
                if config.get("output_file"):
                    model.save(config["output_file"])
                best_model = copy.deepcopy(model)
    trainer.sync()
    if best_model is None:
        best_model = model
    if test_loader is not None:
        evaluate(config, best_model, test_loader)
    return best_model
