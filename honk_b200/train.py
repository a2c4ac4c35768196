"""The inference half of ``utils/train.py``: ``print_eval`` (train.py:41-46) and ``evaluate`` (train.py:56-85).

``evaluate`` keeps the reference's loop (one ``model(model_in)`` per loader batch, accuracy weighted by the batch
size) but argmax, the label comparison and the CrossEntropyLoss run in one kernel on the device
(``honk_eval_metrics``), and a loader may hand raw PCM ``[B, 16000]`` instead of MFCCs: the batched front end then
replaces ``SpeechDataset.collate_fn``'s per-clip ``compute_mfccs`` + ``torch.cat`` (model.py:253-265).
Training (train.py:87-163) is out of scope.
"""
from __future__ import annotations

import torch

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
