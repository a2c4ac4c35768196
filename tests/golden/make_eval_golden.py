#!/usr/bin/env python
"""Golden vectors for the evaluation data path.  Run ONLY in the build container.

Imports the reference's own ``utils/model.py`` and ``utils/train.py`` from ``/root/reference`` (stub modules for the
uninstallable imports, like make_golden.py) and freezes into ``eval_golden.npz``:

* ``SpeechDataset._timeshift_audio`` and ``SpeechDataset.load_audio`` outputs for seeded clips, with Python's
  ``random`` replaced by a scripted source so that every draw (cache bypass, noise window, shift, noise gate,
  gain) is known and stored next to the output;
* ``print_eval``'s accuracy and ``nn.CrossEntropyLoss`` for seeded scores/labels (ties included).

Usage:  python tests/golden/make_eval_golden.py
"""
import io
import os
import sys
import contextlib

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
from make_golden import import_reference_model  # noqa: E402


class Script:
    """Stands in for the ``random`` module inside utils.model: returns the scripted draws in order."""
    def __init__(self, draws):
        self.draws = list(draws)

    def _next(self):
        return self.draws.pop(0)

    def random(self):
        return self._next()

    def randint(self, a, b):
        v = self._next()
        assert a <= v <= b, (a, v, b)
        return v

    def choice(self, seq):
        return seq[self._next()]


def main():
    refmod = import_reference_model()
    import utils.train as reftrain
    rng = np.random.default_rng(77)
    out = {}
    bg = (rng.standard_normal(3 * 16000) * 0.3).astype(np.float32)
    out["bg_noise"] = bg
    cases = []
    #            len   set_type                shift  noise? gain-draw  offset
    specs = [(16000, refmod.DatasetType.TRAIN, -1600, True, 0.73, 100),
             (16000, refmod.DatasetType.TRAIN, 1600, True, 0.11, 31999),
             (9000, refmod.DatasetType.TRAIN, 355, False, None, 7),
             (12345, refmod.DatasetType.TEST, None, True, 0.999, 16000),
             (16000, refmod.DatasetType.TEST, None, False, None, 0),
             (16000, refmod.DatasetType.TRAIN, 0, True, 0.5, 5),
             (1, refmod.DatasetType.TRAIN, -3, True, 0.25, 12)]
    # (clips longer than input_length make the reference's own noise mix fail to broadcast, model.py:303: not a case)
    for k, (n, set_type, shift, noisy, gdraw, off) in enumerate(specs):
        ds = refmod.SpeechDataset.__new__(refmod.SpeechDataset)
        ds.set_type = set_type
        ds.bg_noise_audio = [bg]
        ds.noise_prob = 0.8
        ds.input_length = 16000
        ds.timeshift_ms = 100
        ds._audio_cache = refmod.SimpleCache(10)
        ds._file_cache = refmod.SimpleCache(10)
        x = np.clip(rng.standard_normal(n) * 0.4, -1, 1).astype(np.float32)
        ds._file_cache["clip"] = x
        draws = [0.9 if set_type == refmod.DatasetType.TRAIN else 0.1]   # cache gate (empty cache either way)
        draws += [0, off]                                              # random.choice index, window start
        if set_type == refmod.DatasetType.TRAIN:
            draws += [shift]
        draws += [0.1 if noisy else 0.95]                              # noise gate vs noise_prob 0.8
        if noisy:
            draws += [gdraw]
        refmod.random = Script(draws)
        y = ds.load_audio("clip")
        assert not refmod.random.draws
        out["load_audio/%d/x" % k] = x
        out["load_audio/%d/y" % k] = np.asarray(y)
        cases.append([n, 0 if shift is None else shift, 1 if noisy else 0, (gdraw or 0.0) * 0.1, off])
    out["load_audio/params"] = np.asarray(cases, dtype=np.float64)      # len, shift, noisy, gain a, noise offset
    # silence example: zeros + noise always
    ds = refmod.SpeechDataset.__new__(refmod.SpeechDataset)
    ds.set_type = refmod.DatasetType.TEST
    ds.bg_noise_audio = [bg]
    ds.noise_prob = 0.0
    ds.input_length = 16000
    ds.timeshift_ms = 100
    ds._audio_cache = refmod.SimpleCache(10)
    ds._file_cache = refmod.SimpleCache(10)
    refmod.random = Script([0.1, 0, 4242, 0.99, 0.6])
    out["load_audio/silence/y"] = np.asarray(ds.load_audio("whatever", silence=True))
    out["load_audio/silence/params"] = np.asarray([0.6 * 0.1, 4242])
    import random as _r
    refmod.random = _r
    # _timeshift_audio alone
    ds.timeshift_ms = 100
    x = rng.standard_normal(16000).astype(np.float32)
    out["timeshift/x"] = x
    for sh in (-1600, -1, 0, 1, 1600):
        refmod.random = Script([sh])
        out["timeshift/%d" % sh] = ds._timeshift_audio(x)
    refmod.random = _r

    # print_eval + CrossEntropyLoss
    torch.manual_seed(3)
    scores = torch.randn(257, 12) * 3
    scores[5, 3] = scores[5, 7] = scores[5].max() + 1          # a tie: torch.max returns the first index
    labels = torch.randint(0, 12, (257,))
    labels[:100] = scores[:100].argmax(1)
    loss = torch.nn.CrossEntropyLoss()(scores, labels)
    with contextlib.redirect_stdout(io.StringIO()):
        acc = reftrain.print_eval("test", scores, labels, loss)
    out["eval/scores"] = scores.numpy()
    out["eval/labels"] = labels.numpy()
    out["eval/accuracy"] = np.float64(acc)
    out["eval/loss"] = np.float64(loss.item())
    out["eval/pred"] = torch.max(scores, 1)[1].numpy()
    np.savez_compressed(os.path.join(HERE, "eval_golden.npz"), **out)
    print({k: np.asarray(v).shape for k, v in out.items()})


if __name__ == "__main__":
    main()
