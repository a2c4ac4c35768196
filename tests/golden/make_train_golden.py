#!/usr/bin/env python
"""Golden vectors of the TRAINING step.  Run ONLY in the build container.

    python tests/golden/make_train_golden.py

Imports the reference's own ``utils/model.py`` (as make_golden.py does), puts seeded weights into the reference's
``SpeechResModel`` with a strict ``load_state_dict``, and runs the body of the reference's training loop
(utils/train.py:127-139: model.train(), forward, nn.CrossEntropyLoss, backward, torch.optim.SGD.step with the
reference's momentum / weight_decay defaults, train.py:171-172) for three steps on CPU in float32.  Frozen per case in
``train_golden.npz``: the input batches and labels, the initial state, every step's loss and training-mode scores, the
first step's gradients, and the state_dict after the last step (weights, BatchNorm running statistics,
num_batches_tracked).  The oracle's restatement (oracle.sgd_train_steps) is checked against the same run here.
"""
import json
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, HERE)

from honk_b200 import synth                      # noqa: E402
from oracle import honk_oracle as orc            # noqa: E402
from make_golden import import_reference_model   # noqa: E402

CASES = [  # name, batch, lr, nesterov
    ("res8-narrow", 16, 0.05, False),
    ("res15-narrow", 6, 0.02, True),
    ("res26-narrow", 8, 0.05, False),
    ("res8", 8, 0.1, False),
    # cnn-* (SpeechModel, model.py:123-205): the Dropout keep masks of the reference run are recorded by forward hooks on the
    # reference's own nn.Dropout module (mask = output != 0 where the input is nonzero; where the input is exactly zero the
    # mask has no effect on the forward value or on any gradient that survives the preceding ReLU) and stored bit-packed
    ("cnn-trad-pool2", 4, 0.001, False),      # the reference's default training model (train.py:168): tf_variant, 2 convs, no dnn
    ("cnn-one-stride1", 6, 0.001, True),      # tf_variant, one conv, dnn1 (no ReLU) + dnn2
    ("cnn-tstride4", 6, 0.002, False),        # strided conv1, conv2, lin -> dnn1 (ReLU) -> dnn2
    ("cnn-one-fpool3", 6, 0.002, False),      # pool (1, 3), lin -> dnn1 -> dnn2
]
STEPS = 3


SUB_MIN, SUB_STEP = 20000, 13


def subsample(v):
    v = np.asarray(v)
    return v.reshape(-1)[::SUB_STEP].copy() if v.size > SUB_MIN else v


def cnn_init_state(name, cfg):
    """The initial state of a cnn-* case: seeded He-scaled weights shrunk by 0.3 (raw MFCCs reach +-100, so unshrunk weights
    give logits in the hundreds and a diverging first step)."""
    state = synth.make_state_dict(name, cfg, seed=11)
    return {k: (v * np.float32(0.3) if k.endswith("weight") else v) for k, v in state.items()}


def main():
    refmod = import_reference_model()
    torch.set_num_threads(8)
    torch.manual_seed(1234)                                        # the reference's nn.Dropout draws from torch's generator
    pre = orc.AudioPreprocessorOracle()
    out = {}
    for name, B, lr, nesterov in CASES:
        cfg = dict(refmod.find_config(name))
        cfg["n_labels"] = 12
        is_cnn = name.startswith("cnn")
        clips, labels = synth.make_clips(B * STEPS, start=700000, labelled=True)
        mfcc = pre.compute_mfccs_batch(clips).astype(np.float32)
        state0 = cnn_init_state(name, cfg) if is_cnn else synth.make_state_dict(name, cfg, seed=11, calib_mfcc=mfcc[:8])
        model = (refmod.SpeechModel if is_cnn else refmod.SpeechResModel)(cfg)
        recorded = []                                              # per step: list of (site order) keep masks
        if is_cnn:
            def hook(mod, inp, outp):
                keep = (outp != 0) | (inp[0] == 0)                 # input == 0: "keep" (irrelevant either way)
                recorded[-1].append(keep.detach().numpy().astype(np.uint8))
            model.dropout.register_forward_hook(hook)
        model.load_state_dict({k: torch.from_numpy(np.asarray(v)) for k, v in state0.items()}, strict=True)
        opt = torch.optim.SGD(model.parameters(), lr=lr, nesterov=nesterov, weight_decay=0.00001, momentum=0.9)
        crit = torch.nn.CrossEntropyLoss()
        batches, losses, scores, grads0 = [], [], [], None
        for s in range(STEPS):
            x = torch.from_numpy(mfcc[s * B:(s + 1) * B])
            y = torch.from_numpy(labels[s * B:(s + 1) * B].astype(np.int64))
            batches.append([x.numpy(), y.numpy()])
            model.train()
            opt.zero_grad()
            recorded.append([])
            sc = model(x)
            loss = crit(sc, y)
            loss.backward()
            if s == 0:
                grads0 = {k: p.grad.detach().numpy().copy() for k, p in model.named_parameters()}
            opt.step()
            losses.append(float(loss.detach()))
            scores.append(sc.detach().numpy().copy())
        final = {k: v.detach().numpy().copy() for k, v in model.state_dict().items()}
        if is_cnn:
            # place the recorded masks (in call order) on their sites: conv1, [conv2], [dnn1], [dnn2]
            sites = [0] + ([1] if "conv2_size" in cfg else []) + ([2] if "dnn1_size" in cfg else []) + ([3] if "dnn2_size" in cfg else [])
            for s_, b in enumerate(batches):
                masks = [None] * 4
                for site, mk in zip(sites, recorded[s_]):
                    masks[site] = mk
                b.append(masks)
        # the oracle's restatement against the reference run
        o = orc.sgd_train_steps(state0, cfg, batches, lr=lr, momentum=0.9, weight_decay=0.00001, nesterov=nesterov)
        worst = max(float(np.abs(o["state"][k].astype(np.float64) - final[k]).max()) for k in final)
        print("%-13s losses %s  oracle-vs-reference: state %.2e  loss %.2e" %
              (name, ["%.5f" % v for v in losses], worst, max(abs(a - b) for a, b in zip(o["losses"], losses))))
        assert worst <= 1e-5 and max(abs(a - b) / max(1.0, abs(b)) for a, b in zip(o["losses"], losses)) <= 1e-5
        pfx = name + "/"
        if is_cnn:
            for s_, b in enumerate(batches):
                for site, mk in enumerate(b[2]):
                    if mk is not None:
                        out[pfx + "mask/%d/%d" % (s_, site)] = np.packbits(mk.reshape(-1))
                        out[pfx + "maskshape/%d/%d" % (s_, site)] = np.asarray(mk.shape, dtype=np.int64)
        out[pfx + "config"] = np.frombuffer(json.dumps(cfg).encode(), dtype=np.uint8)
        out[pfx + "hyper"] = np.asarray([lr, 0.9, 0.00001, float(nesterov)], dtype=np.float64)
        out[pfx + "mfcc"] = mfcc
        out[pfx + "labels"] = labels.astype(np.int64)
        out[pfx + "losses"] = np.asarray(losses, dtype=np.float64)
        out[pfx + "scores"] = np.stack(scores)
        # fixtures stay small: the cnn-* initial states are regenerated from their seed by the tests (cnn_init_state below),
        # and tensors of more than SUB_MIN elements are stored as every SUB_STEP-th element of the flattened tensor
        if not is_cnn:
            for k, v in state0.items():
                out[pfx + "init/" + k] = np.asarray(v)
        for k, v in final.items():
            out[pfx + "final/" + k] = subsample(v)
        for k, v in grads0.items():
            out[pfx + "grad0/" + k] = subsample(v)
    path = os.path.join(HERE, "train_golden.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
