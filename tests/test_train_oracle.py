"""The training-step oracle (oracle.sgd_train_steps) against the reference's own run of
utils/train.py:127-139 over utils/model.py's SpeechResModel, frozen in tests/golden/train_golden.npz by
tests/golden/make_train_golden.py.  CPU only."""
import json
import os

import numpy as np
import pytest

from oracle import honk_oracle as orc

CASES = ["res8-narrow", "res15-narrow", "res26-narrow", "res8", "cnn-trad-pool2", "cnn-one-stride1", "cnn-tstride4", "cnn-one-fpool3"]
SUB_MIN, SUB_STEP = 20000, 13          # make_train_golden.py stores large tensors as every 13th element


def sub(v):
    """The view of a tensor that the fixture holds (make_train_golden.subsample)."""
    v = np.asarray(v)
    return v.reshape(-1)[::SUB_STEP] if v.size > SUB_MIN else v


def load_case(golden_dir, name):
    g = np.load(os.path.join(golden_dir, "train_golden.npz"))
    pfx = name + "/"
    case = {"config": json.loads(bytes(g[pfx + "config"]).decode()), "hyper": g[pfx + "hyper"], "mfcc": g[pfx + "mfcc"],
            "labels": g[pfx + "labels"], "losses": g[pfx + "losses"], "scores": g[pfx + "scores"],
            "init": {}, "final": {}, "grad0": {}}
    for k in g.files:
        for part in ("init", "final", "grad0"):
            if k.startswith(pfx + part + "/"):
                case[part][k[len(pfx + part) + 1:]] = g[k]
    steps = len(case["losses"])
    B = len(case["labels"]) // steps
    case["batches"] = [[case["mfcc"][s * B:(s + 1) * B], case["labels"][s * B:(s + 1) * B]] for s in range(steps)]
    if name.startswith("cnn"):
        from honk_b200 import synth
        state = synth.make_state_dict(name, case["config"], seed=11)          # make_train_golden.cnn_init_state
        case["init"] = {k: (v * np.float32(0.3) if k.endswith("weight") else v) for k, v in state.items()}
        for s in range(steps):                                                # the reference run's Dropout keep masks
            masks = [None] * 4
            for site in range(4):
                key = pfx + "mask/%d/%d" % (s, site)
                if key in g.files:
                    shape = tuple(g[pfx + "maskshape/%d/%d" % (s, site)])
                    masks[site] = np.unpackbits(g[key])[:int(np.prod(shape))].reshape(shape)
            case["batches"][s].append(masks)
    return case


@pytest.mark.parametrize("name", CASES)
def test_oracle_training_steps_match_the_reference_run(golden_dir, name):
    c = load_case(golden_dir, name)
    lr, mom, wd, nest = c["hyper"]
    o = orc.sgd_train_steps(c["init"], c["config"], c["batches"], lr=lr, momentum=mom, weight_decay=wd, nesterov=bool(nest))
    assert np.allclose(o["losses"], c["losses"], atol=2e-5, rtol=0)
    for s in range(len(c["losses"])):
        assert np.allclose(o["scores"][s], c["scores"][s], atol=5e-5, rtol=1e-4)
    for k, v in c["grad0"].items():
        assert np.allclose(sub(o["grads"][0][k]), v, atol=1e-6 + 1e-4 * np.abs(v).max(), rtol=0), k
    for k, v in c["final"].items():
        if k.endswith("num_batches_tracked"):
            assert int(o["state"][k]) == int(v) == 100 + len(c["losses"])
        else:
            assert np.allclose(sub(o["state"][k]), v, atol=2e-5, rtol=1e-5), k


def test_sgd_rule_weight_decay_momentum_nesterov():
    """The update rule on its own, against torch.optim.SGD (what train.py:100 constructs), on a tiny res model."""
    import torch
    cfg = dict(n_labels=3, n_layers=2, n_feature_maps=4, use_dilation=False)
    rng = np.random.default_rng(0)
    state = {"conv0.weight": rng.normal(0, 0.5, (4, 1, 3, 3)).astype(np.float32)}
    for i in (1, 2):
        state["conv%d.weight" % i] = rng.normal(0, 0.3, (4, 4, 3, 3)).astype(np.float32)
        state["bn%d.running_mean" % i] = np.zeros(4, np.float32)
        state["bn%d.running_var" % i] = np.ones(4, np.float32)
        state["bn%d.num_batches_tracked" % i] = np.asarray(0, np.int64)
    state["output.weight"] = rng.normal(0, 0.5, (3, 4)).astype(np.float32)
    state["output.bias"] = np.zeros(3, np.float32)
    batches = [(rng.normal(0, 1, (5, 9, 7)).astype(np.float32), rng.integers(0, 3, 5)) for _ in range(4)]
    for nesterov in (False, True):
        o = orc.sgd_train_steps(state, cfg, batches, lr=0.1, momentum=0.8, weight_decay=0.01, nesterov=nesterov)
        params = {k: torch.from_numpy(v.copy()).requires_grad_() for k, v in state.items() if "bn" not in k}
        stats = {k: torch.from_numpy(v.copy()) for k, v in state.items() if "running" in k}
        opt = torch.optim.SGD(params.values(), lr=0.1, momentum=0.8, weight_decay=0.01, nesterov=nesterov)
        for x, y in batches:
            opt.zero_grad()
            torch.nn.functional.cross_entropy(orc.res_train_logits(params, stats, cfg, torch.from_numpy(x)), torch.from_numpy(y)).backward()
            opt.step()
        for k, p in params.items():
            assert np.allclose(o["state"][k], p.detach().numpy(), atol=1e-6), k
        for k, v in stats.items():
            assert np.allclose(o["state"][k], v.numpy(), atol=1e-6), k
