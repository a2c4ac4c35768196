"""Training step on the GPU (reference utils/train.py:127-139) through the C ABI (SgdTrainer ->
honk_trainer_step): against the reference's own run frozen in tests/golden/train_golden.npz, and against the oracle on a
fresh batch.  FP32 on both sides; tolerances are summation-order noise: 1e-4 on the loss, 1e-3 of the largest entry on a
gradient tensor, 2e-4 absolute + 1e-3 relative on the state after three steps."""
import numpy as np
import pytest
import torch

from honk_b200 import synth
from oracle import honk_oracle as orc
from test_train_oracle import CASES, load_case

pytestmark = pytest.mark.gpu


def make_model(name, cfg, state):
    import honk_b200 as hb
    model = hb.find_model(name)(cfg)
    model.load_state_dict({k: torch.from_numpy(np.asarray(v)) for k, v in state.items()}, strict=True)
    return model.cuda()


@pytest.mark.parametrize("name", CASES)
def test_three_steps_match_the_reference_run(golden_dir, name):
    from honk_b200 import train
    c = load_case(golden_dir, name)
    lr, mom, wd, nest = c["hyper"]
    model = make_model(name, c["config"], c["init"])
    tr = train.SgdTrainer(model, lr=lr, momentum=mom, weight_decay=wd, nesterov=bool(nest))
    for s, (x, y) in enumerate(c["batches"]):
        scores, loss = tr.step(torch.from_numpy(x).cuda(), torch.from_numpy(y).cuda())
        assert abs(float(loss) - c["losses"][s]) <= 1e-4, (s, float(loss), c["losses"][s])
        ref = c["scores"][s]
        assert np.abs(scores.cpu().numpy() - ref).max() <= 1e-3 * np.abs(ref).max() + 1e-4, s
        if s == 0:
            for k, v in c["grad0"].items():
                g = tr.read("grad:" + k)
                assert np.abs(g - v).max() <= 1e-3 * np.abs(v).max() + 1e-7, (k, np.abs(g - v).max(), np.abs(v).max())
    assert tr.steps == len(c["batches"])
    tr.sync()
    sd = model.state_dict()
    for k, v in c["final"].items():
        got = sd[k].cpu().numpy()
        if k.endswith("num_batches_tracked"):
            assert int(got) == int(v)
        else:
            assert np.allclose(got, v, atol=2e-4, rtol=1e-3), (k, np.abs(got - v).max())
    # the synced module serves inference with the trained weights (eval mode, running statistics)
    model.eval()
    model.precision = 1
    x = torch.from_numpy(c["batches"][0][0]).cuda()
    want = orc.res_forward({k: v for k, v in c["final"].items()}, c["config"], c["batches"][0][0]).numpy()
    got = model(x).cpu().numpy()
    assert np.abs(got - want).max() <= 2e-3 * np.abs(want).max()


def test_batch64_against_oracle_and_schedule_reset():
    """The reference's batch size (train.py:171) on fresh data, res15 with dilations: four steps against the oracle,
    with the optimiser rebuilt after step 2 like the learning-rate schedule does (train.py:140-143)."""
    import honk_b200 as hb
    from honk_b200 import train
    cfg = dict(hb.find_config("res15-narrow"))
    cfg["n_labels"] = 12
    clips, labels = synth.make_clips(4 * 64, start=810000, labelled=True)
    pre = hb.AudioPreprocessor()
    mfcc = pre.compute_mfccs_batch(torch.from_numpy(clips).cuda()).float()
    state = synth.make_state_dict("res15-narrow", cfg, seed=5, calib_mfcc=mfcc[:8].cpu().numpy())
    model = make_model("res15-narrow", cfg, state)
    tr = train.SgdTrainer(model, lr=0.05, momentum=0.9, weight_decay=1e-5)
    batches = [(mfcc[i * 64:(i + 1) * 64].cpu().numpy(), labels[i * 64:(i + 1) * 64].astype(np.int64)) for i in range(4)]
    o1 = orc.sgd_train_steps(state, cfg, batches[:2], lr=0.05, momentum=0.9, weight_decay=1e-5)
    o2 = orc.sgd_train_steps(o1["state"], cfg, batches[2:], lr=0.01, momentum=0.9, weight_decay=1e-5)   # new SGD: no buffers
    losses = []
    for i, (x, y) in enumerate(batches):
        if i == 2:
            tr.new_optimizer(0.01)
        _, loss = tr.step(torch.from_numpy(x).cuda(), torch.from_numpy(y).cuda())
        losses.append(float(loss))
    assert np.allclose(losses, o1["losses"] + o2["losses"], atol=2e-4), (losses, o1["losses"] + o2["losses"])
    for k, v in o2["momentum"].items():
        assert np.abs(tr.read("momentum:" + k) - v).max() <= 2e-3 * np.abs(v).max() + 1e-7, k
    tr.sync()
    for k, v in o2["state"].items():
        got = model.state_dict()[k].cpu().numpy()
        if k.endswith("num_batches_tracked"):
            assert int(got) == int(v) == 104
        else:
            assert np.allclose(got, v, atol=3e-4, rtol=2e-3), (k, np.abs(got - v).max())


def test_training_loss_falls_and_train_loop_runs(tmp_path, capsys):
    """train() (train.py:87-163) end to end on the class-structured synthetic clips: loss falls, dev accuracy is
    printed, the best model is saved in the reference's bare-state_dict format and reloads."""
    import honk_b200 as hb
    from honk_b200 import train
    cfg = dict(hb.find_config("res8-narrow"))
    cfg.update(n_labels=12, lr=[0.1, 0.02], schedule=[30], n_epochs=4, dev_every=2, momentum=0.9, weight_decay=1e-5,
               use_nesterov=False, output_file=str(tmp_path / "model.pt"), input_file="", model_class=hb.find_model("res8-narrow"))
    clips, labels = synth.make_clips(12 * 64, start=820000, labelled=True)
    pre = hb.AudioPreprocessor()
    mfcc = pre.compute_mfccs_batch(torch.from_numpy(clips).cuda()).float()
    y = torch.from_numpy(labels.astype(np.int64))
    loader = [(mfcc[i * 64:(i + 1) * 64], y[i * 64:(i + 1) * 64]) for i in range(10)]
    dev = [(mfcc[640:704], y[640:704]), (mfcc[704:768], y[704:768])]
    torch.manual_seed(0)
    best = train.train(cfg, loader, dev_loader=dev)
    out = capsys.readouterr().out
    assert "changing learning rate to 0.02" in out and "final dev accuracy" in out and "saving best model..." in out
    first = float(out.split("train step #1 ")[1].split("loss: ")[1].split()[0])
    last = float(out.split("train step #40 ")[1].split("loss: ")[1].split()[0])
    assert last < 0.6 * first, (first, last)
    again = cfg["model_class"](cfg)
    again.load(cfg["output_file"])
    again.cuda().eval()
    best.eval()
    assert torch.allclose(again(dev[0][0]), best(dev[0][0]), atol=1e-5)
    acc = float((best(dev[0][0]).argmax(1).cpu() == dev[0][1]).float().mean())
    assert acc > 0.5, acc


def test_trainer_errors():
    import honk_b200 as hb
    from honk_b200 import train
    cfg = dict(hb.find_config("res8-narrow"))
    cfg["n_labels"] = 12
    model = hb.find_model("res8-narrow")(cfg).cuda()
    tr = train.SgdTrainer(model, lr=0.1)
    x = torch.zeros((4, 101, 40), device="cuda")
    with pytest.raises(IndexError):
        tr.step(x, torch.tensor([0, 1, 12, 3]))
    with pytest.raises(ValueError):
        tr.step(x[:, :100], torch.tensor([0, 1, 2, 3]))
    with pytest.raises(ValueError):
        tr.step(x[:1], torch.tensor([0]))                      # BatchNorm in training mode needs more than one value
    assert tr.steps == 0
    cnn = dict(hb.find_config("cnn-trad-pool2"))
    cnn["n_labels"] = 12
    with pytest.raises(NotImplementedError):
        train.SgdTrainer(hb.find_model("cnn-trad-pool2")(cnn).cuda())
    with pytest.raises(RuntimeError):
        model.train()
        model(x)                                               # the nn.Module itself still only serves inference
