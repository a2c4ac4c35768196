"""Training step on the GPU (reference utils/train.py:127-139) through the C ABI (SgdTrainer ->
honk_trainer_step): against the reference's own run frozen in tests/golden/train_golden.npz, and against the oracle on a
fresh batch.  FP32 on both sides; tolerances are summation-order noise: 1e-4 on the loss, 1e-3 of the largest entry on a
gradient tensor, 2e-4 absolute + 1e-3 relative on the state after three steps."""
import numpy as np
import pytest
import torch

from honk_b200 import synth
from oracle import honk_oracle as orc
from test_train_oracle import CASES, load_case, sub

pytestmark = pytest.mark.gpu


def make_model(name, cfg, state):
    import honk_b200 as hb
    model = hb.find_model(name)(cfg)
    model.load_state_dict({k: torch.from_numpy(np.asarray(v)) for k, v in state.items()}, strict=True)
    return model.cuda()


def rel_err(a, truth):
    return float(np.abs(np.asarray(a, np.float64) - truth).max() / max(np.abs(truth).max(), 1e-30))


def float32_noise(c, truth, trials=6):
    """How far float32 arithmetic itself may land from the float64 result on this case: the float32 oracle on inputs and
    weights perturbed by about one ulp (6e-8 relative, seeded), worst relative error over the gradient tensors of step 0
    and over the final state.  The outcome is bimodal - ~4e-6 when no ReLU changes side, ~1.7e-3 when one activation
    within an ulp of zero does (res26-narrow, batch 8) - which is why a fixed small tolerance cannot be the bar."""
    lr, mom, wd, nest = c["hyper"]
    rng = np.random.default_rng(123)
    g_noise = max(rel_err(v, sub(truth["grads"][0][k])) for k, v in c["grad0"].items())     # the reference's own run
    s_noise = max(rel_err(v, sub(truth["state"][k])) for k, v in c["final"].items() if not k.endswith("num_batches_tracked"))
    for _ in range(trials):
        init = {k: (v * (1 + 6e-8 * rng.standard_normal(v.shape))).astype(np.float32) if k.endswith("weight") else v
                for k, v in c["init"].items()}
        batches = [[(b[0] * (1 + 6e-8 * rng.standard_normal(b[0].shape))).astype(np.float32)] + list(b[1:]) for b in c["batches"]]
        o = orc.sgd_train_steps(init, c["config"], batches, lr=lr, momentum=mom, weight_decay=wd, nesterov=bool(nest))
        g_noise = max(g_noise, max(rel_err(o["grads"][0][k], t) for k, t in truth["grads"][0].items()))
        s_noise = max(s_noise, max(rel_err(o["state"][k], t) for k, t in truth["state"].items() if not k.endswith("num_batches_tracked")))
    return g_noise, s_noise


@pytest.mark.parametrize("name", CASES)
def test_three_steps_match_the_reference_run(golden_dir, name):
    """Loss and scores against the reference's float32 run directly.  Gradients and the final state are judged against
    the float64 oracle with float32's own noise on the case as the yardstick (float32_noise): every tensor must be within
    2 x that noise + 2e-4 - about 2e-4 for the well-conditioned res8 cases, which is where an arithmetic slip shows."""
    from honk_b200 import train
    c = load_case(golden_dir, name)
    lr, mom, wd, nest = c["hyper"]
    truth = orc.sgd_train_steps(c["init"], c["config"], c["batches"], lr=lr, momentum=mom, weight_decay=wd, nesterov=bool(nest),
                                dtype=torch.float64)
    g_noise, s_noise = float32_noise(c, truth)
    model = make_model(name, c["config"], c["init"])
    tr = train.SgdTrainer(model, lr=lr, momentum=mom, weight_decay=wd, nesterov=bool(nest))
    for s, b in enumerate(c["batches"]):
        masks = b[2] if len(b) > 2 else None                       # cnn-*: the reference run's Dropout keep masks
        scores, loss = tr.step(torch.from_numpy(b[0]).cuda(), torch.from_numpy(b[1]).cuda(), dropout_masks=masks)
        assert abs(float(loss) - c["losses"][s]) <= 1e-4 * max(1.0, c["losses"][s]), (s, float(loss), c["losses"][s])
        ref = c["scores"][s]
        assert np.abs(scores.cpu().numpy() - ref).max() <= 1e-3 * np.abs(ref).max() + 1e-4, s
        if s == 0:
            for k in c["grad0"]:
                e_gpu = rel_err(tr.read("grad:" + k), truth["grads"][0][k])
                assert e_gpu <= 2 * g_noise + 2e-4, (k, e_gpu, g_noise)
    assert tr.steps == len(c["batches"])
    tr.sync()
    sd = model.state_dict()
    for k, v in c["final"].items():
        got = sd[k].cpu().numpy()
        if k.endswith("num_batches_tracked"):
            assert int(got) == int(v)
        else:
            e_gpu = rel_err(got, truth["state"][k])
            assert e_gpu <= 2 * s_noise + 2e-4, (k, e_gpu, s_noise)
    # the synced module serves inference with the trained weights (eval mode: running statistics / no Dropout)
    model.eval()
    model.precision = 1
    x = torch.from_numpy(c["batches"][0][0]).cuda()
    want = orc.model_forward(name, {k: v.cpu().numpy() for k, v in sd.items()}, c["config"], c["batches"][0][0], dtype=torch.float64).numpy()
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
    def run(dtype):
        o1 = orc.sgd_train_steps(state, cfg, batches[:2], lr=0.05, momentum=0.9, weight_decay=1e-5, dtype=dtype)
        o2 = orc.sgd_train_steps(o1["state"], cfg, batches[2:], lr=0.01, momentum=0.9, weight_decay=1e-5, dtype=dtype)   # new SGD: no buffers
        return o1["losses"] + o2["losses"], o2
    ref_losses, ref = run(torch.float32)
    true_losses, truth = run(torch.float64)
    losses = []
    for i, (x, y) in enumerate(batches):
        if i == 2:
            tr.new_optimizer(0.01)
        _, loss = tr.step(torch.from_numpy(x).cuda(), torch.from_numpy(y).cuda())
        losses.append(float(loss))
    assert np.allclose(losses, true_losses, atol=2e-4), (losses, true_losses)
    e_ref = max(rel_err(ref["momentum"][k], t) for k, t in truth["momentum"].items())
    for k, t in truth["momentum"].items():
        assert rel_err(tr.read("momentum:" + k), t) <= 5 * e_ref + 2e-4, (k, e_ref)
    tr.sync()
    e_ref = max(rel_err(ref["state"][k], t) for k, t in truth["state"].items() if not k.endswith("num_batches_tracked"))
    for k, t in truth["state"].items():
        got = model.state_dict()[k].cpu().numpy()
        if k.endswith("num_batches_tracked"):
            assert int(got) == int(t) == 104
        else:
            assert rel_err(got, t) <= 5 * e_ref + 2e-4, (k, rel_err(got, t), e_ref)


@pytest.mark.parametrize("name,lrs,fall,min_acc", [("res8-narrow", [0.1, 0.02], 0.65, 0.3), ("cnn-trad-pool2", [0.002, 0.0005], 0.5, 0.6)])
def test_training_loss_falls_and_train_loop_runs(tmp_path, capsys, name, lrs, fall, min_acc):
    """train() (train.py:87-163) end to end on the class-structured synthetic clips: loss falls, dev accuracy is
    printed, the best model is saved in the reference's bare-state_dict format and reloads.  cnn-trad-pool2 is the
    reference's default training model (train.py:168); its Dropout masks are the library's own draws here."""
    import honk_b200 as hb
    from honk_b200 import train
    cfg = dict(hb.find_config(name))
    cfg.update(n_labels=12, lr=lrs, schedule=[30], n_epochs=4, dev_every=2, momentum=0.9, weight_decay=1e-5,
               use_nesterov=False, output_file=str(tmp_path / "model.pt"), input_file="", model_class=hb.find_model(name))
    clips, labels = synth.make_clips(12 * 64, start=820000, labelled=True)
    pre = hb.AudioPreprocessor()
    mfcc = pre.compute_mfccs_batch(torch.from_numpy(clips).cuda()).float()
    y = torch.from_numpy(labels.astype(np.int64))
    loader = [(mfcc[i * 64:(i + 1) * 64], y[i * 64:(i + 1) * 64]) for i in range(10)]
    dev = [(mfcc[640:704], y[640:704]), (mfcc[704:768], y[704:768])]
    torch.manual_seed(0)
    best = train.train(cfg, loader, dev_loader=dev)
    out = capsys.readouterr().out
    assert ("changing learning rate to %s" % lrs[1]) in out and "final dev accuracy" in out and "saving best model..." in out
    losses = [float(out.split("train step #%d " % i)[1].split("loss: ")[1].split()[0]) for i in range(1, 41)]
    first, last = float(np.mean(losses[:4])), float(np.mean(losses[-4:]))
    assert last < fall * first, (first, last)
    again = cfg["model_class"](cfg)
    again.load(cfg["output_file"])
    again.cuda().eval()
    best.eval()
    assert torch.allclose(again(dev[0][0]), best(dev[0][0]), atol=1e-5)
    acc = float((best(dev[0][0]).argmax(1).cpu() == dev[0][1]).float().mean())
    assert acc > min_acc, acc


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
    with pytest.raises(ValueError):
        N_ = __import__("honk_b200._native", fromlist=["x"])
        N_.check(N_.lib().honk_trainer_set_dropout(tr.handle, 1.0, 0))    # probability must stay below 1
    with pytest.raises(RuntimeError):
        model.train()
        model(x)                                               # the nn.Module itself still only serves inference


def test_cnn_dropout_drawn_by_the_library_is_reproducible_and_has_the_right_rate():
    """Without caller masks the library draws the keep masks itself (counter-based hash of seed, step, site, element): two
    trainers with the same seed take identical steps, another seed differs, p = 0 reproduces the mask-free oracle, and with
    the loss still falling on the class-structured clips."""
    import honk_b200 as hb
    from honk_b200 import train
    name = "cnn-tstride8"
    cfg = dict(hb.find_config(name))
    cfg["n_labels"] = 12
    clips, labels = synth.make_clips(3 * 32, start=830000, labelled=True)
    pre = hb.AudioPreprocessor()
    mfcc = pre.compute_mfccs_batch(torch.from_numpy(clips).cuda()).float()
    y = torch.from_numpy(labels.astype(np.int64)).cuda()
    state = synth.make_state_dict(name, cfg, seed=4)
    state = {k: (v * np.float32(0.3) if k.endswith("weight") else v) for k, v in state.items()}

    def run(seed, p=0.5, steps=3):
        c2 = dict(cfg, dropout_prob=p)
        model = make_model(name, c2, state)
        tr = train.SgdTrainer(model, lr=0.002, momentum=0.9, weight_decay=1e-5, dropout_seed=seed)
        losses = [float(tr.step(mfcc[i * 32:(i + 1) * 32], y[i * 32:(i + 1) * 32])[1]) for i in range(steps)]
        return losses, tr.read("conv1.weight"), tr

    l1, w1, _ = run(7)
    l2, w2, _ = run(7)
    l3, w3, _ = run(8)
    assert l1 == l2 and np.array_equal(w1, w2)
    assert l1[0] != l3[0] and not np.array_equal(w1, w3)
    l0, w0, _ = run(7, p=0.0)
    o = orc.sgd_train_steps(state, dict(cfg, dropout_prob=0.0),
                            [(mfcc[i * 32:(i + 1) * 32].cpu().numpy(), labels[i * 32:(i + 1) * 32].astype(np.int64)) for i in range(3)],
                            lr=0.002, momentum=0.9, weight_decay=1e-5)
    assert np.allclose(l0, o["losses"], atol=2e-4), (l0, o["losses"])
    assert rel_err(w0, o["state"]["conv1.weight"].astype(np.float64)) <= 1e-3
    # keep rate of the drawn masks: the first step's scores with p = 0.5 differ from p = 0 but stay finite and of similar scale
    assert np.isfinite(l1).all() and abs(l1[0] - l0[0]) > 1e-4


@pytest.mark.parametrize("momentum,wd,nesterov", [(0.0, 0.0, False), (0.0, 1e-3, False), (0.5, 0.0, True)])
def test_sgd_variants_against_oracle(momentum, wd, nesterov):
    """Plain SGD (no momentum buffer), weight decay alone, Nesterov without weight decay: two steps against the oracle."""
    import honk_b200 as hb
    from honk_b200 import train
    name = "res8-narrow"
    cfg = dict(hb.find_config(name))
    cfg["n_labels"] = 12
    clips, labels = synth.make_clips(2 * 32, start=840000, labelled=True)
    pre = hb.AudioPreprocessor()
    mfcc = pre.compute_mfccs_batch(torch.from_numpy(clips).cuda()).float()
    state = synth.make_state_dict(name, cfg, seed=6, calib_mfcc=mfcc[:8].cpu().numpy())
    batches = [(mfcc[i * 32:(i + 1) * 32].cpu().numpy(), labels[i * 32:(i + 1) * 32].astype(np.int64)) for i in range(2)]
    o = orc.sgd_train_steps(state, cfg, batches, lr=0.05, momentum=momentum, weight_decay=wd, nesterov=nesterov)
    tr = train.SgdTrainer(make_model(name, cfg, state), lr=0.05, momentum=momentum, weight_decay=wd, nesterov=nesterov)
    losses = [float(tr.step(torch.from_numpy(x).cuda(), torch.from_numpy(y).cuda())[1]) for x, y in batches]
    assert np.allclose(losses, o["losses"], atol=1e-4), (losses, o["losses"])
    for k, v in o["state"].items():
        if not k.endswith("num_batches_tracked"):
            assert rel_err(tr.read(k), v.astype(np.float64)) <= 1e-3, k


@pytest.mark.parametrize("name", ["cnn-trad-pool2", "cnn-tstride4"])
def test_cnn_batch64_with_supplied_masks_against_oracle(name):
    """The reference's batch size (train.py:171) for the cnn-* family: two steps with caller-supplied Dropout masks against the
    float64 oracle, float32's own noise on the case (the float32 oracle) as the yardstick like in the res* test."""
    import honk_b200 as hb
    from honk_b200 import train
    cfg = dict(hb.find_config(name))
    cfg["n_labels"] = 12
    B = 64
    clips, labels = synth.make_clips(2 * B, start=850000, labelled=True)
    pre = hb.AudioPreprocessor()
    mfcc = pre.compute_mfccs_batch(torch.from_numpy(clips).cuda()).float()
    state = synth.make_state_dict(name, cfg, seed=8)
    state = {k: (v * np.float32(0.3) if k.endswith("weight") else v) for k, v in state.items()}
    model = make_model(name, cfg, state)
    rng = np.random.default_rng(17)
    h1 = (101 - cfg["conv1_size"][0]) // cfg["conv1_stride"][0] + 1
    w1 = (40 - cfg["conv1_size"][1]) // cfg["conv1_stride"][1] + 1
    hp, wp = h1 // cfg["conv1_pool"][0], w1 // cfg["conv1_pool"][1]
    shapes = [(B, cfg["n_feature_maps1"], h1, w1), (B, cfg["n_feature_maps2"], hp - cfg["conv2_size"][0] + 1, wp - cfg["conv2_size"][1] + 1),
              (B, cfg.get("dnn1_size", 0)), (B, cfg.get("dnn2_size", 0))]
    batches = []
    for i in range(2):
        masks = [rng.integers(0, 2, sh).astype(np.uint8) if sh[-1] > 0 else None for sh in shapes]
        batches.append((mfcc[i * B:(i + 1) * B].cpu().numpy(), labels[i * B:(i + 1) * B].astype(np.int64), masks))
    ref = orc.sgd_train_steps(state, cfg, batches, lr=0.002, momentum=0.9, weight_decay=1e-5)
    truth = orc.sgd_train_steps(state, cfg, batches, lr=0.002, momentum=0.9, weight_decay=1e-5, dtype=torch.float64)
    tr = train.SgdTrainer(model, lr=0.002, momentum=0.9, weight_decay=1e-5)
    losses = [float(tr.step(torch.from_numpy(x).cuda(), torch.from_numpy(y).cuda(), dropout_masks=m)[1]) for x, y, m in batches]
    assert np.allclose(losses, truth["losses"], atol=2e-4, rtol=1e-4), (losses, truth["losses"])
    noise = max(rel_err(ref["state"][k], t) for k, t in truth["state"].items())
    for k, t in truth["state"].items():
        assert rel_err(tr.read(k), t) <= 2 * noise + 2e-4, (k, rel_err(tr.read(k), t), noise)
    noise_g = max(rel_err(ref["momentum"][k], t) for k, t in truth["momentum"].items())
    for k, t in truth["momentum"].items():
        assert rel_err(tr.read("momentum:" + k), t) <= 2 * noise_g + 2e-4, (k, noise_g)
