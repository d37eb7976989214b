"""GPU parity of the training step (forward, mean cross-entropy, backward, Adam) against the oracle
``oracle.dscnn_ref.cross_entropy_train_step`` (float64 autograd restatement of the reference step body,
kws/libs/training.py:285-297 with CrossEntropyLoss :115 and Adam(lr=1e-3) :119-123)."""
import numpy as np
import pytest
import torch

from kws_b200.models import DepthwiseSeparableConv
from kws_b200.training import KWSTrainStep, TrainingConfig, flatten_state_dict
from oracle import dscnn_ref

pytestmark = pytest.mark.gpu


def _setup(B, T, F, seed):
    sd = dscnn_ref.init_state_dict(seed=seed)
    m = DepthwiseSeparableConv()
    m.load_state_dict(sd)
    g = torch.Generator().manual_seed(seed + 7)
    x = torch.randn(B, 1, T, F, generator=g)
    y = torch.randint(0, 12, (B,), generator=g)
    return sd, m, x, y


@pytest.mark.parametrize("B,T,F", [(16, 99, 10), (5, 99, 40), (3, 49, 10)])
def test_gradients_and_loss_match_oracle(B, T, F):
    sd, m, x, y = _setup(B, T, F, seed=B + F)
    ts = KWSTrainStep(m.cuda(), TrainingConfig(), device="cuda:0")
    loss, logits = ts.forward_backward(x.cuda(), y.cuda())
    sd64 = {k: v.double() for k, v in sd.items()}
    ref_loss, ref_grads, _, _ = dscnn_ref.cross_entropy_train_step(sd64, x.double(), y)
    ref_logits = dscnn_ref.forward(sd64, x.double())
    np.testing.assert_allclose(logits.cpu().numpy(), ref_logits.numpy(), rtol=0, atol=2e-5)
    assert abs(float(loss.item()) - float(ref_loss)) < 1e-5
    got = ts.gradients()
    assert list(got.keys()) == list(sd.keys())
    for k, g in got.items():
        r = ref_grads[k].numpy()
        scale = max(np.abs(r).max(), 1e-6)
        np.testing.assert_allclose(g.cpu().numpy(), r, rtol=1e-3, atol=2e-5 * scale, err_msg=k)


def test_adam_steps_match_oracle():
    B, T, F = 8, 99, 10
    sd, m, x, y = _setup(B, T, F, seed=3)
    ts = KWSTrainStep(m.cuda(), TrainingConfig(learning_rate=1e-3), device="cuda:0")
    cur = {k: v.double() for k, v in sd.items()}
    state = None
    losses = []
    for it in range(3):
        loss, _ = ts.step(x.cuda(), y.cuda())
        ref_loss, grads, cur, state = dscnn_ref.cross_entropy_train_step(cur, x.double(), y, state)
        losses.append((float(loss.item()), float(ref_loss)))
        got = ts.state_dict()
        for k in cur:
            a, r, g = got[k].cpu().numpy(), cur[k].numpy(), grads[k].numpy()
            firm = np.abs(g) > 1e-6          # Adam's first updates are ~ lr * sign(g): skip the sign-ambiguous ones
            np.testing.assert_allclose(a[firm], r[firm], rtol=0, atol=2e-5, err_msg=f"{k} step {it}")
            assert np.abs(a - r).max() <= 2.1e-3 * (it + 1)
    for a, r in losses:
        assert abs(a - r) < 2e-4
    assert losses[-1][0] < losses[0][0]      # it learns the fixed batch
    # trained parameters flow back into the module and its inference kernels
    ts.sync_to_model()
    want = dscnn_ref.forward(cur, x.double()).numpy()
    np.testing.assert_allclose(ts.model(x.cuda()).cpu().numpy(), want, rtol=1e-3, atol=2e-4)


def test_flat_layout_is_state_dict_order():
    sd = dscnn_ref.init_state_dict(seed=1)
    flat = flatten_state_dict(sd, "cpu")
    assert flat.numel() == 26444
    m = DepthwiseSeparableConv()
    m.load_state_dict(sd)
    ts = KWSTrainStep(m.cuda(), device="cuda:0")
    np.testing.assert_array_equal(ts.params.cpu().numpy(), flat.numpy())


def test_training_does_not_disturb_the_inference_handle_until_sync():
    """The step trains on its own native handle: the module keeps answering with the weights it holds
    until ``sync_to_model()`` copies the trained parameters back."""
    sd, m, x, y = _setup(8, 99, 10, seed=21)
    m = m.cuda()
    before = m(x.cuda()).cpu().numpy()
    ts = KWSTrainStep(m, TrainingConfig(learning_rate=1e-2), device="cuda:0")
    for _ in range(3):
        ts.step(x, y)
    np.testing.assert_array_equal(m(x.cuda()).cpu().numpy(), before)
    ts.sync_to_model()
    after = m(x.cuda()).cpu().numpy()
    assert np.abs(after - before).max() > 1e-3
    want = dscnn_ref.forward({k: v.cpu().double() for k, v in ts.state_dict().items()}, x.double()).numpy()
    np.testing.assert_allclose(after, want, rtol=1e-3, atol=2e-4)


def test_four_steps_match_the_reference_trainer(golden_dir):
    """One epoch (four batches of 16) of the REFERENCE's own KWSTrainer.train_epoch
    (kws/libs/training.py:266-345), recorded by tests/golden/make_golden.py::golden_trainer: same start, same
    batches -> same epoch loss and the same parameters."""
    import os
    g = np.load(os.path.join(golden_dir, "ref_trainer_epoch.npz"))
    sd0 = {k[4:]: torch.from_numpy(g[k]) for k in g.files if k.startswith("sd0.")}
    m = DepthwiseSeparableConv()
    m.load_state_dict(sd0)
    ts = KWSTrainStep(m.cuda(), TrainingConfig(learning_rate=1e-3), device="cuda:0", check_labels=True)
    xs, ys = torch.from_numpy(g["x"]), torch.from_numpy(g["y"])
    losses = []
    for i in range(xs.shape[0]):
        loss, _ = ts.step(xs[i], ys[i])
        losses.append(float(loss.item()))
    assert abs(np.mean(losses) - float(g["epoch_loss"])) < 2e-4
    got = ts.state_dict()
    for k in sd0:
        a, r = got[k].cpu().numpy(), g["sd1." + k]
        moved = np.abs(r - g["sd0." + k])
        # Adam's early updates are ~ lr * sign(g) per step: parameters whose gradient is rounding noise may
        # step the other way, so the bound is 2 * 4 steps * lr there and tight where the update is firm
        assert np.abs(a - r).max() <= 8.2e-3, k
        firm = moved > 3.5e-3
        if firm.any():
            np.testing.assert_allclose(a[firm], r[firm], rtol=0, atol=2e-4, err_msg=k)
    assert np.mean([np.abs(got[k].cpu().numpy() - g["sd1." + k]).mean() for k in sd0]) < 2e-4


def test_kwstrainer_adapter_and_reference_checkpoint(golden_dir, tmp_path):
    """f3: the KWSTrainer drop-in runs train_epoch / validate / save_checkpoint over a fake loader, loads the
    checkpoint the REFERENCE trainer wrote (tests/golden/ref_trainer_checkpoint.pth: parameters AND Adam
    moments), and a further step from it equals torch.optim.Adam continuing from that checkpoint."""
    import os

    from kws_b200.training import KWSTrainer
    g = np.load(os.path.join(golden_dir, "ref_trainer_epoch.npz"))
    xs, ys = torch.from_numpy(g["x"]), torch.from_numpy(g["y"])
    loader = [(xs[i], ys[i]) for i in range(xs.shape[0])]
    sd0 = {k[4:]: torch.from_numpy(g[k]) for k in g.files if k.startswith("sd0.")}
    m = DepthwiseSeparableConv()
    m.load_state_dict(sd0)
    cfg = TrainingConfig(epochs=1, learning_rate=1e-3, batch_size=16, checkpoint_dir=str(tmp_path))
    tr = KWSTrainer(m, loader, loader, cfg, device="cuda:0")
    loss, acc = tr.train_epoch(0)
    assert abs(loss - float(g["epoch_loss"])) < 2e-4 and 0.0 <= acc <= 100.0
    vloss, vacc = tr.validate(0)
    assert np.isfinite(vloss) and 0.0 <= vacc <= 100.0
    path = tr.save_checkpoint("checkpoint_epoch_1.pth")
    ck = torch.load(path, map_location="cpu", weights_only=False)
    assert set(ck) == {"model_state_dict", "optimizer_state_dict", "metrics", "config", "epoch"}
    # our optimizer state loads into a real torch.optim.Adam over the reference's parameter list
    ref_params = [torch.nn.Parameter(v.clone()) for v in ck["model_state_dict"].values()]
    opt = torch.optim.Adam(ref_params, lr=1e-3)
    opt.load_state_dict(ck["optimizer_state_dict"])
    assert int(opt.state[ref_params[0]]["step"]) == 4

    # now the reference's own checkpoint: resume and take one more step
    m2 = DepthwiseSeparableConv()
    tr2 = KWSTrainer(m2, loader, loader, cfg, device="cuda:0")
    tr2.load_checkpoint(os.path.join(golden_dir, "ref_trainer_checkpoint.pth"))
    got = tr2.step_fn.state_dict()
    for k in sd0:
        np.testing.assert_array_equal(got[k].cpu().numpy(), g["sd1." + k])
    assert tr2.step_fn.step_count == 4
    rck = torch.load(os.path.join(golden_dir, "ref_trainer_checkpoint.pth"), map_location="cpu", weights_only=False)
    params = {k: v.clone().double().requires_grad_(True) for k, v in rck["model_state_dict"].items()}
    opt = torch.optim.Adam(list(params.values()), lr=1e-3)
    osd = rck["optimizer_state_dict"]
    osd = {"state": {i: {kk: (vv.double() if torch.is_tensor(vv) and vv.dim() else vv) for kk, vv in st.items()}
                     for i, st in osd["state"].items()}, "param_groups": osd["param_groups"]}
    opt.load_state_dict(osd)
    torch.nn.functional.cross_entropy(dscnn_ref.forward(params, xs[0].double()), ys[0]).backward()
    opt.step()
    tr2.step_fn.step(xs[0], ys[0])
    got = tr2.step_fn.state_dict()
    for k in sd0:   # fifth Adam step: moments are warm, the update is firm everywhere
        np.testing.assert_allclose(got[k].cpu().numpy(), params[k].detach().numpy(), rtol=0, atol=3e-4, err_msg=k)


def test_out_of_range_label_is_loud():
    sd, m, x, y = _setup(4, 99, 10, seed=5)
    ts = KWSTrainStep(m.cuda(), TrainingConfig(), device="cuda:0")
    y_bad = y.clone()
    y_bad[2] = 12
    loss, _ = ts.forward_backward(x, y_bad)
    assert torch.isnan(loss).item()
    assert all(torch.isfinite(v).all().item() for v in ts.gradients().values())
    from kws_b200.common.errors import ModelError
    ts2 = KWSTrainStep(m, TrainingConfig(), device="cuda:0", check_labels=True)
    with pytest.raises(ModelError):
        ts2.forward_backward(x, y_bad)


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs (NCCL)")
def test_two_rank_nccl_step_equals_single_process(tmp_path):
    """BASELINE config 5: two ranks, UNEQUAL shards (9 + 7 samples) of one 16-sample batch, one NCCL
    all-reduce of the flat gradient buffer per step; after three steps both ranks hold the parameters the
    single-process step computes on the whole batch."""
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    out = tmp_path / "nccl.json"
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
           "--master-port", "29631", os.path.join(root, "tests", "nccl_train_worker.py"), str(out)]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    import json
    res = json.load(open(out))
    # Adam's first steps are ~ lr * sign(g): the few parameters whose gradient is rounding noise may differ by up to
    # 2 * 3 steps * lr between two summation orders; everything else agrees to fp32 rounding
    assert res["world"] == 2 and res["ranks_equal"]
    assert res["grad_rel_diff"] < 1e-5 and res["frac_within_2e-5"] > 0.995 and res["max_abs_diff_vs_single"] < 6.1e-3
