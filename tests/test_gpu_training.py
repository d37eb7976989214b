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
