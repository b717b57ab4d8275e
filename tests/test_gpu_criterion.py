"""Fused RewardCriterion (model.py:7-23) against a plain PyTorch fp32 restatement of
the same op, forward and backward.  Floating point: tolerance 1e-5 relative."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def torch_reference(seq, logprobs, reward):
    """fp32 eager restatement of model.py:12-23."""
    mask = (seq > 0).float()
    mask = torch.cat([mask.new_ones(mask.size(0), 1), mask[:, :-1]], 1)
    out = -logprobs * reward * mask
    return out.sum() / mask.sum()


@pytest.fixture(scope='module')
def cst():
    import __graft_entry__ as graft
    graft.build()
    import cst_captioning_b200
    return cst_captioning_b200


@pytest.mark.parametrize('n,t,dtype', [(7, 5, torch.int64), (1280, 28, torch.int64), (4099, 28, torch.int32),
                                       (33, 40, torch.int64), (5, 3, torch.int64), (1, 1, torch.int32),
                                       (3, 7, torch.int32), (70001, 27, torch.int64), (300000, 28, torch.int64)])
def test_forward_backward_match_torch(cst, n, t, dtype):
    dev = torch.device('cuda:0')
    g = torch.Generator(device='cpu').manual_seed(n * 31 + t)
    seq = torch.randint(0, 50, (n, t), generator=g)
    seq[seq < 8] = 0                                   # plenty of <end>/padding
    seq[0] = 0                                         # an all-padding row: only its first step counts
    seq = seq.to(dev).to(dtype)
    logp = (-torch.rand(n, t, generator=g) * 5).to(dev).requires_grad_(True)
    col = torch.randn(n, generator=g).to(dev)
    crit = cst.RewardCriterion()
    for reward in (col.unsqueeze(1).expand(n, t), col, col.unsqueeze(1).repeat(1, t) + 0.0):
        logp.grad = None
        loss = crit(seq, logp, reward)
        (loss * 3.0).backward()
        got_grad = logp.grad.clone()
        logp.grad = None
        full = reward if reward.dim() == 2 else reward.unsqueeze(1).expand(n, t)
        want = torch_reference(seq.long(), logp, full)
        (want * 3.0).backward()
        np.testing.assert_allclose(loss.item(), want.item(), rtol=1e-5, atol=1e-7)
        np.testing.assert_allclose(got_grad.cpu().numpy(), logp.grad.cpu().numpy(), rtol=1e-5, atol=1e-9)
    # deterministic: the same inputs give the same bits
    a = crit(seq, logp.detach(), col).item()
    b = crit(seq, logp.detach(), col).item()
    assert a == b
    # row-strided views (a slice of a wider buffer) take the row-per-warp kernels; dense (N, T)
    # tensors the element-stream kernels: both must give the same loss and gradient
    wide_seq = torch.zeros(n, t + 3, dtype=seq.dtype, device=dev)
    wide_seq[:, :t] = seq
    wide_lp = torch.zeros(n, t + 5, device=dev)
    wide_lp[:, :t] = logp.detach()
    lp_view = wide_lp[:, :t].requires_grad_(True)
    loss_v = crit(wide_seq[:, :t], lp_view, col)
    g_v, = torch.autograd.grad(loss_v, lp_view)
    lp_d = logp.detach().clone().requires_grad_(True)
    loss_d = crit(seq, lp_d, col)
    g_d, = torch.autograd.grad(loss_d, lp_d)
    np.testing.assert_allclose(loss_v.item(), loss_d.item(), rtol=1e-6, atol=1e-8)
    assert torch.equal(g_v, g_d) or torch.allclose(g_v, g_d, rtol=1e-6, atol=1e-12)


def test_reward_path_feeds_the_criterion(cst):
    """get_cst_reward -> RewardCriterion without the (N, T) reward ever being built."""
    from cst_captioning_b200 import synth
    dev = torch.device('cuda:0')
    rng = np.random.default_rng(2)
    corpus, _ = synth.make_refs(rng, 200, 20)
    scorer = cst.CiderD(df=cst.DfTable.from_rows(torch.from_numpy(corpus).to(dev), use_eos=1))
    b = synth.make_batch(16, 20, 20, seed=3)
    seq = torch.from_numpy(b['sampled']).to(dev).long()
    reward, m, bs = cst.get_cst_reward(seq, torch.from_numpy(b['refs']).to(dev), scorer, expand_feat=1,
                                       seq_per_img=20, scb_captions=20, scb_baseline=2, use_eos=1)
    assert reward.stride(1) == 0 and tuple(reward.shape) == tuple(seq.shape)
    logp = (-torch.rand(seq.shape, device=dev)).requires_grad_(True)
    loss = cst.RewardCriterion()(seq, logp, reward)
    loss.backward()
    want = torch_reference(seq, logp.detach(), reward)
    np.testing.assert_allclose(loss.item(), want.item(), rtol=1e-5, atol=1e-8)
    with pytest.raises(ValueError):
        cst.RewardCriterion()(seq.cpu(), logp.detach().cpu(), reward.cpu())
