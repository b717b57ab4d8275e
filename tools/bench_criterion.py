"""Roofline check of the fused RewardCriterion kernels (HBM-bound streaming)."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import __graft_entry__ as g
g.build()
import cst_captioning_b200 as cst

dev = torch.device('cuda:0')
n, t = 1310720, 28
seq = torch.randint(0, 10000, (n, t), device=dev)
seq[torch.rand(n, t, device=dev) < 0.3] = 0
logp = (-torch.rand(n, t, device=dev)).requires_grad_(True)
col = torch.randn(n, device=dev)
crit = cst.RewardCriterion()
peak = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), 'MEASURED_PEAKS.json')))['hbm_gbs']


def timed(fn, reps=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps


fwd_ms = timed(lambda: crit(seq, logp.detach(), col))
fwd_bytes = n * t * (8 + 4) + n * 4                      # int64 seq + fp32 logp read once, reward column
loss = crit(seq, logp, col)
bwd_ms = timed(lambda: torch.autograd.grad(loss, logp, retain_graph=True))
bwd_bytes = n * t * (8 + 4) + n * 4                      # int64 seq read, fp32 grad written, reward column


def eager():
    mask = (seq > 0).float()
    mask = torch.cat([mask.new_ones(n, 1), mask[:, :-1]], 1)
    return (-logp.detach() * col.unsqueeze(1) * mask).sum() / mask.sum()


eager_ms = timed(eager)
print(json.dumps({'rows': n, 'width': t,
                  'fwd_ms': fwd_ms, 'fwd_GBs': fwd_bytes / fwd_ms / 1e6, 'fwd_frac_of_measured_peak': fwd_bytes / fwd_ms / 1e6 / peak,
                  'bwd_ms': bwd_ms, 'bwd_GBs': bwd_bytes / bwd_ms / 1e6, 'bwd_frac_of_measured_peak': bwd_bytes / bwd_ms / 1e6 / peak,
                  'torch_eager_fwd_ms': eager_ms}))
