"""The callers either side of reduce(): the upstream training loop restated in harness.py (ddp_init.py:185-222)
with this reducer dropped in, and its fused form (reduce_fused + FusedMomentumSGD), against the same loop driven by
the oracle reducer (the reference algorithm as plain torch ops on the GPU)."""
import pytest
import torch

from tests.helpers import rel_err

pytestmark = pytest.mark.gpu


def test_fused_momentum_sgd_matches_torch_ops():
    """fp32 elementwise kernel vs the plain PyTorch fp32 ops of ddp_init.py:210-222; tolerance 1e-6 relative
    (one FMA contraction per element)."""
    from network_distributed_pytorch_b200.train_step import FusedMomentumSGD

    dev = torch.device("cuda", 0)
    torch.manual_seed(0)
    shapes = [(64, 33), (7,), (3, 5, 2), (4099,), (256, 256)]
    p = [torch.randn(s, device=dev) for s in shapes]
    m = [torch.empty(s, device=dev) for s in shapes]
    p_ref, m_ref = [t.clone() for t in p], [t.clone() for t in m]
    opt = FusedMomentumSGD(p, lr=0.01, momentum=0.9)
    for step in range(3):
        g = [torch.randn(s, device=dev) for s in shapes]
        g_ref = [t.clone() for t in g]
        for gr, mr, pr in zip(g_ref, m_ref, p_ref):
            if step == 0:
                mr.copy_(gr)
            else:
                mr.mul_(0.9).add_(gr)
            gr += mr
            pr.add_(gr, alpha=-0.01)
        opt.step(p, g, m)
        torch.cuda.synchronize()
        for a, b in zip(p + g + m, p_ref + g_ref + m_ref):
            assert rel_err(a.cpu().numpy(), b.cpu().numpy()) < 1e-6


@pytest.mark.parametrize("loop", ["upstream", "fused"])
def test_training_loop_matches_oracle_reducer(loop):
    """Three chained steps of Algorithm 2 on a tiny DistilBERT.  Three labels on purpose: with two, the classifier
    gradient has exactly rank 1 (its rows are negatives of each other), P = M Q is singular and the second
    Gram-Schmidt column is normalised rounding noise in the reference and here alike -- nothing to compare.
    Parameters whose update is itself noise (e.g. k_lin.bias, whose true gradient is zero) are skipped."""
    pytest.importorskip("transformers")
    from network_distributed_pytorch_b200 import harness
    from network_distributed_pytorch_b200.reducer import PowerSGDReducer
    from oracle.powersgd_oracle import OracleReducer
    from tests.gpu_common import ensure_group

    ensure_group()
    dev = torch.device("cuda", 0)

    def run(reducer_factory, which):
        torch.manual_seed(123)
        model, vocab, kind = harness.build_model("distilbert-tiny", dev, dropout=0.0, num_labels=3)
        init = [p.detach().clone() for p in model.parameters()]
        gen = torch.Generator(device=dev).manual_seed(9)
        batches = [harness.make_batch(kind, vocab, 4, 32, dev, gen, num_labels=3) for _ in range(3)]
        state = harness.train_steps(model, kind, batches, reducer_factory(), 3, lr=0.05, loop=which, autocast=False)
        return init, [p.detach().clone() for p in model.parameters()], state["losses"]

    init, ref_params, ref_losses = run(lambda: OracleReducer(714, dev, reuse_query=True, rank=2), "upstream")
    _, got_params, got_losses = run(lambda: PowerSGDReducer(714, dev, 0, True, 2), loop)
    assert abs(ref_losses[0] - got_losses[0]) < 1e-6  # same init, same batch
    assert all(abs(a - b) < 2e-5 for a, b in zip(ref_losses, got_losses)), (ref_losses, got_losses)
    compared = 0
    for p0, a, b in zip(init, got_params, ref_params):
        upd = (b - p0).norm().item()
        if upd < 1e-4 * max(p0.norm().item(), 1e-3):
            continue  # the reference's own update of this tensor is noise
        compared += 1
        assert rel_err((a - p0).cpu().numpy(), (b - p0).cpu().numpy()) < 1e-3
    assert compared >= 20
