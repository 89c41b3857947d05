"""tensor_buffer.TensorBuffer on the GPU: the flat layout must be bit-identical to torch.cat of the
flattened inputs (tensor_buffer.py:19) and unpack must be its inverse (tensor_buffer.py:34-36)."""
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("dtype", [torch.float32, torch.float16, torch.bfloat16, torch.float64, torch.int64, torch.uint8])
def test_pack_is_cat_and_unpack_inverts(dtype):
    from network_distributed_pytorch_b200.tensor_buffer import TensorBuffer

    dev = torch.device("cuda", 0)
    shapes = [(7,), (3, 5), (1,), (0,), (2, 3, 4), (1025,)]
    torch.manual_seed(0)
    ts = [(torch.randn(s, device=dev) * 50).to(dtype) for s in shapes]
    tb = TensorBuffer(ts)
    want = torch.cat([t.view(-1) for t in ts])
    assert tb.buffer.dtype == dtype and torch.equal(tb.buffer, want)
    assert len(tb) == len(ts) and tb.nelement() == want.nelement() and tb.element_size() == want.element_size()
    assert tb.bits() == 8 * want.nelement() * want.element_size()
    for i, t in enumerate(ts):
        assert tb[i].shape == t.shape and torch.equal(tb[i], t)
        if t.nelement():  # a view into the flat buffer, not a copy
            assert tb[i].data_ptr() == tb.buffer.data_ptr() + tb._start_idx[i] * tb.element_size()
    outs = [torch.zeros_like(t) for t in ts]
    tb.unpack(outs)
    assert all(torch.equal(a, b) for a, b in zip(outs, ts))
    # pack(new tensors) overwrites the buffer
    ts2 = [t + 1 if t.is_floating_point() else t for t in ts]
    tb.pack(ts2)
    assert torch.equal(tb.buffer, torch.cat([t.view(-1) for t in ts2]))


def test_errors_and_world1_collectives():
    import torch.distributed as dist

    from network_distributed_pytorch_b200 import tensor_buffer as tbm
    from tests.gpu_common import ensure_group

    ensure_group()
    dev = torch.device("cuda", 0)
    with pytest.raises(RuntimeError, match="non-empty"):
        tbm.TensorBuffer([])
    with pytest.raises(RuntimeError, match="CUDA"):
        tbm.TensorBuffer([torch.zeros(3)])
    tb = tbm.TensorBuffer([torch.ones(4, device=dev), torch.ones(2, 2, device=dev)])
    assert tbm.all_reduce(tb.buffer) is None  # world 1: guarded (tensor_buffer.py:59-61)
    out = [torch.empty_like(tb.buffer)]
    tbm.all_gather(out, tb.buffer)
    assert out[0].data_ptr() == tb.buffer.data_ptr()  # tensor_buffer.py:68-69 aliases
    assert dist.get_world_size() == 1


def test_dense_reducer_world1_is_a_copy():
    from network_distributed_pytorch_b200.dense import DenseReducer
    from tests.gpu_common import ensure_group

    ensure_group()
    dev = torch.device("cuda", 0)
    shapes = [(64, 33), (7,), (3, 5, 2)]
    g = [torch.randn(s, device=dev) for s in shapes]
    out = [torch.zeros(s, device=dev) for s in shapes]
    red = DenseReducer(714, dev)
    bits = red.reduce(g, out, None)
    assert bits == 32 * sum(t.nelement() for t in g)
    assert all(torch.equal(a, b) for a, b in zip(g, out))
    g2 = [t * 2 for t in g]
    red.reduce(g2, out, None)
    assert all(torch.equal(a, b) for a, b in zip(g2, out))
