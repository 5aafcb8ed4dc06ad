"""CPU tier for the pieces around the engine-backed API path that need no GPU: network recognition, chain linking of the fused
slots, U8Images semantics, the loader's samplers (row-level and block-level), automatic reshuffling, the evaluation accumulator and
optimizer dispatch."""
import pytest
import torch
from torch import nn
from torch.utils.data import DataLoader

from ray_torch_distributed_checkpoint_b200.data import (DeviceLoader, DistributedSampler, FashionMNISTSynthetic, U8BatchLoader,
                                                        U8Images)
from ray_torch_distributed_checkpoint_b200.data.loader import BlockSampler
from ray_torch_distributed_checkpoint_b200.models import NeuralNetwork, fuse_linear_chains
from ray_torch_distributed_checkpoint_b200.nn import CrossEntropyLoss, EvalAccumulator
from ray_torch_distributed_checkpoint_b200.optim import SGD
from ray_torch_distributed_checkpoint_b200.parallel.engine import match_reference_mlp


def test_reference_network_is_recognised_and_others_are_not():
    spec = match_reference_mlp(NeuralNetwork())
    assert spec == {"prefix": "linear_relu_stack.", "names": [f"linear_relu_stack.{i}" for i in (0, 3, 6)], "p_drop": 0.25}
    assert match_reference_mlp(fuse_linear_chains(NeuralNetwork())) == spec            # also after the module-path rewrite
    assert match_reference_mlp(NeuralNetwork(hidden=256, classes=16, p_drop=0.0))["p_drop"] == 0.0
    assert match_reference_mlp(NeuralNetwork(hidden=1024)) is None                     # hidden width beyond the fused head
    assert match_reference_mlp(NeuralNetwork(classes=12)) is None
    assert match_reference_mlp(nn.Sequential(nn.Linear(8, 8), nn.ReLU())) is None
    no_trailing_relu = nn.Sequential(nn.Flatten(), nn.Sequential(nn.Linear(784, 512), nn.ReLU(), nn.Linear(512, 512), nn.ReLU(), nn.Linear(512, 10)))
    assert match_reference_mlp(no_trailing_relu) is None


def test_fused_slots_are_chained_inside_a_sequential():
    net = fuse_linear_chains(NeuralNetwork())
    s = net.linear_relu_stack
    assert s[0].next is s[3] and s[3].prev is s[0] and s[3].next is s[6] and s[6].prev is s[3] and s[6].next is None
    assert list(net.state_dict()) == list(NeuralNetwork().state_dict())
    x = torch.randn(4, 1, 28, 28)
    net.eval()
    ref = NeuralNetwork()
    ref.load_state_dict(net.state_dict())
    ref.eval()
    assert torch.allclose(net(x), ref(x), atol=1e-6)                                   # CPU fallback of the slots


def test_u8_images_are_bytes_until_computed_on():
    raw = torch.randint(0, 256, (6, 1, 28, 28), dtype=torch.uint8)
    x = U8Images.wrap(raw, 0.5, 0.5)
    assert isinstance(x, torch.Tensor) and x.dtype == torch.uint8 and x.shape == raw.shape and len(x) == 6 and x.dim() == 4
    assert x.raw.data_ptr() == raw.data_ptr() and type(x.raw) is torch.Tensor
    want = (raw.float() / 255 - 0.5) / 0.5
    assert torch.allclose(x.float(), want) and torch.allclose(nn.Flatten()(x), want.flatten(1)) and torch.allclose(x * 2, want * 2)
    assert type(x + 0) is torch.Tensor and x.norm == (0.5, 0.5)


def test_samplers_row_and_block_level():
    for n, w, dl in ((1000, 3, False), (1000, 3, True), (7, 4, False), (1000, 1, False)):
        for r in range(w):
            for sh in (True, False):
                s = DistributedSampler(n, w, r, shuffle=sh, seed=3, drop_last=dl)
                s.set_epoch(2)
                assert s.indices() == s.indices_tensor().tolist()
    seen = set()
    sizes = set()
    for r in range(3):
        b = BlockSampler(1000, 64, 3, r, shuffle=True, seed=1)
        b.set_epoch(4)
        rows = list(b)
        sizes.add(len(rows))
        assert len(rows) == len(b) and max(rows) < 1000
        seen |= set(rows)
    assert seen == set(range(1000)) and len(sizes) == 1                                # every row visited, equal shards
    b0 = BlockSampler(1000, 64, 1, 0, shuffle=True, seed=1)
    e0 = b0.block_starts()
    b0.set_epoch(1)
    assert sorted(e0) == sorted(b0.block_starts()) and e0 != b0.block_starts()       # same blocks, new order


def test_u8_loader_on_cpu_reshuffles_keeps_tail_batch_and_byte_labels():
    ds = FashionMNISTSynthetic(train=False, size=300)
    ld = U8BatchLoader(ds.data, ds.targets, 128, "cpu", shuffle=True, drop_last=False, wrap=True)
    assert ld.transfer == "host_gather" and ld.label_dtype == torch.uint8 and len(ld) == 3
    a = [(x, y) for x, y in ld]
    assert [x.shape[0] for x, _ in a] == [128, 128, 44] and isinstance(a[0][0], U8Images) and a[0][1].dtype == torch.int64
    ya, yb = torch.cat([y for _, y in a]), torch.cat([y for _, y in ld])
    assert sorted(ya.tolist()) == sorted(ds.targets.tolist()) and not torch.equal(ya, yb)   # no set_epoch call: still reshuffled
    ld.close()
    plain = U8BatchLoader(ds.data, ds.targets, 100, "cpu", shuffle=False)
    assert torch.equal(torch.cat([y for _, y in plain]), ds.targets)
    plain.close()


def test_cpu_worker_path_uses_plain_modules_and_flat_optimizer():
    from ray_torch_distributed_checkpoint_b200.train.torch import prepare_data_loader, prepare_model
    ds = FashionMNISTSynthetic(train=False, size=64)
    loader = prepare_data_loader(DataLoader(ds, batch_size=16))
    assert isinstance(loader, DeviceLoader)                                            # no GPU worker: generic device loader
    model = prepare_model(NeuralNetwork())
    assert type(model) is NeuralNetwork
    opt = SGD(model.parameters(), lr=0.1, momentum=0.9)
    assert type(opt).__name__ == "SGD"
    loss_fn = CrossEntropyLoss()
    acc = EvalAccumulator(model)
    first = None
    for X, y in loader:
        loss = loss_fn(model(X), y)
        opt.zero_grad(); loss.backward(); opt.step()
        first = first if first is not None else loss.item()
    model.eval()
    with torch.no_grad():
        for X, y in loader:
            pred = model(X)
            acc.update(loss_fn(pred, y), pred, y)
    val_loss, accuracy = acc.compute()
    assert val_loss == val_loss and 0.0 <= accuracy <= 1.0 and acc.samples == 64 and acc.batches == 4
