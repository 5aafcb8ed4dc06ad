"""CPU tier: sampler parity with torch, bucket plan == measured DDP layout, mini Dataset, checkpoint
byte-compatibility, async writer, Philox reference."""
import os
import zipfile

import numpy as np
import pytest
import torch
from torch.utils.data import DataLoader, TensorDataset

from ray_torch_distributed_checkpoint_b200 import ckpt, data
from ray_torch_distributed_checkpoint_b200.models import NeuralNetwork, mlp_flops_per_sample
from ray_torch_distributed_checkpoint_b200.parallel import buckets


@pytest.mark.parametrize("n,world,shuffle,drop_last", [(103, 4, True, False), (100, 8, False, False), (103, 4, True, True), (5, 8, False, False)])
def test_sampler_matches_torch(n, world, shuffle, drop_last):
    ds = TensorDataset(torch.arange(n))
    for rank in range(world):
        ours = data.DistributedSampler(n, world, rank, shuffle=shuffle, seed=0, drop_last=drop_last)
        ref = torch.utils.data.distributed.DistributedSampler(ds, world, rank, shuffle=shuffle, seed=0, drop_last=drop_last)
        for epoch in (0, 3):
            ours.set_epoch(epoch); ref.set_epoch(epoch)
            assert list(ours) == list(ref)
            assert len(ours) == len(ref)


def test_shard_data_loader_keeps_shuffle_flag():
    ds = TensorDataset(torch.arange(64).float())
    a = data.shard_data_loader(DataLoader(ds, batch_size=8, shuffle=True), 2, 0)
    b = data.shard_data_loader(DataLoader(ds, batch_size=8), 2, 1)
    assert a.sampler.shuffle and not b.sampler.shuffle
    assert len(a) == 4 and [int(x) for x in next(iter(b))[0]] == list(range(1, 16, 2))


def test_bucket_plan_matches_measured_ddp_layout():
    model = NeuralNetwork()
    bks, total = buckets.plan(model.named_parameters())
    assert [b.payload_bytes for b in bks] == [1071144, 1607680]
    assert [s.name for s in bks[0].slots] == ["linear_relu_stack.6.bias", "linear_relu_stack.6.weight",
                                              "linear_relu_stack.3.bias", "linear_relu_stack.3.weight"]
    assert [s.name for s in bks[1].slots] == ["linear_relu_stack.0.bias", "linear_relu_stack.0.weight"]
    for b in bks:
        assert b.offset % 32 == 0 and b.numel % 32 == 0
        for s in b.slots:
            assert s.offset % 32 == 0                                  # 128-byte aligned slots
    assert total == sum(b.numel for b in bks)
    assert sum(p.numel() for p in model.parameters()) == 669706
    assert mlp_flops_per_sample() == 3209216 and mlp_flops_per_sample(training=False) == 1337344
    # torch's own bucketing agrees
    import torch.distributed as dist
    params = list(reversed(list(model.parameters())))
    idx, _ = dist._compute_bucket_assignment_by_size(params, [1024 * 1024, 25 * 1024 * 1024])
    assert [len(i) for i in idx] == [len(b.slots) for b in bks]


def test_model_state_dict_keys_match_reference():
    keys = list(NeuralNetwork().state_dict().keys())
    assert keys == [f"linear_relu_stack.{i}.{k}" for i in (0, 3, 6) for k in ("weight", "bias")]


def test_mini_dataset_map_batches():
    items = [{"features": np.full((1, 28, 28), i, dtype=np.float32), "labels": i % 10} for i in range(37)]
    ds = data.from_items(items)
    assert ds.count() == 37 and set(ds.columns()) == {"features", "labels"}
    seen = []

    class Pred:
        def __call__(self, batch):
            seen.append(batch["features"].shape)
            return {"logits": batch["features"].reshape(len(batch["labels"]), -1)[:, :10], "predicted_values": batch["labels"]}

    out = ds.map_batches(Pred(), batch_size=16, concurrency=1, num_gpus=1).take_all()
    assert seen == [(16, 1, 28, 28), (16, 1, 28, 28), (5, 1, 28, 28)]
    assert len(out) == 37 and out[5]["predicted_values"] == 5 and out[5]["logits"].shape == (10,)
    out2 = ds.map_batches(Pred, batch_size=8, concurrency=3).take_all()      # class + workers: order preserved
    assert [r["predicted_values"] for r in out2] == [i % 10 for i in range(37)]
    df = ds.to_pandas()
    assert df.shape == (37, 2) and isinstance(df.features[0], np.ndarray)
    assert data.DataContext.get_current() is data.DataContext.get_current()


def test_synthetic_fashion_mnist_and_idx_writer(tmp_path):
    ds = data.FashionMNISTSynthetic(train=False, size=64)
    x, y = ds[3]
    assert x.shape == (1, 28, 28) and x.dtype == torch.float32 and -1.0 <= float(x.min()) and float(x.max()) <= 1.0
    assert isinstance(y, int) and 0 <= y < 10 and len(data.get_labels_map()) == 10
    raw = data.write_fashion_mnist_idx(str(tmp_path), n_train=32, n_test=16)
    assert os.path.getsize(os.path.join(raw, "train-images-idx3-ubyte")) == 16 + 32 * 784
    import torchvision
    tv = torchvision.datasets.FashionMNIST(root=str(tmp_path), train=True, download=False)
    assert len(tv) == 32 and tv.data.shape == (32, 28, 28)


def test_u8_loader_cpu_path():
    ds = data.FashionMNISTSynthetic(train=True, size=100, raw=True)
    loader = data.U8BatchLoader(ds.data, ds.targets, 32, "cpu", shuffle=False, normalize=True)
    batches = list(loader)
    assert len(batches) == 3 and batches[0][0].shape == (32, 784) and batches[0][0].dtype == torch.float32
    assert torch.allclose(batches[0][0], ds.normalized()[:32].reshape(32, -1))
    assert torch.equal(batches[1][1], ds.targets[32:64])
    raw = list(data.U8BatchLoader(ds.data, ds.targets, 32, "cpu", shuffle=True, world=2, rank=1, seed=3))
    assert len(raw) == 1 and raw[0][0].dtype == torch.uint8 and raw[0][0].shape == (32, 784)
    ref = data.DistributedSampler(100, 2, 1, shuffle=True, seed=3).indices()[:32]
    assert torch.equal(raw[0][1], ds.targets[ref]) and torch.equal(raw[0][0], ds.data.reshape(100, -1)[ref])


def _ref_state():
    torch.manual_seed(0)
    model = NeuralNetwork()
    opt = torch.optim.SGD(model.parameters(), lr=1e-3, momentum=0.9)
    model(torch.randn(4, 1, 28, 28)).sum().backward()
    opt.step()
    return {"epoch": 0, "model_state_dict": model.state_dict(), "optimizer_state_dict": opt.state_dict(),
            "val_losses": [1.5], "val_accuracy": [0.25]}


def test_atomic_save_is_byte_compatible_with_torch_save(tmp_path):
    state = _ref_state()
    ours, ref = tmp_path / "a" / "latest_model.pt", tmp_path / "b" / "latest_model.pt"
    os.makedirs(ref.parent)
    ckpt.atomic_torch_save(state, str(ours))
    torch.save(state, str(ref))
    za, zb = zipfile.ZipFile(ours), zipfile.ZipFile(ref)
    names = lambda z: sorted(n for n in z.namelist() if not n.endswith("serialization_id"))
    assert names(za) == names(zb) and all(n.startswith("latest_model/") for n in za.namelist())
    for n in names(za):
        assert za.read(n) == zb.read(n), n
    assert len([n for n in za.namelist() if "/data/" in n]) == 12          # 6 params + 6 momentum buffers
    assert abs(os.path.getsize(ours) - 5362905) < 4096
    loaded = torch.load(str(ours), weights_only=True)
    assert loaded["optimizer_state_dict"]["param_groups"][0]["momentum"] == 0.9
    assert not [f for f in os.listdir(ours.parent) if f != "latest_model.pt"]


def test_async_checkpointer_snapshot_is_isolated(tmp_path):
    state = _ref_state()
    w = state["model_state_dict"]["linear_relu_stack.0.weight"]
    expect = w.clone()
    cp = ckpt.AsyncCheckpointer("cpu")
    pending = cp.save(state, str(tmp_path / "ck"), ["latest_model.pt", "best_model.pt"])
    w.add_(1.0)                                    # training continues and overwrites the live tensor
    c = pending.result()
    assert sorted(os.listdir(c.path)) == ["best_model.pt", "latest_model.pt"]
    loaded = torch.load(os.path.join(c.path, "best_model.pt"), weights_only=True)
    assert torch.equal(loaded["model_state_dict"]["linear_relu_stack.0.weight"], expect)
    assert loaded["val_losses"] == [1.5]
    p2 = cp.save(state, str(tmp_path / "ck2"), "latest_model.pt")
    p3 = cp.save(state, str(tmp_path / "ck3"), "latest_model.pt")   # third save recycles slot 0
    ckpt.wait_all()
    assert p2.done() and p3.done()
    cp.close()


def test_philox_reference_statistics_and_determinism():
    from ray_torch_distributed_checkpoint_b200.ops.philox_ref import dropout_keep_mask, philox4x32_10
    # known-answer test of Philox4x32-10 (Random123 test vector: counter 0, key 0)
    out = philox4x32_10(0, 0, 0, 0, 0, 0)
    assert [int(v) for v in out] == [0x6627E8D5, 0xE169C58D, 0xBC57AC4C, 0x9B00DBD8]
    m = dropout_keep_mask(64, 512, 0.25, seed=7, step=1, salt=2)
    assert abs(m.mean() - 0.75) < 0.01
    assert np.array_equal(m, dropout_keep_mask(64, 512, 0.25, seed=7, step=1, salt=2))
    assert not np.array_equal(m, dropout_keep_mask(64, 512, 0.25, seed=7, step=2, salt=2))


def test_parallel_archive_writer_matches_stock_torch_save(tmp_path):
    """ckpt/fast_zip.py: same entries, order and contents as torch.save of the same object; valid CRCs; 64-byte aligned
    payloads; loadable with weights_only=True (also through mmap and with zip64 fields forced)."""
    import struct
    from ray_torch_distributed_checkpoint_b200.ckpt import fast_zip
    state = _ref_state()
    state["extra"] = {"empty": torch.empty(0), "half": torch.randn(33, 7).to(torch.bfloat16), "ints": torch.arange(10)}
    ours, ref = tmp_path / "a" / "latest_model.pt", tmp_path / "b" / "latest_model.pt"
    os.makedirs(ours.parent), os.makedirs(ref.parent)
    n = fast_zip.save(state, str(ours), threads=3, chunk_bytes=1 << 16)
    torch.save(state, str(ref))
    assert n == os.path.getsize(ours)
    za, zb = zipfile.ZipFile(ours), zipfile.ZipFile(ref)
    assert za.testzip() is None
    assert [i.filename for i in za.infolist()] == [i.filename for i in zb.infolist()]
    for i in za.infolist():
        if not i.filename.endswith("serialization_id"):
            assert za.read(i.filename) == zb.read(i.filename), i.filename
    raw = ours.read_bytes()
    for i in za.infolist():
        fnl, exl = struct.unpack("<HH", raw[i.header_offset + 26:i.header_offset + 30])
        assert (i.header_offset + 30 + fnl + exl) % 64 == 0, i.filename
    for kw in ({}, {"mmap": True}):
        got = torch.load(str(ours), weights_only=True, **kw)
        assert torch.equal(got["model_state_dict"]["linear_relu_stack.0.weight"], state["model_state_dict"]["linear_relu_stack.0.weight"])
        assert torch.equal(got["extra"]["half"], state["extra"]["half"]) and got["extra"]["empty"].numel() == 0
        assert got["optimizer_state_dict"]["param_groups"][0]["momentum"] == 0.9
    z64 = tmp_path / "z64.pt"
    fast_zip.save(state, str(z64), force_zip64=True)
    assert zipfile.ZipFile(z64).testzip() is None
    assert torch.equal(torch.load(str(z64), weights_only=True)["extra"]["ints"], state["extra"]["ints"])


def test_big_states_take_the_parallel_writer_and_small_ones_torch_save(tmp_path, monkeypatch):
    from ray_torch_distributed_checkpoint_b200.ckpt import async_writer, fast_zip
    calls = []
    real = fast_zip.save
    monkeypatch.setattr(fast_zip, "save", lambda *a, **k: (calls.append(a[1]), real(*a, **k))[1])
    state = _ref_state()
    ckpt.atomic_torch_save(state, str(tmp_path / "small" / "latest_model.pt"))
    assert not calls                                                        # 5 MB: stock torch.save
    monkeypatch.setattr(async_writer, "FAST_WRITER_MIN_BYTES", 1 << 20)
    cp = ckpt.AsyncCheckpointer("cpu")
    pending = cp.save(state, str(tmp_path / "big"), ["latest_model.pt", "best_model.pt"])
    w = state["model_state_dict"]["linear_relu_stack.0.weight"]
    expect = w.clone()
    w.add_(1.0)                                                             # the snapshot is isolated from later updates
    pending.result()
    assert len(calls) == 2 and all(".latest_model.pt." in c or ".best_model.pt." in c for c in calls)
    for name in ("latest_model.pt", "best_model.pt"):
        got = torch.load(str(tmp_path / "big" / name), weights_only=True)
        assert torch.equal(got["model_state_dict"]["linear_relu_stack.0.weight"], expect)
        assert zipfile.ZipFile(tmp_path / "big" / name).namelist()[0] == name[:-3] + "/data.pkl"
    monkeypatch.setenv("RTDC_CKPT_WRITER", "torch")
    calls.clear()
    ckpt.atomic_torch_save(state, str(tmp_path / "forced" / "latest_model.pt"))
    assert not calls
    cp.close()


def test_parallel_writer_failure_falls_back_to_torch_save(tmp_path, monkeypatch):
    from ray_torch_distributed_checkpoint_b200.ckpt import fast_zip
    monkeypatch.setattr(fast_zip, "save", lambda *a, **k: (_ for _ in ()).throw(RuntimeError("moved")))
    state = _ref_state()
    with pytest.warns(UserWarning, match="parallel checkpoint writer unavailable"):
        ckpt.atomic_torch_save(state, str(tmp_path / "latest_model.pt"), parallel=True)
    assert torch.load(str(tmp_path / "latest_model.pt"), weights_only=True)["epoch"] == 0
