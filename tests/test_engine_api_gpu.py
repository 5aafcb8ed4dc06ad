"""The reference's literal loop (model(X) / loss_fn / zero_grad / backward / optimizer.step) on the engine-backed model
that train.torch.prepare_model returns for the reference network: same trajectory as the engine driven by hand and as a
plain PyTorch fp64 restatement; lazy loss / logits semantics; tail batches; evaluation accumulators; U8Images."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


@pytest.fixture(autouse=True)
def gpu_worker_session():
    """What TorchTrainer.fit() installs in a GPU worker: train.torch.get_device() answers cuda:<local_rank>."""
    from ray_torch_distributed_checkpoint_b200.train import session
    session._set_session(session._Session(session.TrainContext(use_gpu=True), None, None))
    yield
    session._set_session(None)


def _prepared(seed=0, **kw):
    from ray_torch_distributed_checkpoint_b200.models import NeuralNetwork
    from ray_torch_distributed_checkpoint_b200.train.torch import prepare_model
    torch.manual_seed(seed)
    return prepare_model(NeuralNetwork(), engine_kwargs=dict(seed=77, **kw))


def test_literal_loop_equals_engine_by_hand():
    from ray_torch_distributed_checkpoint_b200.models import NeuralNetwork
    from ray_torch_distributed_checkpoint_b200.nn import CrossEntropyLoss
    from ray_torch_distributed_checkpoint_b200.optim import SGD
    from ray_torch_distributed_checkpoint_b200.parallel.engine import MLPTrainEngine
    from ray_torch_distributed_checkpoint_b200.parallel.engine_api import EngineModel, EngineSGD
    model = _prepared()
    assert isinstance(model, EngineModel)
    assert list(model.state_dict())[0] == "linear_relu_stack.0.weight"          # one worker: no 'module.' prefix
    torch.manual_seed(0)
    eng = MLPTrainEngine(NeuralNetwork(), batch_size=320, lr=0.05, momentum=0.9, device="cuda", seed=77, distributed=False)
    assert torch.equal(eng.p_arena, model.core.p_arena)
    loss_fn = CrossEntropyLoss()
    opt = SGD(model.parameters(), lr=0.05, momentum=0.9)                        # parameters() resolves to the engine optimizer
    assert isinstance(opt, EngineSGD)
    g = torch.Generator().manual_seed(5)
    got, want = [], []
    for step in range(6):
        x = torch.randint(0, 256, (320, 1, 28, 28), dtype=torch.uint8, generator=g).cuda()
        y = torch.randint(0, 10, (320,), generator=g).cuda()
        pred = model(x)
        loss = loss_fn(pred, y)
        if step == 2:
            peek = loss.item()                                                  # before step(): one extra forward, same dropout
        opt.zero_grad()
        loss.backward()
        opt.step()
        got.append(loss.item())                                                 # after step(): the loss the step computed
        want.append(eng.step(x.reshape(320, 784), y).item())
        if step == 2:
            assert abs(peek - got[-1]) < 1e-5
    assert got == want
    assert torch.equal(eng.p_arena, model.core.p_arena) and torch.equal(eng.m_arena, model.core.m_arena)
    assert model.core.recent_losses(3) == got[-3:]
    # the wrapped module's parameters are views of the arena
    assert model.module.linear_relu_stack[0].weight.data_ptr() == model.core.W[0].data_ptr()
    sd = opt.state_dict()
    assert sd["param_groups"][0]["momentum"] == 0.9 and set(sd["state"]) == set(range(6))
    # what the fused step cannot express fails loudly
    p1 = model(x)
    loss_fn(p1, y)
    with pytest.raises(NotImplementedError):
        model(x)
    eng.close()
    model.close()


def test_tail_batches_eval_and_materialisation():
    from ray_torch_distributed_checkpoint_b200.models import NeuralNetwork
    from ray_torch_distributed_checkpoint_b200.nn import CrossEntropyLoss, EvalAccumulator
    from ray_torch_distributed_checkpoint_b200.optim import SGD
    model = _prepared(seed=3)
    ref = NeuralNetwork().cuda()
    loss_fn = CrossEntropyLoss()
    opt = SGD(model, lr=0.05, momentum=0.9)
    g = torch.Generator().manual_seed(9)
    for n in (256, 256, 100, 256, 100):                                         # full batches and the epoch's tail batch
        x = torch.randn(n, 1, 28, 28, generator=g).cuda()
        y = torch.randint(0, 10, (n,), generator=g).cuda()
        loss = loss_fn(model(x), y)
        opt.zero_grad(); loss.backward(); opt.step()
        assert loss.item() == loss.item()
    assert sorted(model.core.plans) == [100, 256]
    ref.load_state_dict(model.state_dict())
    ref.eval()
    model.eval()
    acc = EvalAccumulator(model)
    want_loss, want_correct, total = 0.0, 0, 0
    with torch.no_grad():
        for n in (200, 200, 56):
            x = torch.randn(n, 1, 28, 28, generator=g).cuda()
            y = torch.randint(0, 10, (n,), generator=g).cuda()
            pred = model(x)
            loss = loss_fn(pred, y)
            acc.update(loss, pred, y)
            z = ref(x)
            assert abs(loss.item() - F.cross_entropy(z, y).item()) < 1e-4
            assert torch.allclose(pred.materialize(), z, rtol=1e-4, atol=1e-5)   # logits of the eval graph, on demand
            assert int((pred.argmax(1) == y).sum()) == int((z.argmax(1) == y).sum())
            want_loss += F.cross_entropy(z, y).item(); want_correct += int((z.argmax(1) == y).sum()); total += n
        # inference only: no loss asked for
        x = torch.randn(64, 1, 28, 28, generator=g).cuda()
        assert torch.allclose(torch.as_tensor(model(x).materialize()), ref(x), rtol=1e-4, atol=1e-5)
    val_loss, accuracy = acc.compute()
    assert abs(val_loss - want_loss / 3) < 1e-4 and accuracy == want_correct / total
    model.train()
    model.eval()                                                                # a new validation pass starts from zero
    assert model.read_eval() == (0.0, 0)
    model.close()


def test_u8_images_and_generic_module_path():
    from ray_torch_distributed_checkpoint_b200.data import U8Images
    from ray_torch_distributed_checkpoint_b200.models import NeuralNetwork
    from ray_torch_distributed_checkpoint_b200.nn import CrossEntropyLoss
    from ray_torch_distributed_checkpoint_b200.optim import SGD
    from ray_torch_distributed_checkpoint_b200.train.torch import prepare_model
    raw = torch.randint(0, 256, (128, 1, 28, 28), dtype=torch.uint8, device="cuda")
    xb = U8Images.wrap(raw, 0.5, 0.5)
    want = (raw.float() / 255 - 0.5) / 0.5
    assert xb.dtype == torch.uint8 and xb.shape == raw.shape and xb.raw.data_ptr() == raw.data_ptr()
    assert torch.allclose(xb.float(), want) and torch.allclose(torch.flatten(xb, 1), want.flatten(1))
    # a generic model (autograd path over the fused slots) consumes the same batch as normalised float32
    torch.manual_seed(1)
    net = prepare_model(NeuralNetwork(), fuse="modules")
    ref = NeuralNetwork().cuda()
    ref.load_state_dict(net.state_dict())
    net.eval(); ref.eval()
    assert torch.allclose(net(xb), ref(want), rtol=1e-4, atol=1e-5)
    # the engine-backed model takes the bytes; same numbers as the float path
    torch.manual_seed(1)
    eng_model = prepare_model(NeuralNetwork())
    eng_model.eval()
    with torch.no_grad():
        assert torch.allclose(eng_model(xb).materialize(), ref(want), rtol=1e-4, atol=1e-5)
    # training through the module path: chain-linked backward (bit masks + fused column sums) against stock autograd
    torch.manual_seed(2)
    a = prepare_model(NeuralNetwork(p_drop=0.0), fuse="modules")
    b = NeuralNetwork(p_drop=0.0).cuda()
    b.load_state_dict(a.state_dict())
    oa, ob = SGD(a, lr=0.05, momentum=0.9), torch.optim.SGD(b.parameters(), lr=0.05, momentum=0.9)
    la = CrossEntropyLoss()
    for step in range(4):
        x = torch.randn(200, 1, 28, 28, device="cuda")
        y = torch.randint(0, 10, (200,), device="cuda")
        oa.zero_grad(); l1 = la(a(x), y); l1.backward(); oa.step()
        ob.zero_grad(); l2 = F.cross_entropy(b(x), y); l2.backward(); ob.step()
        assert abs(l1.item() - l2.item()) < 1e-4, step
    for (n, p), (_, q) in zip(a.state_dict().items(), b.state_dict().items()):
        assert (p - q).norm() / q.norm() < 2e-4, n
    eng_model.close()


def test_prepare_data_loader_u8_fast_path_matches_reference_pipeline():
    from torch.utils.data import DataLoader
    from ray_torch_distributed_checkpoint_b200.data import FashionMNISTSynthetic, U8BatchLoader, U8Images
    from ray_torch_distributed_checkpoint_b200.train.torch import prepare_data_loader
    ds = FashionMNISTSynthetic(train=False, size=1000)
    loader = prepare_data_loader(DataLoader(ds, batch_size=384))            # sequential sampler, tail batch kept
    assert isinstance(loader, U8BatchLoader) and len(loader) == 3
    seen = 0
    for xb, yb in loader:
        assert isinstance(xb, U8Images) and yb.dtype == torch.int64 and xb.is_cuda
        n = xb.shape[0]
        want_x = torch.stack([ds[i][0] for i in range(seen, seen + n)]).cuda()
        assert torch.allclose(xb.float(), want_x, atol=1e-6)
        assert torch.equal(yb.cpu(), ds.targets[seen:seen + n])
        seen += n
    assert seen == 1000
    shuffled = prepare_data_loader(DataLoader(ds, batch_size=250, shuffle=True))
    labels = torch.cat([yb.cpu() for _, yb in shuffled])
    assert sorted(labels.tolist()) == sorted(ds.targets.tolist()) and not torch.equal(labels, ds.targets)


def test_loader_transfer_modes_agree_with_their_samplers():
    from ray_torch_distributed_checkpoint_b200.data import DistributedSampler, FashionMNISTSynthetic, U8BatchLoader
    ds = FashionMNISTSynthetic(train=False, size=1000)
    flat = ds.data.reshape(1000, -1)
    key = lambda rows: {(bytes(r[:16].tolist()), int(r.to(torch.int64).sum())) for r in rows}
    everything = key(flat)
    for world in (1, 2):
        seen = {"resident": set(), "stream": set(), "host_gather": set()}
        for rank in range(world):
            want = DistributedSampler(1000, world, rank, shuffle=True, seed=3)
            want.set_epoch(1)
            idx = torch.tensor(want.indices())
            for mode in seen:
                ld = U8BatchLoader(ds.data, ds.targets, 192, "cuda", shuffle=True, seed=3, world=world, rank=rank, drop_last=False,
                                   transfer=mode, shuffle_block=64)
                assert ld.transfer == mode
                ld.sampler.set_epoch(1)
                xs, ys = zip(*[(x.cpu(), y.cpu()) for x, y in ld])
                assert [x.shape[0] for x in xs[:-1]] == [192] * (len(xs) - 1) and len(xs) == len(ld)
                x, y = torch.cat(xs), torch.cat(ys)
                if mode != "stream":                                   # exact DistributedSampler order
                    assert torch.equal(x, flat[idx]) and torch.equal(y, ds.targets[idx])
                else:                                                  # block shuffle: rows are real rows with their own labels
                    lut = {k: int(l) for k, l in zip([(bytes(r[:16].tolist()), int(r.to(torch.int64).sum())) for r in flat], ds.targets)}
                    assert all(lut[(bytes(r[:16].tolist()), int(r.to(torch.int64).sum()))] == int(l) for r, l in zip(x, y))
                    assert x.shape[0] == len(ld.sampler)
                seen[mode] |= key(x)
                ld.close()
        for mode, got in seen.items():
            assert got == everything, (mode, world, len(got))          # every sample is visited (padding may repeat a few)
    # no shuffle: storage order
    ld = U8BatchLoader(ds.data, ds.targets, 400, "cuda", shuffle=False, drop_last=False, transfer="stream", shuffle_block=128)
    assert torch.equal(torch.cat([y.cpu() for _, y in ld])[:1000 - 1000 % 128], ds.targets[:1000 - 1000 % 128])


def test_async_snapshot_of_arena_views_is_exact():
    from ray_torch_distributed_checkpoint_b200.ckpt import AsyncCheckpointer
    from ray_torch_distributed_checkpoint_b200.nn import CrossEntropyLoss
    from ray_torch_distributed_checkpoint_b200.optim import SGD
    model = _prepared(seed=5)
    opt = SGD(model, lr=0.05, momentum=0.9)
    x = torch.randn(128, 1, 28, 28, device="cuda"); y = torch.randint(0, 10, (128,), device="cuda")
    loss = CrossEntropyLoss()(model(x), y); loss.backward(); opt.step()
    state = {"model_state_dict": model.state_dict(), "optimizer_state_dict": opt.state_dict(), "extra": torch.arange(5, device="cuda"), "epoch": 3}
    want = {k: v.clone() for k, v in state["model_state_dict"].items()}
    want_m = [v["momentum_buffer"].clone() for v in state["optimizer_state_dict"]["state"].values()]
    ck = AsyncCheckpointer(torch.device("cuda"))
    cpu_state, _slot = ck.snapshot(state)
    loss = CrossEntropyLoss()(model(x), y); loss.backward(); opt.step()          # the live arenas move on; the snapshot must not
    got = cpu_state()
    assert got["epoch"] == 3 and torch.equal(got["extra"], torch.arange(5))
    for k, v in want.items():
        assert not got["model_state_dict"][k].is_cuda and torch.equal(got["model_state_dict"][k], v.cpu()), k
    for i, v in enumerate(want_m):
        assert torch.equal(got["optimizer_state_dict"]["state"][i]["momentum_buffer"], v.cpu())
    assert ck.stats["snapshots"] == 1
    ck.close()
    model.close()
