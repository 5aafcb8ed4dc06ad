"""Multi-GPU checks, one rank per GPU (launched by tests/test_multigpu.py through torch.distributed.run)."""
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def check_collectives(rank, world, dev):
    from ray_torch_distributed_checkpoint_b200.parallel.symm import SymmetricArena
    arena = SymmetricArena(64 << 20, device=dev)
    algos = ["two_shot"] + (["nvls"] if arena.has_multicast else [])
    if rank == 0:
        print(f"[mgpu] world={world} multicast={arena.has_multicast} algos={algos}", flush=True)
    sizes = [4, 1024, 267786, 401920, 1 << 20, (4 << 20) + 12]        # incl. the MLP's two bucket sizes (elements)
    for algo in algos:
        for n in sizes:
            n4 = (n + 3) // 4 * 4
            off, t = 4096, arena.tensor(4096, n4)
            g = torch.Generator(device=dev).manual_seed(1000 * rank + n)
            src = torch.randn(n4, device=dev, generator=g)
            t.copy_(src)
            ref = src.clone()
            dist.all_reduce(ref)
            ref /= world
            torch.cuda.synchronize()
            arena.all_reduce(off, n4, scale=1.0 / world, algo=algo)
            torch.cuda.synchronize()
            assert torch.allclose(t, ref, rtol=1e-5, atol=1e-6), (algo, n, float((t - ref).abs().max()))
            gathered = [torch.empty_like(t) for _ in range(world)]
            dist.all_gather(gathered, t.contiguous())
            assert all(torch.equal(gathered[0], x) for x in gathered), f"{algo} n={n}: replicas differ"
    # one-shot, out of place
    n = 65536
    t = arena.tensor(4096, n)
    src = torch.randn(n, device=dev) + rank
    t.copy_(src)
    ref = src.clone(); dist.all_reduce(ref)
    out = torch.empty(n, device=dev)
    torch.cuda.synchronize()
    arena.all_reduce(4096, n, algo="one_shot", out=out)
    torch.cuda.synchronize()
    assert torch.allclose(out, ref, rtol=1e-5, atol=1e-5)
    assert torch.equal(t, src)                                    # input untouched
    # bf16 payload, fp32 accumulate
    nb = 1 << 18
    tb = arena.tensor(4096, nb, torch.bfloat16)
    srcb = (torch.randn(nb, device=dev) * 0.1).to(torch.bfloat16)
    tb.copy_(srcb)
    refb = srcb.float(); dist.all_reduce(refb); refb /= world
    torch.cuda.synchronize()
    for algo in algos:
        tb.copy_(srcb); torch.cuda.synchronize(); dist.barrier()
        arena.all_reduce(4096, nb, dtype=torch.bfloat16, scale=1.0 / world, algo=algo)
        torch.cuda.synchronize()
        assert torch.allclose(tb.float(), refb, rtol=2e-2, atol=2e-3), algo
    # flag-reuse stress: many back-to-back launches with skewed ranks
    t = arena.tensor(4096, 401920)
    for it in range(300):
        t.fill_(float(rank + it % 7))
        if (it + rank) % 5 == 0:
            torch.cuda._sleep(200000)
        arena.all_reduce(4096, 401920, scale=1.0 / world, algo=algos[it % len(algos)])
        if it % 50 == 49:
            torch.cuda.synchronize()
            expect = sum(float(r + it % 7) for r in range(world)) / world
            assert torch.allclose(t, torch.full_like(t, expect)), (it, float(t[0]), expect)
    # broadcast + barrier
    stage = arena.tensor(4096, 1000)
    stage.fill_(float(rank + 1))
    dst = torch.zeros(1000, device=dev)
    torch.cuda.synchronize(); dist.barrier()
    arena.broadcast(4096, 4000, world - 1, dst)
    arena.barrier()
    torch.cuda.synchronize()
    assert torch.equal(dst, torch.full_like(dst, float(world)))
    dist.barrier()
    arena.close()
    if rank == 0:
        print("[mgpu] collectives ok", flush=True)


def check_ddp(rank, world, dev):
    from ray_torch_distributed_checkpoint_b200.models import NeuralNetwork
    from ray_torch_distributed_checkpoint_b200.parallel.ddp import DistributedDataParallel
    torch.manual_seed(7 + rank)                                  # different init per rank: wrap must broadcast rank 0's
    model = NeuralNetwork().to(dev)
    for m in model.modules():
        if isinstance(m, torch.nn.Dropout):
            m.p = 0.0
    ours = DistributedDataParallel(model, device=dev)
    ref_model = NeuralNetwork().to(dev)
    for m in ref_model.modules():
        if isinstance(m, torch.nn.Dropout):
            m.p = 0.0
    ref_model.load_state_dict({k[len("module."):]: v.clone() for k, v in ours.state_dict().items()})
    ref = torch.nn.parallel.DistributedDataParallel(ref_model, device_ids=[dev.index])
    assert list(ours.state_dict()) == list(ref.state_dict())     # 'module.' prefix
    w0 = [torch.empty_like(ours.param_arena) for _ in range(world)]
    dist.all_gather(w0, ours.param_arena)
    assert all(torch.equal(w0[0], w) for w in w0)
    opt_o = torch.optim.SGD(ours.parameters(), lr=0.1, momentum=0.9)
    opt_r = torch.optim.SGD(ref.parameters(), lr=0.1, momentum=0.9)
    for step in range(4):
        g = torch.Generator(device=dev).manual_seed(100 * step + rank)
        x = torch.randn(64, 1, 28, 28, device=dev, generator=g)
        y = torch.randint(0, 10, (64,), device=dev, generator=g)
        for net, opt in ((ours, opt_o), (ref, opt_r)):
            opt.zero_grad(set_to_none=(step % 2 == 0))           # both zero_grad flavours must work
            torch.nn.functional.cross_entropy(net(x), y).backward()
            opt.step()
        for (n1, p1), (n2, p2) in zip(ours.named_parameters(), ref.named_parameters()):
            assert torch.allclose(p1, p2, rtol=1e-4, atol=1e-6), (step, n1, float((p1 - p2).abs().max()))
    wN = [torch.empty_like(ours.param_arena) for _ in range(world)]
    dist.all_gather(wN, ours.param_arena)
    assert all(torch.equal(wN[0], w) for w in wN), "replicas diverged"
    if rank == 0:
        print("[mgpu] ddp ok", flush=True)


def check_engine(rank, world, dev):
    from ray_torch_distributed_checkpoint_b200.models import NeuralNetwork
    from ray_torch_distributed_checkpoint_b200.parallel.engine import MLPTrainEngine
    B = 512
    finals, moms = {}, {}
    from ray_torch_distributed_checkpoint_b200.parallel.symm import SymmetricArena
    probe = SymmetricArena(1 << 20, device=dev)
    has_mc = probe.has_multicast
    probe.close()
    modes = [("fused", "auto"), ("fused", "two_shot")] + ([("fused", "nvls")] if has_mc else []) + [("merged", "auto"), ("bucketed", "auto")]
    for mode, algo in modes:
        torch.manual_seed(3 + rank)
        eng = MLPTrainEngine(NeuralNetwork(), batch_size=B, lr=0.05, device=dev, seed=11, ar_mode=mode, algo=algo)
        losses = []
        for step in range(6):
            g = torch.Generator(device=dev).manual_seed(50 * step + rank)
            x = torch.randint(0, 256, (B, 784), dtype=torch.uint8, device=dev, generator=g)
            y = torch.randint(0, 10, (B,), device=dev, generator=g)
            losses.append(eng.step(x, y).item())
            if mode == "fused" and step == 3 and rank == 0:
                # one rank alone (a checkpoint on rank 0) completes its momentum out of the owners' slices; no collective
                m3 = [v["momentum_buffer"].clone() for v in eng.optimizer_state_dict()["state"].values()]
                assert all(float(t.abs().sum()) > 0 for t in m3)
        ws = [torch.empty_like(eng.p_arena) for _ in range(world)]
        dist.all_gather(ws, eng.p_arena)
        assert all(torch.equal(ws[0], w) for w in ws), f"engine replicas diverged ({mode}/{algo})"
        assert list(eng.model_state_dict())[0].startswith("module.")
        assert losses[-1] == losses[-1]
        assert eng.core.recent_losses(6) == losses                # every step's loss reached the pinned host ring
        finals[(mode, algo)] = eng.p_arena.clone()
        torch.cuda.synchronize()
        dist.barrier()
        moms[(mode, algo)] = torch.cat([v["momentum_buffer"].reshape(-1) for v in eng.optimizer_state_dict()["state"].values()]).clone()
        torch.cuda.synchronize()
        dist.barrier()
        eng.close()
    # reduce-scatter + sharded SGD + parameter multicast, one all-reduce of the whole arena, one all-reduce per bucket: the
    # same update (the in-switch NVLS reduction need not add in the same order as the P2P kernels, so compare to rounding)
    ref = finals[("merged", "auto")]
    for k, v in finals.items():
        assert torch.allclose(v, ref, rtol=1e-5, atol=1e-7), f"{k} disagrees with the merged all-reduce"
        assert torch.allclose(moms[k], moms[("merged", "auto")], rtol=1e-4, atol=1e-7), f"{k}: gathered momentum disagrees"
    if rank == 0:
        print(f"[mgpu] engine ok (losses {losses[0]:.4f} -> {losses[-1]:.4f}; modes {[m + '/' + a for m, a in modes]})", flush=True)


def check_trainer(rank, world, dev):
    """The reference-shaped program under this launch (SPMD mode of TorchTrainer.fit): engine-backed model, sharded uint8
    loader, per-epoch checkpoint from rank 0, retention, warm start."""
    import importlib
    import tempfile
    os.environ.update(RTDC_TRAIN_SIZE="16384", RTDC_TEST_SIZE="4096", RTDC_SYNTHETIC_DATA="1")
    sys.path.insert(0, os.path.join(ROOT, "examples"))
    import my_module
    importlib.reload(my_module)
    box = [tempfile.mkdtemp(prefix="rtdc_mgpu_trainer_") if rank == 0 else None]
    dist.broadcast_object_list(box, src=0)
    result = my_module.train_fashion_mnist(num_workers=world, use_gpu=True, global_batch_size=256 * world, learning_rate=0.05,
                                           epochs=3, checkpoint_storage_path=box[0])
    assert result.error is None and result.metrics["training_iteration"] == 3
    hist = result.metrics_dataframe
    assert hist["val_loss"].iloc[-1] < hist["val_loss"].iloc[0]
    if rank == 0:
        kept = sorted(d for d in os.listdir(result.path) if d.startswith("checkpoint_"))
        assert kept == ["checkpoint_000001", "checkpoint_000002"], kept
        with result.checkpoint.as_directory() as d:
            sd = torch.load(os.path.join(d, "latest_model.pt"), weights_only=True)
        assert list(sd["model_state_dict"])[0] == "module.linear_relu_stack.0.weight"
        assert all(float(v["momentum_buffer"].abs().sum()) > 0 for v in sd["optimizer_state_dict"]["state"].values())
    dist.barrier()
    again = my_module.train_fashion_mnist(num_workers=world, use_gpu=True, global_batch_size=256 * world, learning_rate=0.05,
                                          epochs=1, checkpoint_storage_path=box[0], checkpoint=result.checkpoint)
    assert again.metrics["val_loss"] < hist["val_loss"].iloc[0]
    if rank == 0:
        pred = my_module.TorchPredictor(result.checkpoint)
        ds = my_module.get_dataloaders(512, val_only=True, as_ray_ds=True)
        out = ds.map_batches(pred, batch_size=512, concurrency=1, num_gpus=1).take_all()
        acc = sum(int(r["predicted_values"] == l) for r, l in zip(out, ds._cols["labels"])) / len(out)
        assert abs(acc - result.metrics["accuracy"]) < 0.03, (acc, result.metrics["accuracy"])
        print(f"[mgpu] trainer ok (val_loss {hist['val_loss'].iloc[0]:.4f} -> {hist['val_loss'].iloc[-1]:.4f}, accuracy {acc:.4f})", flush=True)
    dist.barrier()


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    which = sys.argv[1:] or ["collectives", "ddp", "engine", "trainer"]
    if "collectives" in which:
        check_collectives(rank, world, dev)
    if "ddp" in which:
        check_ddp(rank, world, dev)
    if "engine" in which:
        check_engine(rank, world, dev)
    if "trainer" in which:
        check_trainer(rank, world, dev)
    dist.barrier()
    dist.destroy_process_group()
    if rank == 0:
        print("[mgpu] all ok", flush=True)


if __name__ == "__main__":
    main()
