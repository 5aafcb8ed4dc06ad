"""GPU tier for the module-level API: fused Linear chains under autograd, fused loss and optimizer
drop-ins, and the reference-shaped training program through TorchTrainer on a GPU worker."""
import os
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "examples"))


def test_fused_chain_matches_stock_modules_under_autograd():
    from ray_torch_distributed_checkpoint_b200.models import NeuralNetwork, fuse_linear_chains
    from ray_torch_distributed_checkpoint_b200.nn import CrossEntropyLoss
    from ray_torch_distributed_checkpoint_b200.optim import SGD
    torch.manual_seed(0)
    ref = NeuralNetwork(p_drop=0.0).cuda()
    ours = NeuralNetwork(p_drop=0.0).cuda()
    ours.load_state_dict(ref.state_dict())
    fuse_linear_chains(ours)
    assert list(ours.state_dict()) == list(ref.state_dict())           # keys survive the rewrite
    opt_r = torch.optim.SGD(ref.parameters(), lr=0.05, momentum=0.9)
    opt_o = SGD(ours, lr=0.05, momentum=0.9)
    loss_o = CrossEntropyLoss()
    for step in range(5):
        x = torch.randn(200, 1, 28, 28, device="cuda")
        y = torch.randint(0, 10, (200,), device="cuda")
        opt_r.zero_grad(); lr = torch.nn.functional.cross_entropy(ref(x), y); lr.backward(); opt_r.step()
        opt_o.zero_grad(); lo = loss_o(ours(x), y); lo.backward(); opt_o.step()
        assert abs(lr.item() - lo.item()) < 1e-4, step
    for (n, a), (_, b) in zip(ours.state_dict().items(), ref.state_dict().items()):
        assert (a - b).norm() / b.norm() < 2e-4, n
    sd = opt_o.state_dict()
    ref_sd = opt_r.state_dict()
    assert sd["param_groups"][0]["params"] == ref_sd["param_groups"][0]["params"]
    assert set(sd["state"]) == set(ref_sd["state"])
    # dropout active in train mode, off in eval mode
    net = fuse_linear_chains(NeuralNetwork().cuda())
    x = torch.randn(64, 1, 28, 28, device="cuda")
    net.train(); a = net(x); b = net(x)
    net.eval(); c = net(x); d = net(x)
    assert not torch.equal(a, b) and torch.equal(c, d)


def _gpu_loop(config):
    import my_module
    my_module.train_func_per_worker(config)


def test_trainer_single_gpu_worker_end_to_end(tmp_path, monkeypatch):
    monkeypatch.setenv("RTDC_TRAIN_SIZE", "4096")
    monkeypatch.setenv("RTDC_TEST_SIZE", "1024")
    monkeypatch.setenv("RTDC_SYNTHETIC_DATA", "1")
    import importlib
    import my_module
    importlib.reload(my_module)
    result = my_module.train_fashion_mnist(num_workers=1, use_gpu=True, global_batch_size=256, learning_rate=0.05, epochs=3,
                                           checkpoint_storage_path=str(tmp_path))
    assert result.error is None and result.metrics["training_iteration"] == 3
    hist = result.metrics_dataframe
    assert hist["val_loss"].iloc[-1] < hist["val_loss"].iloc[0]          # it learns
    assert result.metrics["accuracy"] > 0.5
    with result.checkpoint.as_directory() as d:
        sd = torch.load(os.path.join(d, "latest_model.pt"), weights_only=True)
    assert list(sd["model_state_dict"])[0] == "linear_relu_stack.0.weight"      # world_size 1: no 'module.' prefix
    pred = my_module.TorchPredictor(result.checkpoint)
    ds = my_module.get_dataloaders(512, val_only=True, as_ray_ds=True)
    out = ds.map_batches(pred, batch_size=512, concurrency=1, num_gpus=1).take_all()
    acc = sum(int(r["predicted_values"] == l) for r, l in zip(out, ds._cols["labels"])) / len(out)
    assert abs(acc - result.metrics["accuracy"]) < 0.02
