"""Whole-step engine against a plain PyTorch (fp64) restatement of the same MLP / loss / SGD, driven with
the same dropout masks (replayed from the Philox reference)."""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def _ref_step(params, mom, x, y, step, p, seed, lr, mu, first):
    from ray_torch_distributed_checkpoint_b200.ops.philox_ref import dropout_keep_mask
    W1, b1, W2, b2, W3, b3 = [t.clone().requires_grad_(True) for t in params]
    B = x.shape[0]
    k1 = torch.from_numpy(dropout_keep_mask(B, W1.shape[0], p, seed, step=step, salt=1)).to(x.device) if p > 0 else 1.0
    k2 = torch.from_numpy(dropout_keep_mask(B, W2.shape[0], p, seed, step=step, salt=2)).to(x.device) if p > 0 else 1.0
    h1 = torch.relu(x @ W1.T + b1) * k1 / (1 - p)
    h2 = torch.relu(h1 @ W2.T + b2) * k2 / (1 - p)
    z = torch.relu(h2 @ W3.T + b3)
    loss = F.cross_entropy(z, y)
    loss.backward()
    new_p, new_m = [], []
    for t, m in zip((W1, b1, W2, b2, W3, b3), mom):
        g = t.grad
        m2 = g.clone() if first else mu * m + g
        new_m.append(m2)
        new_p.append((t - lr * m2).detach())
    return new_p, new_m, loss.detach()


@pytest.mark.parametrize("use_graph,fused_head", [(False, True), (True, True), (True, False)])
def test_engine_matches_reference_training(use_graph, fused_head):
    from ray_torch_distributed_checkpoint_b200.models import NeuralNetwork
    from ray_torch_distributed_checkpoint_b200.parallel.engine import MLPTrainEngine
    torch.manual_seed(0)
    B, steps, lr, mu, p, seed = 384, 6, 0.05, 0.9, 0.25, 1234
    model = NeuralNetwork()
    eng = MLPTrainEngine(model, batch_size=B, lr=lr, momentum=mu, device="cuda", seed=seed, use_graph=use_graph, distributed=False,
                         fused_head=fused_head)
    sd = model.state_dict()
    params = [sd[f"linear_relu_stack.{i}.{k}"].double().cuda() for i in (0, 3, 6) for k in ("weight", "bias")]
    mom = [torch.zeros_like(t) for t in params]
    g = torch.Generator().manual_seed(1)
    for t in range(steps):
        x = torch.randn(B, 784, generator=g)
        y = torch.randint(0, 10, (B,), generator=g)
        loss = eng.step(x.cuda(), y.cuda())
        params, mom, ref_loss = _ref_step(params, mom, x.double().cuda(), y.cuda(), t, p, seed, lr, mu, t == 0)
        assert abs(loss.item() - ref_loss.item()) < 2e-4 * max(1.0, abs(ref_loss.item())), f"step {t}"
    got = eng.model_state_dict()
    names = [f"linear_relu_stack.{i}.{k}" for i in (0, 3, 6) for k in ("weight", "bias")]
    for n, ref in zip(names, params):
        # fp32 vs fp64 can flip a ReLU/dropout-masked unit that sits at ~0, so compare in norm and bound outliers
        diff = (got[n].double() - ref)
        assert diff.norm() / ref.norm() < 2e-4, (n, float(diff.norm() / ref.norm()))
        assert diff.abs().max() < 1e-3, (n, float(diff.abs().max()))
    osd = eng.optimizer_state_dict()
    assert osd["param_groups"][0]["momentum"] == mu and osd["param_groups"][0]["params"] == list(range(6))
    for i, ref in enumerate(mom):
        diff = osd["state"][i]["momentum_buffer"].double() - ref
        # one ReLU unit flipping between fp32 and fp64 moves a gradient row by ~1/B of its norm
        assert diff.norm() / ref.norm() < 2e-2, (i, float(diff.norm() / ref.norm()))
    if use_graph:
        assert ("train_f32", 0) in eng._graphs and ("train_f32", 1) in eng._graphs
    assert eng.kernel_launches_per_step["train_f32"] >= 10
    eng.close()


def test_engine_eval_u8_and_state_roundtrip():
    from ray_torch_distributed_checkpoint_b200.models import NeuralNetwork
    from ray_torch_distributed_checkpoint_b200.parallel.engine import MLPTrainEngine
    torch.manual_seed(1)
    B = 512
    model = NeuralNetwork().cuda()
    eng = MLPTrainEngine(model, batch_size=B, device="cuda", distributed=False)
    xu8 = torch.randint(0, 256, (B, 28, 28), dtype=torch.uint8, device="cuda")
    y = torch.randint(0, 10, (B,), device="cuda")
    xf = (xu8.float() / 255 - 0.5) / 0.5
    model.eval()
    with torch.no_grad():
        ref_logits = model(xf.unsqueeze(1))
    eng.reset_eval()
    for _ in range(4):                                   # past warm-up, into the captured graph
        eng.eval_batch(xu8, y)
    loss_sum, correct = eng.read_eval()
    ref_loss = F.cross_entropy(ref_logits, y).item()
    assert abs(loss_sum / 4 - ref_loss) < 1e-4
    assert correct == 4 * int((ref_logits.argmax(1) == y).sum())
    assert torch.allclose(eng.predict(xf), ref_logits, rtol=1e-4, atol=1e-5)
    # u8 and float inputs give the same training step
    e2 = MLPTrainEngine(NeuralNetwork(), batch_size=B, device="cuda", distributed=False, seed=7)
    e3 = MLPTrainEngine(NeuralNetwork(), batch_size=B, device="cuda", distributed=False, seed=7)
    e3.load_model_state_dict(e2.model_state_dict())
    l2 = e2.step(xu8, y).item()
    l3 = e3.step(xf, y).item()
    assert abs(l2 - l3) < 1e-5
    # ... including the folded input layer's gradients (dW1 = s*(dh1^T raw) + t*db1 1^T)
    # (a ReLU unit at ~0 may flip between the two roundings and move one sample's contribution, so the norm bound
    # is loose and the median, which such outliers cannot move, is tight)
    for got, ref in zip(e2.dW + e2.db, e3.dW + e3.db):
        err = (got - ref).abs()
        assert err.norm() <= 1e-2 * ref.norm()
        assert err.median() <= 1e-5 * ref.abs().mean()
    e4 = MLPTrainEngine(NeuralNetwork(), batch_size=B, device="cuda", distributed=False, seed=7, fold_input=False)
    e4.load_model_state_dict(e3.model_state_dict())
    e2.step(xu8, y); e3.step(xf, y)
    assert e2.input_mode == "shift" and e4.input_mode == "plain"
    # a normalisation whose shifted pixels are not tf32-exact takes the rowsum-folding route; same step as plain
    kw = dict(batch_size=B, device="cuda", distributed=False, seed=7, input_mean=0.1307, input_std=0.3081)
    e5, e6 = MLPTrainEngine(NeuralNetwork(), **kw), MLPTrainEngine(NeuralNetwork(), fold_input=False, **kw)
    e6.load_model_state_dict(e5.model_state_dict())
    assert e5.input_mode == "fold" and e6.input_mode == "plain"
    assert abs(e5.step(xu8, y).item() - e6.step(xu8, y).item()) < 1e-5
    for got, ref in zip(e5.dW + e5.db, e6.dW + e6.db):
        err = (got - ref).abs()
        assert err.norm() <= 1e-2 * ref.norm()
        assert err.median() <= 1e-5 * ref.abs().mean()
    # checkpoint-format round trip through a bare module
    fresh = e2.copy_to_module(NeuralNetwork())
    assert list(fresh.state_dict()) == list(e2.model_state_dict(prefix=""))
    e3.load_optimizer_state_dict(e2.optimizer_state_dict())
    assert torch.equal(e3.m_arena, e2.m_arena)
    # staged (pipelined) use gives the same training trajectory as passing each batch to step()
    kw = dict(batch_size=B, device="cuda", distributed=False, seed=11)
    ea, eb = MLPTrainEngine(NeuralNetwork(), **kw), MLPTrainEngine(NeuralNetwork(), **kw)
    eb.load_model_state_dict(ea.model_state_dict())
    batches = [(torch.randint(0, 256, (B, 28, 28), dtype=torch.uint8, device="cuda"), torch.randint(0, 10, (B,), device="cuda"))
               for _ in range(6)]
    la = [ea.step(xb, yb).item() for xb, yb in batches]
    lb = []
    eb.stage(*batches[0])
    for i in range(len(batches)):
        if i + 1 < len(batches):
            eb.stage(*batches[i + 1])          # next batch's conversion overlaps this step
        lb.append(eb.step().item())
    assert la == lb
    assert torch.equal(ea.p_arena, eb.p_arena)
    with pytest.raises(RuntimeError):
        eb.step()
    for e in (eng, e2, e3, e4, e5, e6, ea, eb):
        e.close()
