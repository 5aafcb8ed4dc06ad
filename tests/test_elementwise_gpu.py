"""Fused loss / optimizer / normalisation kernels against plain PyTorch fp32 references."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def _ops():
    from ray_torch_distributed_checkpoint_b200 import ops
    return ops


@pytest.mark.parametrize("rows", [1, 37, 4096, 70000])
def test_softmax_ce_forward_backward(rows):
    ops = _ops()
    g = torch.Generator(device="cuda").manual_seed(rows)
    z = torch.zeros(rows, 16, device="cuda")
    z[:, :10] = torch.relu(torch.randn(rows, 10, device="cuda", generator=g) * 3)
    y = torch.randint(0, 10, (rows,), device="cuda", generator=g)
    dz = torch.full_like(z, 5.0)
    correct = torch.zeros((), dtype=torch.int64, device="cuda")
    db = torch.full((10,), 7.0, device="cuda")
    loss = ops.softmax_cross_entropy(z, y, classes=10, dlogits=dz, relu_mask=True, correct_accum=correct, dbias_out=db)
    zr = z[:, :10].clone().requires_grad_(True)
    ref = F.cross_entropy(zr, y)
    ref.backward()
    assert torch.allclose(loss, ref.detach(), rtol=2e-5, atol=1e-6)
    ref_dz = zr.grad * (z[:, :10] > 0)
    assert torch.allclose(dz[:, :10], ref_dz, rtol=1e-4, atol=1e-7 / max(rows, 1) + 1e-9)
    assert (dz[:, 10:] == 0).all()
    assert torch.allclose(db.double(), dz[:, :10].double().sum(0), rtol=1e-4, atol=1e-6)      # fused bias gradient
    assert correct.item() == (z[:, :10].argmax(1) == y).sum().item()
    # determinism + accumulation on a second launch
    loss2 = ops.softmax_cross_entropy(z, y, classes=10, correct_accum=correct)
    assert loss2.item() == loss.item()
    assert correct.item() == 2 * (z[:, :10].argmax(1) == y).sum().item()


def test_softmax_ce_unpadded_32_classes():
    ops = _ops()
    z = torch.randn(513, 32, device="cuda")
    y = torch.randint(0, 32, (513,), device="cuda")
    dz = torch.empty_like(z)
    loss = ops.softmax_cross_entropy(z, y, dlogits=dz)
    zr = z.clone().requires_grad_(True)
    ref = F.cross_entropy(zr, y)
    ref.backward()
    assert torch.allclose(loss, ref.detach(), rtol=2e-5)
    assert torch.allclose(dz, zr.grad, rtol=1e-4, atol=1e-8)


@pytest.mark.parametrize("rows,cols,ld", [(4096, 512, 512), (1000, 10, 16), (333, 784, 784), (5, 3, 7)])
def test_colsum(rows, cols, ld):
    ops = _ops()
    x = torch.randn(rows, ld, device="cuda")
    out = ops.colsum(x, cols=cols)
    ref = x[:, :cols].double().sum(0)
    assert torch.allclose(out.double(), ref, rtol=1e-5, atol=1e-4)
    out2 = ops.colsum(x, cols=cols)
    assert torch.equal(out, out2)


def test_u8_normalize():
    ops = _ops()
    src = torch.randint(0, 256, (1031, 28, 28), dtype=torch.uint8, device="cuda")
    out = ops.u8_normalize(src)
    ref = (src.float() / 255.0 - 0.5) / 0.5
    assert out.shape == src.shape
    assert torch.allclose(out, ref, atol=1e-6)


@pytest.mark.parametrize("n", [669706, 1023, 4])
@pytest.mark.parametrize("momentum,nesterov,wd", [(0.9, False, 0.0), (0.0, False, 0.0), (0.9, True, 1e-2)])
def test_fused_sgd_matches_torch(n, momentum, nesterov, wd):
    ops = _ops()
    p = torch.randn(n, device="cuda")
    ref_p = p.clone().requires_grad_(True)
    opt = torch.optim.SGD([ref_p], lr=1e-2, momentum=momentum, nesterov=nesterov, weight_decay=wd)
    m = torch.zeros(n, device="cuda")
    st = ops.SgdState("cuda", lr=1e-2, momentum=momentum, nesterov=nesterov, weight_decay=wd)
    for i in range(4):
        g = torch.randn(n, device="cuda")
        ref_p.grad = g.clone()
        opt.step()
        ops.sgd_step(p, g, m, st)
        ops.counter_advance(None, st)
        assert torch.allclose(p, ref_p.detach(), rtol=1e-5, atol=1e-6), f"step {i}"
    if momentum:
        assert torch.allclose(m, opt.state[ref_p]["momentum_buffer"], rtol=1e-5, atol=1e-6)


@pytest.mark.parametrize("decoupled", [False, True])
def test_fused_adam_matches_torch(decoupled):
    ops = _ops()
    n = 100003
    p = torch.randn(n, device="cuda")
    ref_p = p.clone().requires_grad_(True)
    cls = torch.optim.AdamW if decoupled else torch.optim.Adam
    opt = cls([ref_p], lr=1e-3, weight_decay=1e-2)
    m = torch.zeros(n, device="cuda"); v = torch.zeros(n, device="cuda")
    st = ops.AdamState("cuda", lr=1e-3, weight_decay=1e-2, decoupled=decoupled)
    for i in range(5):
        g = torch.randn(n, device="cuda")
        ref_p.grad = g.clone()
        opt.step()
        ops.adam_step(p, g, m, v, st)
        assert torch.allclose(p, ref_p.detach(), rtol=2e-5, atol=2e-6), f"step {i}"


@pytest.mark.parametrize("rows,classes", [(1000, 10), (64, 10), (4133, 16), (18944, 10)])
def test_fused_mlp_head_matches_autograd(rows, classes):
    ops = _ops()
    K, p = 512, 0.25
    g = torch.Generator(device="cuda").manual_seed(rows + classes)
    h = torch.relu(torch.randn(rows, K, device="cuda", generator=g)) * (torch.rand(rows, K, device="cuda", generator=g) > p)
    w = torch.randn(classes, K, device="cuda", generator=g) * 0.05
    b = torch.randn(classes, device="cuda", generator=g) * 0.1
    y = torch.randint(0, classes, (rows,), device="cuda", generator=g)
    dh = torch.full((rows, K), 3.0, device="cuda"); dw = torch.empty(classes, K, device="cuda"); db = torch.empty(classes, device="cuda")
    for _ in range(2):
        dbh = torch.empty(K, device="cuda")
        loss = ops.mlp_head_train(h, w, b, y, dh=dh, dw=dw, db=db, mask_scale=1 / (1 - p), db_below=dbh)
    hr = h.double().requires_grad_(True); wr = w.double().requires_grad_(True); br = b.double().requires_grad_(True)
    ref = F.cross_entropy(torch.relu(hr @ wr.T + br), y)
    ref.backward()
    assert abs(loss.item() - ref.item()) < 1e-5 * max(1.0, abs(ref.item()))
    ref_dh = hr.grad * (h > 0) / (1 - p)
    assert torch.allclose(dh.double(), ref_dh, rtol=1e-4, atol=1e-9)
    assert (dw.double() - wr.grad).norm() / wr.grad.norm() < 1e-5
    assert (db.double() - br.grad).norm() / br.grad.norm() < 1e-5
    assert (dbh.double() - ref_dh.sum(0)).norm() / ref_dh.sum(0).norm() < 1e-5
