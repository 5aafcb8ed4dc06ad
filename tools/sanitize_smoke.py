"""Small run of every kernel family for compute-sanitizer (memcheck / racecheck): tiny shapes, few launches."""
import torch
from ray_torch_distributed_checkpoint_b200 import ops

dev = "cuda"
torch.manual_seed(0)
x = torch.randn(300, 64, device=dev); w = torch.randn(48, 64, device=dev); b = torch.randn(48, device=dev)
y = ops.linear_forward(x, w, b, relu=True, drop_p=0.25, seed=1)                       # K-major, 3xTF32, TMA store
y1 = ops.linear_forward(x, w, b, relu=True, precision=ops.PRECISION_TF32)              # single tf32
dy = torch.randn(300, 48, device=dev)
dx = ops.linear_dgrad(dy, w, mask_src=x.abs(), mask_scale=1.0)                         # MN-major B + mask
dw = ops.linear_wgrad(dy, x)                                                          # MN-major both
xb = torch.randn(2048, 64, device=dev); dyb = torch.randn(2048, 48, device=dev)
dwb = ops.linear_wgrad(dyb, xb)                                                        # split-K + reduce kernel
z = torch.zeros(300, 16, device=dev); z[:, :10] = torch.randn(300, 10, device=dev)
lab = torch.randint(0, 10, (300,), device=dev)
dz = torch.empty_like(z); db = torch.empty(10, device=dev)
loss = ops.softmax_cross_entropy(z, lab, classes=10, dlogits=dz, relu_mask=True, dbias_out=db)
cs = ops.colsum(dy)
u8 = torch.randint(0, 256, (37, 784), dtype=torch.uint8, device=dev)
f = ops.u8_normalize(u8)
p = torch.randn(1001, device=dev); g = torch.randn(1001, device=dev); m = torch.zeros(1001, device=dev)
st = ops.SgdState(dev, 0.1, 0.9)
ops.sgd_step(p, g, m, st); ops.counter_advance(None, st)
v = torch.zeros(1001, device=dev); ast = ops.AdamState(dev, 1e-3)
ops.adam_step(p, g, m, v, ast)
# CTA-pair kernel (N >= 128, M > 128): odd number of 128-row tiles (the last pair's second half lies past M), fused column
# sums, exact operand + affine epilogue, split-K with both operands MN-major
w2 = torch.randn(160, 64, device=dev); b2 = torch.randn(160, device=dev)
y2 = ops.linear_forward(x, w2, b2, relu=True, drop_p=0.25, seed=1)
xi = torch.randint(0, 256, (300, 64), device=dev).float() - 127.5
y3 = ops.linear_forward(xi, w2, b2, relu=True, x_exact=True, alpha=2.0 / 255)
dy2 = torch.randn(300, 160, device=dev); w3 = torch.randn(160, 128, device=dev); hmask = torch.randn(300, 128, device=dev)
part = torch.zeros(12, 128, device=dev)
dx2 = ops.linear_dgrad(dy2, w3, mask_src=hmask, mask_scale=4 / 3, colsum_partial=part)
dyc = torch.randn(2048, 256, device=dev); xc = torch.randint(0, 256, (2048, 128), device=dev).float()
dwc = ops.linear_wgrad(dyc, xc, x_exact=True, alpha=0.5, row_bias=dyc.sum(0), row_bias_scale=-1.0)
# fused head, input conversion helpers
hh = torch.relu(torch.randn(100, 64, device=dev)); wh = torch.randn(10, 64, device=dev) * 0.1; bh = torch.zeros(10, device=dev)
labh = torch.randint(0, 10, (100,), device=dev)
dhh = torch.empty_like(hh); dwh = torch.empty_like(wh); dbh = torch.empty(10, device=dev); dbl = torch.empty(64, device=dev)
lh = ops.mlp_head_train(hh, wh, bh, labh, dh=dhh, dw=dwh, db=dbh, db_below=dbl)
fa = ops.u8_affine(u8, shift=-127.5)
fb = ops.fold_input_bias(w2, b2, -1.0)
# whole engine step, staged use included
from ray_torch_distributed_checkpoint_b200.models import NeuralNetwork
from ray_torch_distributed_checkpoint_b200.parallel.engine import MLPTrainEngine
eng = MLPTrainEngine(NeuralNetwork(), batch_size=256, device=dev, distributed=False, use_graph=False)
xe = torch.randint(0, 256, (256, 784), dtype=torch.uint8, device=dev); ye = torch.randint(0, 10, (256,), device=dev)
eng.stage(xe, ye)
for _ in range(2):
    eng.stage(xe, ye)
    le = eng.step()
le = eng.step()
torch.cuda.synchronize()
assert torch.allclose(y3.double(), torch.relu((2.0 / 255) * (xi.double() @ w2.double().T) + b2.double()), rtol=1e-4, atol=1e-4)
assert torch.allclose(dwc.double(), 0.5 * (dyc.double().T @ xc.double()) - dyc.double().sum(0)[:, None], rtol=1e-4, atol=1e-2)
ref = torch.relu(x.double() @ w.double().T + b.double())
assert torch.allclose(y1.double(), ref, rtol=1e-2, atol=1e-2)
assert torch.allclose(dwb.double(), dyb.double().T @ xb.double(), rtol=1e-4, atol=1e-3)
print("sanitize smoke ok", float(loss), float(lh), float(le))
