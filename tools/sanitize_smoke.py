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
torch.cuda.synchronize()
ref = torch.relu(x.double() @ w.double().T + b.double())
assert torch.allclose(y1.double(), ref, rtol=1e-2, atol=1e-2)
assert torch.allclose(dwb.double(), dyb.double().T @ xb.double(), rtol=1e-4, atol=1e-3)
print("sanitize smoke ok", float(loss))
