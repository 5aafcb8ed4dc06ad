"""tcgen05 GEMM (csrc/gemm_tf32.cu) against plain PyTorch references computed in fp64."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _ops():
    from ray_torch_distributed_checkpoint_b200 import ops
    return ops


def _rel_err(got, ref64, scale64):
    return ((got.double() - ref64).abs() / scale64.clamp_min(1e-30)).max().item()


@pytest.mark.parametrize("M,N,K", [(300, 512, 784), (128, 256, 64), (1000, 512, 512), (77, 48, 40)])
@pytest.mark.parametrize("precision,tol", [(3, 4e-6), (1, 3e-3)])
def test_forward_kmajor(M, N, K, precision, tol):
    ops = _ops()
    g = torch.Generator(device="cuda").manual_seed(1)
    x = torch.randn(M, K, device="cuda", generator=g)
    w = torch.randn(N, K, device="cuda", generator=g) * 0.05
    b = torch.randn(N, device="cuda", generator=g)
    y = ops.linear_forward(x, w, b, precision=precision)
    ref = x.double() @ w.double().T + b.double()
    scale = x.double().abs() @ w.double().abs().T + b.double().abs()
    assert _rel_err(y, ref, scale) < tol
    # fp32-equivalent mode must be at least as close to fp64 as torch's own fp32 GEMM, within 4x
    if precision == 3:
        torch_err = _rel_err(torch.addmm(b, x, w.T), ref, scale)
        assert _rel_err(y, ref, scale) < max(4 * torch_err, 1e-6)


def test_forward_relu_dropout_matches_philox():
    ops = _ops()
    from ray_torch_distributed_checkpoint_b200.ops.philox_ref import dropout_keep_mask
    M, N, K, p = 260, 512, 784, 0.25
    g = torch.Generator(device="cuda").manual_seed(2)
    x = torch.randn(M, K, device="cuda", generator=g)
    w = torch.randn(N, K, device="cuda", generator=g) * 0.05
    b = torch.randn(N, device="cuda", generator=g)
    step_t = torch.tensor([7], dtype=torch.int32, device="cuda")
    y = ops.linear_forward(x, w, b, relu=True, drop_p=p, seed=0x1234567890ABCDEF, step_tensor=step_t, salt=3)
    keep = torch.from_numpy(dropout_keep_mask(M, N, p, 0x1234567890ABCDEF, step=7, salt=3)).cuda()
    ref = torch.relu(x.double() @ w.double().T + b.double()) * keep / (1 - p)
    scale = (x.double().abs() @ w.double().abs().T + b.double().abs()) / (1 - p)
    assert _rel_err(y, ref, scale) < 4e-6
    assert (y[~keep] == 0).all()
    frac = keep.float().mean().item()
    assert abs(frac - 0.75) < 0.01
    # a different step draws a different mask
    y2 = ops.linear_forward(x, w, b, relu=True, drop_p=p, seed=0x1234567890ABCDEF, step=8, salt=3)
    assert ((y2 == 0) != (y == 0)).any()


def test_forward_skinny_n_padded_store():
    ops = _ops()
    M, N, K = 515, 10, 512
    g = torch.Generator(device="cuda").manual_seed(3)
    x = torch.randn(M, K, device="cuda", generator=g)
    w = torch.randn(N, K, device="cuda", generator=g) * 0.05
    b = torch.randn(N, device="cuda", generator=g)
    out = torch.full((M, 16), 9.0, device="cuda")
    ops.linear_forward(x, w, b, relu=True, out=out, n_store=16)
    ref = torch.relu(x.double() @ w.double().T + b.double())
    scale = x.double().abs() @ w.double().abs().T + b.double().abs()
    assert _rel_err(out[:, :10], ref, scale) < 4e-6
    assert (out[:, 10:] == 0).all()


@pytest.mark.parametrize("B,n_out,k_in,ld_dy", [(300, 512, 512, 512), (515, 10, 512, 16), (256, 512, 784, 512)])
def test_dgrad_with_mask(B, n_out, k_in, ld_dy):
    ops = _ops()
    g = torch.Generator(device="cuda").manual_seed(4)
    dy = torch.zeros(B, ld_dy, device="cuda")
    dy[:, :n_out] = torch.randn(B, n_out, device="cuda", generator=g)
    w = torch.randn(n_out, k_in, device="cuda", generator=g) * 0.05
    h = torch.relu(torch.randn(B, k_in, device="cuda", generator=g))
    dx = ops.linear_dgrad(dy, w, N_out=n_out, mask_src=h, mask_scale=1.0 / 0.75)
    ref = (dy[:, :n_out].double() @ w.double()) * (h > 0) / 0.75
    scale = (dy[:, :n_out].double().abs() @ w.double().abs()) / 0.75
    assert _rel_err(dx, ref, scale) < 4e-6


@pytest.mark.parametrize("B,n_out,k_in,ld_dy", [(4096, 512, 784, 512), (4096, 512, 512, 512), (2080, 10, 512, 16), (100, 64, 96, 64)])
@pytest.mark.parametrize("splits", [0, 1])
def test_wgrad(B, n_out, k_in, ld_dy, splits):
    ops = _ops()
    g = torch.Generator(device="cuda").manual_seed(5)
    dy = torch.zeros(B, ld_dy, device="cuda")
    dy[:, :n_out] = torch.randn(B, n_out, device="cuda", generator=g)
    x = torch.randn(B, k_in, device="cuda", generator=g)
    ref = dy[:, :n_out].double().T @ x.double()
    scale = dy[:, :n_out].double().abs().T @ x.double().abs()
    for _ in range(2):  # second call reuses the scratch and its self-resetting counters
        dw = ops.linear_wgrad(dy, x, N_out=n_out, splits=splits)
        assert dw.shape == (n_out, k_in)
        # one fp32 accumulator over the whole batch when splits == 1: allow sqrt(K) growth
        assert _rel_err(dw, ref, scale) < (4e-6 if splits == 0 else 1.2e-5)


def test_wgrad_into_bucket_view():
    ops = _ops()
    B, n_out, k_in = 2048, 512, 784
    g = torch.Generator(device="cuda").manual_seed(6)
    dy = torch.randn(B, n_out, device="cuda", generator=g)
    x = torch.randn(B, k_in, device="cuda", generator=g)
    bucket = torch.zeros(512 + n_out * k_in + 128, device="cuda")
    view = bucket[512:512 + n_out * k_in].view(n_out, k_in)
    ops.linear_wgrad(dy, x, out=view)
    ref = dy.double().T @ x.double()
    assert _rel_err(view, ref, dy.double().abs().T @ x.double().abs()) < 4e-6
    assert (bucket[:512] == 0).all() and (bucket[512 + n_out * k_in:] == 0).all()


def test_plan_and_errors():
    ops = _ops()
    plan = ops.gemm_plan(16384, 512, 784)
    assert plan["bn"] == 256 and plan["splits"] == 1
    plan = ops.gemm_plan(512, 784, 16384, b_mn=True)
    assert plan["bn"] % 32 == 0 and plan["splits"] > 1
    x = torch.randn(64, 30, device="cuda")  # leading dimension not a multiple of 4 floats
    w = torch.randn(32, 30, device="cuda")
    with pytest.raises(ops.RtdcError):
        ops.linear_forward(x, w)


@pytest.mark.parametrize("B", [1000, 384, 640])
def test_dgrad_fused_column_sums(B):
    """1000 rows: 8 tiles of 128 (4 CTA pairs); 384 / 640: an odd number of 128-row tiles, so the last CTA pair's
    second half lies past M and must not write its (empty) partial rows beyond the buffer."""
    ops = _ops()
    n_out, k_in = 512, 512
    g = torch.Generator(device="cuda").manual_seed(8)
    dy = torch.randn(B, n_out, device="cuda", generator=g)
    w = torch.randn(n_out, k_in, device="cuda", generator=g) * 0.05
    h = torch.relu(torch.randn(B, k_in, device="cuda", generator=g))
    rows = 4 * ((B + 127) // 128)
    arena = torch.full((rows + 8, k_in), 7.0, device="cuda")
    part = arena[:rows]
    dx = ops.linear_dgrad(dy, w, mask_src=h, mask_scale=4 / 3, colsum_partial=part)
    groups = (B + 31) // 32
    assert torch.allclose(part[:groups].sum(0).double(), dx.double().sum(0), rtol=1e-4, atol=1e-3)
    assert torch.allclose(ops.colsum(part[:groups].contiguous()).double(), dx.double().sum(0), rtol=1e-4, atol=1e-3)
    assert bool((arena[rows:] == 7.0).all()), "partial sums written past the buffer"


@pytest.mark.parametrize("which", ["a", "b", "both"])
def test_exact_operand_and_affine_epilogue(which):
    """Operands whose values are exact in tf32 (raw pixels) skip their correction pass; alpha / row_bias fold an input
    normalisation into the layer. fp32-accurate against float64."""
    from ray_torch_distributed_checkpoint_b200 import ops
    torch.manual_seed(3)
    M, N, K = 300, 520, 784
    s, t = 2.0 / 255.0, -1.0
    ints = lambda *shape: torch.randint(0, 256, shape, device="cuda").float()
    a = ints(M, K) if which in ("a", "both") else torch.randn(M, K, device="cuda")
    b = ints(N, K) if which in ("b", "both") else torch.randn(N, K, device="cuda") * 0.05
    bias = torch.randn(N, device="cuda")
    rb = torch.randn(M, device="cuda")
    ref = s * (a.double() @ b.double().T) + bias.double() + t * rb.double()[:, None]
    scale = (s * (a.double().abs() @ b.double().abs().T)).max().item()
    for splits in (1, 0):      # in-kernel epilogue, then the split-K reduce epilogue
        out = torch.empty(M, N, device="cuda")
        ops.gemm(a, b, out, bias=bias, relu=False, a_exact=which in ("a", "both"), b_exact=which in ("b", "both"), alpha=s,
                 row_bias=rb, row_bias_scale=t, splits=splits)
        assert (out.double() - ref).abs().max().item() < 1.2e-5 * scale, splits
    # the same through split-K (wgrad shape: both operands MN-major, K = batch)
    Kb = 4096
    dy = torch.randn(Kb, 64, device="cuda")
    x = ints(Kb, 200)
    db = dy.sum(0)
    dw = torch.empty(64, 200, device="cuda")
    ops.linear_wgrad(dy, x, out=dw, x_exact=True, alpha=s, row_bias=db, row_bias_scale=t)
    ref = dy.double().T @ (s * x.double() + t)
    scale = (dy.double().abs().T @ (s * x.double())).max().item()
    assert (dw.double() - ref).abs().max().item() < 1.2e-5 * scale
    assert ops.gemm_plan(64, 200, Kb, b_mn=True)["splits"] > 1


def test_fold_input_bias():
    from ray_torch_distributed_checkpoint_b200 import ops
    torch.manual_seed(4)
    for rows, cols in ((512, 784), (7, 33), (64, 1000)):
        w = torch.randn(rows, cols, device="cuda")
        b = torch.randn(rows, device="cuda")
        got = ops.fold_input_bias(w, b, -1.0)
        ref = b.double() - w.double().sum(1)
        assert (got.double() - ref).abs().max().item() < 1e-4
        assert torch.equal(ops.fold_input_bias(w, None, 0.5), ops.fold_input_bias(w, None, 0.5))


def test_u8_affine_widens_exactly():
    from ray_torch_distributed_checkpoint_b200 import ops
    src = torch.arange(256, dtype=torch.uint8, device="cuda").repeat(64)
    assert torch.equal(ops.u8_affine(src), src.float())


@pytest.mark.parametrize("bn", [0, 176, 48])
def test_multi_tile_n_not_multiple_of_32(bn):
    """Several N tiles whose width is not a multiple of the 32-column store box (regression: the last box of a tile
    used to overwrite the first columns of the next tile)."""
    from ray_torch_distributed_checkpoint_b200 import ops
    torch.manual_seed(5)
    M, N, K = 300, 520, 96
    a, b = torch.randn(M, K, device="cuda"), torch.randn(N, K, device="cuda")
    out = torch.full((M, N), float("nan"), device="cuda")
    ops.gemm(a, b, out, bn=bn, splits=1)
    ref = a.double() @ b.double().T
    assert (out.double() - ref).abs().max().item() < 4e-6 * (a.double().abs() @ b.double().abs().T).max().item()


@pytest.mark.parametrize("M,N", [(300, 512), (384, 160), (129, 130), (640, 784)])
def test_output_stays_inside_its_rows(M, N):
    """Canary rows behind the output (and behind the split-K result) stay untouched: the CTA-pair kernel schedules a
    phantom half-tile past M when the number of 128-row tiles is odd, TMA stores clip at the tensor edge."""
    ops = _ops()
    torch.manual_seed(9)
    K = 96
    a, b = torch.randn(M, K, device="cuda"), torch.randn(N, K, device="cuda")
    ld = (N + 3) // 4 * 4
    arena = torch.full((M + 130, ld), 5.0, device="cuda")
    ops.gemm(a, b, arena[:M], N=N, splits=1)
    ref = a.double() @ b.double().T
    assert (arena[:M, :N].double() - ref).abs().max().item() < 4e-6 * (a.double().abs() @ b.double().abs().T).max().item()
    assert bool((arena[M:] == 5.0).all())
    # split-K twin (wgrad shape)
    dy, x = torch.randn(2048, (M + 3) // 4 * 4, device="cuda"), torch.randn(2048, ld, device="cuda")
    arena.fill_(5.0)
    ops.linear_wgrad(dy, x, N_out=M, out=arena[:M])
    assert bool((arena[M:] == 5.0).all())
    assert (arena[:M].double() - dy[:, :M].double().T @ x.double()).abs().max().item() < 1e-3


@pytest.mark.parametrize("B,N", [(300, 512), (1000, 200), (129, 96)])
def test_bit_packed_mask_roundtrip(B, N):
    """A forward GEMM writes (out > 0) as one bit per element; a dgrad consumes those bits instead of the activation and
    must give exactly what the float mask gives."""
    ops = _ops()
    torch.manual_seed(12)
    K = 96
    x, w, b = torch.randn(B, K, device="cuda"), torch.randn(N, K, device="cuda") * 0.1, torch.randn(N, device="cuda")
    words = (N + 31) // 32
    arena = torch.full((B + 4, words + 1), -1, dtype=torch.int32, device="cuda")
    bits = arena[:B, :words]
    h = ops.linear_forward(x, w, b, relu=True, drop_p=0.25, seed=3, mask_bits_out=bits)
    ref_bits = torch.zeros(B, words * 32, dtype=torch.bool, device="cuda")
    ref_bits[:, :N] = h > 0
    weights = (1 << torch.arange(32, device="cuda", dtype=torch.int64))
    packed = (ref_bits.view(B, words, 32).long() * weights).sum(-1)
    assert torch.equal(bits.long() & 0xFFFFFFFF, packed)
    assert bool((arena[B:] == -1).all()) and bool((arena[:, words] == -1).all())
    dy, w2 = torch.randn(B, 64, device="cuda"), torch.randn(64, N, device="cuda") * 0.1
    a = ops.linear_dgrad(dy, w2, mask_src=h, mask_scale=4 / 3)
    c = ops.linear_dgrad(dy, w2, mask_bits=bits, mask_scale=4 / 3)
    assert torch.equal(a, c)
