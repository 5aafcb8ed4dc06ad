"""The allocation-free column-sum entry point, called through the C ABI the way a non-Python host would.

Some users of this library cannot allocate device memory at the moment they need a kernel: a serving
process that gave its whole HBM budget to one arena at start-up, or a control loop where the hidden
device synchronisation of cudaMalloc/cudaFree is off limits. ``rtdc_colsum_noalloc`` is for them: it
keeps its scratch inside the library. This test drives it with ctypes only (no wrapper from ops/),
twice in a row on different data, and checks each result against the ordinary ``rtdc_colsum`` entry
point fed with an explicit scratch buffer. The file name sorts last on purpose: with ``pytest -x`` a
failure here cannot hide the result of any other test.
"""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu


def _lib():
    from ray_torch_distributed_checkpoint_b200.ops import _build
    if not _build.LIB_PATH.exists():
        _build.build()
    lib = C.CDLL(str(_build.LIB_PATH))
    vp, i, sz = C.c_void_p, C.c_int, C.c_size_t
    lib.rtdc_last_error.restype = C.c_char_p
    lib.rtdc_colsum_noalloc.argtypes = [vp, i, i, i, vp, vp]
    lib.rtdc_colsum_noalloc.restype = i
    lib.rtdc_colsum.argtypes = [vp, i, i, i, vp, vp, sz, vp]
    lib.rtdc_colsum.restype = i
    lib.rtdc_colsum_scratch_bytes.argtypes = [i, i]
    lib.rtdc_colsum_scratch_bytes.restype = sz
    return lib


def test_colsum_noalloc_matches_colsum_twice_in_a_row():
    lib = _lib()
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    for call, (rows, cols, ld) in enumerate([(4096, 512, 512), (3001, 1000, 1024)]):
        x = torch.randn(rows, ld, device="cuda") * (call + 1)
        got = torch.full((cols,), float("nan"), device="cuda")
        ref = torch.full((cols,), float("nan"), device="cuda")
        if call == 0:
            # the first call on a device resolves the built-in scratch's address; make it before measuring free memory
            assert lib.rtdc_colsum_noalloc(x.data_ptr(), ld, rows, cols, got.data_ptr(), stream) == 0, lib.rtdc_last_error().decode()
        torch.cuda.synchronize()
        before, free_before = torch.cuda.memory_allocated(), torch.cuda.mem_get_info()[0]
        rc = lib.rtdc_colsum_noalloc(x.data_ptr(), ld, rows, cols, got.data_ptr(), stream)
        assert rc == 0, lib.rtdc_last_error().decode()
        torch.cuda.synchronize()
        assert torch.cuda.memory_allocated() == before          # nothing was taken from torch's allocator ...
        assert torch.cuda.mem_get_info()[0] == free_before      # ... and no cudaMalloc happened inside the library either
        need = lib.rtdc_colsum_scratch_bytes(rows, cols)
        scratch = torch.zeros(need, dtype=torch.uint8, device="cuda")
        rc = lib.rtdc_colsum(x.data_ptr(), ld, rows, cols, ref.data_ptr(), scratch.data_ptr(), need, stream)
        assert rc == 0, lib.rtdc_last_error().decode()
        torch.cuda.synchronize()
        assert torch.allclose(got, ref, rtol=1e-5, atol=1e-4), f"call {call}: max diff {(got - ref).abs().max().item()}"
        assert torch.allclose(got.double(), x[:, :cols].double().sum(0), rtol=1e-5, atol=1e-4 * (call + 1))


def test_colsum_noalloc_rejects_oversize():
    lib = _lib()
    x = torch.zeros(8, 2048, device="cuda")
    out = torch.zeros(2048, device="cuda")
    rc = lib.rtdc_colsum_noalloc(x.data_ptr(), 2048, 8, 1025, out.data_ptr(), C.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc != 0
    msg = lib.rtdc_last_error().decode()
    assert "1024" in msg and "exceed" in msg
