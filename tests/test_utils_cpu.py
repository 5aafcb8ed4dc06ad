"""Host-side helpers that need no GPU."""
import os


def test_cpulist_parsing_and_numa_binding_is_safe_without_gpu():
    from ray_torch_distributed_checkpoint_b200.utils.affinity import _parse_cpulist, bind_to_gpu_numa
    assert _parse_cpulist("0-3,8,10-11\n") == {0, 1, 2, 3, 8, 10, 11}
    assert _parse_cpulist("") == set()
    before = os.sched_getaffinity(0)
    info = bind_to_gpu_numa(0)                      # no CUDA device here: must report, not raise, and change nothing
    assert info["bound"] is False and info["numa_node"] is None
    assert os.sched_getaffinity(0) == before


def test_graphed_step_runs_eagerly_without_cuda():
    """utils.GraphedStep off the GPU (or with enabled=False) is a plain call of the step function."""
    from ray_torch_distributed_checkpoint_b200.utils import GraphedStep
    calls = []
    step = GraphedStep(lambda: calls.append(1) or len(calls), warmup=3, enabled=False)
    assert step.graph is None and step() == 1 and step() == 2 and len(calls) == 2


def test_fused_layernorm_module_rejects_a_wrong_width_like_torch():
    import pytest
    import torch
    from ray_torch_distributed_checkpoint_b200.ops.layernorm import LayerNormBf16
    ln = LayerNormBf16(torch.nn.LayerNorm(16))
    with pytest.raises(RuntimeError):
        ln(torch.randn(2, 24))
