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
