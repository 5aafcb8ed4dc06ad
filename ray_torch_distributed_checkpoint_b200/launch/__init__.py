"""Worker-group launcher: one rank per GPU (or CPU worker) on this node.

Native path: ``lib/rtdc-launch`` (``csrc/launcher_main.cpp``) forks/execs the ranks, sets the
rendezvous environment, pins CPU cores, and tears the whole group down when any rank fails. The
Python fallback does the same with ``subprocess`` when the binary has not been built.
"""
from .group import WorkerGroup, find_free_port

__all__ = ["WorkerGroup", "find_free_port"]
