"""A ``torch.save``-compatible archive written with parallel positional writes.

Why: the asynchronous checkpointer (``ckpt/async_writer.py``) hides a checkpoint behind training only while its writer keeps
up, and for a GPT-2-sized state (1.2 GB of parameters + Adam moments) stock ``torch.save`` is one thread pushing every byte
through a CRC and a buffered ``fwrite``, after a staging copy out of the snapshot buffer. Here PyTorch still
produces the pickle (``torch.serialization._save`` with a record collector standing in for its zip writer, so the object
graph, storage keys and dtypes are exactly what ``torch.load`` expects), but the container is laid out up front — every
record's offset is known from the sizes alone — and the payload goes to the file with ``os.pwrite`` straight out of the
tensors' memory (the pinned snapshot buffer: no staging copy) from ONE writer thread in file order (buffered writes to one
inode serialise in the kernel; several writers only fight over its lock), while a small pool computes the CRC-32 of each
record beside it. Both release the GIL. Measured on the B200 host (1.24 GB state, ``tools/ckpt_writer_probe.py``): 0.34-0.8 s
against 0.58-0.76 s for ``torch.save`` — the box's page-cache write-back (~1.6 GB/s sustained) is the bound either way; what
the checkpointer gains is the staging copy it no longer makes (GPT-2 small, checkpoint every 5 steps: 164 -> 133 ms/step).

The result is a plain zip with the entries, names, order and contents of a stock archive (``<stem>/data.pkl``,
``.format_version``, ``.storage_alignment``, ``byteorder``, ``data/<key>`` stored uncompressed at 64-byte aligned offsets,
``version``, ``.data/serialization_id``): ``torch.load(path, weights_only=True)`` and ``zipfile`` read it like any other; the
reference's loader (my_ray_module.py:255-259) needs nothing else. Differences from the stock container, invisible to
readers: sizes and CRCs sit in the local headers instead of trailing data descriptors.
"""
from __future__ import annotations

import concurrent.futures as _fut
import ctypes
import os
import pickle
import struct
import zlib
from typing import Any, List, Optional, Tuple

import torch

_ALIGN = 64
_CHUNK = 32 << 20
_U32 = 0xFFFFFFFF


class _Record:
    __slots__ = ("name", "data", "ptr", "size", "keep", "crc", "header_off", "data_off", "extra_local")

    def __init__(self, name: str, data: Optional[bytes], ptr: int, size: int, keep: Any):
        self.name, self.data, self.ptr, self.size, self.keep = name, data, ptr, size, keep
        self.crc = 0
        self.header_off = self.data_off = 0
        self.extra_local = b""


class _Collector:
    """Stands in for ``torch._C.PyTorchFileWriter`` inside ``torch.serialization._save``: remembers what would be written."""

    def __init__(self):
        self.records: List[_Record] = []

    def write_record(self, name: str, data: Any, size: int):
        if isinstance(data, str):
            data = data.encode()
        if isinstance(data, (bytes, bytearray, memoryview)):
            b = bytes(data[:size])
            self.records.append(_Record(name, b, 0, len(b), None))
            return
        # an UntypedStorage on the CPU (torch moves device storages to the host before calling)
        if getattr(getattr(data, "device", None), "type", "cpu") != "cpu":
            raise TypeError("expected a CPU storage")
        self.records.append(_Record(name, None, int(data.data_ptr()) if size else 0, int(size), data))

    def write_record_metadata(self, name: str, size: int):          # torch.serialization.skip_data(): not supported here
        raise NotImplementedError("skip_data archives go through torch.save")


def _view(ptr: int, size: int):
    return memoryview((ctypes.c_char * size).from_address(ptr)) if size else memoryview(b"")


def _pwrite_all(fd: int, buf, offset: int):
    mv = memoryview(buf)
    done = 0
    while done < len(mv):
        done += os.pwrite(fd, mv[done:], offset + done)


def _zip64_needed(size: int, offset: int, force: bool) -> bool:
    return force or size >= _U32 or offset >= _U32


def _local_header(name: bytes, crc: int, size: int, extra: bytes, zip64: bool) -> bytes:
    sz = _U32 if zip64 else size
    return struct.pack("<IHHHHHIIIHH", 0x04034B50, 45 if zip64 else 0, 0x0800, 0, 0, 0, crc, sz, sz, len(name), len(extra)) + name + extra


def _central_header(name: bytes, crc: int, size: int, header_off: int, zip64: bool) -> bytes:
    extra = b""
    sz, off = size, header_off
    if zip64:
        extra = struct.pack("<HHQQQ", 0x0001, 24, size, size, header_off)
        sz, off = _U32, _U32
    return struct.pack("<IHHHHHHIIIHHHHHII", 0x02014B50, 45 if zip64 else 0, 45 if zip64 else 0, 0x0800, 0, 0, 0, crc, sz, sz,
                       len(name), len(extra), 0, 0, 0, 0, off) + name + extra


def save(obj: Any, path: str, archive_name: Optional[str] = None, threads: Optional[int] = None, chunk_bytes: int = _CHUNK,
         force_zip64: bool = False, pickle_protocol: int = torch.serialization.DEFAULT_PROTOCOL) -> int:
    """Write ``obj`` to ``path`` as a ``torch.load``-able archive whose root folder is ``archive_name`` (default: the file's
    stem, as ``torch.save`` names it). Tensors must live on the CPU (the checkpointer hands in views of its pinned snapshot
    buffer). Returns the number of bytes written. Raises whatever PyTorch's pickler raises; callers fall back to
    ``torch.save`` on any exception (``atomic_torch_save`` does)."""
    stem = archive_name if archive_name is not None else os.path.splitext(os.path.basename(path))[0]
    col = _Collector()
    torch.serialization._save(obj, col, pickle, pickle_protocol, False)
    recs = col.records
    recs.append(_Record("version", b"3\n", 0, 2, None))
    # one writer + two checksum threads: measured on the B200 host, more threads buy nothing once the page cache throttles
    n_threads = threads if threads is not None else max(2, min(3, (os.cpu_count() or 4)))

    # ---- layout: offsets follow from the sizes alone
    cur = 0
    id_slot = _Record(".data/serialization_id", b"0" * 40, 0, 40, None)      # contents filled in once the CRCs are known
    recs.append(id_slot)
    for r in recs:
        name = f"{stem}/{r.name}".encode()
        r.header_off = cur
        z64 = _zip64_needed(r.size, cur, force_zip64)
        extra = struct.pack("<HHQQ", 0x0001, 16, r.size, r.size) if z64 else b""
        base = cur + 30 + len(name) + len(extra) + 4                        # + the padding field's own header
        pad = (-base) % _ALIGN
        r.extra_local = extra + b"FB" + struct.pack("<H", pad) + b"Z" * pad
        r.data_off = base + pad
        cur = r.data_off + r.size
    cd_off = cur

    fd = os.open(path, os.O_WRONLY | os.O_CREAT | os.O_TRUNC, 0o644)
    try:
        # One writer (this thread) streams the payload in file order: buffered writes to one inode serialise in the kernel anyway,
        # and a single sequential writer keeps them at full speed. The checksums, the other half of the work, run beside it.
        with _fut.ThreadPoolExecutor(max_workers=max(1, n_threads - 1), thread_name_prefix="rtdc-zip-crc") as pool:
            def crc_of(r: _Record):
                r.crc = zlib.crc32(_view(r.ptr, r.size)) & _U32

            big = [r for r in recs if r.data is None and r.size]
            jobs = [pool.submit(crc_of, r) for r in sorted(big, key=lambda r: -r.size)]      # longest first
            for r in big:
                for o in range(0, r.size, chunk_bytes):
                    _pwrite_all(fd, _view(r.ptr + o, min(chunk_bytes, r.size - o)), r.data_off + o)
            for r in recs:
                if r.data is not None:
                    r.crc = zlib.crc32(r.data) & _U32
            for j in jobs:
                j.result()
        # the serialization id: stock archives carry a 40-digit token derived from the record names and checksums
        h = 0
        for r in recs:
            if r is not id_slot:
                h = zlib.crc32(f"{r.name}:{r.crc}:{r.size}".encode(), h)
        id_slot.data = f"{zlib.crc32(stem.encode()) & _U32:020d}{h & _U32:020d}".encode()
        id_slot.crc = zlib.crc32(id_slot.data) & _U32
        central = []
        for r in recs:
            name = f"{stem}/{r.name}".encode()
            z64 = _zip64_needed(r.size, r.header_off, force_zip64)
            _pwrite_all(fd, _local_header(name, r.crc, r.size, r.extra_local, z64), r.header_off)
            if r.data is not None:
                _pwrite_all(fd, r.data, r.data_off)
            central.append(_central_header(name, r.crc, r.size, r.header_off, z64))
        cd = b"".join(central)
        n = len(recs)
        z64_eocd = struct.pack("<IQHHIIQQQQ", 0x06064B50, 44, 45, 45, 0, 0, n, n, len(cd), cd_off)
        locator = struct.pack("<IIQI", 0x07064B50, 0, cd_off + len(cd), 1)
        eocd = struct.pack("<IHHHHIIH", 0x06054B50, 0, 0, min(n, 0xFFFF), min(n, 0xFFFF), min(len(cd), _U32), min(cd_off, _U32), 0)
        tail = cd + z64_eocd + locator + eocd
        _pwrite_all(fd, tail, cd_off)
        total = cd_off + len(tail)
        os.ftruncate(fd, total)
    finally:
        os.close(fd)
    return total


def state_bytes(obj: Any) -> int:
    """Bytes of tensor payload in a nested dict / list / tuple state (decides between this writer and ``torch.save``)."""
    if isinstance(obj, torch.Tensor):
        return obj.numel() * obj.element_size()
    if isinstance(obj, dict):
        return sum(state_bytes(v) for v in obj.values())
    if isinstance(obj, (list, tuple)):
        return sum(state_bytes(v) for v in obj)
    return 0
