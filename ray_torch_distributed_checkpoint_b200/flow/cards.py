"""Card components (``metaflow.cards`` surface used at ``/root/reference/eval_flow.py:15,98-139``):
``Markdown``, ``Table``, ``Image``, ``Artifact``; a card renders to a standalone HTML file in the task
directory. Images are PNG data URIs written by a tiny pure-Python encoder (no matplotlib / graphics
stack needed); ``Image.from_matplotlib`` works when matplotlib happens to be installed."""
from __future__ import annotations

import base64
import html
import io
import struct
import zlib
from typing import Any, List, Optional, Sequence

import numpy as np


def _png_bytes(arr: np.ndarray) -> bytes:
    """Encode a uint8 array [H,W] (gray) or [H,W,3] (RGB) as PNG."""
    a = np.asarray(arr)
    if a.dtype != np.uint8:
        lo, hi = float(a.min()), float(a.max())
        a = ((a - lo) / (hi - lo + 1e-12) * 255.0).astype(np.uint8)
    if a.ndim == 2:
        color, rows = 0, a
    elif a.ndim == 3 and a.shape[2] == 3:
        color, rows = 2, a.reshape(a.shape[0], -1)
    else:
        raise ValueError(f"unsupported image shape {a.shape}")
    h, w = a.shape[0], a.shape[1]
    raw = b"".join(b"\x00" + rows[i].tobytes() for i in range(h))

    def chunk(tag: bytes, data: bytes) -> bytes:
        return struct.pack(">I", len(data)) + tag + data + struct.pack(">I", zlib.crc32(tag + data) & 0xFFFFFFFF)

    return (b"\x89PNG\r\n\x1a\n" + chunk(b"IHDR", struct.pack(">IIBBBBB", w, h, 8, color, 0, 0, 0))
            + chunk(b"IDAT", zlib.compress(raw, 6)) + chunk(b"IEND", b""))


class _Component:
    def render(self) -> str:
        raise NotImplementedError


class Markdown(_Component):
    def __init__(self, text: str = ""):
        self.text = text

    def render(self) -> str:
        out = []
        for line in self.text.splitlines() or [""]:
            level = len(line) - len(line.lstrip("#"))
            if 0 < level <= 6 and line[level:level + 1] == " ":
                out.append(f"<h{level}>{html.escape(line[level + 1:])}</h{level}>")
            else:
                out.append(f"<p>{html.escape(line)}</p>")
        return "\n".join(out)


class Artifact(_Component):
    def __init__(self, obj: Any, name: Optional[str] = None):
        self.obj, self.name = obj, name

    def render(self) -> str:
        label = f"<b>{html.escape(self.name)}</b>: " if self.name else ""
        return f"<pre>{label}{html.escape(repr(self.obj))}</pre>"


class Image(_Component):
    def __init__(self, png: bytes, label: Optional[str] = None, width: Optional[int] = None, svg: Optional[str] = None):
        self.png, self.label, self.width, self.svg = png, label, width, svg

    @classmethod
    def from_array(cls, arr: np.ndarray, label: Optional[str] = None, width: Optional[int] = 112) -> "Image":
        return cls(_png_bytes(arr), label, width)

    @classmethod
    def from_matplotlib(cls, fig, label: Optional[str] = None) -> "Image":
        buf = io.BytesIO()
        fig.savefig(buf, format="png", bbox_inches="tight")
        return cls(buf.getvalue(), label)

    @classmethod
    def bar_chart(cls, categories: Sequence[str], values: Sequence[float], title: str = "", width: int = 360) -> "Image":
        """Horizontal bar chart as inline SVG (the logits panel of the reference's error-analysis card)."""
        vals = [float(v) for v in values]
        vmax = max(max(vals, default=1.0), 1e-9)
        row_h, left = 18, 90
        h = row_h * len(vals) + 24
        parts = [f'<svg xmlns="http://www.w3.org/2000/svg" width="{width}" height="{h}" font-family="sans-serif" font-size="11">']
        if title:
            parts.append(f'<text x="{left}" y="12" font-weight="bold">{html.escape(title)}</text>')
        for i, (c, v) in enumerate(zip(categories, vals)):
            y = 20 + i * row_h
            bw = max(0.0, v) / vmax * (width - left - 50)
            parts.append(f'<text x="0" y="{y + 11}">{html.escape(str(c))}</text>')
            parts.append(f'<rect x="{left}" y="{y}" width="{bw:.1f}" height="{row_h - 4}" fill="#4878a8"/>')
            parts.append(f'<text x="{left + bw + 3:.1f}" y="{y + 11}">{v:.2f}</text>')
        parts.append("</svg>")
        return cls(b"", None, width, svg="".join(parts))

    def render(self) -> str:
        if self.svg is not None:
            return self.svg
        w = f' width="{self.width}"' if self.width else ""
        cap = f"<figcaption>{html.escape(self.label)}</figcaption>" if self.label else ""
        return f'<figure><img{w} style="image-rendering:pixelated" src="data:image/png;base64,{base64.b64encode(self.png).decode()}"/>{cap}</figure>'


class Table(_Component):
    def __init__(self, data: Optional[List[List[Any]]] = None, headers: Optional[List[str]] = None):
        self.data, self.headers = data or [], headers or []

    def render(self) -> str:
        def cell(c):
            return c.render() if isinstance(c, _Component) else html.escape(str(c))
        head = "".join(f"<th>{html.escape(str(h))}</th>" for h in self.headers)
        rows = "\n".join("<tr>" + "".join(f"<td>{cell(c)}</td>" for c in row) + "</tr>" for row in self.data)
        return f'<table border="1" cellspacing="0" cellpadding="4"><thead><tr>{head}</tr></thead><tbody>{rows}</tbody></table>'


class Card:
    def __init__(self, card_id: str, card_type: str = "blank"):
        self.id, self.type = card_id, card_type
        self.components: List[_Component] = []

    def append(self, component: _Component):
        self.components.append(component)

    def extend(self, components):
        self.components.extend(components)

    def render(self, title: str = "") -> str:
        body = "\n".join(c.render() for c in self.components)
        return (f"<!doctype html><html><head><meta charset='utf-8'><title>{html.escape(title)}</title>"
                "<style>body{font-family:sans-serif;margin:1.5em} td{vertical-align:middle}</style></head>"
                f"<body><h2>{html.escape(title)}</h2>\n{body}\n</body></html>")


class CardCollection:
    """``current.card['id']`` / ``current.card.append(...)`` (default card)."""

    def __init__(self):
        self._cards = {}

    def _get(self, card_id: str) -> Card:
        if card_id not in self._cards:
            self._cards[card_id] = Card(card_id)
        return self._cards[card_id]

    def __getitem__(self, card_id: str) -> Card:
        return self._get(card_id)

    def append(self, component):
        self._get("default").append(component)

    def items(self):
        return self._cards.items()
