"""Minimal headless raster canvas + PNG encoder (zlib only): matplotlib is not in the image and there is
no display stack, so charts are drawn straight into a numpy RGB buffer and written as PNG files.
Replaces the role of matplotlib in the reference's chart scripts (reference:
scripts/visualize_results.py:77-276, baseline/baseline_training.py:112-147) -- files only, no window."""
from __future__ import annotations

import struct
import zlib
from typing import Tuple

import numpy as np

Color = Tuple[int, int, int]

# 5x7 bitmap font (digits, upper-case letters and a few symbols), column-major bits
_FONT = {
    "0": "3E5149453E", "1": "00427F4000", "2": "4261514946", "3": "2141454B31", "4": "1814127F10",
    "5": "2745454539", "6": "3C4A494930", "7": "0171090503", "8": "3649494936", "9": "064949291E",
    "A": "7E1111117E", "B": "7F49494936", "C": "3E41414122", "D": "7F4141221C", "E": "7F49494941",
    "F": "7F09090901", "G": "3E4149497A", "H": "7F0808087F", "I": "00417F4100", "J": "2040413F01",
    "K": "7F08142241", "L": "7F40404040", "M": "7F020C027F", "N": "7F0408107F", "O": "3E4141413E",
    "P": "7F09090906", "Q": "3E4151215E", "R": "7F09192946", "S": "4649494931", "T": "01017F0101",
    "U": "3F4040403F", "V": "1F2040201F", "W": "3F4038403F", "X": "6314081463", "Y": "0708700807",
    "Z": "6151494543", " ": "0000000000", ".": "0060600000", "-": "0808080808", "/": "2010080402",
    "%": "2313086462", ":": "0036360000", "(": "001C224100", ")": "0041221C00", "=": "1414141414",
    "+": "08083E0808", ",": "00A0600000", "_": "4040404040", "X2": "0000000000",
}


class Canvas:
    def __init__(self, width: int, height: int, bg: Color = (255, 255, 255)):
        self.w, self.h = width, height
        self.px = np.empty((height, width, 3), dtype=np.uint8)
        self.px[:] = bg

    def rect(self, x0: int, y0: int, x1: int, y1: int, color: Color) -> None:
        x0, x1 = sorted((max(0, x0), min(self.w, x1)))
        y0, y1 = sorted((max(0, y0), min(self.h, y1)))
        self.px[y0:y1, x0:x1] = color

    def line(self, x0: float, y0: float, x1: float, y1: float, color: Color, width: int = 1) -> None:
        n = int(max(abs(x1 - x0), abs(y1 - y0))) + 1
        xs = np.linspace(x0, x1, n).round().astype(int)
        ys = np.linspace(y0, y1, n).round().astype(int)
        for dx in range(-(width // 2), width - width // 2):
            for dy in range(-(width // 2), width - width // 2):
                xx, yy = xs + dx, ys + dy
                ok = (xx >= 0) & (xx < self.w) & (yy >= 0) & (yy < self.h)
                self.px[yy[ok], xx[ok]] = color

    def marker(self, x: float, y: float, color: Color, r: int = 3) -> None:
        self.rect(int(x) - r, int(y) - r, int(x) + r + 1, int(y) + r + 1, color)

    def text(self, x: int, y: int, s: str, color: Color = (0, 0, 0), scale: int = 1) -> None:
        cx = x
        for ch in s.upper():
            cols = _FONT.get(ch, _FONT[" "])
            for ci in range(5):
                bits = int(cols[2 * ci: 2 * ci + 2], 16)
                for bi in range(8):
                    if bits >> bi & 1:
                        self.rect(cx + ci * scale, y + bi * scale, cx + (ci + 1) * scale, y + (bi + 1) * scale, color)
            cx += 6 * scale

    def save(self, path: str) -> None:
        raw = b"".join(b"\x00" + self.px[r].tobytes() for r in range(self.h))

        def chunk(tag: bytes, data: bytes) -> bytes:
            return struct.pack(">I", len(data)) + tag + data + struct.pack(">I", zlib.crc32(tag + data) & 0xFFFFFFFF)

        png = b"\x89PNG\r\n\x1a\n" + chunk(b"IHDR", struct.pack(">IIBBBBB", self.w, self.h, 8, 2, 0, 0, 0)) \
            + chunk(b"IDAT", zlib.compress(raw, 6)) + chunk(b"IEND", b"")
        with open(path, "wb") as f:
            f.write(png)


def read_png_size(path: str) -> Tuple[int, int]:
    with open(path, "rb") as f:
        head = f.read(24)
    assert head[:8] == b"\x89PNG\r\n\x1a\n"
    return struct.unpack(">II", head[16:24])
