#!/usr/bin/env python
"""ppm2png.py IN.ppm OUT.png [DOWNSCALE] — box-downscale a binary PPM (the driver's --format ppm) and write a PNG
(zlib + struct only; no imaging library in this image).  For eyeballing frames brought back through gpurun_out/."""
import struct
import sys
import zlib

import numpy as np


def main():
    src, dst = sys.argv[1], sys.argv[2]
    k = int(sys.argv[3]) if len(sys.argv) > 3 else 1
    raw = open(src, "rb").read()
    parts = raw.split(b"\n", 3)
    assert parts[0] == b"P6"
    w, h = map(int, parts[1].split())
    img = np.frombuffer(parts[3], np.uint8, w * h * 3).reshape(h, w, 3)
    if k > 1:
        img = img[: h // k * k, : w // k * k].reshape(h // k, k, w // k, k, 3).astype(np.float32).mean(axis=(1, 3)).round().astype(np.uint8)
    h, w = img.shape[:2]
    rows = b"".join(b"\0" + img[y].tobytes() for y in range(h))

    def chunk(tag, data):
        return struct.pack(">I", len(data)) + tag + data + struct.pack(">I", zlib.crc32(tag + data) & 0xFFFFFFFF)

    open(dst, "wb").write(b"\x89PNG\r\n\x1a\n" + chunk(b"IHDR", struct.pack(">IIBBBBB", w, h, 8, 2, 0, 0, 0)) +
                          chunk(b"IDAT", zlib.compress(rows, 6)) + chunk(b"IEND", b""))


if __name__ == "__main__":
    main()
