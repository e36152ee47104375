"""Texture / HDRI asset files for the seeded scenes.

The reference decodes textures with the stb_image it vendors (texture.hpp:23,:31).  So that both
sides of a parity test see identical texels, every synthetic texture is written twice from the same
numpy array: once in a format the reference can open (lossless 8-bit PNG, or Radiance RGBE .hdr) and
once as the raw ".zbtex" dump the host's image_texture prefers ("ZBTX", u32 width, u32 height,
u32 is_float, then RGB8 / RGB32F rows, top row first).  For .hdr the float texels are the values an
RGBE decoder reproduces: mantissa * 2**(E - 136) (stb_image.h stbi__hdr_convert).
"""
from __future__ import annotations

import struct
import zlib

import numpy as np


def zbtex_bytes(arr: np.ndarray) -> bytes:
    """(H, W, 3) uint8 or float32 -> .zbtex file contents."""
    assert arr.ndim == 3 and arr.shape[2] == 3
    is_float = arr.dtype != np.uint8
    body = np.ascontiguousarray(arr, np.float32 if is_float else np.uint8).tobytes()
    return b"ZBTX" + struct.pack("<III", arr.shape[1], arr.shape[0], int(is_float)) + body


def read_zbtex(data: bytes) -> np.ndarray:
    assert data[:4] == b"ZBTX"
    w, h, is_float = struct.unpack("<III", data[4:16])
    dt = np.float32 if is_float else np.uint8
    return np.frombuffer(data[16:], dt, w * h * 3).reshape(h, w, 3)


def png_bytes(arr: np.ndarray) -> bytes:
    """(H, W, 3) uint8 -> 8-bit RGB PNG (filter 0, zlib level 6); stb_image decodes it losslessly."""
    assert arr.dtype == np.uint8 and arr.ndim == 3 and arr.shape[2] == 3
    h, w, _ = arr.shape
    raw = b"".join(b"\x00" + arr[y].tobytes() for y in range(h))

    def chunk(tag, payload):
        c = tag + payload
        return struct.pack(">I", len(payload)) + c + struct.pack(">I", zlib.crc32(c) & 0xFFFFFFFF)

    return (b"\x89PNG\r\n\x1a\n" + chunk(b"IHDR", struct.pack(">IIBBBBB", w, h, 8, 2, 0, 0, 0))
            + chunk(b"IDAT", zlib.compress(raw, 6)) + chunk(b"IEND", b""))


def rgbe_encode(rgb: np.ndarray) -> np.ndarray:
    """(H, W, 3) float -> (H, W, 4) uint8 RGBE (shared exponent, truncating mantissas)."""
    rgb = np.maximum(np.asarray(rgb, np.float64), 0.0)
    m = rgb.max(axis=2)
    frac, exp = np.frexp(m)  # m = frac * 2**exp, frac in [0.5, 1)
    scale = np.where(m > 1e-32, 256.0 / np.ldexp(1.0, exp), 0.0)
    out = np.zeros(rgb.shape[:2] + (4,), np.uint8)
    out[..., :3] = np.clip(np.floor(rgb * scale[..., None]), 0, 255).astype(np.uint8)
    out[..., 3] = np.where(m > 1e-32, exp + 128, 0).astype(np.uint8)
    return out


def rgbe_decode(rgbe: np.ndarray) -> np.ndarray:
    """What stb_image's HDR loader returns for these bytes: mantissa * 2**(E - 136), 0 when E == 0."""
    e = rgbe[..., 3].astype(np.int32)
    scale = np.where(e != 0, np.ldexp(np.float32(1.0), e - 136), np.float32(0.0)).astype(np.float32)
    return (rgbe[..., :3].astype(np.float32) * scale[..., None]).astype(np.float32)


def hdr_bytes(rgbe: np.ndarray) -> bytes:
    """(H, W, 4) uint8 RGBE -> Radiance .hdr with flat (non-RLE) scanlines."""
    h, w, _ = rgbe.shape
    first = rgbe[0, 0]
    assert not (first[0] == 2 and first[1] == 2 and first[2] < 128), "first pixel would read as an RLE marker"
    head = b"#?RADIANCE\nFORMAT=32-bit_rle_rgbe\n\n" + f"-Y {h} +X {w}\n".encode()
    return head + np.ascontiguousarray(rgbe).tobytes()


def ldr_pair(name: str, arr: np.ndarray) -> dict:
    """files for an 8-bit texture: NAME (PNG, for the reference) + NAME.zbtex (for the host)."""
    return {name: png_bytes(arr), name + ".zbtex": zbtex_bytes(arr)}


def hdr_pair(name: str, rgb: np.ndarray) -> dict:
    """files for a float texture / HDRI: NAME (.hdr RGBE) + NAME.zbtex holding the decoded floats."""
    rgbe = rgbe_encode(rgb)
    return {name: hdr_bytes(rgbe), name + ".zbtex": zbtex_bytes(rgbe_decode(rgbe))}
