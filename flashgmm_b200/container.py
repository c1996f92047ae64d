"""A byte container for what ``net.compress`` returns.

The reference keeps its result as Python objects -- ``strings = [(bytes, abs_max, zero_bitmap), ...,
[z_bytes]]`` plus a ``shape`` dict (/root/reference/compressai/latent_codecs/hyperprior.py:120-128,
entropy_models/entropy_models.py:867) -- and its bits-per-pixel figure counts the payload bytes only
(eval_ckbd.py:137); nothing in it can be written to a file (examples/codec.py carries lists of byte
strings, not tuples). This module defines the file format; ``pack`` / ``unpack`` are exact inverses and
``unpack``'s output is what ``net.decompress(strings, shape)`` takes.

Layout (all integers are LEB128 varints unless sized; little endian)::

    "FGMC"  u8 version (1)  u8 flags (bit 0: payloads are interleaved-rANS "FGIL" containers, whose own
                                      header carries the segment size and the chunk index)
    varint B            images
    varint P            GMM coding passes per image (2 checkerboard, 10 ELIC)
    shape               the ``shape`` object: tagged nested ints  (0x00 int | 0x01 list n | 0x02 dict n (key: varint len + utf8))
    B x  [ varint len, z payload ]                                  hyper-latent stream of image b
    P x B x [ varint abs_max, varint C, ceil(C/8) bitmap bytes (channel c = bit c%8 of byte c//8),
              varint len, payload ]                                 stream of pass p, image b
    u32 crc32 of everything before it

Side information per GMM stream: 1-2 bytes of ``abs_max``, C/8 bytes of bitmap, 1-3 bytes of length --
counted by ``side_info_bytes`` so that a bpp figure can include what the reference's leaves out.
"""
from __future__ import annotations

import zlib
from typing import Any, Dict, List, Sequence, Tuple

import numpy as np
import torch

MAGIC = b"FGMC"
VERSION = 1
FLAG_INTERLEAVED = 1


class ContainerError(ValueError):
    pass


def _varint(n: int) -> bytes:
    if n < 0:
        raise ContainerError("negative integer in container header")
    out = bytearray()
    while True:
        b = n & 0x7F
        n >>= 7
        out.append(b | (0x80 if n else 0))
        if not n:
            return bytes(out)


class _Reader:
    def __init__(self, blob: bytes):
        self.b, self.i = blob, 0

    def take(self, n: int) -> bytes:
        if n < 0 or self.i + n > len(self.b):
            raise ContainerError("container truncated")
        out = self.b[self.i: self.i + n]
        self.i += n
        return out

    def varint(self) -> int:
        shift = val = 0
        while True:
            (byte,) = self.take(1)
            val |= (byte & 0x7F) << shift
            if not byte & 0x80:
                return val
            shift += 7
            if shift > 63:
                raise ContainerError("varint too long")


def _put_shape(obj) -> bytes:
    if isinstance(obj, dict):
        out = b"\x02" + _varint(len(obj))
        for k, v in obj.items():
            kb = str(k).encode()
            out += _varint(len(kb)) + kb + _put_shape(v)
        return out
    if isinstance(obj, (list, tuple, torch.Size)):
        return b"\x01" + _varint(len(obj)) + b"".join(_put_shape(v) for v in obj)
    return b"\x00" + _varint(int(obj))


def _get_shape(r: _Reader):
    (tag,) = r.take(1)
    if tag == 0:
        return r.varint()
    if tag == 1:
        return tuple(_get_shape(r) for _ in range(r.varint()))
    if tag == 2:
        out = {}
        for _ in range(r.varint()):
            k = r.take(r.varint()).decode()
            out[k] = _get_shape(r)
        return out
    raise ContainerError(f"bad shape tag {tag}")


def _bitmap_bytes(bitmap) -> Tuple[int, bytes]:
    bm = np.asarray(torch.as_tensor(bitmap).detach().cpu().numpy() != 0, np.uint8).reshape(-1)
    return bm.size, np.packbits(bm, bitorder="little").tobytes()


def pack(strings: Sequence, shape: Any, interleaved: bool = False) -> bytes:
    """``out = net.compress(x); blob = pack(out["strings"], out["shape"])``. One image (the reference's structure:
    every GMM entry a tuple) or a batch (every entry a list of B tuples)."""
    *gmm, z = strings
    if not isinstance(z, (list, tuple)) or not all(isinstance(s, (bytes, bytearray)) for s in z):
        raise ContainerError("last entry of `strings` must be the list of hyper-latent byte strings")
    B = len(z)
    per_pass: List[List[Tuple]] = []
    for entry in gmm:
        items = [entry] if isinstance(entry, tuple) else list(entry)
        if len(items) != B:
            raise ContainerError(f"a pass carries {len(items)} streams for {B} images")
        per_pass.append(items)
    out = bytearray(MAGIC + bytes([VERSION, FLAG_INTERLEAVED if interleaved else 0]))
    out += _varint(B) + _varint(len(per_pass)) + _put_shape(shape)
    for s in z:
        out += _varint(len(s)) + bytes(s)
    for items in per_pass:
        for payload, abs_max, bitmap in items:
            c, bits = _bitmap_bytes(bitmap)
            out += _varint(int(abs_max)) + _varint(c) + bits + _varint(len(payload)) + bytes(payload)
    out += zlib.crc32(bytes(out)).to_bytes(4, "little")
    return bytes(out)


def unpack(blob: bytes) -> Tuple[List, Any, Dict[str, Any]]:
    """-> (strings, shape, info): ``net.decompress(strings, shape)``. Bitmaps come back as CPU int64 tensors."""
    if len(blob) < 10 or blob[:4] != MAGIC:
        raise ContainerError("not a FGMC container")
    if zlib.crc32(blob[:-4]).to_bytes(4, "little") != blob[-4:]:
        raise ContainerError("container checksum mismatch")
    r = _Reader(blob[:-4])
    r.take(4)
    version, flags = r.take(2)
    if version != VERSION:
        raise ContainerError(f"container version {version} not supported (this build reads version {VERSION})")
    B, P = r.varint(), r.varint()
    shape = _get_shape(r)
    z = [r.take(r.varint()) for _ in range(B)]
    strings: List[Any] = []
    side = 0
    for _ in range(P):
        items = []
        for _b in range(B):
            i0 = r.i
            abs_max, c = r.varint(), r.varint()
            bits = np.frombuffer(r.take((c + 7) // 8), np.uint8)
            bitmap = torch.from_numpy(np.unpackbits(bits, bitorder="little")[:c].astype(np.int64))
            n = r.varint()
            side += r.i - i0
            items.append((r.take(n), abs_max, bitmap))
        strings.append(items[0] if B == 1 else items)
    if r.i != len(r.b):
        raise ContainerError("trailing bytes in container")
    strings.append(z)
    info = {"version": version, "interleaved": bool(flags & FLAG_INTERLEAVED), "images": B, "passes": P,
            "side_info_bytes": side, "payload_bytes": sum(len(s[0]) for e in strings[:-1] for s in ([e] if isinstance(e, tuple) else e)),
            "hyper_bytes": sum(len(s) for s in z), "total_bytes": len(blob)}
    return strings, shape, info
