"""Animated output and state I/O, host side only.

* `GifEncoder` — the reference's second image target (cem-solver/src/project.rs:196-232): every projection pass appends
  one frame of `frame_size` with `frame_delay` to an animated GIF (the application wires it to `Observer::write_to_gif`,
  cem-app/src/solver/observer.rs:51-57).  The reference delegates the encoding to the `image` crate (NeuQuant palette);
  here a small GIF89a writer with a fixed 6×7×6 colour cube + LZW — the container is standard, the palette choice is not
  part of the solver path (parity unpinned: the reference has no test for its GIF bytes).
* `PngSequence` — the same protocol writing numbered PNG files (zlib from the standard library).
* `save_checkpoint` / `load_checkpoint` — raw SoA field planes + tick/time of a state; the reference has no checkpointing
  (SURVEY §5), this exists so long runs of the B200 engine can be resumed bit-exactly.
"""
from __future__ import annotations

import os
import struct
import zlib
from typing import Callable, Optional

import numpy as np

from .solver import FieldComponent


# ---- GIF89a ---------------------------------------------------------------------------------------------------------------
def _cube_palette() -> np.ndarray:
    r, g, b = np.meshgrid(np.arange(6), np.arange(7), np.arange(6), indexing="ij")
    pal = np.stack([r.ravel() * 51, np.round(g.ravel() * 42.5), b.ravel() * 51], axis=1).astype(np.uint8)   # 252 colours
    return np.concatenate([pal, np.zeros((4, 3), dtype=np.uint8)])


def _quantise(rgba: np.ndarray) -> np.ndarray:
    """Nearest colour of the 6×7×6 cube (index = (r*7 + g)*6 + b)."""
    r = (rgba[..., 0].astype(np.uint16) + 25) // 51
    g = np.round(rgba[..., 1].astype(np.float32) / 42.5).astype(np.uint16)
    b = (rgba[..., 2].astype(np.uint16) + 25) // 51
    return ((r * 7 + g) * 6 + b).astype(np.uint8)


def _lzw(indices: bytes, min_code_size: int = 8) -> bytes:
    """GIF variable-width LZW, codes packed LSB first."""
    clear, end = 1 << min_code_size, (1 << min_code_size) + 1
    out = bytearray()
    acc = nbits = 0

    def emit(code: int, width: int) -> None:
        nonlocal acc, nbits
        acc |= code << nbits
        nbits += width
        while nbits >= 8:
            out.append(acc & 0xFF)
            acc >>= 8
            nbits -= 8

    table: dict = {}
    width, nxt = min_code_size + 1, end + 1
    emit(clear, width)
    prefix = -1
    for byte in indices:
        if prefix < 0:
            prefix = byte
            continue
        key = (prefix << 8) | byte
        code = table.get(key)
        if code is not None:
            prefix = code
            continue
        emit(prefix, width)
        if nxt < 4096:
            table[key] = nxt
            if nxt == (1 << width) and width < 12:
                width += 1
            nxt += 1
        else:
            emit(clear, width)
            table.clear()
            width, nxt = min_code_size + 1, end + 1
        prefix = byte
    if prefix >= 0:
        emit(prefix, width)
    emit(end, width)
    if nbits:
        out.append(acc & 0xFF)
    return bytes(out)


class GifEncoder:
    """`GifEncoder<Writer>` (project.rs:196-232): an image target; each `with_image_buffer` call becomes one frame."""

    def __init__(self, path: str, frame_size: tuple, frame_delay_ms: int = 40):
        self.path = path
        self.frame_size = (int(frame_size[0]), int(frame_size[1]))
        self.delay_cs = max(int(round(frame_delay_ms / 10.0)), 0)
        self.frames = 0
        self._f = open(path, "wb")
        w, h = self.frame_size
        self._f.write(b"GIF89a" + struct.pack("<HHBBB", w, h, 0xF7, 0, 0) + _cube_palette().tobytes())
        self._f.write(b"\x21\xFF\x0BNETSCAPE2.0\x03\x01\x00\x00\x00")   # loop forever

    def size(self) -> tuple:
        return self.frame_size

    def with_image_buffer(self, f: Callable[[np.ndarray], None]) -> None:
        w, h = self.frame_size
        buffer = np.zeros((h, w, 4), dtype=np.uint8)   # a fresh buffer per frame, like the reference (project.rs:223-225)
        f(buffer)
        data = _lzw(_quantise(buffer).tobytes())
        self._f.write(b"\x21\xF9\x04\x04" + struct.pack("<H", self.delay_cs) + b"\x00\x00")
        self._f.write(b"\x2C" + struct.pack("<HHHHB", 0, 0, w, h, 0) + b"\x08")
        for i in range(0, len(data), 255):
            chunk = data[i:i + 255]
            self._f.write(bytes([len(chunk)]) + chunk)
        self._f.write(b"\x00")
        self.frames += 1

    def close(self) -> None:
        if self._f is not None:
            self._f.write(b"\x3B")
            self._f.close()
            self._f = None

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()


def decode_gif(path: str) -> list:
    """Minimal reader of what GifEncoder writes (tests): list of [h, w, 3] uint8 frames."""
    raw = open(path, "rb").read()
    assert raw[:6] == b"GIF89a"
    w, h, flags = struct.unpack("<HHB", raw[6:11])
    pos = 13
    pal = np.frombuffer(raw[pos:pos + 3 * (2 << (flags & 7))], dtype=np.uint8).reshape(-1, 3)
    pos += pal.size
    frames = []
    while raw[pos] != 0x3B:
        if raw[pos] == 0x21:                      # extension: skip its sub-blocks
            pos += 2
            while raw[pos]:
                pos += raw[pos] + 1
            pos += 1
            continue
        assert raw[pos] == 0x2C
        fw, fh = struct.unpack("<HH", raw[pos + 5:pos + 9])
        mcs = raw[pos + 10]
        pos += 11
        data = bytearray()
        while raw[pos]:
            data += raw[pos + 1:pos + 1 + raw[pos]]
            pos += raw[pos] + 1
        pos += 1
        clear, end = 1 << mcs, (1 << mcs) + 1
        table = {i: bytes([i]) for i in range(clear)}
        width, nxt, acc, nbits, prev, out = mcs + 1, end + 1, 0, 0, None, bytearray()
        for byte in data:
            acc |= byte << nbits
            nbits += 8
            while nbits >= width:
                code = acc & ((1 << width) - 1)
                acc >>= width
                nbits -= width
                if code == clear:
                    table = {i: bytes([i]) for i in range(clear)}
                    width, nxt, prev = mcs + 1, end + 1, None
                    continue
                if code == end:
                    break
                entry = table[code] if code in table else prev + prev[:1]
                out += entry
                if prev is not None and nxt < 4096:
                    table[nxt] = prev + entry[:1]
                    nxt += 1
                    if nxt == (1 << width) and width < 12:
                        width += 1
                prev = entry
        frames.append(pal[np.frombuffer(bytes(out[:fw * fh]), dtype=np.uint8)].reshape(fh, fw, 3))
    return frames


# ---- PNG sequence -----------------------------------------------------------------------------------------------------------
def write_png(path: str, rgba: np.ndarray) -> None:
    h, w = rgba.shape[:2]

    def chunk(tag: bytes, body: bytes) -> bytes:
        return struct.pack(">I", len(body)) + tag + body + struct.pack(">I", zlib.crc32(tag + body) & 0xFFFFFFFF)

    rows = np.concatenate([np.zeros((h, 1), dtype=np.uint8), np.ascontiguousarray(rgba, dtype=np.uint8).reshape(h, w * 4)], axis=1)
    with open(path, "wb") as f:
        f.write(b"\x89PNG\r\n\x1a\n" + chunk(b"IHDR", struct.pack(">IIBBBBB", w, h, 8, 6, 0, 0, 0)) +
                chunk(b"IDAT", zlib.compress(rows.tobytes(), 6)) + chunk(b"IEND", b""))


class PngSequence:
    """Image target writing `pattern.format(frame_index)` per projection pass."""

    def __init__(self, pattern: str, frame_size: tuple):
        self.pattern, self.frame_size, self.frames = pattern, (int(frame_size[0]), int(frame_size[1])), 0

    def size(self) -> tuple:
        return self.frame_size

    def with_image_buffer(self, f: Callable[[np.ndarray], None]) -> None:
        buffer = np.zeros((self.frame_size[1], self.frame_size[0], 4), dtype=np.uint8)
        f(buffer)
        write_png(self.pattern.format(self.frames), buffer)
        self.frames += 1


# ---- checkpoints ------------------------------------------------------------------------------------------------------------
_MAGIC_V1 = b"FDTDB200CKPT\x01\x00\x00\x00"
_MAGIC = b"FDTDB200CKPT\x02\x00\x00\x00"
_HEADER = "<6QQdQ"   # nx ny nz z_begin z_end n_psi | tick | time | row pitch of the ψ blob


def save_checkpoint(state, path: str, planes_per_chunk: int = 16) -> None:
    """Header + E planes + H planes ([z][y][x][3] f32, the caller's nx-wide lattice) of this handle's slab, streamed
    `planes_per_chunk` at a time, then the CPML ψ state.  The ψ blob is the handle's own (row-pitched) layout, so the header
    records the pitch and `load_checkpoint` refuses a handle that stores its rows differently."""
    nx, ny, _ = state.lattice
    zb, ze = state.z_begin, state.z_end
    state.sync()
    tmp = path + ".tmp"
    with open(tmp, "wb") as f:
        psi = state.read_pml_state()   # CPML ψ arrays (empty without a PML)
        f.write(_MAGIC + struct.pack(_HEADER, nx, ny, state.lattice[2], zb, ze, psi.size, state.tick(), state.time(), state.row_pitch))
        for comp in (FieldComponent.E, FieldComponent.H):
            for z in range(zb, ze, planes_per_chunk):
                z1 = min(z + planes_per_chunk, ze)
                f.write(state.read_field(comp, (0, 0, z), (nx, ny, z1)).tobytes())
        f.write(psi.tobytes())
    os.replace(tmp, path)


def load_checkpoint(state, path: str, planes_per_chunk: int = 16) -> None:
    """Restores fields, tick and time into a state created for the same lattice/slab (materials, PML profiles, sources and probes
    belong to the instance and are not part of the checkpoint; the PML's ψ state is).  Everything that can be wrong with the file
    — magic, lattice, slab, ψ count, ψ row pitch, length — is checked BEFORE the first byte is written into the state, so a
    mismatch or a truncated file leaves the state as it was.  Slab-decomposed runs refresh their halos afterwards."""
    nx, ny, nz = state.lattice
    size = os.path.getsize(path)
    with open(path, "rb") as f:
        magic = f.read(len(_MAGIC))
        if magic not in (_MAGIC, _MAGIC_V1):
            raise ValueError(f"{path}: not a fdtd_b200 checkpoint")
        if magic == _MAGIC:
            hx, hy, hz, zb, ze, n_psi, tick, time, pitch = struct.unpack(_HEADER, f.read(struct.calcsize(_HEADER)))
        else:   # version 1 had no pitch field: its ψ blob is only valid for a handle with unpadded rows
            hx, hy, hz, zb, ze, n_psi, tick, time = struct.unpack("<6QQd", f.read(64))
            pitch = hx
        if (hx, hy, hz, zb, ze) != (nx, ny, nz, state.z_begin, state.z_end):
            raise ValueError(f"{path}: checkpoint of lattice {(hx, hy, hz)} slab {(zb, ze)} does not match this state")
        have_psi = state.read_pml_state().size
        if n_psi != have_psi:
            raise ValueError(f"{path}: checkpoint carries {n_psi} ψ values, this state's PML has {have_psi}")
        if n_psi and pitch != state.row_pitch:
            raise ValueError(f"{path}: the ψ state was saved with rows of {pitch} cells, this handle stores rows of {state.row_pitch} "
                             "(row padding / FDTD_B200_NO_ROW_PADDING / forced generic kernels differ between the two handles)")
        body = 2 * 3 * 4 * nx * ny * (ze - zb) + 4 * n_psi
        if size != f.tell() + body:
            raise ValueError(f"{path}: truncated or oversized checkpoint ({size} bytes, expected {f.tell() + body})")
        for comp in (FieldComponent.E, FieldComponent.H):
            for z in range(zb, ze, planes_per_chunk):
                z1 = min(z + planes_per_chunk, ze)
                n = (z1 - z) * ny * nx * 3
                buf = np.frombuffer(f.read(4 * n), dtype=np.float32)
                if buf.size != n:
                    raise ValueError(f"{path}: truncated checkpoint")
                state.write_field(comp, buf, (0, 0, z), (nx, ny, z1))
        if n_psi:
            state.write_pml_state(np.frombuffer(f.read(4 * n_psi), dtype=np.float32))
    state.set_time(int(tick), float(time))
