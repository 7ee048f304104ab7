"""Host-side output targets of fdtd_b200/io.py: GIF container round trip, PNG validity."""
import struct
import zlib

import numpy as np

from fdtd_b200 import io as fio


def _frames(n, w, h, seed=0):
    rng = np.random.default_rng(seed)
    out = []
    for k in range(n):
        img = np.zeros((h, w, 4), dtype=np.uint8)
        img[..., 0] = (np.arange(w)[None, :] * 7 + k * 30) % 256
        img[..., 1] = (np.arange(h)[:, None] * 5) % 256
        img[..., 2] = rng.integers(0, 256, (h, w))        # noise: exercises LZW table resets on larger frames
        img[..., 3] = 255
        out.append(img)
    return out


def test_gif_round_trip(tmp_path):
    for (w, h, n) in ((33, 17, 3), (160, 120, 2), (1, 1, 1)):
        path = str(tmp_path / f"a{w}.gif")
        frames = _frames(n, w, h)
        with fio.GifEncoder(path, (w, h), frame_delay_ms=50) as enc:
            assert enc.size() == (w, h)
            for img in frames:
                enc.with_image_buffer(lambda buf, img=img: buf.__setitem__(Ellipsis, img))
            assert enc.frames == n
        got = fio.decode_gif(path)
        assert len(got) == n
        pal = fio._cube_palette()
        for g, img in zip(got, frames):
            assert np.array_equal(g, pal[fio._quantise(img)])                 # lossless after palette quantisation
            assert np.abs(g.astype(int) - img[..., :3].astype(int)).max() <= 26   # cube spacing 51 / 42.5
        raw = open(path, "rb").read()
        assert raw[:6] == b"GIF89a" and raw[-1] == 0x3B and struct.unpack("<HH", raw[6:10]) == (w, h)


def test_png_sequence(tmp_path):
    seq = fio.PngSequence(str(tmp_path / "f{:03d}.png"), (20, 10))
    frames = _frames(2, 20, 10)
    for img in frames:
        seq.with_image_buffer(lambda buf, img=img: buf.__setitem__(Ellipsis, img))
    for k, img in enumerate(frames):
        raw = open(tmp_path / f"f{k:03d}.png", "rb").read()
        assert raw[:8] == b"\x89PNG\r\n\x1a\n"
        pos, idat, ihdr = 8, b"", None
        while pos < len(raw):
            n, tag = struct.unpack(">I4s", raw[pos:pos + 8])
            body = raw[pos + 8:pos + 8 + n]
            assert struct.unpack(">I", raw[pos + 8 + n:pos + 12 + n])[0] == zlib.crc32(tag + body) & 0xFFFFFFFF
            if tag == b"IHDR":
                ihdr = struct.unpack(">IIBBBBB", body)
            if tag == b"IDAT":
                idat += body
            pos += 12 + n
        assert ihdr == (20, 10, 8, 6, 0, 0, 0)
        rows = np.frombuffer(zlib.decompress(idat), dtype=np.uint8).reshape(10, 1 + 80)
        assert not rows[:, 0].any() and np.array_equal(rows[:, 1:].reshape(10, 20, 4), img)
