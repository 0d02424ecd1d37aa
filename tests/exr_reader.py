"""TEST INFRASTRUCTURE: minimal OpenEXR scan-line reader (NO_COMPRESSION / ZIPS / ZIP, HALF / FLOAT / UINT
channels) for the reference's golden renders under resources/sunsky/test_data/renders/ -- cv2 cannot open the
ten-band spectral files. Only `tests/gen_reference_fixtures.py` needs it at generation time;
the tests read the committed npz.

File layout (OpenEXR 2 single-part scan-line): magic 20000630, version/flags, attributes
(name\\0 type\\0 size value) up to an empty name, an offset table with one uint64 per chunk, chunks of
(y:int32, size:int32, payload). ZIP payloads are zlib streams of the block's bytes after the byte-delta predictor
and the even/odd interleave; inside a block every scan line stores its channels one after the other in
alphabetical channel order."""
import struct
import zlib

import numpy as np

_PIXEL = {0: (np.uint32, 4), 1: (np.float16, 2), 2: (np.float32, 4)}
_LINES = {0: 1, 2: 1, 3: 16}                # compression -> scan lines per chunk


def _cstr(buf, pos):
    end = buf.index(b"\0", pos)
    return buf[pos:end].decode("latin1"), end + 1


def _undo_zip(raw):
    data = np.frombuffer(zlib.decompress(raw), dtype=np.uint8)
    # predictor: d[i] = d[i-1] + t[i] - 128 (mod 256)
    t = (np.cumsum(np.concatenate([data[:1].astype(np.int64), data[1:].astype(np.int64) - 128])) & 0xFF).astype(np.uint8)
    half = (t.size + 1) // 2
    out = np.empty_like(t)
    out[0::2] = t[:half]
    out[1::2] = t[half:]
    return out.tobytes()


def read_exr(path):
    """-> (dict channel name -> float32 array (height, width), [channel names in file order])"""
    buf = open(path, "rb").read()
    magic, version = struct.unpack_from("<II", buf, 0)
    if magic != 20000630:
        raise ValueError(f"{path}: not an OpenEXR file")
    if version & 0x1E00:
        raise ValueError(f"{path}: tiled / multi-part / deep files are not supported")
    pos = 8
    attrs = {}
    while True:
        name, pos = _cstr(buf, pos)
        if not name:
            break
        typ, pos = _cstr(buf, pos)
        (size,) = struct.unpack_from("<i", buf, pos)
        pos += 4
        attrs[name] = (typ, buf[pos:pos + size])
        pos += size
    channels = []
    cb = attrs["channels"][1]
    p = 0
    while cb[p] != 0:
        cname, p = _cstr(cb, p)
        ptype, _plinear, xs, ys = struct.unpack_from("<iB3xii", cb, p)
        p += 16
        if xs != 1 or ys != 1:
            raise ValueError("subsampled channels are not supported")
        channels.append((cname, ptype))
    compression = attrs["compression"][1][0]
    if compression not in _LINES:
        raise ValueError(f"{path}: compression {compression} is not supported")
    x0, y0, x1, y1 = struct.unpack("<4i", attrs["dataWindow"][1])
    width, height = x1 - x0 + 1, y1 - y0 + 1
    lines = _LINES[compression]
    n_chunks = (height + lines - 1) // lines
    offsets = struct.unpack_from(f"<{n_chunks}Q", buf, pos)
    out = {c: np.zeros((height, width), np.float32) for c, _ in channels}
    line_bytes = sum(_PIXEL[t][1] for _, t in channels) * width
    for off in offsets:
        y, size = struct.unpack_from("<ii", buf, off)
        raw = buf[off + 8:off + 8 + size]
        n_lines = min(lines, y1 - y + 1)
        want = line_bytes * n_lines
        data = raw if (compression == 0 or size == want) else _undo_zip(raw)
        if len(data) != want:
            raise ValueError(f"{path}: chunk at y={y} has {len(data)} bytes, expected {want}")
        q = 0
        for ly in range(n_lines):
            for cname, ptype in channels:                       # already alphabetical in a valid file
                dt, sz = _PIXEL[ptype]
                out[cname][y - y0 + ly] = np.frombuffer(data, dtype=dt, count=width, offset=q).astype(np.float32)
                q += sz * width
    return out, [c for c, _ in channels]
