"""Minimal OpenEXR 2 scan-line writer for generated scene assets (float environment maps,
scenes/gen_scenes.py) and the reader tests (tests/test_exr_reader.py): single
part, HALF or FLOAT channels, compression NONE / RLE / ZIPS / ZIP, either line order, any data
window origin.  Follows the published file layout (magic 20000630, attribute list, offset
table, chunks of { y, size, data }; ZIP / RLE data = even/odd byte split, byte delta, deflate or
run lengths; a chunk that does not shrink is stored raw)."""
import struct
import zlib

import numpy as np

COMP = {"none": 0, "rle": 1, "zips": 2, "zip": 3}


def _attr(name, typ, payload):
    return name.encode() + b"\0" + typ.encode() + b"\0" + struct.pack("<i", len(payload)) + payload


def _rle(data):
    out = bytearray()
    i, n = 0, len(data)
    while i < n:
        j = i + 1
        while j < n and data[j] == data[i] and j - i < 128:
            j += 1
        if j - i >= 3:
            out += struct.pack("b", j - i - 1) + bytes([data[i]])
            i = j
            continue
        j = i
        while j < n and j - i < 127 and not (j + 2 < n and data[j] == data[j + 1] == data[j + 2]):
            j += 1
        out += struct.pack("b", -(j - i)) + bytes(data[i:j])
        i = j
    return bytes(out)


def _pack(raw, comp):
    if comp == "none":
        return raw
    a = np.frombuffer(raw, np.uint8)
    t = np.concatenate([a[0::2], a[1::2]]).astype(np.int32)
    d = t.copy()
    d[1:] = (t[1:] - t[:-1] + 128 + 256) & 255
    pre = d.astype(np.uint8).tobytes()
    packed = _rle(pre) if comp == "rle" else zlib.compress(pre, 6)
    return packed if len(packed) < len(raw) else raw


def write_exr(path, channels, comp="none", origin=(0, 0), decreasing=False):
    """channels: dict name -> 2-D array (h, w) of float16 or float32; row 0 = top scan line."""
    names = sorted(channels)
    h, w = channels[names[0]].shape
    x0, y0 = origin
    chlist = b""
    for n in names:
        t = {np.dtype(np.float16): 1, np.dtype(np.float32): 2}[channels[n].dtype]
        chlist += n.encode() + b"\0" + struct.pack("<iB3xii", t, 0, 1, 1)
    chlist += b"\0"
    box = struct.pack("<4i", x0, y0, x0 + w - 1, y0 + h - 1)
    head = struct.pack("<ii", 20000630, 2)
    head += _attr("channels", "chlist", chlist)
    head += _attr("compression", "compression", bytes([COMP[comp]]))
    head += _attr("dataWindow", "box2i", box) + _attr("displayWindow", "box2i", box)
    head += _attr("lineOrder", "lineOrder", bytes([1 if decreasing else 0]))
    head += _attr("pixelAspectRatio", "float", struct.pack("<f", 1.0))
    head += _attr("screenWindowCenter", "v2f", struct.pack("<ff", 0.0, 0.0))
    head += _attr("screenWindowWidth", "float", struct.pack("<f", 1.0))
    head += b"\0"
    lpb = 16 if comp == "zip" else 1
    chunks = []
    for r0 in range(0, h, lpb):
        raw = b""
        for r in range(r0, min(h, r0 + lpb)):
            for n in names:
                raw += np.ascontiguousarray(channels[n][r]).astype(channels[n].dtype.newbyteorder("<")).tobytes()
        data = _pack(raw, comp)
        chunks.append(struct.pack("<ii", y0 + r0, len(data)) + data)
    order = list(range(len(chunks)))
    stored = order[::-1] if decreasing else order  # file order of the chunks
    pos = len(head) + 8 * len(chunks)
    offsets = [0] * len(chunks)
    body = b""
    for k in stored:
        offsets[k] = pos + len(body)
        body += chunks[k]
    # the offset table lists chunks in increasing y for either line order
    with open(path, "wb") as f:
        f.write(head + b"".join(struct.pack("<Q", o) for o in offsets) + body)
