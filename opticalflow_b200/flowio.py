"""On-disk flow formats of the reference drivers: byte-exact writers / readers.

  write_flo / writeFlowFile / read_flow     Middlebury .flo   /root/reference/script_pwc.py:12-27,
                                            pwc_extract_flow.py:46-56, data_processing.py:17-29
  encode_kitti / decode_kitti               u = (R - 2^15)/64, v = (G - 2^15)/64, valid = B != 0
                                            /root/reference/inference_kitti.py:23-52, inference.py:88-102,266-282
  write_kitti_png / read_kitti_png          16-bit RGB PNG container (zlib from the standard library; no cv2 / PIL needed)

These are host-side byte formats (files); the arithmetic on either side of them that touches whole frames on the
hot path (x20, resize, unpad) is on the GPU in prepost.py.
"""
import struct
import zlib

import numpy as np

TAG_FLOAT = 202021.25          # "PIEH" as float32, little endian


def write_flo(filename, flow_uv):
    """Middlebury .flo: float32 tag 202021.25, int32 W, int32 H, then H*W*(u,v) float32 rows (pwc_extract_flow.py:46-56)."""
    flow_uv = np.asarray(flow_uv)
    if flow_uv.ndim != 3 or flow_uv.shape[2] != 2:
        raise ValueError("write_flo: flow must have two bands ([H,W,2])")
    h, w, _ = flow_uv.shape
    with open(filename, "wb") as f:
        f.write(struct.pack("<f", TAG_FLOAT))
        f.write(struct.pack("<i", w))
        f.write(struct.pack("<i", h))
        f.write(np.ascontiguousarray(flow_uv, dtype="<f4").tobytes())


def writeFlowFile(filename, uv):
    """Name and behaviour of script_pwc.py:12-27."""
    write_flo(filename, uv)


def read_flow(flow_file):
    """.flo reader (data_processing.py:17-29): returns [H,W,2] float32; ValueError on a wrong tag."""
    with open(flow_file, "rb") as f:
        head = f.read(12)
        if len(head) < 12 or struct.unpack("<f", head[:4])[0] != TAG_FLOAT:
            raise ValueError("Invalid .flo file")
        w, h = struct.unpack("<ii", head[4:12])
        data = np.frombuffer(f.read(8 * w * h), dtype="<f4", count=2 * w * h)
    return data.reshape(h, w, 2).astype(np.float32)


def encode_kitti(flow, valid=None):
    """[H,W,2] float flow (+ optional [H,W] validity) -> uint16 [H,W,3] in R,G,B order: R = u*64 + 2^15, G = v*64 + 2^15
    (truncated like numpy's astype(uint16)), B = valid (inference.py:266-277)."""
    flow = np.asarray(flow, dtype=np.float32)
    u = (flow[..., 0] * 64.0 + 2 ** 15).astype(np.uint16)
    v = (flow[..., 1] * 64.0 + 2 ** 15).astype(np.uint16)
    b = np.ones(u.shape, np.uint16) if valid is None else (np.asarray(valid) != 0).astype(np.uint16)
    return np.stack([u, v, b], axis=-1)


def decode_kitti(rgb16):
    """uint16 [H,W,3] (R,G,B) -> (flow [H,W,2] float32, valid [H,W] bool): inference_kitti.py:39-52."""
    rgb16 = np.asarray(rgb16)
    if rgb16.ndim != 3 or rgb16.shape[2] != 3 or rgb16.dtype != np.uint16:
        raise RuntimeError(f"KITTI flow PNG must be uint16 [H,W,3], got {rgb16.dtype} {rgb16.shape}")
    u = (rgb16[..., 0].astype(np.float32) - 32768.0) / 64.0
    v = (rgb16[..., 1].astype(np.float32) - 32768.0) / 64.0
    return np.stack([u, v], axis=-1).astype(np.float32), rgb16[..., 2] != 0


def _chunk(tag, data):
    return struct.pack(">I", len(data)) + tag + data + struct.pack(">I", zlib.crc32(tag + data) & 0xFFFFFFFF)


def write_png16(filename, rgb16, level=6):
    """Minimal 16-bit RGB PNG writer (colour type 2, bit depth 16, filter 0, no interlace)."""
    rgb16 = np.ascontiguousarray(rgb16, dtype=np.uint16)
    h, w, c = rgb16.shape
    if c != 3:
        raise ValueError("write_png16: [H,W,3] expected")
    raw = np.empty((h, 1 + w * 6), np.uint8)
    raw[:, 0] = 0
    raw[:, 1:] = rgb16.astype(">u2").view(np.uint8).reshape(h, w * 6)
    png = b"\x89PNG\r\n\x1a\n" + _chunk(b"IHDR", struct.pack(">IIBBBBB", w, h, 16, 2, 0, 0, 0))
    png += _chunk(b"IDAT", zlib.compress(raw.tobytes(), level)) + _chunk(b"IEND", b"")
    with open(filename, "wb") as f:
        f.write(png)


def read_png16(filename):
    """16-bit RGB, non-interlaced PNG reader (all five row filters) -> uint16 [H,W,3] in R,G,B order."""
    with open(filename, "rb") as f:
        buf = f.read()
    if buf[:8] != b"\x89PNG\r\n\x1a\n":
        raise RuntimeError(f"{filename}: not a PNG file")
    pos, idat, hdr = 8, [], None
    while pos < len(buf):
        n, tag = struct.unpack(">I", buf[pos:pos + 4])[0], buf[pos + 4:pos + 8]
        body = buf[pos + 8:pos + 8 + n]
        pos += 12 + n
        if tag == b"IHDR":
            hdr = struct.unpack(">IIBBBBB", body)
        elif tag == b"IDAT":
            idat.append(body)
        elif tag == b"IEND":
            break
    if hdr is None or hdr[2:] != (16, 2, 0, 0, 0):
        raise RuntimeError(f"{filename}: expected a 16-bit RGB non-interlaced PNG, header {hdr}")
    w, h = hdr[0], hdr[1]
    bpp, stride = 6, w * 6
    raw = np.frombuffer(zlib.decompress(b"".join(idat)), np.uint8).reshape(h, 1 + stride)
    out = np.zeros((h, stride), np.uint8)
    prev = np.zeros(stride, np.int32)
    for y in range(h):
        ft, line = int(raw[y, 0]), raw[y, 1:].astype(np.int32)
        if ft == 0:
            cur = line
        elif ft == 2:
            cur = (line + prev) & 255
        elif ft == 1:                                   # Sub: running sum per byte lane of a pixel
            cur = (np.cumsum(line.reshape(w, bpp), axis=0) & 255).reshape(-1)
        else:                                           # Average / Paeth: sequential in x
            cur = np.zeros(stride, np.int32)
            for i in range(stride):
                a = cur[i - bpp] if i >= bpp else 0
                b = prev[i]
                if ft == 3:
                    pred = (a + b) >> 1
                else:
                    c = prev[i - bpp] if i >= bpp else 0
                    p = a + b - c
                    pa, pb, pc = abs(p - a), abs(p - b), abs(p - c)
                    pred = a if (pa <= pb and pa <= pc) else (b if pb <= pc else c)
                cur[i] = (line[i] + pred) & 255
        out[y] = cur
        prev = cur
    return out.view(">u2").astype(np.uint16).reshape(h, w, 3)


def write_kitti_png(filename, flow, valid=None):
    """Role of save_flow (inference.py:266-282), NOT byte-compatible with it: the reference stacks [valid, v, u] (R = valid,
    B = u) and hands a uint16 RGB array to PIL, which cannot encode it.  This writer emits the standard KITTI layout
    (R = u, G = v, B = valid, 16 bit: u16 = flow*64 + 2^15), the one load_flow_kitti_png (inference_kitti.py:23-52) and
    read_kitti_png below decode; write -> read round-trips (tests/test_prepost.py)."""
    write_png16(filename, encode_kitti(flow, valid))


def read_kitti_png(filename):
    """load_flow_kitti_png of inference_kitti.py:23-52: (flow [H,W,2] float32, valid [H,W] bool)."""
    return decode_kitti(read_png16(filename))
