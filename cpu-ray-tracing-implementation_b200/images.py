"""Image ingest for the opt-in texture path: float (un-quantised) images and bilinear lookups.

The reference turns every image into 8-bit at load time (src/image.h:33-67 -> convert_to_bytes, :99-117) and picks the nearest
texel (src/texture.h:67-74); that path is what `Scene` files carry by default and what parity is judged on.  This module feeds the
backend's two extensions (include/b200rt.h: b200rt_image.bytes_per_pixel 12 / 16, B200RT_TEX_PICTURE_LINEAR):

  read_exr(path)                  -> float32 (H, W, C) from a scan-line OpenEXR file (NONE / ZIPS / ZIP, HALF or FLOAT channels);
                                     numpy + zlib only: what tinyexr's LoadEXR (src/image.h:54) would deliver as `fdata`
  add_image(scene, pixels)        -> image row; uint8 (H, W, 3|4) = the reference's bdata, float32 (H, W, 3|4) = un-quantised
  set_picture_filter(scene, row, linear)   switch a picture texture between nearest (reference) and bilinear lookups
"""
import struct
import zlib

import numpy as np

from . import _abi as abi


def _exr_header(buf):
    if buf[:4] != bytes([0x76, 0x2F, 0x31, 0x01]):
        raise ValueError("not an OpenEXR file")
    version = struct.unpack("<I", buf[4:8])[0]
    if version & 0x200 or version & 0x1000:
        raise ValueError("tiled / multi-part OpenEXR files are not supported")
    pos, attrs = 8, {}
    while buf[pos] != 0:
        e = buf.index(b"\0", pos)
        name = buf[pos:e].decode("latin1")
        pos = e + 1
        e = buf.index(b"\0", pos)
        typ = buf[pos:e].decode("latin1")
        pos = e + 1
        (size,) = struct.unpack("<i", buf[pos:pos + 4])
        attrs[name] = (typ, buf[pos + 4:pos + 4 + size])
        pos += 4 + size
    return attrs, pos + 1


def _undo_zip(raw, expected):
    """OpenEXR ZIP(S): zlib, then the byte predictor, then de-interleaving of the two halves."""
    data = np.frombuffer(zlib.decompress(raw), np.uint8).astype(np.int32)
    if len(data) != expected:
        raise ValueError("OpenEXR block: unexpected decompressed size")
    data[1:] -= 128
    data = (np.cumsum(data) & 0xFF).astype(np.uint8)  # t[i] = t[i-1] + d[i] - 128
    half = (len(data) + 1) // 2
    out = np.empty(len(data), np.uint8)
    out[0::2] = data[:half]
    out[1::2] = data[half:]
    return out.tobytes()


def read_exr(path):
    """-> float32 array (H, W, C) with C = 3 (R, G, B) or 4 (R, G, B, A)."""
    buf = open(path, "rb").read()
    attrs, pos = _exr_header(buf)
    x0, y0, x1, y1 = struct.unpack("<4i", attrs["dataWindow"][1])
    W, H = x1 - x0 + 1, y1 - y0 + 1
    comp = attrs["compression"][1][0]
    if comp not in (0, 2, 3):
        raise ValueError(f"OpenEXR compression {comp} is not supported (NONE, ZIPS and ZIP are)")
    lines_per_block = {0: 1, 2: 1, 3: 16}[comp]
    chans, ch, p = [], attrs["channels"][1], 0
    while ch[p] != 0:
        e = ch.index(b"\0", p)
        name = ch[p:e].decode("latin1")
        ptype, = struct.unpack("<i", ch[e + 1:e + 5])
        xs, ys = struct.unpack("<2i", ch[e + 9:e + 17])
        if ptype not in (1, 2) or xs != 1 or ys != 1:
            raise ValueError("OpenEXR: only HALF / FLOAT channels without subsampling are supported")
        chans.append((name, ptype))
        p = e + 17
    n_blocks = (H + lines_per_block - 1) // lines_per_block
    offs = np.frombuffer(buf, "<u8", n_blocks, pos)
    planes = {name: np.zeros((H, W), np.float32) for name, _ in chans}
    row_bytes = sum(W * (2 if t == 1 else 4) for _, t in chans)
    for o in offs:
        y, nbytes = struct.unpack("<2i", buf[o:o + 8])
        lines = min(lines_per_block, y1 - y + 1)
        raw = buf[o + 8:o + 8 + nbytes]
        if comp and nbytes < lines * row_bytes:
            raw = _undo_zip(raw, lines * row_bytes)
        q = 0
        for k in range(lines):
            for name, t in chans:  # channels are stored in alphabetical order, one run per scan line
                n = W * (2 if t == 1 else 4)
                planes[name][y - y0 + k] = np.frombuffer(raw, "<f2" if t == 1 else "<f4", W, q).astype(np.float32)
                q += n
    order = [c for c in ("R", "G", "B", "A") if c in planes]
    if order[:3] != ["R", "G", "B"]:
        raise ValueError(f"OpenEXR: channels {sorted(planes)} (need R, G, B)")
    return np.stack([planes[c] for c in order], axis=-1)


def add_image(scene, pixels):
    """Appends an image to the scene description.  uint8 -> the reference's 8-bit rows (3 or 4 bytes per pixel, read as byte / 256);
    float32 -> 12 or 16 bytes per pixel, read as they are.  Returns the image row."""
    px = np.ascontiguousarray(pixels)
    if px.ndim != 3 or px.shape[2] not in (3, 4) or px.dtype not in (np.uint8, np.float32):
        raise ValueError("pixels must be (H, W, 3|4) uint8 or float32")
    blob = np.asarray(scene.image_bytes, np.uint8)
    pad = (-len(blob)) % 16
    blob = np.concatenate([blob, np.zeros(pad, np.uint8)])
    row = np.zeros(1, abi.image_dtype)
    row["width"], row["height"] = px.shape[1], px.shape[0]
    row["bytes_per_pixel"] = px.shape[2] * px.dtype.itemsize
    row["has_data"] = 1
    row["offset"] = len(blob)
    scene.image_bytes = np.concatenate([blob, px.reshape(-1).view(np.uint8)])
    scene.images = np.concatenate([scene.images, row])
    return len(scene.images) - 1


def float_copy_of_image(scene, image_row):
    """The 8-bit image `image_row` again as a float image holding byte / 256 exactly (every such value is a binary32): nearest
    lookups on the copy must equal the reference path bit for bit.  Returns the new image row."""
    im = scene.images[image_row]
    n = int(im["width"]) * int(im["height"]) * int(im["bytes_per_pixel"])
    b = np.asarray(scene.image_bytes[int(im["offset"]):int(im["offset"]) + n], np.uint8).reshape(int(im["height"]), int(im["width"]), int(im["bytes_per_pixel"]))
    return add_image(scene, (b.astype(np.float32) / np.float32(256.0)))


def set_picture_filter(scene, tex_row, linear):
    t = scene.textures
    if t["kind"][tex_row] not in (abi.TEX_PICTURE, abi.TEX_PICTURE_LINEAR):
        raise ValueError("texture row is not a picture texture")
    t["kind"][tex_row] = abi.TEX_PICTURE_LINEAR if linear else abi.TEX_PICTURE
