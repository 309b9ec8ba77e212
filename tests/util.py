"""Shared helpers for the tests: seeded corpus (encoded with the reference's libjpeg-turbo), oracle access, reorderings."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from nvimagecodec_b200.corpus import synth_image  # noqa: E402
from oracle import port, ref  # noqa: E402

GOLDEN = os.path.join(ROOT, "tests", "golden")


def have_ref():
    return ref.available()


def make_jpeg(seed, w, h, samp=(2, 2), quality=90, restart_interval=0, restart_in_rows=0, gray=False, kind="photo", **kw):
    img = synth_image(seed, w, h, 1 if gray else 3, kind)
    return ref.ljt_encode(img, quality, samp, restart_interval=restart_interval, restart_in_rows=restart_in_rows, **kw)


def mcu_order_coefficients(data):
    """oracle (jpeg_read_coefficients) coefficients re-ordered into the plugin's MCU scan order: int16[blocks*64]"""
    import nvimagecodec_b200 as nb
    info = nb.parse_header(data)  # (the plain-C port's header reader knows baseline streams only)
    ncomp = info.num_components
    hs = [info.h_samp[0], 1, 1] if ncomp == 3 else [1]
    vs = [info.v_samp[0], 1, 1] if ncomp == 3 else [1]
    hmax, vmax = max(hs), max(vs)
    mx = (info.width + 8 * hmax - 1) // (8 * hmax)
    my = (info.height + 8 * vmax - 1) // (8 * vmax)
    comps = [ref.ljt_coefficients(data, c, mx * hs[c], my * vs[c]) for c in range(ncomp)]
    parts = []
    for c in range(ncomp):
        a = comps[c].reshape(my, vs[c], mx, hs[c], 64).transpose(0, 2, 1, 3, 4).reshape(my * mx, vs[c] * hs[c], 64)
        parts.append(a)
    return np.concatenate(parts, axis=1).reshape(-1)


def unstuff_reference(scan):
    """python statement of the unstuffing rule (drop FF00 stuffing, FF fill, RSTn) -> (bytes, marker positions)"""
    out = bytearray()
    marks = []
    n = len(scan)
    for p in range(n):
        cur = scan[p]
        pv = scan[p - 1] if p else 0
        nx = scan[p + 1] if p + 1 < n else 0
        if cur == 0 and pv == 0xFF:
            continue
        if cur == 0xFF and nx == 0xFF:
            continue
        if cur == 0xFF and 0xD0 <= nx <= 0xD7:
            marks.append(len(out))
            continue
        if 0xD0 <= cur <= 0xD7 and pv == 0xFF:
            continue
        out.append(cur)
    return bytes(out), marks


def apply_orientation(img, rotated, flip_x, flip_y):
    """nvimgcodecOrientation_t applied to an (H, W, C) array: rotate counter-clockwise by `rotated` degrees, then mirror
    horizontally (flip_x) / vertically (flip_y) -- the meaning src/parsers/exif_orientation.h:37-60 gives the EXIF values
    (e.g. EXIF 6 "rotate 90 CW" = {270, False, False}, EXIF 5 "transpose" = {90, False, True})"""
    out = np.rot90(img, k=(rotated // 90) % 4, axes=(0, 1))
    if flip_x:
        out = out[:, ::-1]
    if flip_y:
        out = out[::-1]
    return np.ascontiguousarray(out)


# the eight EXIF orientations as (rotated, flip_x, flip_y), in the order of the tag's values 1..8
EXIF_ORIENTATIONS = [(0, False, False), (0, True, False), (180, False, False), (0, False, True), (90, False, True), (270, False, False),
                     (270, False, True), (90, False, False)]
