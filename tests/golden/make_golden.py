"""Regenerates tests/golden/* from the reference tree (run in the build container where /root/reference exists).

Known answers come from libjpeg-turbo's own regression suite (external/libjpeg-turbo/CMakeLists.txt:1273-1345): cjpeg/djpeg
-dct int on testimages/testorig.ppm.  We store the *JPEG inputs* produced by the reference's libjpeg-turbo (a few KB each)
and golden.json with the MD5s of (a) those JPEGs (encoder known answers) and (b) their decoded PPM/PGM (decoder known
answers), plus SHA-256 of reference-plugin RGB output for a seeded synthetic corpus covering every sampling mode."""
import hashlib
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
from nvimagecodec_b200.corpus import synth_image  # noqa: E402
from oracle import ref  # noqa: E402

LJT = "/root/reference/external/libjpeg-turbo"
# MD5 constants quoted from external/libjpeg-turbo/CMakeLists.txt (8-bit, non-"JPEG 7/8 API" section, lines 1273-1345)
LJT_MD5 = {
    "MD5_JPEG_440_ISLOW": "538bc02bd4b4658fd85de6ece6cbeda6", "MD5_PPM_440_ISLOW": "11e7eab7ef7ef3276934bb7e7b6bb377",
    "MD5_JPEG_420_ISLOW": "9a68f56bc76e466aa7e52f415d0f4a5f",
    "MD5_JPEG_GRAY_ISLOW": "72b51f894b8f4a10b3ee3066770aa38d", "MD5_PPM_GRAY_ISLOW": "8d3596c56eace32f205deccc229aa5ed",
    "MD5_PPM_GRAY_ISLOW_RGB": "116424ac07b79e5e801f00508eab48ec",
}


def main():
    raw = open(os.path.join(LJT, "testimages", "testorig.ppm"), "rb").read()
    img = np.frombuffer(raw.split(b"\n", 3)[3], dtype=np.uint8).reshape(149, 227, 3)
    files = {
        "testorig_440_islow.jpg": ref.ljt_encode(img, 75, (1, 2)),   # cjpeg -sample 1x2 -dct int
        "testorig_420_islow.jpg": ref.ljt_encode(img, 75, (2, 2)),   # cjpeg -sample 2x2 -dct int (== jpegtran of testimgari)
        "testorig_gray_islow.jpg": ref.ljt_encode(img, 75, (0, 0)),  # cjpeg -gray -dct int
        "testorig.jpg": open(os.path.join(LJT, "testimages", "testorig.jpg"), "rb").read(),
    }
    assert hashlib.md5(files["testorig_440_islow.jpg"]).hexdigest() == LJT_MD5["MD5_JPEG_440_ISLOW"]
    assert hashlib.md5(files["testorig_420_islow.jpg"]).hexdigest() == LJT_MD5["MD5_JPEG_420_ISLOW"]
    assert hashlib.md5(files["testorig_gray_islow.jpg"]).hexdigest() == LJT_MD5["MD5_JPEG_GRAY_ISLOW"]
    for name, data in files.items():
        open(os.path.join(HERE, name), "wb").write(data)
    open(os.path.join(HERE, "testorig.ppm"), "wb").write(raw)  # source image of the encoder known answers
    dec = ref.Session.shared().decoder(options=":fancy_upsampling=1")
    golden = {"ljt_md5": LJT_MD5, "decoded_md5": {}, "synthetic_sha256": []}
    for name, data in files.items():
        rgb, st, _ = dec.decode_host([data])
        golden["decoded_md5"][name] = {"ppm_rgb": hashlib.md5(b"P6\n227 149\n255\n" + rgb[0].tobytes()).hexdigest()}
        if "gray" in name:
            y, _, _ = dec.decode_host([data], fmt=ref.P_Y)
            golden["decoded_md5"][name]["pgm_y"] = hashlib.md5(b"P5\n227 149\n255\n" + y[0].tobytes()).hexdigest()
    assert golden["decoded_md5"]["testorig_440_islow.jpg"]["ppm_rgb"] == LJT_MD5["MD5_PPM_440_ISLOW"]
    assert golden["decoded_md5"]["testorig_gray_islow.jpg"]["pgm_y"] == LJT_MD5["MD5_PPM_GRAY_ISLOW"]
    assert golden["decoded_md5"]["testorig_gray_islow.jpg"]["ppm_rgb"] == LJT_MD5["MD5_PPM_GRAY_ISLOW_RGB"]
    # seeded synthetic corpus: (seed, w, h, samp, quality, restart_interval, gray, fancy) -> sha256 of reference RGB
    cases = [(11, 227, 149, (2, 2), 90, 0, False), (12, 131, 77, (2, 1), 85, 0, False), (13, 131, 77, (1, 2), 85, 0, False),
             (14, 100, 100, (1, 1), 95, 7, False), (15, 333, 211, (2, 2), 75, 5, False), (16, 201, 99, (1, 1), 85, 0, True),
             (17, 17, 9, (2, 2), 90, 0, False), (18, 640, 426, (2, 2), 75, 0, False)]
    d0 = ref.Session.shared().decoder(options=":fancy_upsampling=0")
    for seed, w, h, samp, q, ri, gray in cases:
        data = ref.ljt_encode(synth_image(seed, w, h, 1 if gray else 3), q, samp, restart_interval=ri)
        entry = {"seed": seed, "w": w, "h": h, "samp": list(samp), "quality": q, "restart_interval": ri, "gray": gray,
                 "jpeg_sha256": hashlib.sha256(data).hexdigest()}
        for fancy, dd in ((1, dec), (0, d0)):
            rgb, st, _ = dd.decode_host([data])
            entry["rgb_sha256_fancy%d" % fancy] = hashlib.sha256(rgb[0].tobytes()).hexdigest()
        golden["synthetic_sha256"].append(entry)
    json.dump(golden, open(os.path.join(HERE, "golden.json"), "w"), indent=1)
    print("wrote", sorted(files), "golden.json")


if __name__ == "__main__":
    main()
