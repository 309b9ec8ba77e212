"""nvimagecodec_b200 -- B200-native baseline-JPEG codec path for nvImageCodec.

Product pieces:
  lib/libb200jpeg.so                C ABI (include/b200jpeg.h): CUDA sm_100a decode / encode of JPEG batches
  extensions/libb200jpeg_ext.so     nvImageCodec extension module (nvimgcodecExtensionModuleEntry) wrapping the same code
  decoder.Decoder / encoder.Encoder Python host mirror of nvimgcodec.Decoder / nvimgcodec.Encoder for this path
"""
from . import _capi  # noqa: F401
from ._capi import EXT_DIR, EXT_PATH, LIB_PATH, parse_header, status_string  # noqa: F401


def __getattr__(name):
    if name in ("Decoder", "Image", "MultiDeviceDecoder"):
        from . import decoder
        return getattr(decoder, name)
    if name == "Encoder":
        from . import encoder
        return encoder.Encoder
    raise AttributeError(name)
