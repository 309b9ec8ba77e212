/* oracle/jpeg_oracle.h -- TEST INFRASTRUCTURE, NOT PRODUCT.
 * Plain-C restatement of the reference's baseline-JPEG decode path (libjpeg_turbo_ext -> libjpeg-turbo 3.0.3,
 * ISLOW IDCT, fancy or box upsampling, YCbCr->RGB).  Parity status: PINNED -- checked in tests/ against
 * (a) libjpeg-turbo's own golden MD5s (external/libjpeg-turbo/CMakeLists.txt:1273-1345) and
 * (b) the compiled reference (oracle/_ref) on the seeded corpus.  See jpeg_oracle.c for per-function citations. */
#ifndef JPEG_ORACLE_H
#define JPEG_ORACLE_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

enum { JO_FMT_I_RGB = 0, JO_FMT_I_BGR = 1, JO_FMT_Y = 2, JO_FMT_P_RGB = 3, JO_FMT_P_BGR = 4 };

/* out[0..7] = width, height, ncomp, hs0, vs0, restart_interval, SOF marker (0xC0/0xC1), entropy-coded bytes */
int jo_info(const uint8_t* jpg, size_t len, int* out);

/* Full decode.  Interleaved: out[y*stride + x*3 + c]; planar: plane p at out + p*stride*height; Y: 1 channel. */
int jo_decode(const uint8_t* jpg, size_t len, int fmt, int fancy_upsampling, uint8_t* out, int64_t stride);

/* Quantised coefficients of one component, natural order, over the MCU-padded block grid (same layout as
 * refh_ljt_read_coefficients).  dims[0..1] = blocks per row, block rows. */
int jo_read_coefficients(const uint8_t* jpg, size_t len, int comp, int16_t* out, size_t cap_blocks, int* dims);

/* Component plane after dequant + IDCT (MCU-padded: dims[0] x dims[1] samples, pitch = dims[0]). */
int jo_read_plane(const uint8_t* jpg, size_t len, int comp, uint8_t* out, size_t cap, int* dims);

/* Stage-level helpers for kernel unit tests */
void jo_idct_islow(const int16_t* coef, const uint16_t* quant_natural, uint8_t* out, int out_stride);

/* ---- encode mirror (jcparam/jccolor/jcsample/jfdctint/jcdctmgr/jchuff/jcmarker restated) ---- */
/* pixels: interleaved RGB (channels=3) or gray (channels=1). hs/vs: luma sampling factors (1 or 2).
 * Returns bytes written (<= cap) or -1. Standard tables, baseline, JFIF header, optional DRI. */
long jo_encode(const uint8_t* pixels, int width, int height, int channels, int64_t stride, int quality, int hs, int vs,
    int restart_interval, uint8_t* out, size_t cap);

#ifdef __cplusplus
}
#endif
#endif
