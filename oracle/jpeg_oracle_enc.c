/* oracle/jpeg_oracle_enc.c -- TEST INFRASTRUCTURE. Encode-mirror restatement (filled in with the encoder milestone). */
#include "jpeg_oracle.h"
long jo_encode(const uint8_t* pixels, int width, int height, int channels, int64_t stride, int quality, int hs, int vs,
    int restart_interval, uint8_t* out, size_t cap)
{
    (void)pixels; (void)width; (void)height; (void)channels; (void)stride; (void)quality; (void)hs; (void)vs;
    (void)restart_interval; (void)out; (void)cap;
    return -1;
}
