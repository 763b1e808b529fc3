// bitcopy.h — appending bit ranges to an Arrow bitmap under construction (validity bitmaps and Boolean values of
// the host batch coalescer, coalesce.inc).  Plain C++ (no CUDA) so tests/test_bitcopy_cpu.py can check it with g++.
// Contract: bits [0, pos) of dst are final; everything at or above bit `pos` is unspecified before the call and,
// after it, bits [pos, pos + n) are set and the remaining bits of the last touched byte are zero.
#pragma once
#include <cstdint>
#include <cstring>

namespace dfgpu {
inline void bits_fill_ones(uint8_t *dst, int64_t pos, int64_t n) {
    if (n <= 0) return;
    if (pos & 7) dst[pos >> 3] &= (uint8_t)((1u << (pos & 7)) - 1);
    while (n > 0 && (pos & 7)) { dst[pos >> 3] |= (uint8_t)(1u << (pos & 7)); pos++; n--; }
    if (n >= 8) { memset(dst + (pos >> 3), 0xff, (size_t)(n >> 3)); pos += n & ~(int64_t)7; n &= 7; }
    if (n > 0) dst[pos >> 3] = (uint8_t)((1u << n) - 1);
}
// append bits [soff, soff + n) of src at bit position pos of dst
inline void bits_append(uint8_t *dst, int64_t pos, const uint8_t *src, int64_t soff, int64_t n) {
    if (n <= 0) return;
    if (pos & 7) dst[pos >> 3] &= (uint8_t)((1u << (pos & 7)) - 1);
    while (n > 0 && (pos & 7)) {  // head: up to the next byte boundary of dst
        if ((src[soff >> 3] >> (soff & 7)) & 1) dst[pos >> 3] |= (uint8_t)(1u << (pos & 7));
        pos++; soff++; n--;
    }
    if (n <= 0) return;
    uint8_t *d = dst + (pos >> 3);
    const uint8_t *s = src + (soff >> 3);
    const int sh = (int)(soff & 7);
    const int64_t nbytes = (n + 7) >> 3;
    if (sh == 0) {
        memcpy(d, s, (size_t)nbytes);
    } else {
        const int64_t sbytes = (sh + n + 7) >> 3;  // source bytes the range spans (never read past them)
        for (int64_t i = 0; i < nbytes; i++) {
            unsigned lo = (unsigned)s[i] >> sh;
            unsigned hi = i + 1 < sbytes ? (unsigned)s[i + 1] << (8 - sh) : 0u;
            d[i] = (uint8_t)(lo | hi);
        }
    }
    if (n & 7) d[nbytes - 1] &= (uint8_t)((1u << (n & 7)) - 1);  // bits past the end stay zero
}
}  // namespace dfgpu
