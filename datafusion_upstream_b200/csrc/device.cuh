// device.cuh — device-side helpers: bitmap access, vectorised loads, the row hash
// ("dfhash v1") with the structure of the reference's create_hashes
// (common/src/hash_utils.rs:38-41, 92-132, 182-210, 357-416), and 128-bit arithmetic.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "common.h"

namespace dfgpu {

#define DF_HD __host__ __device__ __forceinline__
#define DF_D __device__ __forceinline__

DF_HD bool bit_get(const uint8_t *b, int64_t i) { return (b[i >> 3] >> (i & 7)) & 1; }
DF_HD int type_width_dev(int id) {
    switch (id) {
    case T_I8: case T_U8: return 1;
    case T_I16: case T_U16: return 2;
    case T_I32: case T_U32: case T_F32: case T_DATE32: case T_TIME32: return 4;
    case T_DEC128: return 16;
    default: return 8;
    }
}

DF_HD uint64_t mix64(uint64_t x) {
    x ^= x >> 30; x *= 0xbf58476d1ce4e5b9ULL;
    x ^= x >> 27; x *= 0x94d049bb133111ebULL;
    x ^= x >> 31;
    return x;
}
DF_HD uint64_t h_u64(uint64_t v, uint64_t seed) { return mix64(v + 0x9e3779b97f4a7c15ULL + seed); }
DF_HD uint64_t h_u128(uint64_t lo, uint64_t hi, uint64_t seed) {
    return mix64(h_u64(lo, seed) ^ (hi * 0xff51afd7ed558ccdULL));
}
// combine_hashes (hash_utils.rs:38-41)
DF_HD uint64_t combine_hashes(uint64_t l, uint64_t r) { return (17ULL * 37ULL + l) * 37ULL + r; }

// n bytes (1..8) at an arbitrary address as a little-endian word, read with one or two ALIGNED 64-bit loads
// (Utf8 values start at any byte; a byte loop costs eight loads per word).  An aligned word that holds at least one
// byte of the range lies in the same page as that byte, so neither load can fault; bytes outside the range are
// shifted or masked away.  CG: read through L2 (data another SM may have just published).
template <bool CG = false> DF_D uint64_t load_le(const uint8_t *p, int n) {
    const uintptr_t a = (uintptr_t)p;
    const uint64_t *w = (const uint64_t *)(a & ~(uintptr_t)7);
    const int o = (int)(a & 7), sh = o * 8;
    uint64_t v = (CG ? __ldcg(w) : *w) >> sh;
    if (o + n > 8) v |= (CG ? __ldcg(w + 1) : w[1]) << (64 - sh);  // o >= 1 here
    if (n < 8) v &= (1ULL << (8 * n)) - 1;
    return v;
}
DF_D uint64_t h_bytes(const uint8_t *p, int64_t len, uint64_t seed) {
    uint64_t h = h_u64((uint64_t)len, seed);
    int64_t i = 0;
    for (; i + 8 <= len; i += 8) h = mix64(h ^ load_le(p + i, 8));
    if (i < len) h = mix64(h ^ load_le(p + i, (int)(len - i)));
    return h;
}
// String ids inside packed key images (groups.h: kind 2 fields).  A string of <= 3 bytes IS its id - the length and the
// bytes themselves, top bit set - so it needs no table (the inline rule of ArrowBytesMap, physical-expr/src/binary_map.rs:
// 381-393, at the width of a 4-byte field); longer strings get dense ids < 2^31 from the column's interner.
static const uint32_t STRID_INLINE = 0x80000000u;
static const int STRID_INLINE_MAX = 3;
DF_D uint32_t strid_inline(const uint8_t *p, int len) { return STRID_INLINE | ((uint32_t)len << 24) | (len ? (uint32_t)load_le(p, len) : 0u); }

// a[0, len) == b[0, len), eight bytes per step; CG: b is read through L2
template <bool CG = false> DF_D bool bytes_equal(const uint8_t *a, const uint8_t *b, int64_t len) {
    int64_t k = 0;
    for (; k + 8 <= len; k += 8)
        if (load_le(a + k, 8) != load_le<CG>(b + k, 8)) return false;
    return k == len || load_le(a + k, (int)(len - k)) == load_le<CG>(b + k, (int)(len - k));
}

// ---- column access ----------------------------------------------------------------------------
DF_D bool col_valid(const ColView &c, int64_t i) { return !c.validity || bit_get(c.validity, c.vbit + i); }

// widened 64-bit pattern of a fixed-width (<= 8 byte) value: signed ints sign-extended, unsigned and
// float bit patterns zero-extended, bool 0/1 — the integer the hash and the packed key are built from.
DF_D uint64_t col_bits64(const ColView &c, int64_t i) {
    switch (c.type) {
    case T_BOOL: return bit_get((const uint8_t *)c.values, c.dbit + i) ? 1 : 0;
    case T_I8: return (uint64_t)(int64_t)((const int8_t *)c.values)[i];
    case T_I16: return (uint64_t)(int64_t)((const int16_t *)c.values)[i];
    case T_I32: case T_DATE32: case T_TIME32: return (uint64_t)(int64_t)((const int32_t *)c.values)[i];
    case T_I64: case T_DATE64: case T_TS: case T_TIME64: case T_U64: case T_F64: return ((const uint64_t *)c.values)[i];
    case T_U8: return ((const uint8_t *)c.values)[i];
    case T_U16: return ((const uint16_t *)c.values)[i];
    case T_U32: case T_F32: return ((const uint32_t *)c.values)[i];
    }
    return 0;
}
DF_D int64_t dict_index(const ColView &c, int64_t i) {
    switch (c.width) {
    case 1: return c.index_signed ? (int64_t)((const int8_t *)c.values)[i] : (int64_t)((const uint8_t *)c.values)[i];
    case 2: return c.index_signed ? (int64_t)((const int16_t *)c.values)[i] : (int64_t)((const uint16_t *)c.values)[i];
    case 4: return c.index_signed ? (int64_t)((const int32_t *)c.values)[i] : (int64_t)((const uint32_t *)c.values)[i];
    default: return ((const int64_t *)c.values)[i];
    }
}
DF_D void col_bytes(const ColView &c, int64_t i, const uint8_t *&p, int64_t &len) {
    if (c.large) {
        const int64_t *o = (const int64_t *)c.offsets;
        p = (const uint8_t *)c.values + o[i];
        len = o[i + 1] - o[i];
    } else {
        const int32_t *o = (const int32_t *)c.offsets;
        p = (const uint8_t *)c.values + o[i];
        len = o[i + 1] - o[i];
    }
}
// hash of one non-null, non-dictionary value
DF_D uint64_t col_hash_value(const ColView &c, int64_t i, uint64_t seed) {
    if (c.type >= T_UTF8 && c.type <= T_LARGE_BINARY) {
        const uint8_t *p; int64_t len;
        col_bytes(c, i, p, len);
        return h_bytes(p, len, seed);
    }
    if (c.type == T_DEC128) {
        const uint64_t *v = (const uint64_t *)c.values + 2 * i;
        return h_u128(v[0], v[1], seed);
    }
    return h_u64(col_bits64(c, i), seed);
}

// ---- 128-bit integers as (lo, hi) pairs -------------------------------------------------------
struct u128_t {
    uint64_t lo, hi;
};
DF_HD u128_t add128(u128_t a, u128_t b) {
    u128_t r;
    r.lo = a.lo + b.lo;
    r.hi = a.hi + b.hi + (r.lo < a.lo ? 1 : 0);
    return r;
}
DF_HD bool lt128_signed(u128_t a, u128_t b) {
    if ((int64_t)a.hi != (int64_t)b.hi) return (int64_t)a.hi < (int64_t)b.hi;
    return a.lo < b.lo;
}
// exact wrapping 128-bit accumulate into global memory with two 64-bit atomics
// (add_wrapping: physical-expr/src/aggregate/sum.rs:137; integer addition is associative, every
// carry out of the low word is forwarded exactly once).
DF_D void atomic_add128(uint64_t *dst, u128_t v) {
    unsigned long long old = atomicAdd((unsigned long long *)dst, (unsigned long long)v.lo);
    uint64_t carry = (old + v.lo < old) ? 1 : 0;
    uint64_t hi = v.hi + carry;
    if (hi) atomicAdd((unsigned long long *)dst + 1, (unsigned long long)hi);
}
// 128-bit compare-and-swap (sm_90+): returns true when the swap happened; `expected` is updated.
DF_D bool atomic_cas128(uint64_t *addr, u128_t &expected, u128_t desired) {
    uint64_t olo, ohi;
    asm volatile(
        "{\n .reg .b128 c, d, o;\n mov.b128 c, {%2,%3};\n mov.b128 d, {%4,%5};\n"
        " atom.global.cas.b128 o, [%6], c, d;\n mov.b128 {%0,%1}, o;\n}\n"
        : "=l"(olo), "=l"(ohi)
        : "l"(expected.lo), "l"(expected.hi), "l"(desired.lo), "l"(desired.hi), "l"(addr)
        : "memory");
    bool ok = (olo == expected.lo && ohi == expected.hi);
    expected.lo = olo;
    expected.hi = ohi;
    return ok;
}

// streaming (read-once) loads that do not pollute L1
DF_D uint64_t ld_stream_u64(const uint64_t *p) {
    uint64_t v;
    asm volatile("ld.global.nc.L1::no_allocate.u64 %0, [%1];" : "=l"(v) : "l"(p));
    return v;
}
DF_D uint4 ld_stream_u128(const void *p) {
    uint4 v;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
                 : "l"(p));
    return v;
}
DF_D uint32_t ld_volatile_u32(const uint32_t *p) {
    uint32_t v;
    asm volatile("ld.volatile.global.u32 %0, [%1];" : "=r"(v) : "l"(p));
    return v;
}
DF_D uint64_t ld_volatile_u64(const uint64_t *p) {
    uint64_t v;
    asm volatile("ld.volatile.global.u64 %0, [%1];" : "=l"(v) : "l"(p));
    return v;
}

}  // namespace dfgpu
