// ops.h — the primitive state operations every GroupsAccumulator on the path decomposes into
// (physical-expr/src/aggregate/{sum,count,average,min_max}.rs + groups_accumulator/prim_op.rs).
#pragma once
#include "common.h"

namespace dfgpu {

enum OpKind : int32_t {
    OP_COUNT = 1,      // state += 1 for every accepted row          (count.rs:110-133)
    OP_ADD_COUNT = 2,  // state += value (merge of partial counts)    (count.rs:135-160, average.rs:483-493)
    OP_ADD_I64 = 3,    // wrapping 64-bit integer sum                 (sum.rs:137, prim_op.rs:94-110)
    OP_ADD_F64 = 4,    // IEEE f64 sum (inputs cast to f64)           (sum.rs:137)
    OP_ADD_I128 = 5,   // wrapping 128-bit integer sum (Decimal128)
    OP_MIN_I64 = 6, OP_MAX_I64 = 7, OP_MIN_U64 = 8, OP_MAX_U64 = 9,
    OP_MIN_F64 = 10, OP_MAX_F64 = 11,  // strict compares, NaN never replaces (min_max.rs:102-139)
    OP_MIN_I128 = 12, OP_MAX_I128 = 13,
    OP_AND_U64 = 14, OP_OR_U64 = 15, OP_XOR_U64 = 16  // BIT_AND/OR/XOR, BOOL_AND/OR (bit_and_or_xor.rs:123-126, bool_op.rs:92-101)
};

struct OpDesc {
    int32_t kind;
    int32_t has_val;   // 0: COUNT(*) — every row counts
    int32_t has_filt;
    int32_t woff;      // first word of this op inside a privatised group record
    int32_t nwords;    // 1, or 2 for 128-bit state
    int32_t seen_bit;  // bit in the record's "seen" word, -1 when the op does not track NullState
    int32_t val_is_f32, pad;
    ColView val, filt;
    uint64_t *state;   // [max_groups * nwords]
    uint8_t *seen;     // [max_groups] NullState.seen_values (accumulate.rs:126-233) or nullptr
    uint64_t init_lo, init_hi;
};

// order-preserving map double -> u64 (so float MIN/MAX can use integer atomics)
#ifdef __CUDACC__
__host__ __device__ __forceinline__ uint64_t f64_to_ordered(double d) {
    uint64_t b;
    memcpy(&b, &d, 8);
    return (b >> 63) ? ~b : (b | 0x8000000000000000ULL);
}
__host__ __device__ __forceinline__ double ordered_to_f64(uint64_t u) {
    uint64_t b = (u >> 63) ? (u & 0x7fffffffffffffffULL) : ~u;
    double d;
    memcpy(&d, &b, 8);
    return d;
}
#endif

}  // namespace dfgpu
