// agg.cu — grouped hash aggregation on the device: GroupsAccumulator update/merge kernels,
// emit kernels, and (agg_host.inc) the GroupedHashAggregateStream-shaped operator
// (physical-plan/src/aggregates/row_hash.rs:281-662).
//
// Two accumulate paths, chosen ON THE DEVICE from the current number of groups:
//   PRIVATE  (#groups * record size fits shared memory): every thread owns a private copy of all
//            group records in shared memory — no atomics, no bank conflicts (record word w of
//            thread t lives at [w][t]); one warp-reduced flush per block.  This is the Q1 shape
//            (4 groups, 6M+ rows) where global or shared atomics would serialise.
//   GLOBAL   one atomic per (row, state word) into HBM-resident state arrays indexed by group id.
// Both read each input column once, R rows per thread per op so R independent 128-bit loads are
// in flight per thread.
#include <functional>
#include <algorithm>
#include <cfloat>
#include <chrono>

#include "device.cuh"
#include "fastagg.h"
#include "groups.h"
#include "ops.h"

namespace dfgpu {

static const int ACC_T = 128;  // threads per block
static const int ACC_R = 4;    // rows per thread per tile
static const int MAX_WPG = 96; // max 64-bit words in a privatised group record

__device__ __forceinline__ bool op_accept(const OpDesc &op, int64_t row) {
    // NullState::accumulate: non-null AND (no filter OR filter == Some(true)) (accumulate.rs:147-231)
    if (op.has_filt) {
        if (!col_valid(op.filt, row)) return false;
        if (!bit_get((const uint8_t *)op.filt.values, op.filt.dbit + row)) return false;
    }
    if (op.has_val && !col_valid(op.val, row)) return false;
    return true;
}
__device__ __forceinline__ double load_as_f64(const ColView &c, int64_t i) {
    switch (c.type) {
    case T_F64: return ((const double *)c.values)[i];
    case T_F32: return (double)((const float *)c.values)[i];
    case T_U8: case T_U16: case T_U32: case T_U64: return (double)col_bits64(c, i);
    default: return (double)(int64_t)col_bits64(c, i);
    }
}
__device__ __forceinline__ int op_category(int kind) {
    if (kind == OP_COUNT) return 0;
    if (kind == OP_ADD_I128 || kind == OP_MIN_I128 || kind == OP_MAX_I128) return 1;
    if (kind == OP_ADD_F64 || kind == OP_MIN_F64 || kind == OP_MAX_F64) return 2;
    return 3;
}

// a <- a (op) b on raw (non-mapped) state words
__device__ __forceinline__ void combine_raw(int kind, uint64_t &lo, uint64_t &hi, uint64_t blo, uint64_t bhi) {
    switch (kind) {
    case OP_COUNT: case OP_ADD_COUNT: case OP_ADD_I64: lo += blo; break;
    case OP_ADD_F64: lo = (uint64_t)__double_as_longlong(__longlong_as_double((long long)lo) + __longlong_as_double((long long)blo)); break;
    case OP_ADD_I128: { u128_t r = add128(u128_t{lo, hi}, u128_t{blo, bhi}); lo = r.lo; hi = r.hi; break; }
    case OP_MIN_I64: if ((int64_t)lo > (int64_t)blo) lo = blo; break;
    case OP_MAX_I64: if ((int64_t)lo < (int64_t)blo) lo = blo; break;
    case OP_MIN_U64: if (lo > blo) lo = blo; break;
    case OP_MAX_U64: if (lo < blo) lo = blo; break;
    case OP_MIN_F64: if (__longlong_as_double((long long)lo) > __longlong_as_double((long long)blo)) lo = blo; break;
    case OP_MAX_F64: if (__longlong_as_double((long long)lo) < __longlong_as_double((long long)blo)) lo = blo; break;
    case OP_MIN_I128: if (lt128_signed(u128_t{blo, bhi}, u128_t{lo, hi})) { lo = blo; hi = bhi; } break;
    case OP_MAX_I128: if (lt128_signed(u128_t{lo, hi}, u128_t{blo, bhi})) { lo = blo; hi = bhi; } break;
    case OP_AND_U64: lo &= blo; break;
    case OP_OR_U64: lo |= blo; break;
    case OP_XOR_U64: lo ^= blo; break;
    }
}

__device__ __forceinline__ void apply_global(const OpDesc &op, uint32_t g, uint64_t lo, uint64_t hi) {
    uint64_t *st = op.state + (uint64_t)g * op.nwords;
    switch (op.kind) {
    case OP_COUNT: case OP_ADD_COUNT: case OP_ADD_I64: atomicAdd((unsigned long long *)st, (unsigned long long)lo); break;
    case OP_ADD_F64: atomicAdd((double *)st, __longlong_as_double((long long)lo)); break;
    case OP_ADD_I128: atomic_add128(st, u128_t{lo, hi}); break;
    case OP_MIN_I64: atomicMin((long long *)st, (long long)lo); break;
    case OP_MAX_I64: atomicMax((long long *)st, (long long)lo); break;
    case OP_MIN_U64: atomicMin((unsigned long long *)st, (unsigned long long)lo); break;
    case OP_MAX_U64: atomicMax((unsigned long long *)st, (unsigned long long)lo); break;
    case OP_MIN_F64: { double d = __longlong_as_double((long long)lo); if (d == d) atomicMin((unsigned long long *)st, (unsigned long long)f64_to_ordered(d)); break; }
    case OP_MAX_F64: { double d = __longlong_as_double((long long)lo); if (d == d) atomicMax((unsigned long long *)st, (unsigned long long)f64_to_ordered(d)); break; }
    case OP_AND_U64: atomicAnd((unsigned long long *)st, (unsigned long long)lo); break;
    case OP_OR_U64: atomicOr((unsigned long long *)st, (unsigned long long)lo); break;
    case OP_XOR_U64: atomicXor((unsigned long long *)st, (unsigned long long)lo); break;
    case OP_MIN_I128: case OP_MAX_I128: {
        u128_t v{lo, hi};
        u128_t cur{0, 0};
        atomic_cas128(st, cur, cur);  // atomic 128-bit read (two 64-bit loads could tear)
        for (;;) {
            bool better = op.kind == OP_MIN_I128 ? lt128_signed(v, cur) : lt128_signed(cur, v);
            if (!better) break;
            if (atomic_cas128(st, cur, v)) break;  // cur refreshed on failure
        }
        break;
    }
    }
}

template <bool PRIVATE>
__global__ void __launch_bounds__(ACC_T) k_accumulate(const OpDesc *__restrict__ ops, int nops, const uint32_t *__restrict__ gids,
                                                      TableView tv, int64_t n, uint32_t gp_limit, int wpg) {
    extern __shared__ uint64_t sm[];
    __shared__ uint64_t winit[MAX_WPG];
    const uint32_t G = *tv.ngroups;
    if (PRIVATE != (G <= gp_limit)) return;  // the other launch handles this chunk
    const int tid = threadIdx.x;
    if (PRIVATE) {
        if (tid == 0) {
            for (int o = 0; o < nops; o++) {
                winit[ops[o].woff] = ops[o].init_lo;
                if (ops[o].nwords == 2) winit[ops[o].woff + 1] = ops[o].init_hi;
            }
            winit[wpg - 1] = 0;  // seen word
        }
        __syncthreads();
        for (uint32_t idx = tid; idx < G * wpg * ACC_T; idx += ACC_T) sm[idx] = winit[(idx / ACC_T) % wpg];
        // each thread only ever touches its own column [.][tid]; no barrier needed before the main loop
    }
    const int64_t tile_rows = (int64_t)ACC_T * ACC_R;
    for (int64_t tile = blockIdx.x; tile * tile_rows < n; tile += gridDim.x) {
        const int64_t base = tile * tile_rows + tid;
        uint32_t g[ACC_R];
#pragma unroll
        for (int k = 0; k < ACC_R; k++) {
            int64_t row = base + (int64_t)k * ACC_T;
            g[k] = row < n ? resolve_gid(tv, gids[row]) : GID_NONE;
        }
        for (int o = 0; o < nops; o++) {
            const OpDesc op = ops[o];
            uint64_t lo[ACC_R], hi[ACC_R];
            bool acc[ACC_R];
#pragma unroll
            for (int k = 0; k < ACC_R; k++) {
                acc[k] = g[k] != GID_NONE && op_accept(op, base + (int64_t)k * ACC_T);
                lo[k] = 1; hi[k] = 0;
            }
            const int cat = op_category(op.kind);
            if (!PRIVATE && op.kind == OP_COUNT) {
                // COUNT: warp-level pre-aggregation before the global atomic — the lanes of a warp that hit the same
                // group elect one lane, which adds their number (hot keys: one atomic per warp instead of one per row)
#pragma unroll
                for (int k = 0; k < ACC_R; k++) {
                    const unsigned m = __ballot_sync(0xffffffffu, acc[k]);
                    if (!acc[k]) continue;
                    const unsigned peers = __match_any_sync(m, g[k]);
                    if ((tid & 31) == __ffs(peers) - 1) {
                        atomicAdd((unsigned long long *)(op.state + g[k]), (unsigned long long)__popc(peers));
                        if (op.seen) op.seen[g[k]] = 1;
                    }
                }
                continue;
            }
            if (cat == 1) {
#pragma unroll
                for (int k = 0; k < ACC_R; k++)
                    if (acc[k]) {
                        uint4 v = ld_stream_u128((const uint8_t *)op.val.values + 16 * (base + (int64_t)k * ACC_T));
                        lo[k] = (uint64_t)v.x | ((uint64_t)v.y << 32);
                        hi[k] = (uint64_t)v.z | ((uint64_t)v.w << 32);
                    }
            } else if (cat == 2) {
#pragma unroll
                for (int k = 0; k < ACC_R; k++)
                    if (acc[k]) lo[k] = (uint64_t)__double_as_longlong(load_as_f64(op.val, base + (int64_t)k * ACC_T));
            } else if (cat == 3) {
#pragma unroll
                for (int k = 0; k < ACC_R; k++)
                    if (acc[k]) lo[k] = col_bits64(op.val, base + (int64_t)k * ACC_T);
            }
#pragma unroll
            for (int k = 0; k < ACC_R; k++) {
                if (!acc[k]) continue;
                if (PRIVATE) {
                    uint64_t *rec = &sm[((uint64_t)g[k] * wpg + op.woff) * ACC_T + tid];
                    uint64_t a = rec[0], b = op.nwords == 2 ? rec[ACC_T] : 0;
                    if (op.seen_bit >= 0) sm[((uint64_t)g[k] * wpg + wpg - 1) * ACC_T + tid] |= 1ULL << op.seen_bit;
                    if ((op.kind == OP_MIN_F64 || op.kind == OP_MAX_F64)) {
                        double d = __longlong_as_double((long long)lo[k]);
                        if (!(d == d)) continue;  // NaN never replaces (strict compare is false)
                    }
                    combine_raw(op.kind, a, b, lo[k], hi[k]);
                    rec[0] = a;
                    if (op.nwords == 2) rec[ACC_T] = b;
                } else {
                    apply_global(op, g[k], lo[k], hi[k]);
                    if (op.seen) op.seen[g[k]] = 1;
                }
            }
        }
    }
    if (PRIVATE) {
        __syncthreads();
        const int warp = tid >> 5, lane = tid & 31, nw = ACC_T / 32;
        for (uint32_t p = warp; p < G * (uint32_t)nops; p += nw) {
            uint32_t gi = p / nops;
            const OpDesc op = ops[p % nops];
            const uint64_t *rec = &sm[((uint64_t)gi * wpg + op.woff) * ACC_T];
            uint64_t lo = op.init_lo, hi = op.init_hi;
            for (int c = lane; c < ACC_T; c += 32) combine_raw(op.kind, lo, hi, rec[c], op.nwords == 2 ? rec[ACC_T + c] : 0);
#pragma unroll
            for (int o = 16; o; o >>= 1) {
                uint64_t olo = __shfl_xor_sync(0xffffffffu, lo, o), ohi = __shfl_xor_sync(0xffffffffu, hi, o);
                combine_raw(op.kind, lo, hi, olo, ohi);
            }
            if (lane == 0) {
                bool is_add = op.kind <= OP_ADD_I128;
                if (!(is_add && lo == 0 && hi == 0)) apply_global(op, gi, lo, hi);
            }
        }
        for (uint32_t gi = warp; gi < G; gi += nw) {
            const uint64_t *rec = &sm[((uint64_t)gi * wpg + wpg - 1) * ACC_T];
            uint64_t w = 0;
            for (int c = lane; c < ACC_T; c += 32) w |= rec[c];
#pragma unroll
            for (int o = 16; o; o >>= 1) w |= __shfl_xor_sync(0xffffffffu, w, o);
            for (int o = lane; o < nops; o += 32)
                if (ops[o].seen_bit >= 0 && ops[o].seen && ((w >> ops[o].seen_bit) & 1)) ops[o].seen[gi] = 1;
        }
    }
}

// ---- state init / emit kernels -------------------------------------------------------------------
__global__ void k_init_state(uint64_t *state, int nwords, uint64_t lo, uint64_t hi, uint64_t from, uint64_t to) {
    uint64_t i = from + blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
    for (; i < to; i += (uint64_t)gridDim.x * blockDim.x) {
        state[i * nwords] = lo;
        if (nwords == 2) state[i * nwords + 1] = hi;
    }
}
// ---- emit: state -> output columns ---------------------------------------------------------------
// Every output buffer of an emit is produced by ONE launch: the host records one EmitJob per buffer
// (values of a state, AVG results, validity bitmaps from seen bytes), uploads the array once, and
// blockIdx.y picks the job.  (One launch + one allocation per buffer cost ~90 us of host time for
// the 11 columns of a Q1 result - with 4 output rows.)
enum { EJ_VALUES = 0, EJ_AVG_F64 = 1, EJ_AVG_DEC128 = 2, EJ_BITMAP = 3, EJ_BOOL_VALUES = 4 };
struct EmitJob {
    int32_t kind, op_kind, out_type, pad;
    const uint64_t *a;      // state words (values / sums)
    const uint64_t *b;      // counts (AVG)
    const uint8_t *seen;    // seen bytes (AVG, bitmap)
    void *dst;
    uint64_t factor_lo, factor_hi, lim_lo, lim_hi;  // AVG Decimal128: 10^(s_out - s_in), 10^precision - 1
    uint32_t *err;
};
DF_D void ej_bitmap(const uint8_t *__restrict__ seen, int64_t G, uint32_t *__restrict__ words) {
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    int64_t nround = (G + 31) / 32 * 32;
    for (; i < nround; i += (int64_t)gridDim.x * blockDim.x) {
        bool v = i < G && seen[i] != 0;
        uint32_t m = __ballot_sync(0xffffffffu, v);
        if ((threadIdx.x & 31) == 0) words[i >> 5] = m;
    }
}
// state words of a Boolean accumulator (0 / 1) -> bit-packed Boolean values (bool_op.rs:105-121)
DF_D void ej_bool_values(const uint64_t *__restrict__ state, int64_t G, uint32_t *__restrict__ words) {
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    int64_t nround = (G + 31) / 32 * 32;
    for (; i < nround; i += (int64_t)gridDim.x * blockDim.x) {
        bool v = i < G && state[i] != 0;
        uint32_t m = __ballot_sync(0xffffffffu, v);
        if ((threadIdx.x & 31) == 0) words[i >> 5] = m;
    }
}
// state words -> output values of type `out_type` (narrowing / un-mapping)
DF_D void ej_values(const uint64_t *__restrict__ state, int kind, int out_type, int64_t G, uint8_t *__restrict__ out) {
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    for (; i < G; i += (int64_t)gridDim.x * blockDim.x) {
        if (kind == OP_ADD_I128 || kind == OP_MIN_I128 || kind == OP_MAX_I128) {
            ((uint64_t *)out)[2 * i] = state[2 * i];
            ((uint64_t *)out)[2 * i + 1] = state[2 * i + 1];
            continue;
        }
        uint64_t v = state[i];
        if (kind == OP_MIN_F64 || kind == OP_MAX_F64) {
            double d = ordered_to_f64(v);
            if (out_type == T_F32) ((float *)out)[i] = (float)d; else ((double *)out)[i] = d;
            continue;
        }
        switch (type_width_dev(out_type)) {
        case 1: out[i] = (uint8_t)v; break;
        case 2: ((uint16_t *)out)[i] = (uint16_t)v; break;
        case 4: ((uint32_t *)out)[i] = (uint32_t)v; break;
        default: ((uint64_t *)out)[i] = v; break;
        }
    }
}
// AVG evaluate, Float64: sum / count (average.rs:165)
DF_D void ej_avg_f64(const uint64_t *__restrict__ sums, const uint64_t *__restrict__ counts, const uint8_t *__restrict__ seen, int64_t G,
                     double *__restrict__ out) {
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    for (; i < G; i += (int64_t)gridDim.x * blockDim.x)
        out[i] = seen[i] ? __longlong_as_double((long long)sums[i]) / (double)counts[i] : 0.0;
}
// AVG evaluate, Decimal128: DecimalAverager::avg (aggregate/utils.rs:108-124) —
// sum.mul_checked(10^(s_out - s_in)) then truncating division by count, then precision check.
DF_D void ej_avg_dec128(const uint64_t *__restrict__ sums, const uint64_t *__restrict__ counts, const uint8_t *__restrict__ seen, int64_t G,
                        __int128 factor, __int128 lim, uint64_t *__restrict__ out, uint32_t *__restrict__ err) {
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    for (; i < G; i += (int64_t)gridDim.x * blockDim.x) {
        __int128 r = 0;
        if (seen[i]) {
            unsigned __int128 us = ((unsigned __int128)sums[2 * i + 1] << 64) | sums[2 * i];
            __int128 s = (__int128)us;
            // checked multiply: |s| * factor must stay below 2^127
            bool neg = s < 0;
            unsigned __int128 mag = neg ? (unsigned __int128)(-(s + 1)) + 1 : (unsigned __int128)s;
            unsigned __int128 f = (unsigned __int128)factor;
            unsigned __int128 maxmag = neg ? ((unsigned __int128)1 << 127) : (((unsigned __int128)1 << 127) - 1);
            if (mag != 0 && f > maxmag / mag) { atomicExch(err, 1u); continue; }
            __int128 value = s * factor;
            __int128 c = (__int128)counts[i];
            __int128 nv = value / c;  // div_wrapping == truncation toward zero
            if (nv > lim || nv < -lim) { atomicExch(err, 1u); continue; }
            r = nv;
        }
        out[2 * i] = (uint64_t)(unsigned __int128)r;
        out[2 * i + 1] = (uint64_t)((unsigned __int128)r >> 64);
    }
}
__global__ void __launch_bounds__(256) k_emit_jobs(const EmitJob *__restrict__ jobs, int64_t G) {
    const EmitJob j = jobs[blockIdx.y];
    switch (j.kind) {
    case EJ_VALUES: ej_values(j.a, j.op_kind, j.out_type, G, (uint8_t *)j.dst); break;
    case EJ_AVG_F64: ej_avg_f64(j.a, j.b, j.seen, G, (double *)j.dst); break;
    case EJ_AVG_DEC128: {
        __int128 factor = (__int128)(((unsigned __int128)j.factor_hi << 64) | j.factor_lo);
        __int128 lim = (__int128)(((unsigned __int128)j.lim_hi << 64) | j.lim_lo);
        ej_avg_dec128(j.a, j.b, j.seen, G, factor, lim, (uint64_t *)j.dst, j.err);
        break;
    }
    case EJ_BOOL_VALUES: ej_bool_values(j.a, G, (uint32_t *)j.dst); break;
    default: ej_bitmap(j.seen, G, (uint32_t *)j.dst); break;
    }
}

}  // namespace dfgpu

#include "agg_host.inc"
