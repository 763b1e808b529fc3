// fastagg.cu — kernels and host driver of the two-level aggregation path (see fastagg.h).
//
// Semantics reproduced per row (all bit-exact with the table path in agg.cu):
//   * which rows update an accumulator: NullState::accumulate (accumulate.rs:147-231);
//   * SUM add_wrapping (sum.rs:137), COUNT (count.rs:110-160), AVG sum+count (average.rs:443-509),
//     MIN/MAX strict compares with T::MAX / T::MIN starting values (min_max.rs:102-139);
//   * NULL group keys form a group (primitive.rs:118-122, row.rs:97).
// Integer state arithmetic is associative, so the order in which partial states are merged cannot
// change a result bit; Float64 sums differ from the row-order sum only within the 1e-9 contract.
#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <mutex>
#include <set>

#include "device.cuh"
#include "fastagg.h"

namespace dfgpu {

static const int FA_T1 = 512;   // level-1 threads per CTA (2 CTAs / SM)
static const int FA_R1 = 4;     // rows per thread per tile
static const int FA_T2 = 512;   // level-2 threads per CTA (2 CTAs / SM: one merges while the other initialises / emits)
static const int FA_KMAX = 1024; // largest per-warp chunk of the record store (records)
static const int FA_CSTRIDE = 32; // scatter cursors live 128 B apart: same-line atomics serialise in L2

struct FaStore {
    uint64_t *recs;
    unsigned long long *cursor;
    uint64_t cap;
    uint32_t *hist;        // [FA_P]
    uint64_t *side_state;  // [2][SW]: group of the NULL key, group of the key image ~0
    uint32_t *side_exists; // [2]
    uint32_t *err;
};
struct FaOut {
    uint64_t *keys;
    uint64_t *state[FA_MAX_OPS];
    uint8_t *seen[FA_MAX_OPS];  // indexed by seen bit
    unsigned long long *gcount;
    uint64_t *ovf;
    unsigned long long *ovf_cursor;
    uint64_t ovf_cap;
    uint32_t *err;
};

// ---- per-row helpers ----------------------------------------------------------------------------
DF_D uint64_t fa_hash(uint64_t key) { return h_u64(key, 0); }

// key image of one row: identical bytes to k_pack_keys (groups.cu) so GroupTable::emit_keys can
// unpack it.  side = 0 for the NULL key of a single 8-byte column, 1 for the image ~0, else -1.
DF_D void fa_load_key(const FaPlan &pl, int64_t row, uint64_t &key, int &side) {
    side = -1;
    if (pl.direct) {
        const ColView &c = pl.keys[0].v;
        if (!col_valid(c, row)) { key = 0; side = 0; return; }
        key = ((const uint64_t *)c.values)[row];
    } else {
        key = pl.key_or;
        for (int k = 0; k < pl.nkeys; k++) {
            const FaKeyCol &kc = pl.keys[k];
            if (kc.null_bit >= 0 && !col_valid(kc.v, row)) {
                key |= 1ULL << (8 * pl.null_off + kc.null_bit);
                continue;
            }
            uint64_t v;  // raw little-endian bytes of the value, zero-extended
            if (kc.kind == 3) {  // short string: its id is the string itself (NULL is STRID_NULL in the field)
                const uint8_t *p; int64_t len;
                if (!col_valid(kc.v, row)) v = STRID_NULL;
                else { col_bytes(kc.v, row, p, len); v = strid_inline(p, (int)len); }
            }
            else if (kc.kind == 1) v = bit_get((const uint8_t *)kc.v.values, kc.v.dbit + row) ? 1 : 0;
            else if (kc.w == 1) v = ((const uint8_t *)kc.v.values)[row];
            else if (kc.w == 2) v = ((const uint16_t *)kc.v.values)[row];
            else if (kc.w == 4) v = ((const uint32_t *)kc.v.values)[row];
            else v = ((const uint64_t *)kc.v.values)[row];
            key |= v << (8 * kc.off);
        }
    }
    if (key == FA_EMPTY) side = 1;
}

DF_D bool fa_accept(const FaOp &op, int64_t row) {
    if (op.has_filt) {
        if (!col_valid(op.filt, row)) return false;
        if (!bit_get((const uint8_t *)op.filt.values, op.filt.dbit + row)) return false;
    }
    if (op.has_val && !col_valid(op.val, row)) return false;
    return true;
}
DF_D double fa_as_f64(const ColView &c, int64_t i) {
    switch (c.type) {
    case T_F64: return ((const double *)c.values)[i];
    case T_F32: return (double)((const float *)c.values)[i];
    case T_U8: case T_U16: case T_U32: case T_U64: return (double)col_bits64(c, i);
    default: return (double)(int64_t)col_bits64(c, i);
    }
}
// Lift an accepted row into the state representation.  Returns false when the value must not be
// combined (NaN under MIN/MAX: the strict compare is never true) — the seen bit is still set.
DF_D bool fa_lift_kv(int kind, const ColView &val, int64_t row, uint64_t &lo, uint64_t &hi) {
    hi = 0;
    switch (kind) {
    case OP_COUNT: lo = 1; return true;
    case OP_ADD_F64: lo = (uint64_t)__double_as_longlong(fa_as_f64(val, row)); return true;
    case OP_ADD_I128: {
        uint4 v = ld_stream_u128((const uint8_t *)val.values + 16 * row);
        lo = (uint64_t)v.x | ((uint64_t)v.y << 32);
        hi = (uint64_t)v.z | ((uint64_t)v.w << 32);
        return true;
    }
    case OP_MIN_F64: case OP_MAX_F64: {
        double d = fa_as_f64(val, row);
        if (!(d == d)) return false;
        lo = f64_to_ordered(d);
        return true;
    }
    default: lo = col_bits64(val, row); return true;  // ADD_COUNT, ADD_I64, MIN/MAX I64/U64
    }
}
DF_D bool fa_lift(const FaOp &op, int64_t row, uint64_t &lo, uint64_t &hi) { return fa_lift_kv(op.kind, op.val, row, lo, hi); }

// ---- combining into shared memory (native 32-bit atomics with exact carry forwarding) ----------
DF_D void sm_add_limbs(uint32_t *p0, uint32_t *p1, uint32_t *p2, uint32_t *p3, uint64_t lo, uint64_t hi, int nlimbs) {
    uint32_t v[4] = {(uint32_t)lo, (uint32_t)(lo >> 32), (uint32_t)hi, (uint32_t)(hi >> 32)};
    uint32_t *p[4] = {p0, p1, p2, p3};
    uint32_t c = 0;
#pragma unroll
    for (int i = 0; i < 4; i++) {
        if (i >= nlimbs) break;
        uint64_t t = (uint64_t)v[i] + c;
        uint32_t a = (uint32_t)t;
        c = (uint32_t)(t >> 32);
        if (a) {
            uint32_t old = atomicAdd(p[i], a);
            c += (old + a < old) ? 1u : 0u;
        }
    }
}
// state word(s) of one slot: w0 (and w1 = w0 + stride for 128-bit state)
DF_D void sm_combine(int kind, uint64_t *w0, uint64_t *w1, uint64_t lo, uint64_t hi) {
    switch (kind) {
    case OP_COUNT: case OP_ADD_COUNT: case OP_ADD_I64:
        sm_add_limbs((uint32_t *)w0, (uint32_t *)w0 + 1, nullptr, nullptr, lo, 0, 2);
        break;
    case OP_ADD_I128:
        sm_add_limbs((uint32_t *)w0, (uint32_t *)w0 + 1, (uint32_t *)w1, (uint32_t *)w1 + 1, lo, hi, 4);
        break;
    case OP_ADD_F64: atomicAdd((double *)w0, __longlong_as_double((long long)lo)); break;
    case OP_MIN_I64: atomicMin((long long *)w0, (long long)lo); break;
    case OP_MAX_I64: atomicMax((long long *)w0, (long long)lo); break;
    case OP_MIN_U64: case OP_MIN_F64: atomicMin((unsigned long long *)w0, (unsigned long long)lo); break;
    case OP_MAX_U64: case OP_MAX_F64: atomicMax((unsigned long long *)w0, (unsigned long long)lo); break;
    }
}
// same operation on a state word in global memory
DF_D void gl_combine(int kind, uint64_t *w, uint64_t lo, uint64_t hi) {
    switch (kind) {
    case OP_COUNT: case OP_ADD_COUNT: case OP_ADD_I64: if (lo) atomicAdd((unsigned long long *)w, (unsigned long long)lo); break;
    case OP_ADD_I128: atomic_add128(w, u128_t{lo, hi}); break;
    case OP_ADD_F64: atomicAdd((double *)w, __longlong_as_double((long long)lo)); break;
    case OP_MIN_I64: atomicMin((long long *)w, (long long)lo); break;
    case OP_MAX_I64: atomicMax((long long *)w, (long long)lo); break;
    case OP_MIN_U64: case OP_MIN_F64: atomicMin((unsigned long long *)w, (unsigned long long)lo); break;
    case OP_MAX_U64: case OP_MAX_F64: atomicMax((unsigned long long *)w, (unsigned long long)lo); break;
    }
}
// plain (non-atomic) combine of two states
DF_D void reg_combine(int kind, uint64_t &lo, uint64_t &hi, uint64_t blo, uint64_t bhi) {
    switch (kind) {
    case OP_COUNT: case OP_ADD_COUNT: case OP_ADD_I64: lo += blo; break;
    case OP_ADD_I128: { u128_t r = add128(u128_t{lo, hi}, u128_t{blo, bhi}); lo = r.lo; hi = r.hi; break; }
    case OP_ADD_F64: lo = (uint64_t)__double_as_longlong(__longlong_as_double((long long)lo) + __longlong_as_double((long long)blo)); break;
    case OP_MIN_I64: if ((int64_t)blo < (int64_t)lo) lo = blo; break;
    case OP_MAX_I64: if ((int64_t)blo > (int64_t)lo) lo = blo; break;
    case OP_MIN_U64: case OP_MIN_F64: if (blo < lo) lo = blo; break;
    case OP_MAX_U64: case OP_MAX_F64: if (blo > lo) lo = blo; break;
    }
}

// ---- shared-memory hash table -------------------------------------------------------------------
// Slots are grouped in buckets of two 64-bit key images (16 bytes, read with one 128-bit load), so a
// warp never walks a divergent probe loop slot by slot.  Nothing is ever deleted and a key is only
// ever inserted into the first free slot of its probe order, so a key that is present sits before
// the first empty slot of that order.
DF_D ulonglong2 fa_ld_bucket(const uint64_t *p) {
    ulonglong2 v;
    asm volatile("ld.volatile.shared.v2.u64 {%0, %1}, [%2];" : "=l"(v.x), "=l"(v.y) : "r"((uint32_t)__cvta_generic_to_shared(p)));
    return v;
}
// one bucket: slot of `key` (inserting it into a free slot if absent), -1 when the bucket is held by other keys
DF_D int fa_bucket_find_insert(uint64_t *hkeys, uint32_t b, uint64_t key) {
    const ulonglong2 kk = fa_ld_bucket(hkeys + 2 * b);
    if (kk.x == key) return (int)(2 * b);
    if (kk.y == key) return (int)(2 * b + 1);
    if (kk.x == FA_EMPTY) {
        unsigned long long old = atomicCAS((unsigned long long *)&hkeys[2 * b], (unsigned long long)FA_EMPTY, (unsigned long long)key);
        if (old == FA_EMPTY || old == key) return (int)(2 * b);
    }
    if (kk.x == FA_EMPTY || kk.y == FA_EMPTY) {  // (a lost race for slot 0 may have left slot 1 free)
        unsigned long long old = atomicCAS((unsigned long long *)&hkeys[2 * b + 1], (unsigned long long)FA_EMPTY, (unsigned long long)key);
        if (old == FA_EMPTY || old == key) return (int)(2 * b + 1);
    }
    return -1;
}
// Level 1 (a cache of hot keys: a miss is spilled, so one bucket is all that is looked at).
DF_D int fa_cache_find_insert(uint64_t *hkeys, int S, uint64_t key, uint64_t h) {
    return fa_bucket_find_insert(hkeys, __umulhi((uint32_t)h, (uint32_t)(S >> 1)), key);
}
DF_D void fa_init_table(const FaPlan &pl, uint64_t *hkeys, uint64_t *hstate, int S, int stride, int tid, int nthreads) {
    for (int s = tid; s < stride; s += nthreads) hkeys[s] = FA_EMPTY;
    for (int o = 0; o < pl.nops; o++) {
        const FaOp &op = pl.ops[o];
        for (int s = tid; s < stride; s += nthreads) {
            hstate[(size_t)op.woff * stride + s] = op.init_lo;
            if (op.nwords == 2) hstate[(size_t)(op.woff + 1) * stride + s] = op.init_hi;
        }
    }
    if (pl.seen_word >= 0)
        for (int s = tid; s < stride; s += nthreads) hstate[(size_t)pl.seen_word * stride + s] = 0;
    (void)S;
}

// ---- level 1: per-CTA pre-aggregation + spill ---------------------------------------------------
// The tile loop has no block-wide barrier: every warp reserves space for its spilled records in
// private chunks of K records (one global atomic per chunk); what is left of a chunk when a tile
// does not fit, and at the end, is filled with hole records (key image FA_EMPTY, skipped by the
// scatter).
__global__ void __launch_bounds__(FA_T1, 2) k_fa_l1(const FaPlan *__restrict__ plan_g, int64_t n, FaStore st, int S, int K,
                                                    const uint32_t *__restrict__ mode_flag, uint32_t my_mode) {
    if (mode_flag && *mode_flag != my_mode) return;
    extern __shared__ __align__(16) uint8_t fa_smem[];
    __shared__ uint32_t s_cnt, s_side[2];
    __shared__ unsigned long long s_base;
    FaPlan &pl = *(FaPlan *)fa_smem;
    const int tid = threadIdx.x, lane = tid & 31;
    for (int i = tid; i < (int)(sizeof(FaPlan) / 4); i += FA_T1) ((uint32_t *)&pl)[i] = ((const uint32_t *)plan_g)[i];
    if (tid == 0) { s_cnt = 0; s_side[0] = 0; s_side[1] = 0; }
    __syncthreads();
    const int stride = S + 2;
    const int SW = pl.SW, RW = pl.RW, nops = pl.nops;
    uint64_t *hkeys = (uint64_t *)(fa_smem + ((sizeof(FaPlan) + 15) & ~(size_t)15));
    uint64_t *hstate = hkeys + stride;
    uint32_t *hist = (uint32_t *)(hstate + (size_t)SW * stride);
    fa_init_table(pl, hkeys, hstate, S, stride, tid, FA_T1);
    for (int p = tid; p < FA_P; p += FA_T1) hist[p] = 0;
    __syncthreads();

    uint64_t w_pos = 0, w_end = 0;  // this warp's current chunk of the record store (warp-uniform)
    bool can_write = true;
    const int64_t tile_rows = (int64_t)FA_T1 * FA_R1;
    // a single 8-byte key column without a validity bitmap: the keys of the next tile are loaded
    // into registers before this tile is worked on
    const bool kdirect = pl.direct && pl.keys[0].v.validity == nullptr;
    const uint64_t *const kvals = (const uint64_t *)pl.keys[0].v.values;
    uint64_t nkey[FA_R1];
    auto load_next_keys = [&](int64_t tile) {
#pragma unroll
        for (int k = 0; k < FA_R1; k++) {
            const int64_t row = tile * tile_rows + tid + (int64_t)k * FA_T1;
            nkey[k] = (kdirect && row < n) ? ld_stream_u64(kvals + row) : 0;
        }
    };
    load_next_keys(blockIdx.x);
    for (int64_t tile = blockIdx.x; tile * tile_rows < n; tile += gridDim.x) {
        const int64_t base = tile * tile_rows + tid;
        uint64_t key[FA_R1];
        int slot[FA_R1];  // >= 0: table slot (S, S+1 = side groups); -1: spill; -2: past the end
        int nspill = 0;
#pragma unroll
        for (int k = 0; k < FA_R1; k++) key[k] = nkey[k];
        load_next_keys(tile + gridDim.x);
#pragma unroll
        for (int k = 0; k < FA_R1; k++) {
            int64_t row = base + (int64_t)k * FA_T1;
            slot[k] = -2;
            if (row < n) {
                int side = -1;
                if (kdirect) { if (key[k] == FA_EMPTY) side = 1; }
                else fa_load_key(pl, row, key[k], side);
                if (side >= 0) {
                    slot[k] = S + side;
                    s_side[side] = 1;
                } else {
                    slot[k] = fa_cache_find_insert(hkeys, S, key[k], fa_hash(key[k]));
                    if (slot[k] < 0) nspill++;
                }
            }
        }
        // reserve space for this warp's spilled records
        uint64_t rec0 = 0;
        if (__any_sync(0xffffffffu, nspill)) {
            int incl = nspill;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                int t = __shfl_up_sync(0xffffffffu, incl, o);
                if (lane >= o) incl += t;
            }
            const uint64_t m = (uint64_t)__shfl_sync(0xffffffffu, incl, 31);
            if (w_pos + m > w_end) {
                if (can_write)
                    for (uint64_t r = w_pos + lane; r < w_end; r += 32) st.recs[r * RW] = FA_EMPTY;
                unsigned long long b = 0;
                if (lane == 0) b = atomicAdd(st.cursor, (unsigned long long)K);
                b = __shfl_sync(0xffffffffu, b, 0);
                if (b + (uint64_t)K > st.cap) {
                    if (lane == 0) atomicExch(st.err, 1u);
                    can_write = false;
                }
                w_pos = b;
                w_end = b + (uint64_t)K;
            }
            rec0 = w_pos + (uint64_t)(incl - nspill);
            w_pos += m;
        }
        uint32_t seenmask[FA_R1];
#pragma unroll
        for (int k = 0; k < FA_R1; k++) seenmask[k] = 0;
        for (int o = 0; o < nops; o++) {
            const FaOp &op = pl.ops[o];
            uint64_t r = rec0;
#pragma unroll
            for (int k = 0; k < FA_R1; k++) {
                if (slot[k] == -2) continue;
                const int64_t row = base + (int64_t)k * FA_T1;
                uint64_t lo = op.init_lo, hi = op.init_hi;
                bool acc = fa_accept(op, row);
                bool comb = acc && fa_lift(op, row, lo, hi);
                if (slot[k] >= 0) {
                    if (acc && op.seen_bit >= 0) seenmask[k] |= 1u << op.seen_bit;
                    if (comb) sm_combine(op.kind, &hstate[(size_t)op.woff * stride + slot[k]], &hstate[(size_t)(op.woff + 1) * stride + slot[k]], lo, hi);
                } else {
                    if (acc && op.seen_bit >= 0) seenmask[k] |= 1u << op.seen_bit;
                    if (!comb) { lo = op.init_lo; hi = op.init_hi; }
                    if (can_write) {
                        st.recs[r * RW + 1 + op.woff] = lo;
                        if (op.nwords == 2) st.recs[r * RW + 2 + op.woff] = hi;
                    }
                    r++;
                }
            }
        }
        {
            uint64_t r = rec0;
#pragma unroll
            for (int k = 0; k < FA_R1; k++) {
                if (slot[k] == -2) continue;
                if (slot[k] >= 0) {
                    if (pl.seen_word >= 0 && seenmask[k]) {
                        uint32_t *sw = (uint32_t *)&hstate[(size_t)pl.seen_word * stride + slot[k]];
                        if ((*(volatile uint32_t *)sw & seenmask[k]) != seenmask[k]) atomicOr(sw, seenmask[k]);
                    }
                } else {
                    if (can_write) {
                        st.recs[r * RW] = key[k];
                        if (pl.seen_word >= 0) st.recs[r * RW + 1 + pl.seen_word] = seenmask[k];
                    }
                    atomicAdd(&hist[fa_hash(key[k]) >> (64 - FA_PBITS)], 1u);
                    r++;
                }
            }
        }
    }
    if (can_write)  // the rest of the warp's last chunk
        for (uint64_t r = w_pos + lane; r < w_end; r += 32) st.recs[r * RW] = FA_EMPTY;
    __syncthreads();
    // ---- flush the table as partial-state records -----------------------------------------------
    {
        int mine = 0;
        for (int s = tid; s < S; s += FA_T1) mine += hkeys[s] != FA_EMPTY;
#pragma unroll
        for (int o = 16; o; o >>= 1) mine += __shfl_xor_sync(0xffffffffu, mine, o);
        if (lane == 0 && mine) atomicAdd(&s_cnt, (uint32_t)mine);
        __syncthreads();
        if (tid == 0) {
            s_base = atomicAdd(st.cursor, (unsigned long long)s_cnt);
            if (s_base + s_cnt > st.cap) atomicExch(st.err, 1u);
            s_cnt = 0;
        }
        __syncthreads();
        const int rounds = (S + FA_T1 - 1) / FA_T1;
        for (int it = 0; it < rounds; it++) {
            int s = it * FA_T1 + tid;
            bool occ = s < S && hkeys[s] != FA_EMPTY;
            uint32_t m = __ballot_sync(0xffffffffu, occ);
            uint32_t off = 0;
            if (lane == 0 && m) off = atomicAdd(&s_cnt, (uint32_t)__popc(m));
            off = __shfl_sync(0xffffffffu, off, 0);
            if (occ) {
                uint64_t r = s_base + off + __popc(m & ((1u << lane) - 1));
                if (r < st.cap) {
                    uint64_t kk = hkeys[s];
                    st.recs[r * RW] = kk;
                    for (int w = 0; w < SW; w++) st.recs[r * RW + 1 + w] = hstate[(size_t)w * stride + s];
                    atomicAdd(&hist[fa_hash(kk) >> (64 - FA_PBITS)], 1u);
                }
            }
        }
    }
    __syncthreads();
    for (int p = tid; p < FA_P; p += FA_T1)
        if (hist[p]) atomicAdd(&st.hist[p], hist[p]);
    // side groups: merge this CTA's partial state into the two global side records
    for (int sd = 0; sd < 2; sd++) {
        if (!s_side[sd]) continue;
        if (tid == 0) st.side_exists[sd] = 1;
        for (int o = tid; o < nops; o += FA_T1) {
            const FaOp &op = pl.ops[o];
            uint64_t lo = hstate[(size_t)op.woff * stride + S + sd];
            uint64_t hi = op.nwords == 2 ? hstate[(size_t)(op.woff + 1) * stride + S + sd] : 0;
            gl_combine(op.kind, &st.side_state[(size_t)sd * SW + op.woff], lo, hi);
        }
        if (tid == 0 && pl.seen_word >= 0) {
            uint64_t w = hstate[(size_t)pl.seen_word * stride + S + sd];
            if (w) atomicOr((unsigned long long *)&st.side_state[(size_t)sd * SW + pl.seen_word], (unsigned long long)w);
        }
    }
}

}  // namespace dfgpu

#include "fastagg_reg.inc"
#include "fastagg_l2.inc"
#include "fastagg_part.inc"
#include "fastagg_host.inc"
#include "fastagg_xchg.inc"
