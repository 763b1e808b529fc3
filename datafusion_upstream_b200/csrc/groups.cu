// groups.cu — GroupTable kernels and host logic (see groups.h for the design).
#include "groups.h"

#include <algorithm>

#include "device.cuh"

namespace dfgpu {

static const uint64_t U64_EMPTY = ~0ULL;
static const uint64_t ROW_SEED = 0x5eed5eed5eed5eedULL;

// ------------------------------------------------------------------------------------------------
// side slots: index cap + 0 = NULL key group, cap + 1 = key whose bit pattern equals U64_EMPTY
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t side_group(const TableView &t, int which, bool find_only) {
    uint64_t s = t.cap + which;
    if (t.layout == 0) {
        uint64_t *kp = &t.slots[2 * s];
        uint32_t *gp = (uint32_t *)&t.slots[2 * s + 1];
        uint64_t k = *(volatile uint64_t *)kp;
        if (k == U64_EMPTY) {
            if (find_only) return GID_NONE;
            uint64_t old = atomicCAS((unsigned long long *)kp, (unsigned long long)U64_EMPTY, 0ULL);
            if (old == U64_EMPTY) {
                uint32_t g = atomicAdd(t.ngroups, 1u);
                if (g >= t.max_groups) { *t.overflow = 1; return GID_NONE; }
                t.gkeys[g] = which ? U64_EMPTY : 0ULL;
                *(volatile uint32_t *)gp = g;
                return g;
            }
        }
        uint32_t g = *(volatile uint32_t *)gp;
        return g == GID_UNSET ? (GID_PENDING | (uint32_t)s) : g;
    } else {
        uint64_t *wp = &t.slots[s];
        uint64_t e = *(volatile uint64_t *)wp;
        if (e == 0) {
            if (find_only) return GID_NONE;
            uint64_t want = (1ULL << 32) | GID_PENDING;
            uint64_t old = atomicCAS((unsigned long long *)wp, 0ULL, (unsigned long long)want);
            if (old == 0) {
                uint32_t g = atomicAdd(t.ngroups, 1u);
                if (g >= t.max_groups) { *t.overflow = 1; return GID_NONE; }
                if (t.glen) { t.glen[g] = 0; t.goff[g] = 0; }
                __threadfence();
                *(volatile uint64_t *)wp = (1ULL << 32) | g;
                return g;
            }
            e = old;
        }
        uint32_t pl = (uint32_t)e;
        return (pl & GID_PENDING) ? (GID_PENDING | (uint32_t)s) : pl;
    }
}

// ------------------------------------------------------------------------------------------------
// U64 layout: GroupValuesPrimitive::intern (group_values/primitive.rs:112-150)
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_intern_u64(TableView t, const uint64_t *__restrict__ keys,
                                                    const uint8_t *__restrict__ validity, int64_t vbit,
                                                    const uint8_t *__restrict__ excl, int64_t n, uint32_t *__restrict__ out,
                                                    int find_only, int null_policy) {
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; i < n; i += stride) {
        if (excl && excl[i]) { out[i] = GID_NONE; continue; }
        if (validity && !bit_get(validity, vbit + i)) {
            out[i] = null_policy == NULLS_EXCLUDED ? GID_NONE : side_group(t, 0, find_only);
            continue;
        }
        uint64_t key = ld_stream_u64(keys + i);
        if (key == U64_EMPTY) { out[i] = side_group(t, 1, find_only); continue; }
        uint64_t s = (t.force_collisions ? 0 : mix64(key)) & t.mask;
        uint32_t res;
        for (;;) {
            uint64_t *kp = &t.slots[2 * s];
            uint64_t k = *kp;
            if (k == U64_EMPTY) {
                if (find_only) { res = GID_NONE; break; }
                uint64_t old = atomicCAS((unsigned long long *)kp, (unsigned long long)U64_EMPTY, (unsigned long long)key);
                if (old == U64_EMPTY) {
                    uint32_t g = atomicAdd(t.ngroups, 1u);  // dense id (primitive.rs:136-140)
                    if (g >= t.max_groups) { *t.overflow = 1; res = GID_NONE; break; }
                    t.gkeys[g] = key;
                    *(volatile uint32_t *)&t.slots[2 * s + 1] = g;
                    res = g;
                    break;
                }
                k = old;
            }
            if (k == key) {
                uint32_t g = *(const uint32_t *)&t.slots[2 * s + 1];
                res = g == GID_UNSET ? (GID_PENDING | (uint32_t)s) : g;
                break;
            }
            s = (s + 1) & t.mask;
        }
        out[i] = res;
    }
}

__global__ void k_rehash_u64(TableView t, uint32_t G, uint32_t skip0, uint32_t skip1) {
    uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;
    for (; g < G; g += gridDim.x * blockDim.x) {
        if (g == skip0 || g == skip1) continue;
        uint64_t key = t.gkeys[g];
        uint64_t s = (t.force_collisions ? 0 : mix64(key)) & t.mask;
        for (;;) {
            uint64_t old = atomicCAS((unsigned long long *)&t.slots[2 * s], (unsigned long long)U64_EMPTY, (unsigned long long)key);
            if (old == U64_EMPTY) { *(uint32_t *)&t.slots[2 * s + 1] = g; break; }
            s = (s + 1) & t.mask;
        }
    }
}

// ------------------------------------------------------------------------------------------------
// ROW / VAR layout: GroupValuesRows::intern (row.rs:94-146), ArrowBytesMap::insert_if_new
// (binary_map.rs:333-472)
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint64_t hash_row_words(const uint8_t *p, int W) {
    uint64_t h = ROW_SEED;
    for (int k = 0; k < W; k += 8) h = mix64(h ^ *(const uint64_t *)(p + k));
    return h;
}
__device__ __forceinline__ bool words_equal(const uint8_t *a, const uint8_t *b, int W) {
    for (int k = 0; k < W; k += 8)
        if (*(const uint64_t *)(a + k) != *(const uint64_t *)(b + k)) return false;
    return true;
}

// One key (kp, klen) of row i -> group id (or a pending reference): probe, insert on a free slot, compare on a tag match.
template <bool VAR>
__device__ __forceinline__ uint32_t intern_one(const TableView &t, const uint8_t *__restrict__ packed, const ColView &str, int64_t i,
                                               const uint8_t *kp, int64_t klen, int W, int find_only) {
    uint64_t h = t.force_collisions ? 0 : (VAR ? h_bytes(kp, klen, ROW_SEED) : hash_row_words(kp, W));
    uint32_t tag = (uint32_t)(h >> 32) | 1u;
    uint64_t s = h & t.mask;
    for (;;) {
        uint64_t e = t.slots[s];
        if (e == 0) {
            if (find_only) return GID_NONE;
            uint64_t want = ((uint64_t)tag << 32) | (GID_PENDING | (uint32_t)i);
            uint64_t old = atomicCAS((unsigned long long *)&t.slots[s], 0ULL, (unsigned long long)want);
            if (old == 0) {
                uint32_t g = atomicAdd(t.ngroups, 1u);
                if (g >= t.max_groups) { *t.overflow = 1; return GID_NONE; }
                if (VAR) {
                    // heap entries start 8-byte aligned (the length is rounded up): whole-word copy here, one aligned load
                    // per word when the entry is compared or re-hashed later
                    unsigned long long pos = atomicAdd(t.heap_used, (unsigned long long)((klen + 7) & ~(int64_t)7));
                    for (int64_t k = 0; k < klen; k += 8)
                        *(uint64_t *)(t.heap + pos + k) = load_le(kp + k, (int)(klen - k < 8 ? klen - k : 8));
                    t.goff[g] = (int64_t)pos;
                    t.glen[g] = (int32_t)klen;
                } else {
                    uint8_t *dst = t.grows + (uint64_t)g * W;
                    for (int k = 0; k < W; k += 8) *(uint64_t *)(dst + k) = *(const uint64_t *)(kp + k);
                }
                __threadfence();
                *(volatile uint64_t *)&t.slots[s] = ((uint64_t)tag << 32) | g;
                return g;
            }
            e = old;
        }
        if ((uint32_t)(e >> 32) == tag) {
            uint32_t pl = (uint32_t)e;
            bool eq;
            if (pl & GID_PENDING) {  // compare with the inserting row of this batch (immutable input)
                int64_t j = pl & 0x7fffffffu;
                if (VAR) {
                    const uint8_t *op; int64_t ol;
                    col_bytes(str, j, op, ol);
                    eq = ol == klen && bytes_equal(kp, op, klen);
                } else {
                    eq = words_equal(kp, packed + j * W, W);
                }
            } else if (VAR) {
                // the key store is written by other SMs during this kernel: bypass the (non-coherent) L1
                eq = __ldcg(&t.glen[pl]) == (int32_t)klen;
                if (eq) eq = bytes_equal<true>(kp, t.heap + __ldcg(&t.goff[pl]), klen);
            } else {
                const uint8_t *sp = t.grows + (uint64_t)pl * W;
                eq = true;
                for (int k = 0; k < W && eq; k += 8) eq = __ldcg((const uint64_t *)(sp + k)) == *(const uint64_t *)(kp + k);
            }
            if (eq) return (pl & GID_PENDING) ? (GID_PENDING | (uint32_t)s) : pl;
        }
        s = (s + 1) & t.mask;
    }
}

// VAR: rows are handled in tiles of IR_TILE.  Empty strings, NULLs and excluded rows finish in the first sweep (no
// loops); the rows that carry bytes are first compacted into a shared-memory list and then processed one per thread,
// so a warp hashes 32 strings at a time instead of the few non-empty ones among 32 consecutive rows (ClickBench's
// SearchPhrase is empty in 70 % of the rows: ncu showed 5.2 active lanes per instruction before the compaction).
static const int IR_R = 8, IR_TILE = 256 * IR_R;
template <bool VAR>
__global__ void __launch_bounds__(256) k_intern_rows(TableView t, const uint8_t *__restrict__ packed, ColView str,
                                                     const uint8_t *__restrict__ excl, int64_t n, uint32_t *__restrict__ out,
                                                     int find_only, int null_policy) {
    const int W = t.W;
    if (!VAR) {
        int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
        const int64_t stride = (int64_t)gridDim.x * blockDim.x;
        for (; i < n; i += stride) {
            if (excl && excl[i]) { out[i] = GID_NONE; continue; }
            out[i] = intern_one<false>(t, packed, str, i, packed + i * W, W, W, find_only);
        }
        return;
    }
    __shared__ uint32_t s_list[IR_TILE];
    __shared__ uint32_t s_cnt;
    const int tid = threadIdx.x, lane = tid & 31;
    for (int64_t tile0 = (int64_t)blockIdx.x * IR_TILE; tile0 < n; tile0 += (int64_t)gridDim.x * IR_TILE) {
        if (tid == 0) s_cnt = 0;
        __syncthreads();
#pragma unroll 2
        for (int k = 0; k < IR_R; k++) {
            const int64_t i = tile0 + (int64_t)k * 256 + tid;
            bool defer = false;
            if (i < n) {
                if (excl && excl[i]) {
                    out[i] = GID_NONE;
                } else if (!col_valid(str, i)) {  // single null entry (binary_map.rs:360-373)
                    out[i] = null_policy == NULLS_EXCLUDED ? GID_NONE : side_group(t, 0, find_only);
                } else {
                    const uint8_t *kp;
                    int64_t klen;
                    col_bytes(str, i, kp, klen);
                    if (klen == 0) out[i] = intern_one<true>(t, packed, str, i, kp, 0, W, find_only);
                    else defer = true;
                }
            }
            const unsigned m = __ballot_sync(0xffffffffu, defer);
            if (m) {
                uint32_t base = 0;
                if (lane == 0) base = atomicAdd(&s_cnt, (uint32_t)__popc(m));
                base = __shfl_sync(0xffffffffu, base, 0);
                if (defer) s_list[base + __popc(m & ((1u << lane) - 1))] = (uint32_t)(i - tile0);
            }
        }
        __syncthreads();
        const uint32_t cnt = s_cnt;
        for (uint32_t j = tid; j < cnt; j += 256) {
            const int64_t i = tile0 + s_list[j];
            const uint8_t *kp;
            int64_t klen;
            col_bytes(str, i, kp, klen);
            out[i] = intern_one<true>(t, packed, str, i, kp, klen, W, find_only);
        }
        __syncthreads();
    }
}

template <bool VAR> __global__ void k_rehash_rows(TableView t, uint32_t G, uint32_t skip0) {
    uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;
    for (; g < G; g += gridDim.x * blockDim.x) {
        if (g == skip0) continue;
        uint64_t h;
        if (t.force_collisions) h = 0;
        else if (VAR) h = h_bytes(t.heap + t.goff[g], t.glen[g], ROW_SEED);
        else h = hash_row_words(t.grows + (uint64_t)g * t.W, t.W);
        uint32_t tag = (uint32_t)(h >> 32) | 1u;
        uint64_t s = h & t.mask;
        uint64_t want = ((uint64_t)tag << 32) | g;
        for (;;) {
            uint64_t old = atomicCAS((unsigned long long *)&t.slots[s], 0ULL, (unsigned long long)want);
            if (old == 0) break;
            s = (s + 1) & t.mask;
        }
    }
}

__global__ void k_resolve_ids(TableView t, uint32_t *ids, int64_t n) {
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    for (; i < n; i += (int64_t)gridDim.x * blockDim.x) ids[i] = resolve_gid(t, ids[i]);
}

// ------------------------------------------------------------------------------------------------
// key packing: fixed-width columns (+ per-column string ids) -> one fixed-width row per input row
// (stands in for RowConverter::convert_columns, row.rs:97; only equality matters)
// ------------------------------------------------------------------------------------------------
struct PackCol {
    ColView v;            // kind 0/1
    const uint32_t *ids;  // kind 2
    int32_t kind, off, w, null_bit;
};
struct PackPlan {
    PackCol c[MAX_KEY_COLS];
    int32_t ncols, W, null_off;
};

__global__ void __launch_bounds__(256) k_pack_keys(PackPlan p, int64_t n, uint8_t *__restrict__ out, uint8_t *__restrict__ excl,
                                                   int null_policy) {
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; i < n; i += stride) {
        uint8_t *row = out + i * p.W;
        for (int k = 0; k < p.W; k += 8) *(uint64_t *)(row + k) = 0;
        bool any_null = false, miss = false;
        for (int c = 0; c < p.ncols; c++) {
            const PackCol &pc = p.c[c];
            if (pc.kind == 2) {
                uint32_t id = pc.ids[i];
                if (id == GID_NONE) miss = true;            // find-only: string absent from the table
                else if (id == STRID_NULL) any_null = true;  // NULL is encoded in the id field itself
                *(uint32_t *)(row + pc.off) = id;
                continue;
            }
            if (pc.null_bit >= 0 && !col_valid(pc.v, i)) {
                any_null = true;
                row[p.null_off + (pc.null_bit >> 3)] |= (uint8_t)(1u << (pc.null_bit & 7));
                continue;
            }
            if (pc.kind == 1) {
                row[pc.off] = bit_get((const uint8_t *)pc.v.values, pc.v.dbit + i) ? 1 : 0;
            } else if (pc.w == 16) {
                const uint64_t *s = (const uint64_t *)pc.v.values + 2 * i;
                *(uint64_t *)(row + pc.off) = s[0];
                *(uint64_t *)(row + pc.off + 8) = s[1];
            } else if (pc.w == 8) {
                *(uint64_t *)(row + pc.off) = ((const uint64_t *)pc.v.values)[i];
            } else if (pc.w == 4) {
                *(uint32_t *)(row + pc.off) = ((const uint32_t *)pc.v.values)[i];
            } else if (pc.w == 2) {
                *(uint16_t *)(row + pc.off) = ((const uint16_t *)pc.v.values)[i];
            } else {
                row[pc.off] = ((const uint8_t *)pc.v.values)[i];
            }
        }
        if (excl) excl[i] = (miss || (null_policy == NULLS_EXCLUDED && any_null)) ? 1 : 0;
    }
}

// rows [G x stride bytes] -> one output column (values + validity words)
__global__ void k_unpack_col(const uint8_t *__restrict__ rows, int stride_bytes, int off, int w, int kind, int null_bit, int null_off,
                             int64_t G, uint8_t *__restrict__ out, uint32_t *__restrict__ valid_words, uint32_t *__restrict__ bool_words) {
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    int64_t nround = (G + 31) / 32 * 32;
    for (; i < nround; i += stride) {
        bool valid = false, bv = false;
        if (i < G) {
            const uint8_t *row = rows + i * stride_bytes;
            valid = null_bit < 0 || !((row[null_off + (null_bit >> 3)] >> (null_bit & 7)) & 1);
            if (kind == 1) bv = valid && row[off] != 0;
            else if (w == 16) { ((uint64_t *)out)[2 * i] = *(const uint64_t *)(row + off); ((uint64_t *)out)[2 * i + 1] = *(const uint64_t *)(row + off + 8); }
            else if (w == 8) ((uint64_t *)out)[i] = *(const uint64_t *)(row + off);
            else if (w == 4) ((uint32_t *)out)[i] = *(const uint32_t *)(row + off);
            else if (w == 2) ((uint16_t *)out)[i] = *(const uint16_t *)(row + off);
            else out[i] = row[off];
        }
        uint32_t m = __ballot_sync(0xffffffffu, valid), b = __ballot_sync(0xffffffffu, bv);
        if ((threadIdx.x & 31) == 0) {
            if (valid_words) valid_words[i >> 5] = m;
            if (bool_words) bool_words[i >> 5] = b;
        }
    }
}

// Dictionary column -> per-row string ids via the ids of the dictionary values
__global__ void k_map_dict_ids(ColView idx, const uint32_t *__restrict__ dict_ids, int64_t n, uint32_t *__restrict__ out,
                               uint32_t null_id) {
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    for (; i < n; i += (int64_t)gridDim.x * blockDim.x) out[i] = col_valid(idx, i) ? dict_ids[dict_index(idx, i)] : null_id;
}
__global__ void k_ids_to_i64(const uint32_t *ids, int64_t n, int64_t *out) {
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    for (; i < n; i += (int64_t)gridDim.x * blockDim.x) out[i] = ids[i] == STRID_NULL ? -1 : (int64_t)ids[i];
}
__global__ void k_iota_g(void *dst, int width, int64_t n) {
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    for (; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        if (width == 1) ((uint8_t *)dst)[i] = (uint8_t)i;
        else if (width == 2) ((uint16_t *)dst)[i] = (uint16_t)i;
        else if (width == 4) ((uint32_t *)dst)[i] = (uint32_t)i;
        else ((uint64_t *)dst)[i] = (uint64_t)i;
    }
}
__global__ void k_var_lengths(const int32_t *glen, int64_t G, int64_t *out) {
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    for (; i < G; i += (int64_t)gridDim.x * blockDim.x) out[i] = glen[i];
}
__global__ void k_var_gather(const uint8_t *__restrict__ heap, const int64_t *__restrict__ goff, const int32_t *__restrict__ glen,
                             const int64_t *__restrict__ starts, int64_t G, int64_t total, uint8_t *__restrict__ data,
                             void *__restrict__ offsets, int large) {
    int lane = threadIdx.x & 31;
    int64_t w = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
    int64_t nw = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t g = w; g < G; g += nw) {
        int64_t st = starts[g];
        if (lane == 0) {
            if (large) ((int64_t *)offsets)[g] = st; else ((int32_t *)offsets)[g] = (int32_t)st;
            if (g == G - 1) { if (large) ((int64_t *)offsets)[G] = total; else ((int32_t *)offsets)[G] = (int32_t)total; }
        }
        const uint8_t *src = heap + goff[g];
        int len = glen[g];
        for (int k = lane; k < len; k += 32) data[st + k] = src[k];
    }
}
__global__ void k_clear_bit(uint32_t *words, uint32_t bit) { words[bit >> 5] &= ~(1u << (bit & 31)); }
__global__ void k_fill_u32(uint32_t *p, uint32_t v, int64_t n) {
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    for (; i < n; i += (int64_t)gridDim.x * blockDim.x) p[i] = v;
}

}  // namespace dfgpu

#include "groups_host.inc"
