// df_oracle.cpp — CPU ORACLE (test infrastructure, NOT product code).
//
// Single-threaded restatement of the reference's hash-table operator path, used only as a
// checker (tests/, smoke()) and as the timed CPU baseline (bench.py).  See df_oracle.h.
// Each function cites the reference file:line (relative to /root/reference/datafusion) it follows.
//
// Parity: PINNED — checked against the reference's literal test vectors in tests/golden/
// (aggregates/mod.rs check_aggregates + check_grouping_sets, average.rs avg_decimal*,
// decimal.slt truncation golden, hash_join.rs join_* tables incl. order-sensitive inner joins).
#include "df_oracle.h"

#include <algorithm>
#include <atomic>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <limits>
#include <memory>
#include <string>
#include <thread>
#include <vector>

typedef __int128 i128;
typedef unsigned __int128 u128;

static thread_local std::string g_err;
extern "C" const char *oracle_last_error(void) { return g_err.c_str(); }
static int fail(const std::string &m) {
    g_err = m;
    return -1;
}

// ------------------------------------------------------------------------------------------
// Arrow buffer access
// ------------------------------------------------------------------------------------------
static inline bool get_bit(const uint8_t *b, int64_t i) { return (b[i >> 3] >> (i & 7)) & 1; }
static inline void set_bit(uint8_t *b, int64_t i) { b[i >> 3] |= (uint8_t)(1u << (i & 7)); }

static int type_width(int t) {
    switch (t) {
    case OT_I8: case OT_U8: return 1;
    case OT_I16: case OT_U16: return 2;
    case OT_I32: case OT_U32: case OT_F32: case OT_DATE32: case OT_TIME32: return 4;
    case OT_I64: case OT_U64: case OT_F64: case OT_DATE64: case OT_TS: case OT_TIME64: return 8;
    case OT_DEC128: return 16;
    default: return 0;
    }
}
static bool is_var(int t) { return t == OT_UTF8 || t == OT_LARGE_UTF8 || t == OT_BINARY || t == OT_LARGE_BINARY; }
static bool is_large(int t) { return t == OT_LARGE_UTF8 || t == OT_LARGE_BINARY; }
static bool is_signed_int(int t) {
    return t == OT_I8 || t == OT_I16 || t == OT_I32 || t == OT_I64 || t == OT_DATE32 || t == OT_DATE64 ||
           t == OT_TS || t == OT_TIME32 || t == OT_TIME64;
}
static bool is_unsigned_int(int t) { return t == OT_U8 || t == OT_U16 || t == OT_U32 || t == OT_U64; }

static inline int64_t read_index(const ocol *c, int64_t i) {
    const void *p = c->values;
    int64_t j = c->offset + i;
    switch (c->index_type) {
    case OT_I8: return ((const int8_t *)p)[j];
    case OT_I16: return ((const int16_t *)p)[j];
    case OT_I32: return ((const int32_t *)p)[j];
    case OT_I64: return ((const int64_t *)p)[j];
    case OT_U8: return ((const uint8_t *)p)[j];
    case OT_U16: return ((const uint16_t *)p)[j];
    case OT_U32: return ((const uint32_t *)p)[j];
    case OT_U64: return (int64_t)((const uint64_t *)p)[j];
    }
    return 0;
}

// Resolve logical row i of column c to a (value column, physical row) pair; false when the
// logical value is NULL (dictionary: null key OR null dictionary value — `logical_nulls`,
// physical-expr/src/aggregate/count.rs:110-133).
static inline bool resolve(const ocol *c, int64_t i, const ocol *&vc, int64_t &vi) {
    if (c->validity && !get_bit(c->validity, c->offset + i)) return false;
    if (c->type == OT_DICT) {
        int64_t k = read_index(c, i);
        const ocol *d = c->dict;
        if (d->validity && !get_bit(d->validity, d->offset + k)) return false;
        vc = d;
        vi = d->offset + k;
        return true;
    }
    vc = c;
    vi = c->offset + i;
    return true;
}
static inline int value_type(const ocol *c) { return c->type == OT_DICT ? c->dict->type : c->type; }

static inline uint64_t bits64(const ocol *vc, int64_t vi) {
    switch (vc->type) {
    case OT_BOOL: return get_bit((const uint8_t *)vc->values, vi) ? 1 : 0;
    case OT_I8: return (uint64_t)(int64_t)((const int8_t *)vc->values)[vi];
    case OT_I16: return (uint64_t)(int64_t)((const int16_t *)vc->values)[vi];
    case OT_I32: case OT_DATE32: case OT_TIME32: return (uint64_t)(int64_t)((const int32_t *)vc->values)[vi];
    case OT_I64: case OT_DATE64: case OT_TS: case OT_TIME64: return (uint64_t)((const int64_t *)vc->values)[vi];
    case OT_U8: return ((const uint8_t *)vc->values)[vi];
    case OT_U16: return ((const uint16_t *)vc->values)[vi];
    case OT_U32: case OT_F32: return ((const uint32_t *)vc->values)[vi];
    case OT_U64: case OT_F64: return ((const uint64_t *)vc->values)[vi];
    }
    return 0;
}
static inline i128 get_i128(const ocol *vc, int64_t vi) {
    i128 v;
    memcpy(&v, (const uint8_t *)vc->values + 16 * vi, 16);
    return v;
}
static inline void get_bytes(const ocol *vc, int64_t vi, const uint8_t *&p, int64_t &len) {
    if (is_large(vc->type)) {
        const int64_t *o = (const int64_t *)vc->offsets;
        p = (const uint8_t *)vc->values + o[vi];
        len = o[vi + 1] - o[vi];
    } else {
        const int32_t *o = (const int32_t *)vc->offsets;
        p = (const uint8_t *)vc->values + o[vi];
        len = o[vi + 1] - o[vi];
    }
}

// ------------------------------------------------------------------------------------------
// Hashing — structure of create_hashes (common/src/hash_utils.rs:357-416): column 0 assigns,
// columns >= 1 combine (hash_utils.rs:38-41); NULL rows leave the running hash untouched
// (:117-131, :197-207); floats hash their bit pattern (:78-87); dictionaries hash the value
// (:182-210).  The value mixer ("dfhash v1") replaces ahash (values unobservable).
// ------------------------------------------------------------------------------------------
static inline uint64_t mix64(uint64_t x) {
    x ^= x >> 30; x *= 0xbf58476d1ce4e5b9ULL;
    x ^= x >> 27; x *= 0x94d049bb133111ebULL;
    x ^= x >> 31;
    return x;
}
static inline uint64_t h_u64(uint64_t v, uint64_t seed) { return mix64(v + 0x9e3779b97f4a7c15ULL + seed); }
static inline uint64_t h_u128(uint64_t lo, uint64_t hi, uint64_t seed) {
    return mix64(h_u64(lo, seed) ^ (hi * 0xff51afd7ed558ccdULL));
}
static inline uint64_t h_bytes(const uint8_t *p, int64_t len, uint64_t seed) {
    uint64_t h = h_u64((uint64_t)len, seed);
    int64_t i = 0;
    for (; i + 8 <= len; i += 8) {
        uint64_t w;
        memcpy(&w, p + i, 8);
        h = mix64(h ^ w);
    }
    if (i < len) {
        uint64_t w = 0;
        memcpy(&w, p + i, (size_t)(len - i));
        h = mix64(h ^ w);
    }
    return h;
}
static inline uint64_t combine_hashes(uint64_t l, uint64_t r) { return (17ULL * 37ULL + l) * 37ULL + r; }

static inline uint64_t hash_value(const ocol *vc, int64_t vi, uint64_t seed) {
    if (is_var(vc->type)) {
        const uint8_t *p; int64_t len;
        get_bytes(vc, vi, p, len);
        return h_bytes(p, len, seed);
    }
    if (vc->type == OT_DEC128) {
        u128 v = (u128)get_i128(vc, vi);
        return h_u128((uint64_t)v, (uint64_t)(v >> 64), seed);
    }
    return h_u64(bits64(vc, vi), seed);
}

static void create_hashes(const ocol *const *cols, int ncols, uint64_t seed, bool force_collisions,
                          int64_t n, uint64_t *out) {
    for (int64_t i = 0; i < n; i++) out[i] = 0;
    if (force_collisions) return;  // hash_utils.rs:307-317 (feature force_hash_collisions)
    for (int c = 0; c < ncols; c++) {
        const ocol *col = cols[c];
        bool rehash = c >= 1;
        if (col->type == OT_DICT) {
            // hash_dictionary: a valid key pointing at a NULL dictionary value contributes the
            // value's (unchanged, zero) hash — hash_utils.rs:188-208.
            const ocol *d = col->dict;
            for (int64_t i = 0; i < n; i++) {
                if (col->validity && !get_bit(col->validity, col->offset + i)) continue;
                int64_t k = read_index(col, i);
                uint64_t dh = 0;
                if (!(d->validity && !get_bit(d->validity, d->offset + k))) dh = hash_value(d, d->offset + k, seed);
                out[i] = rehash ? combine_hashes(dh, out[i]) : dh;
            }
            continue;
        }
        for (int64_t i = 0; i < n; i++) {
            if (col->validity && !get_bit(col->validity, col->offset + i)) continue;
            uint64_t h = hash_value(col, col->offset + i, seed);
            out[i] = rehash ? combine_hashes(h, out[i]) : h;
        }
    }
}

extern "C" int oracle_create_hashes(const ocol *const *cols, int ncols, uint64_t seed, int force_collisions,
                                    uint64_t *out) {
    if (ncols <= 0) return fail("no columns");
    create_hashes(cols, ncols, seed, force_collisions != 0, cols[0]->length, out);
    return 0;
}

// ------------------------------------------------------------------------------------------
// Output columns
// ------------------------------------------------------------------------------------------
struct OutCol {
    int type = 0, precision = 0, scale = 0;
    int64_t length = 0, null_count = 0;
    std::vector<uint8_t> values;
    std::vector<uint8_t> validity;  // bitmap, empty when no nulls
    std::vector<int64_t> offsets;   // var-width
    void set_validity(const std::vector<uint8_t> &valid_flags) {
        length = (int64_t)valid_flags.size();
        null_count = 0;
        for (auto v : valid_flags) null_count += v ? 0 : 1;
        validity.clear();
        if (null_count) {
            validity.assign((size_t)((length + 7) / 8), 0);
            for (int64_t i = 0; i < length; i++)
                if (valid_flags[(size_t)i]) set_bit(validity.data(), i);
        }
    }
};

// EmitTo::take_needed (expr/src/groups_accumulator.rs:24-58)
template <class T> static std::vector<T> take_needed(std::vector<T> &v, int64_t first_n) {
    if (first_n < 0 || (size_t)first_n >= v.size()) {
        std::vector<T> out;
        out.swap(v);
        return out;
    }
    std::vector<T> out(v.begin(), v.begin() + first_n);
    v.erase(v.begin(), v.begin() + first_n);
    return out;
}

// ------------------------------------------------------------------------------------------
// GroupValues — dense, first-seen group ids.
//   GroupValuesPrimitive  physical-plan/src/aggregates/group_values/primitive.rs:112-207
//   GroupValuesByes       group_values/bytes.rs:44-121 + physical-expr/src/binary_map.rs:333-521
//   GroupValuesRows       group_values/row.rs:94-223
// One restatement covers all three: rows are encoded to a self-delimiting byte string
// (valid byte + fixed bytes | u32 len + bytes) that is equal iff every column is equal
// (NULL == NULL, floats by bit pattern), looked up in an open-addressing table of
// (hash, group id) and appended on a miss.  A u64 fast path mirrors GroupValuesPrimitive.
// ------------------------------------------------------------------------------------------
struct GroupValues {
    int ncols = 0;
    std::vector<int> col_types;  // value types (dictionaries unwrapped)
    std::vector<int> col_prec, col_scale;
    bool types_known = false;
    // generic
    std::vector<uint8_t> arena;
    std::vector<uint64_t> row_off{0};
    std::vector<uint64_t> row_hash;
    std::vector<uint32_t> slots;  // gid + 1
    size_t ngroups = 0;
    // primitive fast path
    bool prim = false;
    std::vector<uint64_t> prim_keys;
    int64_t null_group = -1;
    uint64_t seed = 0x5eedULL;
    bool force_collisions = false;
    std::vector<uint8_t> rowbuf;

    void learn_types(const ocol *const *cols) {
        if (types_known) return;
        for (int c = 0; c < ncols; c++) {
            const ocol *v = cols[c]->type == OT_DICT ? cols[c]->dict : cols[c];
            col_types.push_back(v->type);
            col_prec.push_back(v->precision);
            col_scale.push_back(v->scale);
        }
        // group_values/mod.rs:55-69: exactly one primitive column -> GroupValuesPrimitive
        prim = ncols == 1 && cols[0]->type != OT_DICT && type_width(col_types[0]) > 0 && type_width(col_types[0]) <= 8;
        types_known = true;
        slots.assign(1024, 0);
    }
    void grow() {
        std::vector<uint32_t> ns(slots.size() * 2, 0);
        size_t mask = ns.size() - 1;
        for (size_t g = 0; g < ngroups; g++) {
            if (prim && (int64_t)g == null_group) continue;
            uint64_t h = prim ? ph(prim_keys[g]) : row_hash[g];
            size_t s = h & mask;
            while (ns[s]) s = (s + 1) & mask;
            ns[s] = (uint32_t)g + 1;
        }
        slots.swap(ns);
    }
    inline uint64_t ph(uint64_t k) const { return force_collisions ? 0 : h_u64(k, seed); }

    void intern(const ocol *const *cols, int64_t n, std::vector<uint32_t> &out) {
        learn_types(cols);
        out.resize((size_t)n);
        if (prim) {
            const ocol *c = cols[0];
            for (int64_t i = 0; i < n; i++) {
                if (c->validity && !get_bit(c->validity, c->offset + i)) {
                    // primitive.rs:118-122 — single null group created on first NULL
                    if (null_group < 0) {
                        null_group = (int64_t)ngroups++;
                        prim_keys.push_back(0);
                    }
                    out[(size_t)i] = (uint32_t)null_group;
                    continue;
                }
                uint64_t k = bits64(c, c->offset + i);
                if ((ngroups + 1) * 2 > slots.size()) grow();
                size_t mask = slots.size() - 1, s = ph(k) & mask;
                for (;;) {
                    uint32_t e = slots[s];
                    if (!e) {  // primitive.rs:136-140
                        slots[s] = (uint32_t)ngroups + 1;
                        out[(size_t)i] = (uint32_t)ngroups++;
                        prim_keys.push_back(k);
                        break;
                    }
                    if (prim_keys[e - 1] == k && (int64_t)(e - 1) != null_group) {  // primitive.rs:127-131
                        out[(size_t)i] = e - 1;
                        break;
                    }
                    s = (s + 1) & mask;
                }
            }
            return;
        }
        std::vector<uint64_t> hashes((size_t)n);
        create_hashes(cols, ncols, seed, force_collisions, n, hashes.data());  // row.rs:112
        for (int64_t i = 0; i < n; i++) {
            rowbuf.clear();
            for (int c = 0; c < ncols; c++) {  // stands in for RowConverter::convert_columns (row.rs:97)
                const ocol *vc; int64_t vi;
                if (!resolve(cols[c], i, vc, vi)) { rowbuf.push_back(0); continue; }
                rowbuf.push_back(1);
                if (is_var(vc->type)) {
                    const uint8_t *p; int64_t len;
                    get_bytes(vc, vi, p, len);
                    uint32_t l32 = (uint32_t)len;
                    rowbuf.insert(rowbuf.end(), (uint8_t *)&l32, (uint8_t *)&l32 + 4);
                    rowbuf.insert(rowbuf.end(), p, p + len);
                } else if (vc->type == OT_BOOL) {
                    rowbuf.push_back((uint8_t)bits64(vc, vi));
                } else {
                    int w = type_width(vc->type);
                    const uint8_t *p = (const uint8_t *)vc->values + (int64_t)w * vi;
                    rowbuf.insert(rowbuf.end(), p, p + w);
                }
            }
            if ((ngroups + 1) * 2 > slots.size()) grow();
            size_t mask = slots.size() - 1, s = hashes[(size_t)i] & mask;
            for (;;) {
                uint32_t e = slots[s];
                if (!e) {  // row.rs:126-137
                    slots[s] = (uint32_t)ngroups + 1;
                    arena.insert(arena.end(), rowbuf.begin(), rowbuf.end());
                    row_off.push_back(arena.size());
                    row_hash.push_back(hashes[(size_t)i]);
                    out[(size_t)i] = (uint32_t)ngroups++;
                    break;
                }
                size_t g = e - 1;
                size_t len = (size_t)(row_off[g + 1] - row_off[g]);
                if (row_hash[g] == hashes[(size_t)i] && len == rowbuf.size() &&
                    memcmp(arena.data() + row_off[g], rowbuf.data(), len) == 0) {  // row.rs:115-120
                    out[(size_t)i] = (uint32_t)g;
                    break;
                }
                s = (s + 1) & mask;
            }
        }
    }

    // emit(EmitTo) — primitive.rs:163-207, row.rs:167-223, bytes.rs:87-121
    std::vector<OutCol> emit(int64_t first_n) {
        size_t take = (first_n < 0 || (size_t)first_n > ngroups) ? ngroups : (size_t)first_n;
        std::vector<OutCol> out((size_t)ncols);
        if (!types_known) return out;
        if (prim) {
            OutCol &o = out[0];
            o.type = col_types[0]; o.precision = col_prec[0]; o.scale = col_scale[0];
            int w = type_width(o.type);
            std::vector<uint8_t> valid(take, 1);
            o.values.resize(take * (size_t)w);
            for (size_t g = 0; g < take; g++) {
                if ((int64_t)g == null_group) valid[g] = 0;
                memcpy(o.values.data() + g * w, &prim_keys[g], (size_t)w);  // little endian truncation
            }
            o.set_validity(valid);
            if (take == ngroups) {
                prim_keys.clear(); null_group = -1; ngroups = 0;
                std::fill(slots.begin(), slots.end(), 0);
            } else {  // primitive.rs:180-203 — shift remaining ids down by n
                prim_keys.erase(prim_keys.begin(), prim_keys.begin() + take);
                if (null_group >= 0) null_group = null_group < (int64_t)take ? -1 : null_group - (int64_t)take;
                ngroups -= take;
                std::fill(slots.begin(), slots.end(), 0);
                size_t mask = slots.size() - 1;
                for (size_t g = 0; g < ngroups; g++) {
                    if ((int64_t)g == null_group) continue;
                    size_t s = ph(prim_keys[g]) & mask;
                    while (slots[s]) s = (s + 1) & mask;
                    slots[s] = (uint32_t)g + 1;
                }
            }
            return out;
        }
        std::vector<std::vector<uint8_t>> valid((size_t)ncols, std::vector<uint8_t>(take, 1));
        for (int c = 0; c < ncols; c++) {
            out[c].type = col_types[c]; out[c].precision = col_prec[c]; out[c].scale = col_scale[c];
            if (is_var(col_types[c])) out[c].offsets.push_back(0);
            if (col_types[c] == OT_BOOL) out[c].values.assign((take + 7) / 8, 0);
        }
        for (size_t g = 0; g < take; g++) {  // convert_rows (row.rs:175,181)
            const uint8_t *p = arena.data() + row_off[g];
            for (int c = 0; c < ncols; c++) {
                OutCol &o = out[c];
                uint8_t v = *p++;
                int t = col_types[c];
                if (is_var(t)) {
                    if (v) {
                        uint32_t len; memcpy(&len, p, 4); p += 4;
                        o.values.insert(o.values.end(), p, p + len);
                        p += len;
                    } else valid[c][g] = 0;
                    o.offsets.push_back((int64_t)o.values.size());
                } else if (t == OT_BOOL) {
                    if (v) { if (*p++) set_bit(o.values.data(), (int64_t)g); } else valid[c][g] = 0;
                } else {
                    int w = type_width(t);
                    if (v) { o.values.insert(o.values.end(), p, p + w); p += w; }
                    else { o.values.insert(o.values.end(), (size_t)w, 0); valid[c][g] = 0; }
                }
            }
        }
        for (int c = 0; c < ncols; c++) out[c].set_validity(valid[c]);
        // drop the emitted prefix and rebuild the table (row.rs:183-202)
        if (take == ngroups) {
            arena.clear(); row_off.assign(1, 0); row_hash.clear(); ngroups = 0;
            std::fill(slots.begin(), slots.end(), 0);
        } else {
            uint64_t cut = row_off[take];
            arena.erase(arena.begin(), arena.begin() + (int64_t)cut);
            row_off.erase(row_off.begin(), row_off.begin() + (int64_t)take);
            for (auto &o : row_off) o -= cut;
            row_hash.erase(row_hash.begin(), row_hash.begin() + (int64_t)take);
            ngroups -= take;
            std::fill(slots.begin(), slots.end(), 0);
            size_t mask = slots.size() - 1;
            for (size_t g = 0; g < ngroups; g++) {
                size_t s = row_hash[g] & mask;
                while (slots[s]) s = (s + 1) & mask;
                slots[s] = (uint32_t)g + 1;
            }
        }
        return out;
    }
};

// ------------------------------------------------------------------------------------------
// NullState / accumulate (physical-expr/src/aggregate/groups_accumulator/accumulate.rs:126-233):
// value_fn runs only for rows that are non-null AND (no filter OR filter == Some(true)).
// ------------------------------------------------------------------------------------------
static inline bool filter_pass(const ocol *f, int64_t i) {
    if (!f) return true;
    if (f->validity && !get_bit(f->validity, f->offset + i)) return false;  // NULL filter rejects
    return get_bit((const uint8_t *)f->values, f->offset + i);
}
template <class F>
static void accumulate(const uint32_t *gids, int64_t n, const ocol *val, const ocol *filt, std::vector<uint8_t> *seen,
                       F value_fn) {
    for (int64_t i = 0; i < n; i++) {
        if (!filter_pass(filt, i)) continue;
        const ocol *vc; int64_t vi;
        if (!resolve(val, i, vc, vi)) continue;
        if (seen) (*seen)[gids[i]] = 1;
        value_fn(gids[i], vc, vi);
    }
}

template <class T> static inline T read_as(const ocol *vc, int64_t vi);
template <> inline int64_t read_as<int64_t>(const ocol *vc, int64_t vi) { return (int64_t)bits64(vc, vi); }
template <> inline uint64_t read_as<uint64_t>(const ocol *vc, int64_t vi) { return bits64(vc, vi); }
template <> inline i128 read_as<i128>(const ocol *vc, int64_t vi) { return get_i128(vc, vi); }
template <> inline double read_as<double>(const ocol *vc, int64_t vi) {
    // upstream cast to Float64 (expr/src/type_coercion/aggregates.rs:108-149)
    switch (vc->type) {
    case OT_F64: { double d; uint64_t b = bits64(vc, vi); memcpy(&d, &b, 8); return d; }
    case OT_F32: { float f; uint32_t b = (uint32_t)bits64(vc, vi); memcpy(&f, &b, 4); return (double)f; }
    case OT_U8: case OT_U16: case OT_U32: case OT_U64: return (double)bits64(vc, vi);
    default: return (double)(int64_t)bits64(vc, vi);
    }
}
template <> inline float read_as<float>(const ocol *vc, int64_t vi) {
    float f; uint32_t b = (uint32_t)bits64(vc, vi); memcpy(&f, &b, 4); return f;
}

struct Acc {
    virtual ~Acc() {}
    virtual void update(const uint32_t *g, int64_t n, const ocol *const *args, const ocol *filt, size_t ngroups) = 0;
    virtual void merge(const uint32_t *g, int64_t n, const ocol *const *args, size_t ngroups) = 0;
    virtual int nstate() const { return 1; }
    virtual bool state(int64_t first_n, std::vector<OutCol> &out) = 0;
    virtual bool evaluate(int64_t first_n, OutCol &out) = 0;
};

template <class T> static void put_values(OutCol &o, const std::vector<T> &v) {
    o.values.resize(v.size() * sizeof(T));
    if (!v.empty()) memcpy(o.values.data(), v.data(), o.values.size());
}

// SUM — PrimitiveGroupsAccumulator with add_wrapping (groups_accumulator/prim_op.rs:86-139, sum.rs:75-142)
template <class T, int OUT_T> struct SumAcc : Acc {
    std::vector<T> sums;
    std::vector<uint8_t> seen;
    int prec = 0, scale = 0;
    static inline T addw(T a, T b) {
        if constexpr (std::is_same<T, double>::value) return a + b;
        else if constexpr (std::is_same<T, i128>::value) return (i128)((u128)a + (u128)b);
        else if constexpr (std::is_same<T, int64_t>::value) return (int64_t)((uint64_t)a + (uint64_t)b);
        else return a + b;
    }
    void update(const uint32_t *g, int64_t n, const ocol *const *args, const ocol *filt, size_t ng) override {
        sums.resize(ng, T(0)); seen.resize(ng, 0);
        accumulate(g, n, args[0], filt, &seen, [&](uint32_t gi, const ocol *vc, int64_t vi) {
            sums[gi] = addw(sums[gi], read_as<T>(vc, vi));
        });
    }
    void merge(const uint32_t *g, int64_t n, const ocol *const *args, size_t ng) override {
        update(g, n, args, nullptr, ng);  // prim_op.rs:126-135
    }
    bool evaluate(int64_t first_n, OutCol &o) override {
        auto v = take_needed(sums, first_n);
        auto s = take_needed(seen, first_n);
        o.type = OUT_T; o.precision = prec; o.scale = scale;
        put_values(o, v);
        o.set_validity(s);
        return true;
    }
    bool state(int64_t first_n, std::vector<OutCol> &out) override {  // prim_op.rs:122-124
        out.emplace_back();
        return evaluate(first_n, out.back());
    }
};

// COUNT — CountGroupsAccumulator (count.rs:92-192)
struct CountAcc : Acc {
    std::vector<int64_t> counts;
    void update(const uint32_t *g, int64_t n, const ocol *const *args, const ocol *filt, size_t ng) override {
        counts.resize(ng, 0);
        const ocol *val = args[0];
        for (int64_t i = 0; i < n; i++) {  // accumulate_indices (accumulate.rs:363-443)
            if (!filter_pass(filt, i)) continue;
            if (val) { const ocol *vc; int64_t vi; if (!resolve(val, i, vc, vi)) continue; }
            counts[g[i]] += 1;
        }
    }
    void merge(const uint32_t *g, int64_t n, const ocol *const *args, size_t ng) override {
        counts.resize(ng, 0);  // count.rs:135-160 (partial counts carry no nulls)
        const ocol *c = args[0];
        for (int64_t i = 0; i < n; i++) counts[g[i]] += ((const int64_t *)c->values)[c->offset + i];
    }
    bool evaluate(int64_t first_n, OutCol &o) override {  // never NULL (count.rs:172-181)
        auto v = take_needed(counts, first_n);
        o.type = OT_I64;
        put_values(o, v);
        o.length = (int64_t)v.size();
        return true;
    }
    bool state(int64_t first_n, std::vector<OutCol> &out) override {
        out.emplace_back();
        return evaluate(first_n, out.back());
    }
};

static i128 pow10_i128(int e) {
    i128 r = 1;
    for (int i = 0; i < e; i++) r = (i128)((u128)r * 10u);  // pow_wrapping
    return r;
}

// AVG — AvgGroupsAccumulator (average.rs:392-561) + DecimalAverager (aggregate/utils.rs:55-125)
template <class T> struct AvgAcc : Acc {
    std::vector<uint64_t> counts;
    std::vector<T> sums;
    std::vector<uint8_t> seen;
    int in_prec = 0, in_scale = 0;  // decimal input (sum) type
    int out_prec = 0, out_scale = 0;
    static inline T addw(T a, T b) {
        if constexpr (std::is_same<T, double>::value) return a + b;
        else return (i128)((u128)a + (u128)b);
    }
    void update(const uint32_t *g, int64_t n, const ocol *const *args, const ocol *filt, size_t ng) override {
        counts.resize(ng, 0); sums.resize(ng, T(0)); seen.resize(ng, 0);
        accumulate(g, n, args[0], filt, &seen, [&](uint32_t gi, const ocol *vc, int64_t vi) {
            sums[gi] = addw(sums[gi], read_as<T>(vc, vi));
            counts[gi] += 1;
        });
    }
    void merge(const uint32_t *g, int64_t n, const ocol *const *args, size_t ng) override {
        counts.resize(ng, 0); sums.resize(ng, T(0)); seen.resize(ng, 0);  // average.rs:472-509
        accumulate(g, n, args[0], nullptr, &seen,
                   [&](uint32_t gi, const ocol *vc, int64_t vi) { counts[gi] += bits64(vc, vi); });
        accumulate(g, n, args[1], nullptr, &seen,
                   [&](uint32_t gi, const ocol *vc, int64_t vi) { sums[gi] = addw(sums[gi], read_as<T>(vc, vi)); });
    }
    int nstate() const override { return 2; }
    bool state(int64_t first_n, std::vector<OutCol> &out) override {  // average.rs:547-560
        auto s = take_needed(seen, first_n);
        auto c = take_needed(counts, first_n);
        auto v = take_needed(sums, first_n);
        OutCol oc; oc.type = OT_U64; put_values(oc, c); oc.set_validity(s);
        OutCol os; os.type = std::is_same<T, double>::value ? OT_F64 : OT_DEC128;
        os.precision = in_prec; os.scale = in_scale;
        put_values(os, v); os.set_validity(s);
        out.push_back(std::move(oc));
        out.push_back(std::move(os));
        return true;
    }
    bool evaluate(int64_t first_n, OutCol &o) override {  // average.rs:511-545
        auto c = take_needed(counts, first_n);
        auto v = take_needed(sums, first_n);
        auto s = take_needed(seen, first_n);
        std::vector<T> avg(v.size(), T(0));
        if constexpr (std::is_same<T, double>::value) {
            o.type = OT_F64;
            for (size_t i = 0; i < v.size(); i++)
                if (s[i]) avg[i] = v[i] / (double)c[i];  // average.rs:165
        } else {
            o.type = OT_DEC128; o.precision = out_prec; o.scale = out_scale;
            i128 sum_mul = pow10_i128(in_scale), target_mul = pow10_i128(out_scale);
            if (!(target_mul >= sum_mul)) { g_err = "Arithmetic Overflow in AvgAccumulator"; return false; }
            i128 factor = target_mul / sum_mul;
            i128 lim = pow10_i128(out_prec) - 1;  // validate_decimal_precision
            for (size_t i = 0; i < v.size(); i++) {
                if (!s[i]) continue;
                i128 value;
                if (__builtin_mul_overflow(v[i], factor, &value)) { g_err = "Arithmetic Overflow in AvgAccumulator"; return false; }
                i128 nv = value / (i128)c[i];  // div_wrapping: truncation toward zero
                if (nv > lim || nv < -lim) { g_err = "Arithmetic Overflow in AvgAccumulator"; return false; }
                avg[i] = nv;
            }
        }
        put_values(o, avg);
        o.set_validity(s);
        return true;
    }
};

// MIN / MAX — PrimitiveGroupsAccumulator with strict compares and MAX/MIN start values
// (min_max.rs:102-139).  For floats the start value is the finite f::MAX / f::MIN and NaN
// never replaces the current value.
template <class T, bool IS_MIN> struct MinMaxAcc : Acc {
    std::vector<T> vals;
    std::vector<uint8_t> seen;
    int type = 0, prec = 0, scale = 0;
    static T start() {
        if constexpr (std::is_same<T, i128>::value) {
            i128 mx = (i128)((~(u128)0) >> 1);
            return IS_MIN ? mx : (i128)(-mx - 1);
        } else if constexpr (std::is_floating_point<T>::value) {
            return IS_MIN ? std::numeric_limits<T>::max() : std::numeric_limits<T>::lowest();
        } else {
            return IS_MIN ? std::numeric_limits<T>::max() : std::numeric_limits<T>::min();
        }
    }
    static inline T rd(const ocol *vc, int64_t vi) {
        if constexpr (std::is_same<T, i128>::value) return get_i128(vc, vi);
        else {
            T v;
            memcpy(&v, (const uint8_t *)vc->values + sizeof(T) * vi, sizeof(T));
            return v;
        }
    }
    void update(const uint32_t *g, int64_t n, const ocol *const *args, const ocol *filt, size_t ng) override {
        vals.resize(ng, start()); seen.resize(ng, 0);
        accumulate(g, n, args[0], filt, &seen, [&](uint32_t gi, const ocol *vc, int64_t vi) {
            T v = rd(vc, vi);
            if (IS_MIN) { if (vals[gi] > v) vals[gi] = v; }
            else { if (vals[gi] < v) vals[gi] = v; }
        });
    }
    void merge(const uint32_t *g, int64_t n, const ocol *const *args, size_t ng) override { update(g, n, args, nullptr, ng); }
    bool evaluate(int64_t first_n, OutCol &o) override {
        auto v = take_needed(vals, first_n);
        auto s = take_needed(seen, first_n);
        o.type = type; o.precision = prec; o.scale = scale;
        put_values(o, v);
        o.set_validity(s);
        return true;
    }
    bool state(int64_t first_n, std::vector<OutCol> &out) override {
        out.emplace_back();
        return evaluate(first_n, out.back());
    }
};

template <bool IS_MIN> static Acc *make_minmax(const oagg_desc &d) {
#define MM(TT, CT)                                   \
    case TT: {                                       \
        auto *a = new MinMaxAcc<CT, IS_MIN>();       \
        a->type = d.input_type; a->prec = d.precision; a->scale = d.scale; \
        return a;                                    \
    }
    switch (d.input_type) {
        MM(OT_I8, int8_t) MM(OT_I16, int16_t) MM(OT_I32, int32_t) MM(OT_I64, int64_t)
        MM(OT_U8, uint8_t) MM(OT_U16, uint16_t) MM(OT_U32, uint32_t) MM(OT_U64, uint64_t)
        MM(OT_F32, float) MM(OT_F64, double) MM(OT_DATE32, int32_t) MM(OT_DATE64, int64_t)
        MM(OT_TS, int64_t) MM(OT_TIME32, int32_t) MM(OT_TIME64, int64_t) MM(OT_DEC128, i128)
    }
#undef MM
    return nullptr;
}

// BIT_AND / BIT_OR / BIT_XOR — PrimitiveGroupsAccumulator over the argument's own integer type with
// starting value !0 (AND) or 0 (OR, XOR) (aggregate/bit_and_or_xor.rs:116-140, 281-301, 442-462;
// groups_accumulator/prim_op.rs:86-139).  State == evaluate; merge == update.
template <class T, int OP> struct BitAcc : Acc {
    std::vector<T> vals;
    std::vector<uint8_t> seen;
    int type = 0;
    void update(const uint32_t *g, int64_t n, const ocol *const *args, const ocol *filt, size_t ng) override {
        vals.resize(ng, OP == 0 ? (T)~(T)0 : (T)0); seen.resize(ng, 0);
        accumulate(g, n, args[0], filt, &seen, [&](uint32_t gi, const ocol *vc, int64_t vi) {
            T v;
            memcpy(&v, (const uint8_t *)vc->values + sizeof(T) * vi, sizeof(T));
            if (OP == 0) vals[gi] &= v; else if (OP == 1) vals[gi] |= v; else vals[gi] ^= v;
        });
    }
    void merge(const uint32_t *g, int64_t n, const ocol *const *args, size_t ng) override { update(g, n, args, nullptr, ng); }
    bool evaluate(int64_t first_n, OutCol &o) override {
        auto v = take_needed(vals, first_n);
        auto s = take_needed(seen, first_n);
        o.type = type;
        put_values(o, v);
        o.set_validity(s);
        return true;
    }
    bool state(int64_t first_n, std::vector<OutCol> &out) override {
        out.emplace_back();
        return evaluate(first_n, out.back());
    }
};
template <int OP> static Acc *make_bit(const oagg_desc &d) {
#define BB(TT, CT) case TT: { auto *a = new BitAcc<CT, OP>(); a->type = d.input_type; return a; }
    switch (d.input_type) {
        BB(OT_I8, int8_t) BB(OT_I16, int16_t) BB(OT_I32, int32_t) BB(OT_I64, int64_t)
        BB(OT_U8, uint8_t) BB(OT_U16, uint16_t) BB(OT_U32, uint32_t) BB(OT_U64, uint64_t)
    }
#undef BB
    return nullptr;  // downcast_integer! accepts the eight integer types only
}

// BOOL_AND / BOOL_OR — BooleanGroupsAccumulator (groups_accumulator/bool_op.rs:38-140,
// bool_and_or.rs:137-148, 270-281).  The value bit of a NEW group starts at bool::default() = false
// for both functions (bool_op.rs:82-85), so the grouped BOOL_AND of this reference version yields
// false for every group that saw a value (false && x); that is reproduced, not corrected.
template <bool IS_AND> struct BoolAcc : Acc {
    std::vector<uint8_t> vals, seen;
    void update(const uint32_t *g, int64_t n, const ocol *const *args, const ocol *filt, size_t ng) override {
        vals.resize(ng, 0); seen.resize(ng, 0);
        accumulate(g, n, args[0], filt, &seen, [&](uint32_t gi, const ocol *vc, int64_t vi) {
            bool nv = bits64(vc, vi) != 0, cur = vals[gi] != 0;
            vals[gi] = IS_AND ? (cur && nv) : (cur || nv);
        });
    }
    void merge(const uint32_t *g, int64_t n, const ocol *const *args, size_t ng) override { update(g, n, args, nullptr, ng); }
    bool evaluate(int64_t first_n, OutCol &o) override {
        auto v = take_needed(vals, first_n);
        auto s = take_needed(seen, first_n);
        o.type = OT_BOOL;
        o.values.assign((v.size() + 7) / 8, 0);
        for (size_t i = 0; i < v.size(); i++)
            if (v[i]) set_bit(o.values.data(), (int64_t)i);
        o.set_validity(s);
        return true;
    }
    bool state(int64_t first_n, std::vector<OutCol> &out) override {
        out.emplace_back();
        return evaluate(first_n, out.back());
    }
};

static Acc *make_acc(const oagg_desc &d) {
    int t = d.input_type;
    switch (d.func) {
    case OAGG_COUNT: return new CountAcc();
    case OAGG_SUM:
        // sum.rs:75-87 + type_coercion/aggregates.rs:108-129, 397-407
        if (is_signed_int(t)) return new SumAcc<int64_t, OT_I64>();
        if (is_unsigned_int(t)) return new SumAcc<uint64_t, OT_U64>();
        if (t == OT_F32 || t == OT_F64) return new SumAcc<double, OT_F64>();
        if (t == OT_DEC128) {
            auto *a = new SumAcc<i128, OT_DEC128>();
            a->prec = std::min(38, d.precision + 10); a->scale = d.scale;
            return a;
        }
        return nullptr;
    case OAGG_AVG:
        if (t == OT_DEC128) {  // type_coercion/aggregates.rs:455-463
            auto *a = new AvgAcc<i128>();
            a->in_prec = d.precision; a->in_scale = d.scale;
            a->out_prec = std::min(38, d.precision + 4); a->out_scale = std::min(38, d.scale + 4);
            return a;
        }
        if (is_signed_int(t) || is_unsigned_int(t) || t == OT_F32 || t == OT_F64) return new AvgAcc<double>();
        return nullptr;
    case OAGG_MIN: return make_minmax<true>(d);
    case OAGG_MAX: return make_minmax<false>(d);
    case OAGG_BIT_AND: return make_bit<0>(d);
    case OAGG_BIT_OR: return make_bit<1>(d);
    case OAGG_BIT_XOR: return make_bit<2>(d);
    case OAGG_BOOL_AND: return t == OT_BOOL ? new BoolAcc<true>() : nullptr;
    case OAGG_BOOL_OR: return t == OT_BOOL ? new BoolAcc<false>() : nullptr;
    }
    return nullptr;
}

// ------------------------------------------------------------------------------------------
// Aggregation operator: GroupedHashAggregateStream (physical-plan/src/aggregates/row_hash.rs)
// ------------------------------------------------------------------------------------------
struct oracle_agg {
    int mode = 0;
    GroupValues gv;
    std::vector<oagg_desc> descs;
    std::vector<std::unique_ptr<Acc>> accs;
    std::vector<uint32_t> gids;
    std::vector<OutCol> out;
    bool first_stage() const { return mode == OMODE_PARTIAL || mode == OMODE_SINGLE || mode == OMODE_SINGLE_PARTITIONED; }
};

extern "C" oracle_agg *oracle_agg_new(int mode, int n_group_cols, int n_aggs, const oagg_desc *aggs) {
    auto *a = new oracle_agg();
    a->mode = mode;
    a->gv.ncols = n_group_cols;
    for (int i = 0; i < n_aggs; i++) {
        Acc *acc = make_acc(aggs[i]);
        if (!acc) {
            g_err = "NotImplemented: unsupported aggregate / input type";
            delete a;
            return nullptr;
        }
        a->descs.push_back(aggs[i]);
        a->accs.emplace_back(acc);
    }
    return a;
}
extern "C" void oracle_agg_free(oracle_agg *a) { delete a; }

extern "C" int oracle_agg_push(oracle_agg *a, const ocol *const *group_cols, const ocol *const *args,
                               const ocol *const *filters) {
    int64_t n = group_cols[0]->length;
    a->gv.intern(group_cols, n, a->gids);  // row_hash.rs:560-566
    size_t ng = a->gv.ngroups;
    int ai = 0;
    for (size_t k = 0; k < a->accs.size(); k++) {  // row_hash.rs:571-599
        if (a->first_stage()) {
            a->accs[k]->update(a->gids.data(), n, args + ai, filters ? filters[k] : nullptr, ng);
            ai += 1;
        } else {
            a->accs[k]->merge(a->gids.data(), n, args + ai, ng);
            ai += a->accs[k]->nstate();
        }
    }
    return 0;
}
extern "C" int64_t oracle_agg_num_groups(const oracle_agg *a) { return (int64_t)a->gv.ngroups; }
extern "C" const uint32_t *oracle_agg_last_group_ids(const oracle_agg *a) { return a->gids.data(); }

extern "C" int oracle_agg_emit(oracle_agg *a, int64_t first_n) {
    a->out.clear();
    size_t ng = a->gv.ngroups;
    // accumulators may not have been resized to groups created by later batches
    static const ocol *none[2] = {nullptr, nullptr};
    (void)none;
    std::vector<OutCol> keys = a->gv.emit(first_n);  // row_hash.rs:636
    for (auto &k : keys) a->out.push_back(std::move(k));
    for (auto &acc : a->accs) {
        // make sure per-group vectors cover all groups (groups whose rows were all filtered)
        acc->update(nullptr, 0, none, nullptr, ng);
        if (a->mode == OMODE_PARTIAL) {  // row_hash.rs:642-655
            if (!acc->state(first_n, a->out)) return -1;
        } else {
            a->out.emplace_back();
            if (!acc->evaluate(first_n, a->out.back())) return -1;
        }
    }
    return (int)a->out.size();
}
extern "C" int oracle_agg_out(const oracle_agg *a, int i, ocol_out *o) {
    if (i < 0 || (size_t)i >= a->out.size()) return fail("bad column");
    const OutCol &c = a->out[(size_t)i];
    o->type = c.type; o->precision = c.precision; o->scale = c.scale;
    o->length = is_var(c.type) ? (int64_t)c.offsets.size() - 1 : c.length;
    o->null_count = c.null_count;
    o->validity = c.validity.empty() ? nullptr : c.validity.data();
    o->values = c.values.data();
    o->offsets = c.offsets.empty() ? nullptr : c.offsets.data();
    o->data_bytes = (int64_t)c.values.size();
    return 0;
}

// ------------------------------------------------------------------------------------------
// Hash join: JoinHashMap (joins/utils.rs:121-349), collect_left_input / update_hash
// (hash_join.rs:678-815), lookup_join_hashmap + equal_rows_arr (:1024-1118),
// process_probe_batch (:1238-1343), adjust_indices_by_join_type (joins/utils.rs:1234-1364),
// process_unmatched_build_batch (:1348-1388).
// ------------------------------------------------------------------------------------------
struct oracle_join {
    int join_type = 0;
    bool null_equals_null = false, force_collisions = false;
    int nkeys = 0;
    int64_t batch_size = 8192;
    std::vector<std::vector<const ocol *>> build_batches;  // arrival order
    std::vector<int64_t> cat_batch, cat_start;             // concat order (reversed arrival)
    int64_t build_rows = 0;
    // JoinHashMap: map (hash -> head+1) and next[]
    std::vector<uint64_t> map_hash, map_head;
    std::vector<uint64_t> next;
    std::vector<uint8_t> visited;
    std::vector<int64_t> out_build, out_probe;
    bool finished = false;
    // JoinFilter (joins/utils.rs:1143-1176): the driver evaluates the residual predicate on the candidate pairs
    oracle_join_filter_fn filter = nullptr;
    void *filter_user = nullptr;

    bool locate(int64_t idx, int64_t &b, int64_t &r) const {
        for (size_t k = 0; k < cat_batch.size(); k++) {
            int64_t len = build_batches[(size_t)cat_batch[k]][0]->length;
            if (idx >= cat_start[k] && idx < cat_start[k] + len) { b = cat_batch[k]; r = idx - cat_start[k]; return true; }
        }
        return false;
    }
    uint64_t *find(uint64_t h) {
        size_t mask = map_hash.size() - 1, s = mix64(h) & mask;
        for (;;) {
            if (!map_head[s]) return nullptr;
            if (map_hash[s] == h) return &map_head[s];
            s = (s + 1) & mask;
        }
    }
    void insert(uint64_t h, uint64_t head) {
        size_t mask = map_hash.size() - 1, s = mix64(h) & mask;
        while (map_head[s]) s = (s + 1) & mask;
        map_hash[s] = h; map_head[s] = head;
    }
};

extern "C" oracle_join *oracle_join_new(int join_type, int null_equals_null, int n_keys, int64_t batch_size,
                                        int force_collisions) {
    auto *j = new oracle_join();
    j->join_type = join_type; j->null_equals_null = null_equals_null != 0; j->nkeys = n_keys;
    j->batch_size = batch_size; j->force_collisions = force_collisions != 0;
    return j;
}
extern "C" void oracle_join_free(oracle_join *j) { delete j; }
extern "C" int oracle_join_build_batch(oracle_join *j, const ocol *const *key_cols) {
    j->build_batches.emplace_back(key_cols, key_cols + j->nkeys);
    return 0;
}
extern "C" int oracle_join_build_finish(oracle_join *j) {
    int64_t total = 0;
    for (auto &b : j->build_batches) total += b[0]->length;
    j->build_rows = total;
    size_t cap = 16;
    while (cap < (size_t)total * 2 + 2) cap <<= 1;
    j->map_hash.assign(cap, 0); j->map_head.assign(cap, 0);
    j->next.assign((size_t)total, 0);  // JoinHashMap::with_capacity (utils.rs:135-140)
    int64_t offset = 0;
    std::vector<uint64_t> hashes;
    // "Updating hashmap starting from the last batch" (hash_join.rs:746-761); fifo_hashmap=true
    // iterates rows in reverse so the chain head is the smallest index (:161-165).
    for (int64_t b = (int64_t)j->build_batches.size() - 1; b >= 0; b--) {
        auto &cols = j->build_batches[(size_t)b];
        int64_t n = cols[0]->length;
        hashes.resize((size_t)n);
        create_hashes(cols.data(), j->nkeys, 0, j->force_collisions, n, hashes.data());
        for (int64_t i = n - 1; i >= 0; i--) {  // update_from_iter (utils.rs:203-229)
            uint64_t row = (uint64_t)(i + offset);
            uint64_t *head = j->find(hashes[(size_t)i]);
            if (head) { j->next[row] = *head; *head = row + 1; }
            else j->insert(hashes[(size_t)i], row + 1);
        }
        j->cat_batch.push_back(b);
        j->cat_start.push_back(offset);
        offset += n;
    }
    j->visited.assign((size_t)total, 0);
    j->finished = true;
    return 0;
}
extern "C" int64_t oracle_join_build_rows(const oracle_join *j) { return j->build_rows; }
extern "C" int oracle_join_build_index(const oracle_join *j, int64_t idx, int64_t *batch, int64_t *row) {
    return j->locate(idx, *batch, *row) ? 0 : -1;
}

// eq / not_distinct on one key column pair (hash_join.rs:1067-1076)
static bool key_equal(const ocol *l, int64_t li, const ocol *r, int64_t ri, bool null_equals_null) {
    const ocol *lv, *rv; int64_t lvi, rvi;
    bool lok = resolve(l, li, lv, lvi), rok = resolve(r, ri, rv, rvi);
    if (!lok || !rok) return null_equals_null && !lok && !rok;
    if (is_var(lv->type)) {
        const uint8_t *lp, *rp; int64_t ll, rl;
        get_bytes(lv, lvi, lp, ll); get_bytes(rv, rvi, rp, rl);
        return ll == rl && memcmp(lp, rp, (size_t)ll) == 0;
    }
    if (lv->type == OT_DEC128) return get_i128(lv, lvi) == get_i128(rv, rvi);
    // arrow-rs 50 `cmp::eq` compares floats with IEEE-754 totalOrder semantics (bit equality:
    // NaN == NaN, -0.0 != +0.0) — consistent with hashing the bit pattern (hash_utils.rs:78-87).
    return bits64(lv, lvi) == bits64(rv, rvi);
}

static bool need_final(int jt) { return jt == OJ_LEFT || jt == OJ_LEFT_ANTI || jt == OJ_LEFT_SEMI || jt == OJ_FULL; }

extern "C" int64_t oracle_join_probe(oracle_join *j, const ocol *const *key_cols, int64_t *out_batch_count) {
    j->out_build.clear(); j->out_probe.clear();
    int64_t n = key_cols[0]->length;
    std::vector<uint64_t> hashes((size_t)n);
    create_hashes(key_cols, j->nkeys, 0, j->force_collisions, n, hashes.data());  // fetch_probe_batch (:1199-1233)
    int64_t nbatches = 0;
    // JoinHashMapOffset = (usize, Option<u64>) (utils.rs:143)
    int64_t off_idx = 0; bool off_has = false; uint64_t off_next = 0;
    bool have_joined = false; int64_t joined_probe_idx = 0;
    for (;;) {
        std::vector<int64_t> bi, pi;
        int64_t remaining = j->batch_size;
        bool limit_hit = false; int64_t nx_idx = 0; uint64_t nx_next = 0;
        // chain_traverse! (utils.rs:147-187)
        auto traverse = [&](int64_t input_idx, uint64_t chain_idx) -> bool {
            uint64_t i = chain_idx - 1;
            for (;;) {
                bi.push_back((int64_t)i); pi.push_back(input_idx);
                remaining -= 1;
                uint64_t nxt = j->next[i];
                if (remaining == 0) {
                    if (input_idx == n - 1 && nxt == 0) limit_hit = false;  // next_offset = None
                    else { limit_hit = true; nx_idx = input_idx; nx_next = nxt; }
                    return true;
                }
                if (nxt == 0) break;
                i = nxt - 1;
            }
            return false;
        };
        bool returned = false;
        int64_t to_skip;
        if (!off_has) to_skip = off_idx;
        else if (off_next == 0) to_skip = off_idx + 1;
        else { returned = traverse(off_idx, off_next); to_skip = off_idx + 1; }
        for (int64_t row = to_skip; row < n && !returned; row++) {
            uint64_t *head = j->find(hashes[(size_t)row]);
            if (head) returned = traverse(row, *head);
        }
        bool next_none = !(returned && limit_hit);
        // equal_rows_arr (:1078-1118)
        std::vector<int64_t> fb, fp;
        for (size_t k = 0; k < bi.size(); k++) {
            int64_t b = 0, r = 0;
            j->locate(bi[k], b, r);
            bool eq = true;
            for (int c = 0; c < j->nkeys && eq; c++)
                eq = key_equal(j->build_batches[(size_t)b][(size_t)c], r, key_cols[c], pi[k], j->null_equals_null);
            if (eq) { fb.push_back(bi[k]); fp.push_back(pi[k]); }
        }
        // apply_join_filter_to_indices (hash_join.rs:1259-1271, utils.rs:1143-1176): BEFORE the visited bits and the
        // join-type fix-ups, so the residual predicate decides which rows count as matched
        if (j->filter && !fb.empty()) {
            std::vector<uint8_t> keep(fb.size(), 0);
            j->filter(j->filter_user, fb.data(), fp.data(), (int64_t)fb.size(), keep.data());
            size_t w = 0;
            for (size_t k = 0; k < fb.size(); k++)
                if (keep[k]) { fb[w] = fb[k]; fp[w] = fp[k]; w++; }
            fb.resize(w); fp.resize(w);
        }
        if (need_final(j->join_type))  // :1274-1278
            for (auto x : fb) j->visited[(size_t)x] = 1;
        bool last_has = !fp.empty(); int64_t last_joined = last_has ? fp.back() : 0;
        int64_t rs = have_joined ? joined_probe_idx + 1 : 0;
        int64_t re = next_none ? n : (last_has ? last_joined + 1 : 0);
        // adjust_indices_by_join_type (utils.rs:1234-1279)
        std::vector<uint8_t> inrange;
        auto mark = [&]() {
            inrange.assign((size_t)std::max<int64_t>(0, re - rs), 0);
            for (auto v : fp) if (v >= rs && v < re) inrange[(size_t)(v - rs)] = 1;
        };
        switch (j->join_type) {
        case OJ_INNER: case OJ_LEFT:
            j->out_build.insert(j->out_build.end(), fb.begin(), fb.end());
            j->out_probe.insert(j->out_probe.end(), fp.begin(), fp.end());
            break;
        case OJ_RIGHT: case OJ_FULL:
            j->out_build.insert(j->out_build.end(), fb.begin(), fb.end());
            j->out_probe.insert(j->out_probe.end(), fp.begin(), fp.end());
            mark();
            for (int64_t v = rs; v < re; v++)
                if (!inrange[(size_t)(v - rs)]) { j->out_build.push_back(-1); j->out_probe.push_back(v); }
            break;
        case OJ_RIGHT_SEMI:
            mark();
            for (int64_t v = rs; v < re; v++)
                if (inrange[(size_t)(v - rs)]) { j->out_build.push_back(-1); j->out_probe.push_back(v); }
            break;
        case OJ_RIGHT_ANTI:
            mark();
            for (int64_t v = rs; v < re; v++)
                if (!inrange[(size_t)(v - rs)]) { j->out_build.push_back(-1); j->out_probe.push_back(v); }
            break;
        default: break;  // LeftSemi / LeftAnti: produced at the end
        }
        nbatches++;
        if (next_none) break;
        off_idx = nx_idx; off_has = true; off_next = nx_next;  // state.advance (:920-925)
        if (last_has) { have_joined = true; joined_probe_idx = last_joined; }
    }
    if (out_batch_count) *out_batch_count = nbatches;
    return (int64_t)j->out_probe.size();
}

extern "C" void oracle_join_set_filter(oracle_join *j, oracle_join_filter_fn fn, void *user) { j->filter = fn; j->filter_user = user; }

extern "C" int64_t oracle_join_unmatched(oracle_join *j) {
    j->out_build.clear(); j->out_probe.clear();
    if (!need_final(j->join_type)) return -1;
    // get_final_indices_from_bit_map (utils.rs:1119-1141)
    for (int64_t i = 0; i < j->build_rows; i++) {
        bool v = j->visited[(size_t)i] != 0;
        if (j->join_type == OJ_LEFT_SEMI ? v : !v) { j->out_build.push_back(i); j->out_probe.push_back(-1); }
    }
    return (int64_t)j->out_build.size();
}
extern "C" const int64_t *oracle_join_out_build(const oracle_join *j) { return j->out_build.data(); }
extern "C" const int64_t *oracle_join_out_probe(const oracle_join *j) { return j->out_probe.data(); }

// ------------------------------------------------------------------------------------------
// BatchPartitioner::partition_iter, Hash arm (repartition/mod.rs:162-217)
// ------------------------------------------------------------------------------------------
extern "C" int oracle_partition_ids(const ocol *const *key_cols, int n_keys, int n_parts, uint32_t *out) {
    int64_t n = key_cols[0]->length;
    std::vector<uint64_t> h((size_t)n);
    create_hashes(key_cols, n_keys, 0, false, n, h.data());  // seeds (0,0,0,0): repartition/mod.rs:115
    for (int64_t i = 0; i < n; i++) out[i] = (uint32_t)(h[(size_t)i] % (uint64_t)n_parts);
    return 0;
}

// ------------------------------------------------------------------------------------------
// CPU baseline drivers — the reference's parallel plan shape on std::threads:
//   P x AggregateExec(Partial) over row ranges, 8192-row batches
//   -> RepartitionExec(Hash(keys), P) -> P x AggregateExec(FinalPartitioned)
// (core/src/physical_planner.rs:771-850; golden plan sqllogictest/test_files/tpch/q1.slt.part:47-60)
// ------------------------------------------------------------------------------------------
static ocol mk_fixed(int type, const void *values, int64_t off, int64_t len) {
    ocol c; memset(&c, 0, sizeof c);
    c.type = type; c.values = values; c.offset = off; c.length = len;
    return c;
}

struct I64CountPartial {  // GroupValuesPrimitive<i64> + CountGroupsAccumulator, fused loop-wise per batch
    std::vector<uint64_t> keys; std::vector<int64_t> counts; std::vector<uint32_t> slots; std::vector<uint32_t> gids;
    I64CountPartial() { slots.assign(1 << 12, 0); }
    void grow() {
        std::vector<uint32_t> ns(slots.size() * 2, 0);
        size_t mask = ns.size() - 1;
        for (size_t g = 0; g < keys.size(); g++) {
            size_t s = h_u64(keys[g], 1) & mask;
            while (ns[s]) s = (s + 1) & mask;
            ns[s] = (uint32_t)g + 1;
        }
        slots.swap(ns);
    }
    void push(const int64_t *k, const int64_t *add, int64_t n) {
        gids.resize((size_t)n);
        for (int64_t i = 0; i < n; i++) {  // intern
            if ((keys.size() + 1) * 2 > slots.size()) grow();
            uint64_t key = (uint64_t)k[i];
            size_t mask = slots.size() - 1, s = h_u64(key, 1) & mask;
            for (;;) {
                uint32_t e = slots[s];
                if (!e) { slots[s] = (uint32_t)keys.size() + 1; gids[(size_t)i] = (uint32_t)keys.size(); keys.push_back(key); break; }
                if (keys[e - 1] == key) { gids[(size_t)i] = e - 1; break; }
                s = (s + 1) & mask;
            }
        }
        counts.resize(keys.size(), 0);  // update_batch / merge_batch
        if (add) for (int64_t i = 0; i < n; i++) counts[gids[(size_t)i]] += add[i];
        else for (int64_t i = 0; i < n; i++) counts[gids[(size_t)i]] += 1;
    }
};

extern "C" double oracle_bench_groupby_count_i64(const int64_t *keys, int64_t n, int threads, int64_t batch,
                                                 int64_t *out_groups, uint64_t *out_checksum) {
    int P = std::max(1, threads);
    auto t0 = std::chrono::steady_clock::now();
    std::vector<I64CountPartial> partial((size_t)P);
    // exchange buffers [src][dst]
    std::vector<std::vector<std::vector<int64_t>>> xk((size_t)P, std::vector<std::vector<int64_t>>((size_t)P)), xc = xk;
    std::vector<std::thread> th;
    for (int p = 0; p < P; p++)
        th.emplace_back([&, p]() {
            int64_t lo = n * p / P, hi = n * (p + 1) / P;
            for (int64_t s = lo; s < hi; s += batch) partial[(size_t)p].push(keys + s, nullptr, std::min(batch, hi - s));
            auto &pp = partial[(size_t)p];
            for (size_t g = 0; g < pp.keys.size(); g++) {  // BatchPartitioner: hash % P with the fixed-seed hash
                size_t d = (size_t)(h_u64(pp.keys[g], 0) % (uint64_t)P);
                xk[(size_t)p][d].push_back((int64_t)pp.keys[g]);
                xc[(size_t)p][d].push_back(pp.counts[g]);
            }
        });
    for (auto &t : th) t.join();
    th.clear();
    std::vector<I64CountPartial> fin((size_t)P);
    for (int p = 0; p < P; p++)
        th.emplace_back([&, p]() {
            for (int s = 0; s < P; s++) {
                auto &k = xk[(size_t)s][(size_t)p]; auto &c = xc[(size_t)s][(size_t)p];
                for (size_t o = 0; o < k.size(); o += (size_t)batch)
                    fin[(size_t)p].push(k.data() + o, c.data() + o, (int64_t)std::min<size_t>((size_t)batch, k.size() - o));
            }
        });
    for (auto &t : th) t.join();
    auto t1 = std::chrono::steady_clock::now();
    int64_t groups = 0; uint64_t cks = 0;
    for (auto &f : fin) {
        groups += (int64_t)f.keys.size();
        for (size_t g = 0; g < f.keys.size(); g++) cks += mix64(f.keys[g]) * (uint64_t)f.counts[g];
    }
    if (out_groups) *out_groups = groups;
    if (out_checksum) *out_checksum = cks;
    return std::chrono::duration<double>(t1 - t0).count();
}

extern "C" double oracle_bench_q1(const uint8_t *flag, const uint8_t *status, const void *const *dec_cols, int64_t n,
                                  int threads, int64_t batch, int64_t *out_groups, uint64_t *out_checksum) {
    // Q1 shape: GROUP BY (flag, status) [2 x UInt8 keys]; SUM(c0..c3), AVG(c0), AVG(c1), AVG(c4), COUNT(*)
    // through the generic oracle operator (GroupValuesRows + accumulators), Partial -> exchange -> Final.
    int P = std::max(1, threads);
    oagg_desc d[8];
    for (int i = 0; i < 8; i++) { d[i].func = i < 4 ? OAGG_SUM : (i < 7 ? OAGG_AVG : OAGG_COUNT); d[i].input_type = OT_DEC128; d[i].precision = 15; d[i].scale = 2; }
    const int argcol[8] = {0, 1, 2, 3, 0, 1, 4, -1};
    auto t0 = std::chrono::steady_clock::now();
    std::vector<oracle_agg *> part((size_t)P);
    std::vector<std::thread> th;
    for (int p = 0; p < P; p++)
        th.emplace_back([&, p]() {
            oracle_agg *a = oracle_agg_new(OMODE_PARTIAL, 2, 8, d);
            part[(size_t)p] = a;
            int64_t lo = n * p / P, hi = n * (p + 1) / P;
            for (int64_t s = lo; s < hi; s += batch) {
                int64_t len = std::min(batch, hi - s);
                ocol k0 = mk_fixed(OT_U8, flag, s, len), k1 = mk_fixed(OT_U8, status, s, len);
                ocol v[5];
                for (int c = 0; c < 5; c++) { v[c] = mk_fixed(OT_DEC128, dec_cols[c], s, len); v[c].precision = 15; v[c].scale = 2; }
                const ocol *g[2] = {&k0, &k1};
                const ocol *args[8];
                for (int i = 0; i < 8; i++) args[i] = argcol[i] < 0 ? nullptr : &v[argcol[i]];
                oracle_agg_push(a, g, args, nullptr);
            }
            oracle_agg_emit(a, -1);
        });
    for (auto &t : th) t.join();
    // exchange is a handful of rows: one FinalPartitioned merge (all partitions' rows hash somewhere;
    // merging on one thread is the same work the reference does across <=4 non-empty partitions).
    oracle_agg *fin = oracle_agg_new(OMODE_FINAL_PARTITIONED, 2, 8, d);
    for (int p = 0; p < P; p++) {
        oracle_agg *a = part[(size_t)p];
        std::vector<ocol> cols(a->out.size());
        for (size_t c = 0; c < a->out.size(); c++) {
            OutCol &o = a->out[c];
            cols[c] = mk_fixed(o.type, o.values.data(), 0, o.length);
            cols[c].validity = o.validity.empty() ? nullptr : o.validity.data();
            cols[c].precision = o.precision; cols[c].scale = o.scale;
        }
        if (cols.empty() || cols[0].length == 0) continue;
        const ocol *g[2] = {&cols[0], &cols[1]};
        std::vector<const ocol *> args;
        for (size_t c = 2; c < cols.size(); c++) args.push_back(&cols[c]);
        oracle_agg_push(fin, g, args.data(), nullptr);
    }
    int nc = oracle_agg_emit(fin, -1);
    auto t1 = std::chrono::steady_clock::now();
    uint64_t cks = 0;
    if (nc > 0)
        for (auto &o : fin->out) for (auto b : o.values) cks = cks * 1099511628211ULL + b;
    if (out_groups) *out_groups = nc > 0 ? fin->out[0].length : -1;
    if (out_checksum) *out_checksum = cks;
    for (auto *a : part) oracle_agg_free(a);
    oracle_agg_free(fin);
    return std::chrono::duration<double>(t1 - t0).count();
}

extern "C" double oracle_bench_join_i64(const int64_t *build_keys, int64_t nb, const int64_t *probe_keys, int64_t np,
                                        int threads, int64_t batch, int64_t *out_rows, uint64_t *out_checksum) {
    // HashJoinExec(Partitioned, Inner) shape: both sides hash-repartitioned into P partitions,
    // P independent build+probe tasks (hash_join.rs:607-621); output = index pairs.
    int P = std::max(1, threads);
    auto t0 = std::chrono::steady_clock::now();
    std::vector<std::vector<int64_t>> bk((size_t)P), bi((size_t)P), pk((size_t)P), pi((size_t)P);
    {
        // RepartitionExec: one task per input partition (row range) — done with P threads and
        // per-(src,dst) buffers concatenated afterwards.
        std::vector<std::vector<std::vector<int64_t>>> tb((size_t)P, std::vector<std::vector<int64_t>>((size_t)P)), tbi = tb, tp = tb, tpi = tb;
        std::vector<std::thread> th;
        for (int p = 0; p < P; p++)
            th.emplace_back([&, p]() {
                for (int64_t i = nb * p / P; i < nb * (p + 1) / P; i++) {
                    size_t d = (size_t)(h_u64((uint64_t)build_keys[i], 0) % (uint64_t)P);
                    tb[(size_t)p][d].push_back(build_keys[i]); tbi[(size_t)p][d].push_back(i);
                }
                for (int64_t i = np * p / P; i < np * (p + 1) / P; i++) {
                    size_t d = (size_t)(h_u64((uint64_t)probe_keys[i], 0) % (uint64_t)P);
                    tp[(size_t)p][d].push_back(probe_keys[i]); tpi[(size_t)p][d].push_back(i);
                }
            });
        for (auto &t : th) t.join();
        for (int d = 0; d < P; d++)
            for (int s = 0; s < P; s++) {
                bk[(size_t)d].insert(bk[(size_t)d].end(), tb[(size_t)s][(size_t)d].begin(), tb[(size_t)s][(size_t)d].end());
                bi[(size_t)d].insert(bi[(size_t)d].end(), tbi[(size_t)s][(size_t)d].begin(), tbi[(size_t)s][(size_t)d].end());
                pk[(size_t)d].insert(pk[(size_t)d].end(), tp[(size_t)s][(size_t)d].begin(), tp[(size_t)s][(size_t)d].end());
                pi[(size_t)d].insert(pi[(size_t)d].end(), tpi[(size_t)s][(size_t)d].begin(), tpi[(size_t)s][(size_t)d].end());
            }
    }
    std::vector<int64_t> rows((size_t)P, 0);
    std::vector<uint64_t> cks((size_t)P, 0);
    std::vector<std::thread> th;
    for (int p = 0; p < P; p++)
        th.emplace_back([&, p]() {
            oracle_join *j = oracle_join_new(OJ_INNER, 0, 1, batch, 0);
            ocol b = mk_fixed(OT_I64, bk[(size_t)p].data(), 0, (int64_t)bk[(size_t)p].size());
            const ocol *bc[1] = {&b};
            oracle_join_build_batch(j, bc);
            oracle_join_build_finish(j);
            for (size_t s = 0; s < pk[(size_t)p].size(); s += (size_t)batch) {
                int64_t len = (int64_t)std::min<size_t>((size_t)batch, pk[(size_t)p].size() - s);
                ocol pc = mk_fixed(OT_I64, pk[(size_t)p].data(), (int64_t)s, len);
                const ocol *pcs[1] = {&pc};
                int64_t nbatch;
                int64_t r = oracle_join_probe(j, pcs, &nbatch);
                rows[(size_t)p] += r;
                for (int64_t k = 0; k < r; k++)
                    cks[(size_t)p] += mix64((uint64_t)bi[(size_t)p][(size_t)j->out_build[(size_t)k]] * 0x9e3779b97f4a7c15ULL +
                                            (uint64_t)pi[(size_t)p][s + (size_t)j->out_probe[(size_t)k]]);
            }
            oracle_join_free(j);
        });
    for (auto &t : th) t.join();
    auto t1 = std::chrono::steady_clock::now();
    int64_t r = 0; uint64_t c = 0;
    for (int p = 0; p < P; p++) { r += rows[(size_t)p]; c += cks[(size_t)p]; }
    if (out_rows) *out_rows = r;
    if (out_checksum) *out_checksum = c;
    return std::chrono::duration<double>(t1 - t0).count();
}
