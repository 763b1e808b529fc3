// groups.h — GroupTable: the device-resident open-addressing hash table that interns group /
// join keys to dense 32-bit ids.  It plays the role of the reference's three GroupValues
// implementations (physical-plan/src/aggregates/group_values/{primitive,bytes,row}.rs) and of the
// key side of JoinHashMap (joins/utils.rs:121-140).
//
// Layouts (chosen from the key schema):
//   KM_U64_DIRECT  one 8-byte column: slot = {key u64, gid u32}; NULL and the EMPTY bit pattern
//                  live in two side slots (cf. `null_group`, primitive.rs:118-122)
//   KM_U64_PACKED  columns packed into <= 8 bytes (incl. null bits), same slot layout
//   KM_ROW         packed fixed-width row of W bytes: slot = [tag32 | payload32], keys in a dense
//                  store indexed by gid (cf. GroupValuesRows, row.rs:94-146)
//   KM_VAR         one Utf8/Binary column: slot as KM_ROW, bytes in an append-only heap
//                  (cf. ArrowBytesMap, physical-expr/src/binary_map.rs:333-472)
// Utf8/Binary columns inside multi-column keys are first interned per column (their id becomes a
// 4-byte packed field); Dictionary columns intern their dictionary values once and map indices
// (hash_dictionary, common/src/hash_utils.rs:182-210).
//
// Insertion is lock-free and never spins: a CAS publishes either the key itself (U64) or a
// reference to the inserting row of the CURRENT batch (ROW/VAR, immutable input), so concurrent
// readers can always compare; a reader that races the winner before its dense id is written gets
// a "pending" slot reference (bit 31) that consumers resolve with resolve_gid() after the kernel.
#pragma once
#include "common.h"

namespace dfgpu {

static const uint32_t GID_PENDING = 0x80000000u;
static const uint32_t GID_NONE = 0x7fffffffu;  // excluded row (NULL join key) / find-only miss
static const uint32_t GID_UNSET = 0xffffffffu;
static const uint32_t STRID_NULL = 0xffffffffu;
static const int MAX_KEY_COLS = 16;

struct TableView {
    uint64_t *slots;
    uint64_t mask, cap;
    uint32_t *ngroups;
    uint32_t max_groups;
    uint32_t *overflow;
    int32_t layout;  // 0: two words per slot (U64), 1: one word per slot (ROW / VAR)
    int32_t force_collisions;
    uint64_t *gkeys;  // U64 key store [max_groups]
    uint8_t *grows;   // ROW key store [max_groups * W]
    int32_t W;
    uint8_t *heap;    // VAR byte heap
    unsigned long long *heap_used;
    int64_t *goff;    // VAR [max_groups]
    int32_t *glen;    // VAR [max_groups]
};

#ifdef __CUDACC__
__device__ __forceinline__ uint32_t resolve_gid(const TableView &t, uint32_t g) {
    if (g != GID_NONE && (g & GID_PENDING)) {
        uint64_t s = g & 0x7fffffffu;
        g = t.layout == 0 ? ((const uint32_t *)t.slots)[s * 4 + 2] : (uint32_t)t.slots[s];
    }
    return g;
}
#endif

enum KeyMode { KM_U64_DIRECT = 0, KM_U64_PACKED = 1, KM_ROW = 2, KM_VAR = 3 };
enum NullPolicy { NULLS_ARE_KEYS = 0, NULLS_EXCLUDED = 1 };

class GroupTable {
  public:
    // trust_nullability: columns declared NOT NULL get no null bit in the packed key image (arrow-rs rejects a batch
    // with nulls in a non-nullable field, RecordBatch::try_new), so e.g. (Int32 NOT NULL, Utf8) packs into 64 bits
    // pack_single_string: a lone Utf8/Binary key is laid out as a packed image holding its 4-byte string id (interned
    // in the column's own table) instead of KM_VAR, which makes it eligible for the two-level aggregation path
    GroupTable(Ctx *ctx, const std::vector<TypeInfo> &key_types, bool trust_nullability = false, bool pack_single_string = false);
    ~GroupTable();

    // Intern rows [0, n) of `cols` (one per key column, already sliced to the chunk).  Writes one
    // id per row into out (device, n u32; may hold pending refs — resolve with view()).
    // find_only: never insert; misses yield GID_NONE (join probe).
    void intern(const std::vector<const DevCol *> &cols, int64_t row_off, int64_t n, uint32_t *out, bool find_only,
                NullPolicy nulls);
    TableView view() const { return tv_; }
    int64_t num_groups_sync();     // exact (synchronises the stream)
    uint32_t max_groups() const { return tv_.max_groups; }
    const uint32_t *ngroups_dev() const { return tv_.ngroups; }
    std::vector<ColRef> emit_keys(int64_t G);  // key columns in gid order (types = key_types)
    void reset();
    void clear_shrink();   // reset + release every buffer (capacity 0)
    size_t device_bytes() const;
    KeyMode mode() const { return mode_; }
    void check_overflow();

    // --- hooks for the two-level aggregation path (fastagg.h) ---------------------------------
    struct KeyLayout { int kind, off, w, null_bit; bool dict; };
    std::vector<KeyLayout> key_layout() const {
        std::vector<KeyLayout> v;
        for (auto &c : cols_) v.push_back(KeyLayout{c.kind, c.off, c.w, c.null_bit, c.type.id == T_DICT});
        return v;
    }
    int null_off() const { return null_off_; }
    // Take ownership of a dense key store produced elsewhere (U64 layouts only): emit_keys(G) then
    // reads these images.  null_gid = id of the NULL-key group of a DIRECT table, or -1.
    // per-row ids of the Utf8/Binary key column `c` of a packed layout (interned in that column's own table;
    // STRID_NULL for NULL): lets the two-level path treat the column as a 4-byte field of the key image
    BufRef string_field_ids(size_t c, const DevCol &col, int64_t row_off, int64_t n) { return string_ids(cols_[c], col, row_off, n, false); }
    // shortest / longest value and first offset of rows [row_off, row_off + n) of a Utf8 / Binary key column -> dev_out3[0..2]
    void string_field_lengths(const DevCol &col, int64_t row_off, int64_t n, unsigned long long *dev_out3);
    void adopt_dense(const BufRef &keys, int64_t null_gid) {
        gkeys_ = keys;
        tv_.gkeys = keys ? keys->as<uint64_t>() : nullptr;
        adopted_ = true;
        adopted_null_gid_ = null_gid;
    }

  private:
    struct KeyCol {
        TypeInfo type;      // declared type (may be dictionary)
        int kind;           // 0 fixed, 1 bool, 2 string id
        int off = 0, w = 0;
        int null_bit = -1;
        std::unique_ptr<GroupTable> strs;  // per-column string interner (kind 2)
    };
    void ensure(int64_t n, int64_t var_bytes);
    void note_inserts();
    void grow(uint64_t new_max);
    void alloc_table(uint64_t new_max);
    BufRef string_ids(KeyCol &kc, const DevCol &col, int64_t row_off, int64_t n, bool find_only);

    Ctx *ctx_;
    std::vector<KeyCol> cols_;
    KeyMode mode_;
    int W_ = 8, null_off_ = 0, null_bytes_ = 0;
    TableView tv_;
    BufRef slots_, gkeys_, grows_, heap_, goff_, glen_, scalars_;  // scalars: ngroups, overflow, heap_used
    int64_t groups_ub_ = 0;
    int64_t heap_cap_ = 0, heap_ub_ = 0;
    // Lagged exact counters: after every inserting intern launch the 16 bytes of device scalars (group count, overflow
    // flag, heap bytes) are copied to pinned host memory and an event is recorded.  ensure() then waits for THAT event
    // only - not for the kernels queued behind it (level 1 of the same chunk, accumulate) - so the host can prepare
    // the next chunk while the device still works on this one.  Nothing but intern kernels changes these counters.
    uint32_t *count_host_ = nullptr;   // pinned: [0] groups, [1] overflow, [2..3] heap bytes used
    cudaEvent_t count_ev_ = nullptr;
    bool count_pending_ = false;
    bool adopted_ = false;
    int64_t adopted_null_gid_ = -1;
};

}  // namespace dfgpu
