// fastagg.h — two-level, cardinality-oblivious grouped aggregation for keys whose packed image fits
// 64 bits (GroupValuesPrimitive and short GroupValuesRows keys: group_values/primitive.rs:112-150,
// row.rs:94-146) with accumulators that decompose into the state words of ops.h.
//
// It is the reference's own Partial -> hash repartition -> Final plan
// (core/src/physical_planner.rs:814-849) applied INSIDE one GPU, at SM scale:
//   level 1  every CTA pre-aggregates its rows in a shared-memory hash table (or, for <= 4 groups,
//            in registers); rows whose key finds no room are spilled as single-row partial
//            states.  At the end the table is flushed as partial-state records.  A histogram over
//            the top hash bits of every record is kept on the side.
//   scatter  records are radix-partitioned by those hash bits (count -> scan -> scatter).
//   level 2  one CTA per partition merges its records in a shared-memory hash table and appends
//            the finished groups (key image, state words, seen flags) to dense output arrays.
// No table in HBM, no growth / rehash, no host round trip per batch, no global-atomic contention on
// hot keys; HBM sees only streaming reads and writes.  Partitions that hold more distinct keys
// than a level-2 table are handled by re-partitioning their overflow on the next hash bits.
#pragma once
#include "groups.h"
#include "ops.h"

namespace dfgpu {

static const int FA_MAX_OPS = 24;
static const int FA_MAX_KEYCOLS = 8;
static const int FA_PBITS = 13;
static const int FA_P = 1 << FA_PBITS;     // partitions per level
static const uint64_t FA_EMPTY = ~0ULL;    // free-slot marker; the real key ~0 lives in side group 1

struct FaKeyCol {
    ColView v;
    int32_t kind, off, w, null_bit;  // kind 0: fixed width, 1: bool, 3: Utf8 / Binary of <= 3 bytes read in place (inline string id)
};
struct FaOp {
    ColView val, filt;
    int32_t kind, nwords, woff, seen_bit, has_val, has_filt, pad0, pad1;
    uint64_t init_lo, init_hi;  // identity in the state representation of the HBM state arrays
};
// One launch's description: how to build the key image of a row and how to lift a row into each
// accumulator.  Lives in device memory; every CTA copies it to shared memory.
struct FaPlan {
    FaKeyCol keys[FA_MAX_KEYCOLS];
    FaOp ops[FA_MAX_OPS];
    int32_t nkeys, direct, null_off, nops;
    int32_t SW, RW, seen_word, pad;  // state words / record words (1 + SW) / index of the seen word or -1
    uint64_t key_or;                 // constant bits of the image (the tag and length of constant-length inline strings)
};

// register mode of level 1 (fastagg_reg.inc)
static const int FA_G0 = 4;         // groups held in registers
static const int FA_MAXROLES = 16;  // distinct accumulators of a launch: one warp role each
#ifndef FA_RR_V
#define FA_RR_V 4
#endif
static const int FA_RR = FA_RR_V;   // rows per thread per tile
struct FaRegPlan {
    int32_t nacc, slabs;
    int32_t op[FA_MAXROLES];      // logical op that defines accumulator i
    uint32_t seen[FA_MAXROLES];   // seen bits of every logical op computed by accumulator i
    int32_t woff[FA_MAXROLES];    // word offset in the CTA's partial record
    int32_t alias[FA_MAX_OPS];    // logical op -> word offset of the accumulator that computes it
    int32_t general_keys, pad;    // test / A-B hook (DFGPU_FA_NOK8=1): key phase without the one-byte-key form
};

// pieces of a partition's records after the record exchange between ranks (level 2 input)
static const int FA_MAXSEG = 64;
struct FaSegs {
    const uint64_t *seg_off = nullptr;  // [nseg][stride] record offsets into the receive buffer
    const uint32_t *seg_cnt = nullptr;  // [nseg][stride] record counts
    int nseg = 0, stride = 0;
};

struct FaOpSpec {
    int kind = 0, nwords = 1;
    int seen_bit = -1;  // ops sharing a NullState share the bit
    uint64_t init_lo = 0, init_hi = 0;
};

struct FaResult {
    int64_t G = 0;
    BufRef keys;                // [G] key images
    std::vector<BufRef> state;  // per op: [G * nwords]
    std::vector<BufRef> seen;   // per seen bit: [G] bytes
    int64_t null_gid = -1;      // group id of the NULL key of a single-column 8-byte key
    int64_t empty_gid = -1;     // group id of the key whose image is ~0 (kept outside the hash tables)
};

class FastAgg {
  public:
    FastAgg(Ctx *ctx, const std::vector<FaOpSpec> &ops);
    ~FastAgg();
    static bool kind_supported(int kind) { return kind != OP_MIN_I128 && kind != OP_MAX_I128 && kind < OP_AND_U64; }
    int SW() const { return SW_; }
    int seen_word() const { return seen_word_; }
    const std::vector<FaOpSpec> &ops() const { return ops_; }
    int nbits() const { return nbits_; }

    // Aggregate rows [0, n) described by `plan` (keys/ops column views already positioned at the
    // first row; kind/nwords/woff/seen_bit/init fields are filled in here).
    void push(FaPlan &plan, int64_t n, int64_t algo_bytes);
    // Size the record store once for a batch of `rows` rows pushed in `launches` chunks.
    void reserve(int64_t rows, int64_t launches);
    // Merge everything pushed so far into dense groups and leave the store empty.
    FaResult finalize();
    // Collective (every rank of the context's communicator): partition the records pushed so far by
    // the top hash bits, send every rank the partitions it owns (rank r owns FA_P / world consecutive
    // partitions) and keep what arrives; finalize() then merges only the owned partitions.  This is
    // Partial -> RepartitionExec(Hash) -> FinalPartitioned (physical_planner.rs:814-849) fused at the
    // level of partial-state records: no local final merge, no emit / re-ingest of state batches.
    void exchange();
    // The same exchange between N operators of ONE process and context ("virtual ranks", SURVEY 8e: how the reference tests
    // its 4-way partitioned plans in one process, hash_join.rs:1505-1550): ranks[r] plays rank r of a world of ranks.size().
    static void exchange_virtual(const std::vector<FastAgg *> &ranks);
    // finalize() + re-insert the groups as partial-state records; returns the exact group count.
    int64_t compact();
    size_t device_bytes() const;
    // what the sample of the first push said (tests, metrics): 1 cache path, 2 partition path, 0 not decided yet
    int path_mode() const { return mode_; }
    bool count_only_op() const { return count_only(); }
    size_t record_bytes() const { return (size_t)RW_ * 8; }

  private:
    void ensure_capacity(int64_t more);
    void reinsert(const FaResult &r);
    // partition path (fastagg_part.inc): COUNT(*)-shaped operators whose keys no per-CTA cache can hold
    bool part_eligible(const FaPlan &plan) const;
    void decide_mode(const FaPlan &plan, int64_t n);
    void push_part(const FaPlan &plan, int64_t n, int64_t algo_bytes);
    FaResult finalize_part();
    void reset_store();
    struct Piece { BufRef keys, offs, hist; int64_t rows = 0; int64_t arena_off = -1; /* >= 0: keys lie in the send arena (partition.cu) */ };
    // record exchange in phases, so that real ranks (collectives between the phases) and virtual ranks (loops) share them
    struct Xchg;
    bool count_only() const { return ops_.size() == 1 && ops_[0].kind == 1 /* OP_COUNT */ && nbits_ == 0 && SW_ == 1; }
    size_t xchg_meta_bytes() const;
    size_t xchg_small_off() const;
    void xchg_finish_small(Xchg &x, const BufRef &meta_all_dev);   // every rank's records rode in the meta blocks: keep the owned ones
    void xchg_prepare(Xchg &x, int W, int me);                                      // local: partition order, meta block on the device
    size_t xchg_plan(Xchg &x, const std::vector<uint8_t> &meta_all);                // from everybody's meta: what goes where; bytes my window needs
    void xchg_push(Xchg &x, void *const *windows);                                  // my runs -> the owners' windows
    void xchg_finish(Xchg &x, void *my_window, bool window_borrowed, const BufRef &meta_all_dev);  // pieces of what arrived; side groups
    // pull model (count-only operators): nothing is copied; the merge reads every source's pieces where they lie.
    // src[s * 3 + k] = stream k of rank s as this rank addresses it (peer mapping / plain pointer), recv = elements per piece
    void xchg_finish_pull(Xchg &x, const uint64_t *const *src, const int64_t *recv, const BufRef &meta_all_dev);
    BufRef arena_buf(size_t bytes, int64_t *arena_off);   // from the send arena when this operator may use it, else from the pool
    void arena_refresh();                                 // the arena moved: re-derive the pieces' pointers
    bool remote_loop_ = false;    // DFGPU_FA_REMOTE_LOOP=1
    bool xpull_ = false;          // the exchange that ran was a pull between real ranks: finalize ends with the closing barrier
    std::vector<BufRef> xhold_;        // received window (virtual ranks), gathered meta, piece tables: alive until the merge ran
    BufRef xp_offs_, xp_hist_;         // [npieces][FA_P] of the received pieces (count-only operators)
    std::vector<const uint64_t *> xp_base_;
    std::vector<int> xp_stride_;
    std::vector<Piece> pieces_;
    BufRef hot_keys_, hot_cnt_;
    int mode_ = 0;               // 0 undecided, 1 cache path (level 1), 2 partition path
    int64_t sample_distinct_ = -1, sample_rows_ = 0, sample_hot_ = 0;
    int part_force_ = -1;        // DFGPU_FA_PART: 1 always (when eligible), 0 never, -1 by sample
    int64_t reserve_hint_ = 0;
    Ctx *ctx_;
    std::vector<FaOpSpec> ops_;
    int SW_ = 0, RW_ = 0, seen_word_ = -1, nbits_ = 0;
    int S1_ = 0, S2max_ = 0;
    size_t smem1_ = 0, smem2_ = 0, smem2x_ = 0;
    int S2xmax_ = 0;
    BufRef recs_, scalars_, hist_, side_, plan_dev_, regplan_dev_, mode_flag_;
    // after exchange(): the received records, their pieces per owned partition, and how many there are
    BufRef xrecs_, xseg_off_, xseg_cnt_;
    int64_t xR_ = -1;
    int xfirst_ = 0, xcount_ = 0;  // owned partitions [xfirst_, xfirst_ + xcount_)
    int xW_ = 1;                   // ranks of the exchange that ran
    int64_t cap_ = 0, used_ub_ = 0;
    bool reg_mode_allowed_ = true;
    int reg_stages_ = 4;
    int64_t single_max_ = 8192;  // record sets up to this size are first merged by one CTA, unpartitioned
};

}  // namespace dfgpu
