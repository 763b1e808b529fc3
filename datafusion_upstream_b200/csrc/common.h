// common.h — host-side plumbing shared by every translation unit of libdfgpu:
// status/exception bridge, context (device, stream, memory pools, NCCL), device buffers,
// Arrow type descriptors, device-resident columns and RecordBatches.
#pragma once
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

#include <atomic>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <map>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <functional>
#include <condition_variable>
#include <vector>

#include "../../include/dfgpu.h"

namespace dfgpu {

// ------------------------------------------------------------------------------------------------
// errors: C++ exceptions inside, status codes + thread-local message at the ABI
// ------------------------------------------------------------------------------------------------
struct Error {
    dfgpu_status code;
    std::string msg;
};
[[noreturn]] inline void fail(dfgpu_status code, const std::string &msg) { throw Error{code, msg}; }
void set_last_error(const std::string &m);
const std::string &last_error();
void *pinned_alloc(size_t n);
void pinned_free(void *p);

#define DF_CUDA(expr)                                                                              \
    do {                                                                                           \
        cudaError_t _e = (expr);                                                                   \
        if (_e != cudaSuccess)                                                                     \
            ::dfgpu::fail(DFGPU_CUDA, std::string("CUDA error: ") + cudaGetErrorString(_e) + " at " + \
                                          __FILE__ + ":" + std::to_string(__LINE__));              \
    } while (0)

// ------------------------------------------------------------------------------------------------
// types
// ------------------------------------------------------------------------------------------------
enum TypeId : int32_t {
    T_BOOL = 1, T_I8, T_I16, T_I32, T_I64, T_U8, T_U16, T_U32, T_U64, T_F32, T_F64, T_DATE32, T_DATE64,
    T_TS, T_TIME32, T_TIME64, T_DEC128, T_UTF8, T_LARGE_UTF8, T_BINARY, T_LARGE_BINARY, T_DICT
};

struct TypeInfo {
    int32_t id = 0;
    int32_t precision = 0, scale = 0;
    std::string format;  // Arrow C format string (kept verbatim for pass-through types)
    std::string name;
    bool nullable = true;
    bool dict_ordered = false;
    int32_t index_id = 0;             // T_DICT: type id of the indices
    std::shared_ptr<TypeInfo> value;  // T_DICT: value type

    int width() const;  // fixed byte width, 0 for bool / var-width / dict
    bool is_var() const { return id == T_UTF8 || id == T_LARGE_UTF8 || id == T_BINARY || id == T_LARGE_BINARY; }
    bool is_large() const { return id == T_LARGE_UTF8 || id == T_LARGE_BINARY; }
    bool is_signed_int() const {
        return id == T_I8 || id == T_I16 || id == T_I32 || id == T_I64 || id == T_DATE32 || id == T_DATE64 ||
               id == T_TS || id == T_TIME32 || id == T_TIME64;
    }
    bool is_unsigned_int() const { return id == T_U8 || id == T_U16 || id == T_U32 || id == T_U64; }
    bool is_float() const { return id == T_F32 || id == T_F64; }
    const TypeInfo &value_type() const { return id == T_DICT ? *value : *this; }
    std::string describe() const;
};
int type_width(int32_t id);
TypeInfo type_from_schema(const ArrowSchema *s);
TypeInfo make_type(int32_t id, const std::string &name, bool nullable = true, int precision = 0, int scale = 0);
std::string format_of(const TypeInfo &t);

struct Schema {
    std::vector<TypeInfo> fields;
};
Schema schema_from_arrow(const ArrowSchema *s);  // "+s" struct schema -> fields
void schema_to_arrow(const Schema &s, ArrowSchema *out);

// ------------------------------------------------------------------------------------------------
// context
// ------------------------------------------------------------------------------------------------
struct PinnedPool {
    std::mutex mu;
    std::multimap<size_t, void *> free_list;
    std::map<void *, size_t> sizes;
    size_t cached = 0;
    void *alloc(size_t bytes);
    void release(void *p);
    void trim();
};

// Size-class cache of device blocks in front of the stream-ordered allocator.  An operator step
// allocates the same multi-GB scratch buffers (record store, partition buffer, outputs) every
// time; cudaMallocAsync takes milliseconds when its pool has to carve or map such a block, so
// freed blocks >= 256 KiB are kept here and handed back without a driver call.  Safe because all
// work of a context is ordered on its one stream.
struct BufCache {
    std::mutex mu;
    std::multimap<size_t, void *> free_list;
    size_t cached = 0, limit = (size_t)48 << 30;
    static size_t size_class(size_t n);  // 0: not cached
    void *take(size_t cls);
    bool give(size_t cls, void *p);      // false: cache full, caller frees
    void trim(cudaStream_t st);
};

}  // namespace dfgpu

struct dfgpu_ctx {
    dfgpu_options opts;
    int device = 0;
    cudaStream_t stream = nullptr;
    cudaStream_t copy_stream = nullptr;
    int sm_count = 148;
    size_t smem_optin = 0;
    int64_t launches = 0;
    int64_t h2d_bytes = 0, d2h_bytes = 0, nvlink_bytes = 0;
    void *nccl_comm = nullptr;  // ncclComm_t
    std::string nccl_error;     // sticky: an asynchronous NCCL error aborted the communicator (ncclCommAbort); every later collective fails with it
    // Receive window of the peer-memory exchange (partition.cu): cudaMalloc'd (IPC-exportable), mapped by every
    // peer of the communicator; peers[r] = this process' mapping of rank r's window.
    struct PeerWindow {
        void *buf = nullptr;
        size_t bytes = 0;
        uint64_t gen = 0;
        int state = 0;  // 0 untried, 1 in use, -1 unavailable on some rank: NCCL send/recv instead
        struct Peer { uint64_t gen = 0; void *ptr = nullptr; } peers[64];
        void *scratch = nullptr;  // small device staging for the handle all-gather / barrier
        // send arena only (pull model): bump allocator owned by one operator at a time
        size_t used = 0;
        const void *owner = nullptr;
        std::vector<void *> retired;  // outgrown buffers some peer may still have mapped: freed at the next handle exchange
    } win, swin;  // win: receive window (peers store into it); swin: send arena (peers load from it)
    dfgpu::PinnedPool pinned;
    dfgpu::BufCache bufs;
    // DFGPU_GUARD=1 (debug): every DevBuf gets a 256-byte red zone either side, checked when it is freed; see batch.cu
    int guard = 0;
    unsigned long long *guard_words = nullptr;  // [0] = corrupted red zones seen, [1] = size of the first offender
    uint64_t *flush_dummy = nullptr;
    bool import_pinned_pending = false;   // batch_import queued a copy out of pinned host memory (batch.cu upload())
    bool import_defer_wait = false;       // the caller of batch_import waits for those copies itself (pipelined push, agg_host.inc)
    cudaEvent_t import_event = nullptr;
    std::atomic<int> refs{1};  // the creator's handle + every batch / operator created on this ctx
    // optional per-kernel CUDA-event timing (bench.py roofline): name -> (events, launches)
    bool profiling = false;
    struct ProfRec {
        const char *name;
        cudaEvent_t start, stop;
        int64_t algo_bytes;
    };
    std::vector<ProfRec> prof;
};

namespace dfgpu {
typedef dfgpu_ctx Ctx;
void ctx_release(Ctx *c);  // drops one reference; tears the context down with the last one
// Reference-counted context pointer: handles may be destroyed in any order by the host language
// (garbage collectors, Rust drop order) without dangling.
struct CtxHold {
    Ctx *c = nullptr;
    CtxHold() {}
    CtxHold(const CtxHold &o) : c(o.c) { if (c) c->refs.fetch_add(1); }
    CtxHold &operator=(const CtxHold &o) { return *this = o.c; }
    CtxHold &operator=(Ctx *n) {
        if (n) n->refs.fetch_add(1);
        if (c) ctx_release(c);
        c = n;
        return *this;
    }
    ~CtxHold() { if (c) ctx_release(c); }
    operator Ctx *() const { return c; }
    Ctx *operator->() const { return c; }
};

// Scoped CUDA-event timer around one kernel launch (active only while ctx->profiling): the events
// are recorded on the launching stream, so the elapsed time is that kernel's device duration.
struct KTimer {
    Ctx *ctx;
    size_t idx = 0;
    bool on;
    KTimer(Ctx *c, const char *name, int64_t algo_bytes = 0) : ctx(c), on(c->profiling) {
        if (!on) return;
        dfgpu_ctx::ProfRec r{name, nullptr, nullptr, algo_bytes};
        cudaEventCreate(&r.start);
        cudaEventCreate(&r.stop);
        cudaEventRecord(r.start, c->stream);
        idx = c->prof.size();
        c->prof.push_back(r);
    }
    ~KTimer() {
        if (on) cudaEventRecord(ctx->prof[idx].stop, ctx->stream);
    }
};

// NVTX range per ABI call and per operator phase (nsys / ncu --nvtx show them; free when no tool is attached)
struct NvtxRange {
    explicit NvtxRange(const char *name);
    ~NvtxRange();
};

// DFGPU_TRACE=1: host-side timeline on stderr (microseconds since the previous trace point, any translation unit)
inline void df_trace(const char *label) {
    static const bool on = getenv("DFGPU_TRACE") != nullptr;
    if (!on) return;
    static auto last = std::chrono::steady_clock::now();
    auto now = std::chrono::steady_clock::now();
    fprintf(stderr, "[dfgpu trace] %-36s +%8.1f us\n", label, std::chrono::duration<double, std::micro>(now - last).count());
    last = now;
}

// communicator wrappers (partition.cu; NCCL is dlopen'd there): byte counts, ctx->stream
void guard_poll(Ctx *ctx);  // DFGPU_GUARD=1: fail with DFGPU_INTERNAL if a freed buffer had a damaged red zone
void comm_poll(Ctx *ctx);  // world > 1: raise (and make sticky) an asynchronous NCCL error; aborts the communicator
void comm_all_gather(Ctx *ctx, const void *send, void *recv, size_t bytes_per_rank);
void comm_group_start();
void comm_group_end();
void comm_send(Ctx *ctx, const void *p, size_t bytes, int peer);
void comm_recv(Ctx *ctx, void *p, size_t bytes, int peer);
// Peer-memory exchange.  peer_window_prepare is collective: makes this rank's receive window at least `need`
// bytes, maps every peer's window and returns the mappings in peer_ptrs[world] (own window included); false, on
// every rank alike, when peer mapping is unavailable.  peer_push copies n runs (src on this device, dst in any
// mapped window) with one kernel; peer_barrier orders the pushes of all ranks before what follows on the stream.
bool peer_window_prepare(Ctx *ctx, size_t need, void **peer_ptrs);
void peer_push(Ctx *ctx, int n, const void *const *src, void *const *dst, const size_t *bytes);
void peer_barrier(Ctx *ctx);
void peer_window_destroy(Ctx *ctx);
// Send arena (pull model of the record exchange).  An operator places what its peers will read - its partitioned key
// pieces - in this rank's arena: cudaMalloc'd (IPC-exportable), kept across operators, grown locally.  arena_alloc is
// local: bytes at a 256-byte aligned offset, or false when the arena belongs to another live operator (the caller then
// uses pool memory and the push model); *moved is set when the arena was re-allocated and its live prefix copied, i.e.
// pointers into it must be re-derived from arena_base() + offset.  arena_map is collective: gens[r] = arena generation
// of rank r (gathered by the caller); handles travel only when some rank's arena grew since they were last exchanged.
bool arena_alloc(Ctx *ctx, const void *owner, size_t bytes, size_t *off, bool *moved);
void arena_release(Ctx *ctx, const void *owner);
void *arena_base(Ctx *ctx);
uint64_t arena_gen(Ctx *ctx);
bool arena_map(Ctx *ctx, const uint64_t *gens, void **peer_ptrs);

// device buffer allocated from the stream-ordered pool of ctx->stream
struct DevBuf {
    Ctx *ctx = nullptr;
    void *ptr = nullptr;
    size_t bytes = 0;
    size_t cls = 0;  // size class when the block came from / returns to the context's BufCache
    DevBuf() {}
    std::shared_ptr<DevBuf> parent;  // set for views: the memory belongs to (and lives as long as) parent
    DevBuf(Ctx *c, size_t n, bool zero = false);
    ~DevBuf();
    DevBuf(const DevBuf &) = delete;
    DevBuf &operator=(const DevBuf &) = delete;
    template <class T> T *as() const { return (T *)ptr; }
};
typedef std::shared_ptr<DevBuf> BufRef;
inline BufRef make_buf(Ctx *c, size_t n, bool zero = false) { return std::make_shared<DevBuf>(c, n, zero); }
// non-owning window [off, off + n) of another buffer
// non-owning handle on device memory managed elsewhere (the peer window)
inline BufRef make_borrowed(Ctx *ctx, void *ptr, size_t n) {
    BufRef v = std::make_shared<DevBuf>();
    v->ctx = ctx;
    v->ptr = ptr;
    v->bytes = n;
    v->parent = std::make_shared<DevBuf>();  // (a parent marks the memory as not ours to free)
    return v;
}
inline BufRef make_view(const BufRef &parent, size_t off, size_t n) {
    BufRef v = std::make_shared<DevBuf>();
    v->ctx = parent->ctx;
    v->ptr = (uint8_t *)parent->ptr + off;
    v->bytes = n;
    v->parent = parent;
    return v;
}

// POD view handed to kernels
struct ColView {
    const uint8_t *validity;  // bitmap or nullptr
    const void *values;       // fixed: element 0; bool: bitmap; var: data base; dict: indices
    const void *offsets;      // var: offsets of element 0
    int64_t vbit;             // bit offset of row 0 in validity
    int64_t dbit;             // bool: bit offset of row 0 in values
    int32_t type;             // TypeId of this physical array (T_DICT for dictionary)
    int32_t width;            // fixed width (index width for dict)
    int32_t large;            // var: 64-bit offsets
    int32_t index_signed;     // dict
};

// device-resident Arrow array
struct DevCol {
    TypeInfo type;
    int64_t length = 0;
    int64_t null_count = 0;
    const uint8_t *validity = nullptr;
    int64_t vbit = 0;
    const void *values = nullptr;
    int64_t dbit = 0;
    const void *offsets = nullptr;
    int64_t data_bytes = 0;          // var-width: bytes referenced by this column
    std::shared_ptr<DevCol> dict;    // dictionary values
    std::vector<BufRef> owned;       // keeps device memory alive
    std::shared_ptr<void> keepalive; // e.g. parent batch for slices

    ColView view() const;
};
typedef std::shared_ptr<DevCol> ColRef;

}  // namespace dfgpu

struct dfgpu_batch {
    dfgpu::CtxHold ctx;
    int64_t num_rows = 0;
    std::vector<dfgpu::ColRef> cols;
    dfgpu::Schema schema() const {
        dfgpu::Schema s;
        for (auto &c : cols) s.fields.push_back(c->type);
        return s;
    }
};

namespace dfgpu {
typedef dfgpu_batch Batch;

// ---- batch.cu ----
Batch *batch_import(Ctx *ctx, const ArrowArray *arr, const Schema &schema, bool on_device);
void batch_export(Ctx *ctx, const Batch *b, ArrowArray *out, ArrowSchema *out_schema);
Batch *batch_concat(Ctx *ctx, Batch *const *batches, int n);
// Host-side CoalesceBatches (coalesce.inc): packs small host Arrow batches into a pinned arena, one H2D per buffer per flush.
class HostCoalescer {
  public:
    HostCoalescer(Ctx *ctx, const Schema &schema, int64_t target_rows);
    ~HostCoalescer();
    HostCoalescer(const HostCoalescer &) = delete;
    HostCoalescer &operator=(const HostCoalescer &) = delete;
    static bool supports(const Schema &schema);   // fixed-width, Boolean and Utf8/Binary columns; no dictionaries
    int64_t rows() const { return rows_; }
    int64_t target() const { return target_; }
    void set_target(int64_t rows) { target_ = rows; }   // takes effect with the next append (the arenas grow on demand)
    int64_t flushes() const { return flushes_; }
    bool fits(const ArrowArray *batch) const;     // false: flush first (i32 offsets of a Utf8 column would overflow)
    void append(const ArrowArray *batch);         // copies; the caller may release the batch afterwards
    // n batches at once: when every column is fixed-width and free of nulls (the common shape of an evaluated aggregate
    // input) the copies into the arena are spread over host threads - a single thread packs ~10 GB/s, PCIe takes 55
    void append_many(const ArrowArray *const *batches, int n);
    Batch *flush();                               // nullptr when empty
  private:
    struct Col;
    Ctx *ctx_;
    Schema schema_;
    int64_t target_, rows_ = 0, flushes_ = 0;
    Col *sets_[2] = {nullptr, nullptr};
    cudaEvent_t ev_[2] = {nullptr, nullptr};
    bool pending_[2] = {false, false};
    int cur_ = 0;
};
Batch *batch_slice(Ctx *ctx, const Batch *b, int64_t off, int64_t len);
Batch *batch_filter(Ctx *ctx, const Batch *b, int predicate_col);
Batch *batch_eval_decimal(Ctx *ctx, const Batch *b, const dfgpu_expr_node *nodes, int n_nodes, const char *name);  // FilterExec::batch_filter (filter.rs:315-327)
// gather rows (compute::take); idx < 0 => NULL row.  idx is a device array of n int64.
ColRef col_take(Ctx *ctx, const DevCol &src, const int64_t *idx, int64_t n, bool idx_may_be_null);
// the same for several columns (each with its own index array): plain fixed-width columns share one launch
// packed: per column an alternative source (first != nullptr): the column's field inside row-wise packed rows and the row size
std::vector<ColRef> cols_take(Ctx *ctx, const std::vector<const DevCol *> &srcs, const std::vector<const int64_t *> &idx, int64_t n,
                              const std::vector<bool> &idx_may_be_null, const std::vector<std::pair<const uint8_t *, int>> *packed = nullptr);
struct PackedRows {
    BufRef rows;            // [n][row_bytes], empty when packing does not apply
    int row_bytes = 0;
    std::vector<int> off;   // per input column: offset of its field in a row, -1 = not packed
};
PackedRows pack_rows(Ctx *ctx, const std::vector<ColRef> &cols, int64_t n);
ColRef col_nulls(Ctx *ctx, const TypeInfo &t, int64_t n);
ColRef col_decode_dictionary(Ctx *ctx, const DevCol &src);  // Dictionary -> plain value type
int64_t count_nulls(Ctx *ctx, const uint8_t *validity, int64_t vbit, int64_t n);
// view of rows [off, off+len) of a column view
inline ColView view_at(ColView v, int64_t off) {
    v.vbit += off;
    if (v.type == T_BOOL) v.dbit += off;
    else if (v.offsets) v.offsets = (const uint8_t *)v.offsets + off * (v.large ? 8 : 4);
    else v.values = (const uint8_t *)v.values + off * v.width;
    return v;
}
// exclusive prefix sum of n int64 (device -> device); returns total via *total_host if non-null
void exclusive_scan_i64(Ctx *ctx, const int64_t *in, int64_t *out, int64_t n, int64_t *total_host);

// launch helper: counts launches for bench.py "gpu_launches"
inline void note_launch(Ctx *ctx, int n = 1) { ctx->launches += n; }
inline int grid_for(int64_t n, int block, Ctx *ctx, int max_waves = 64) {
    int64_t g = (n + block - 1) / block;
    int64_t cap = (int64_t)ctx->sm_count * max_waves;
    if (g > cap) g = cap;
    if (g < 1) g = 1;
    return (int)g;
}

}  // namespace dfgpu
