// partition.cu — RepartitionExec(Hash) on the device: BatchPartitioner::partition_iter
// (physical-plan/src/repartition/mod.rs:162-217) as a stable multi-split, and the exchange itself
// (the reference's per-partition channels, repartition/mod.rs:442-580, distributor_channels.rs)
// as grouped ncclSend/ncclRecv all-to-all over NVLink.
#include <functional>
#include <dlfcn.h>
#include <nccl.h>

#include <algorithm>
#include <mutex>

#include "device.cuh"
#include "groups.h"

namespace dfgpu {
void create_hashes_device(Ctx *ctx, const std::vector<const DevCol *> &cols, int64_t row_off, int64_t n, uint64_t seed, uint64_t *out);

static const int PART_T = 256;
static const int MAX_PARTS = 1024;

// pass 1: per-block histogram of hash % P over a contiguous row range
__global__ void __launch_bounds__(PART_T) k_part_count(const uint64_t *__restrict__ hashes, int64_t n, int64_t rows_per_block, int P,
                                                       uint32_t *__restrict__ pid, int64_t *__restrict__ hist /* [P][nblocks] */) {
    extern __shared__ uint32_t cnt[];
    for (int p = threadIdx.x; p < P; p += PART_T) cnt[p] = 0;
    __syncthreads();
    int64_t lo = (int64_t)blockIdx.x * rows_per_block, hi = min(n, lo + rows_per_block);
    for (int64_t i = lo + threadIdx.x; i < hi; i += PART_T) {
        uint32_t p = (uint32_t)(hashes[i] % (uint64_t)P);  // repartition/mod.rs:184-187
        pid[i] = p;
        atomicAdd(&cnt[p], 1u);
    }
    __syncthreads();
    for (int p = threadIdx.x; p < P; p += PART_T) hist[(int64_t)p * gridDim.x + blockIdx.x] = cnt[p];
}
// pass 2: stable scatter of row indices; rows keep input order inside each partition
__global__ void __launch_bounds__(PART_T) k_part_scatter(const uint32_t *__restrict__ pid, int64_t n, int64_t rows_per_block, int P,
                                                         const int64_t *__restrict__ starts /* [P][nblocks] */, int64_t *__restrict__ perm) {
    extern __shared__ int64_t sm64[];
    int64_t *base = sm64;                                 // [P] running cursor of this block
    uint32_t *wcnt = (uint32_t *)(sm64 + P);              // [warps][P]
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = PART_T / 32;
    for (int p = threadIdx.x; p < P; p += PART_T) base[p] = starts[(int64_t)p * gridDim.x + blockIdx.x];
    for (int k = threadIdx.x; k < nw * P; k += PART_T) wcnt[k] = 0;
    __syncthreads();
    int64_t lo = (int64_t)blockIdx.x * rows_per_block, hi = min(n, lo + rows_per_block);
    for (int64_t s = lo; s < hi; s += PART_T) {
        int64_t i = s + threadIdx.x;
        bool in = i < hi;
        uint32_t p = in ? pid[i] : 0xffffffffu;
        uint32_t peers = __match_any_sync(0xffffffffu, p);
        uint32_t rank = __popc(peers & ((1u << lane) - 1));
        if (in && rank == 0) wcnt[warp * P + p] = __popc(peers);
        __syncthreads();
        if (in) {
            int64_t off = base[p] + rank;
            for (int w = 0; w < warp; w++) off += wcnt[w * P + p];
            perm[off] = i;
        }
        __syncthreads();
        for (int q = threadIdx.x; q < P; q += PART_T) {
            uint32_t t = 0;
            for (int w = 0; w < nw; w++) { t += wcnt[w * P + q]; wcnt[w * P + q] = 0; }
            base[q] += t;
        }
        __syncthreads();
    }
}

// returns perm (row indices grouped by partition, stable) and host offsets[P + 1]
static BufRef partition_perm(Ctx *ctx, const Batch &b, const int32_t *key_cols, int n_keys, int P, std::vector<int64_t> &offsets) {
    if (P <= 0 || P > MAX_PARTS) fail(DFGPU_NOT_IMPLEMENTED, "number of hash partitions must be in 1..1024");
    std::vector<const DevCol *> keys;
    for (int i = 0; i < n_keys; i++) {
        if (key_cols[i] < 0 || key_cols[i] >= (int)b.cols.size()) fail(DFGPU_PLAN, "partition key column out of range");
        keys.push_back(b.cols[key_cols[i]].get());
    }
    int64_t n = b.num_rows;
    offsets.assign((size_t)P + 1, 0);
    BufRef perm = make_buf(ctx, (size_t)n * 8 + 16);
    if (n == 0) return perm;
    cudaStream_t st = ctx->stream;
    BufRef hashes = make_buf(ctx, (size_t)n * 8), pid = make_buf(ctx, (size_t)n * 4);
    create_hashes_device(ctx, keys, 0, n, 0, hashes->as<uint64_t>());  // seeds (0,0,0,0): repartition/mod.rs:115
    int nblocks = (int)std::min<int64_t>((n + 4095) / 4096, (int64_t)ctx->sm_count * 8);
    int64_t rpb = (n + nblocks - 1) / nblocks;
    rpb = (rpb + PART_T - 1) / PART_T * PART_T;
    nblocks = (int)((n + rpb - 1) / rpb);
    size_t hn = (size_t)P * nblocks;
    BufRef hist = make_buf(ctx, hn * 8), starts = make_buf(ctx, hn * 8);
    k_part_count<<<nblocks, PART_T, P * 4, st>>>(hashes->as<uint64_t>(), n, rpb, P, pid->as<uint32_t>(), hist->as<int64_t>());
    note_launch(ctx);
    exclusive_scan_i64(ctx, hist->as<int64_t>(), starts->as<int64_t>(), (int64_t)hn, nullptr);
    size_t smem = (size_t)P * 8 + (size_t)(PART_T / 32) * P * 4;
    if (smem > 48 * 1024) DF_CUDA(cudaFuncSetAttribute(k_part_scatter, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    k_part_scatter<<<nblocks, PART_T, smem, st>>>(pid->as<uint32_t>(), n, rpb, P, starts->as<int64_t>(), perm->as<int64_t>());
    note_launch(ctx);
    DF_CUDA(cudaGetLastError());
    // partition p starts at starts[p][0]
    std::vector<int64_t> first((size_t)P);
    DF_CUDA(cudaMemcpy2DAsync(first.data(), 8, starts->ptr, (size_t)nblocks * 8, 8, (size_t)P, cudaMemcpyDeviceToHost, st));
    DF_CUDA(cudaStreamSynchronize(st));
    for (int p = 0; p < P; p++) offsets[(size_t)p] = first[(size_t)p];
    offsets[(size_t)P] = n;
    return perm;
}

void partition_hash(Ctx *ctx, const Batch &b, const int32_t *key_cols, int n_keys, int P, Batch **out_parts) {
    std::vector<int64_t> off;
    BufRef perm;
    {
        KTimer kt(ctx, "stage:partition_perm");
        perm = partition_perm(ctx, b, key_cols, n_keys, P, off);
    }
    KTimer kt(ctx, "stage:partition_take");
    for (int p = 0; p < P; p++) {
        out_parts[p] = nullptr;
        int64_t cnt = off[(size_t)p + 1] - off[(size_t)p];
        if (cnt == 0) continue;  // empty sub-batches are not produced (repartition/mod.rs:196-214)
        std::unique_ptr<Batch> o(new Batch());
        o->ctx = ctx;
        o->num_rows = cnt;
        for (auto &c : b.cols) o->cols.push_back(col_take(ctx, *c, perm->as<int64_t>() + off[(size_t)p], cnt, false));
        out_parts[p] = o.release();
    }
}

// ------------------------------------------------------------------------------------------------
// NCCL (loaded lazily so single-GPU use has no NCCL dependency)
// ------------------------------------------------------------------------------------------------
struct NcclApi {
    void *lib = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId *) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*CommAbort)(ncclComm_t) = nullptr;
    ncclResult_t (*CommGetAsyncError)(ncclComm_t, ncclResult_t *) = nullptr;
    ncclResult_t (*Send)(const void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Recv)(void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    ncclResult_t (*AllGather)(const void *, void *, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
    const char *(*GetErrorString)(ncclResult_t) = nullptr;
};
static NcclApi &nccl() {
    static NcclApi api;
    if (!api.lib) {
        const char *names[] = {"libnccl.so.2", "libnccl.so"};
        for (auto nm : names) {
            api.lib = dlopen(nm, RTLD_NOW | RTLD_GLOBAL);
            if (api.lib) break;
        }
        if (!api.lib) fail(DFGPU_NCCL, std::string("cannot load libnccl: ") + dlerror());
#define LOAD(f) \
    *(void **)(&api.f) = dlsym(api.lib, "nccl" #f); \
    if (!api.f) fail(DFGPU_NCCL, "libnccl lacks nccl" #f);
        LOAD(GetUniqueId) LOAD(CommInitRank) LOAD(CommDestroy) LOAD(CommAbort) LOAD(CommGetAsyncError) LOAD(Send) LOAD(Recv) LOAD(GroupStart) LOAD(GroupEnd)
        LOAD(AllGather) LOAD(GetErrorString)
#undef LOAD
    }
    return api;
}
#define DF_NCCL(expr)                                                                                        \
    do {                                                                                                     \
        ncclResult_t _r = (expr);                                                                            \
        if (_r != ncclSuccess) fail(DFGPU_NCCL, std::string("NCCL error: ") + nccl().GetErrorString(_r));   \
    } while (0)

void ctx_init_nccl(Ctx *ctx) {
    if (!ctx->opts.nccl_unique_id) fail(DFGPU_PLAN, "world > 1 requires a ncclUniqueId");
    ncclUniqueId id;
    memcpy(&id, ctx->opts.nccl_unique_id, sizeof id);
    ncclComm_t comm;
    DF_NCCL(nccl().CommInitRank(&comm, ctx->opts.world, id, ctx->opts.rank));
    ctx->nccl_comm = comm;
}
void ctx_destroy_nccl(Ctx *ctx) {
    peer_window_destroy(ctx);
    if (ctx->nccl_comm) nccl().CommDestroy((ncclComm_t)ctx->nccl_comm);
    ctx->nccl_comm = nullptr;
}
// Asynchronous failures (a peer died, a link error) surface through ncclCommGetAsyncError, not through the enqueue calls.
// Checked before every collective of this context: on an error the communicator is aborted - which unblocks kernels that
// would otherwise wait for the dead peer for ever - and the error sticks to the context, so the operator fails with a
// DFGPU_NCCL status instead of hanging (the reference's analogue: a failed input stream ends the plan with its Err).
static void comm_check(Ctx *ctx) {
    if (!ctx->nccl_error.empty()) fail(DFGPU_NCCL, ctx->nccl_error);
    if (!ctx->nccl_comm) fail(DFGPU_NCCL, "NCCL error: no communicator (world > 1 requires a ncclUniqueId)");
    ncclResult_t async = ncclSuccess;
    ncclResult_t q = nccl().CommGetAsyncError((ncclComm_t)ctx->nccl_comm, &async);
    if (q == ncclSuccess && (async == ncclSuccess || async == ncclInProgress)) return;
    ctx->nccl_error = std::string("NCCL error (asynchronous): ") + nccl().GetErrorString(q != ncclSuccess ? q : async) + "; communicator aborted";
    nccl().CommAbort((ncclComm_t)ctx->nccl_comm);
    ctx->nccl_comm = nullptr;
    fail(DFGPU_NCCL, ctx->nccl_error);
}

void comm_poll(Ctx *ctx) {
    if (ctx->opts.world > 1) comm_check(ctx);
}

// thin communicator wrappers for the other translation units (fastagg.cu): byte-typed, on ctx->stream
void comm_all_gather(Ctx *ctx, const void *send, void *recv, size_t bytes_per_rank) {
    comm_check(ctx);
    DF_NCCL(nccl().AllGather(send, recv, bytes_per_rank, ncclChar, (ncclComm_t)ctx->nccl_comm, ctx->stream));
}
void comm_group_start() { DF_NCCL(nccl().GroupStart()); }
void comm_group_end() { DF_NCCL(nccl().GroupEnd()); }
void comm_send(Ctx *ctx, const void *p, size_t bytes, int peer) {
    if (!ctx->nccl_error.empty() || !ctx->nccl_comm) comm_check(ctx);
    DF_NCCL(nccl().Send(p, bytes, ncclChar, peer, (ncclComm_t)ctx->nccl_comm, ctx->stream));
    if (peer != ctx->opts.rank) ctx->nvlink_bytes += (int64_t)bytes;
}
void comm_recv(Ctx *ctx, void *p, size_t bytes, int peer) {
    if (!ctx->nccl_error.empty() || !ctx->nccl_comm) comm_check(ctx);
    DF_NCCL(nccl().Recv(p, bytes, ncclChar, peer, (ncclComm_t)ctx->nccl_comm, ctx->stream));
}

// ---- peer-memory exchange ---------------------------------------------------------------------------
// NVSwitch gives every GPU full-bandwidth loads / stores into every peer's HBM.  The record exchange of the
// fused sharded aggregation (fastagg_host.inc) therefore does not go through ncclSend / ncclRecv: every rank
// exports one receive window (cudaIpcGetMemHandle), maps the windows of its peers once, and a single copy
// kernel stores each peer's run straight into that peer's window.  NCCL only carries the 80-byte handles and
// the two barriers.
struct PushDesc {
    const void *src[64];
    void *dst[64];
    unsigned long long vec[64];  // vectors (16 or 8 bytes) per run
    int n;
};
template <class V>
__global__ void __launch_bounds__(512) k_peer_push(PushDesc d) {
    // runs are taken in turn by the whole grid with coalesced vector stores; CTAs start on different peers so all
    // links are busy at once
    for (int r = 0; r < d.n; r++) {
        const int rr = (r + blockIdx.x) % d.n;
        const V *__restrict__ s = (const V *)d.src[rr];
        V *__restrict__ t = (V *)d.dst[rr];
        const unsigned long long nv = d.vec[rr];
        for (unsigned long long i = blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x; i < nv; i += (unsigned long long)gridDim.x * blockDim.x)
            t[i] = __ldcs(s + i);
    }
}
void peer_push(Ctx *ctx, int n, const void *const *src, void *const *dst, const size_t *bytes) {
    PushDesc d;
    memset(&d, 0, sizeof d);
    bool v16 = true;
    for (int i = 0; i < n; i++)
        if (bytes[i]) v16 = v16 && !(bytes[i] & 15) && !((uintptr_t)src[i] & 15) && !((uintptr_t)dst[i] & 15);
    for (int i = 0; i < n; i++) {
        if (!bytes[i]) continue;
        if ((bytes[i] & 7) || ((uintptr_t)src[i] & 7) || ((uintptr_t)dst[i] & 7)) fail(DFGPU_INTERNAL, "peer push: runs must be 8-byte aligned");
        d.src[d.n] = src[i];
        d.dst[d.n] = dst[i];
        d.vec[d.n] = bytes[i] / (v16 ? 16 : 8);
        d.n++;
    }
    if (!d.n) return;
    if (v16) k_peer_push<uint4><<<ctx->sm_count * 4, 512, 0, ctx->stream>>>(d);
    else k_peer_push<uint2><<<ctx->sm_count * 4, 512, 0, ctx->stream>>>(d);
    note_launch(ctx);
    DF_CUDA(cudaGetLastError());
}
static void window_scratch(dfgpu_ctx::PeerWindow &w) {
    if (w.scratch) return;
    if (cudaMalloc(&w.scratch, 8192) != cudaSuccess || cudaMemset(w.scratch, 0, 8192) != cudaSuccess)   // [0,1024) barrier, [1024,..) handles
        fail(DFGPU_CUDA, "peer window: cannot allocate the staging buffer");
}
void peer_barrier(Ctx *ctx) {
    auto &w = ctx->win;
    window_scratch(w);
    comm_all_gather(ctx, w.scratch, (uint8_t *)w.scratch + 256, 8);
}
static void window_destroy(dfgpu_ctx::PeerWindow &w) {
    for (auto &p : w.peers) if (p.ptr && p.ptr != w.buf) cudaIpcCloseMemHandle(p.ptr);
    for (auto &p : w.peers) p = {};
    for (void *r : w.retired) cudaFree(r);
    w.retired.clear();
    if (w.buf) cudaFree(w.buf);
    if (w.scratch) cudaFree(w.scratch);
    w.buf = w.scratch = nullptr;
    w.bytes = 0;
}
void peer_window_destroy(Ctx *ctx) {
    window_destroy(ctx->win);
    window_destroy(ctx->swin);
}

// ---- send arena -------------------------------------------------------------------------------------
void *arena_base(Ctx *ctx) { return ctx->swin.buf; }
uint64_t arena_gen(Ctx *ctx) { return ctx->swin.gen; }
bool arena_alloc(Ctx *ctx, const void *owner, size_t bytes, size_t *off, bool *moved) {
    auto &w = ctx->swin;
    *moved = false;
    if (w.state < 0 || ctx->opts.world <= 1 || ctx->opts.world > 64) return false;
    if (const char *e = getenv("DFGPU_EXCHANGE")) if (strcmp(e, "pull")) return false;   // "nccl" / "push": the push model
    if (w.owner && w.owner != owner) return false;
    if (!w.owner) w.used = 0;
    const size_t at = (w.used + 255) & ~(size_t)255;
    if (at + bytes > w.bytes) {
        // grow locally; peers learn the new handle at the next exchange (the generation travels in the meta block)
        const size_t nb = ((at + bytes) + (at + bytes) / 2 + ((size_t)1 << 20)) & ~(size_t)255;
        void *nbuf = nullptr;
        if (cudaMalloc(&nbuf, nb) != cudaSuccess) { cudaGetLastError(); return false; }
        if (w.buf) {
            if (w.owner && w.used) {
                DF_CUDA(cudaMemcpyAsync(nbuf, w.buf, w.used, cudaMemcpyDeviceToDevice, ctx->stream));
                *moved = true;
            }
            DF_CUDA(cudaStreamSynchronize(ctx->stream));
            w.retired.push_back(w.buf);   // a peer may still have it mapped (not reading it: the closing barrier of the last exchange has passed)
        }
        w.buf = nbuf;
        w.bytes = nb;
        w.gen++;
    }
    w.owner = owner;
    *off = at;
    w.used = at + bytes;
    return true;
}
void arena_release(Ctx *ctx, const void *owner) {
    auto &w = ctx->swin;
    if (w.owner == owner) { w.owner = nullptr; w.used = 0; }
}
// Collective: every rank publishes the IPC handle and generation of its buffer of window `w` and maps the others'.
// `ok` = this rank is able to take part; the result (false: peer mapping is unavailable, for good) is alike everywhere.
static bool window_exchange_handles(Ctx *ctx, dfgpu_ctx::PeerWindow &w, bool ok, void **peer_ptrs) {
    const int W = ctx->opts.world, me = ctx->opts.rank;
    cudaStream_t st = ctx->stream;
    struct Msg { cudaIpcMemHandle_t h; uint64_t gen; uint64_t ok; };  // 80 bytes
    static_assert(sizeof(Msg) == 80, "handle message layout");
    window_scratch(w);
    Msg mine;
    memset(&mine, 0, sizeof mine);
    if (ok && w.buf && cudaIpcGetMemHandle(&mine.h, w.buf) != cudaSuccess) { cudaGetLastError(); ok = false; }
    mine.gen = w.gen;
    mine.ok = ok ? 1 : 0;
    uint8_t *stage = (uint8_t *)w.scratch + 1024;             // [Msg mine][Msg all[W]]
    std::vector<Msg> all((size_t)W);
    auto gather = [&]() {
        DF_CUDA(cudaMemcpyAsync(stage, &mine, sizeof mine, cudaMemcpyHostToDevice, st));
        comm_all_gather(ctx, stage, stage + sizeof(Msg), sizeof(Msg));
        DF_CUDA(cudaMemcpyAsync(all.data(), stage + sizeof(Msg), sizeof(Msg) * W, cudaMemcpyDeviceToHost, st));
        DF_CUDA(cudaStreamSynchronize(st));
    };
    gather();  // (also the pre-barrier: a peer's window is free once that peer got here)
    for (int r = 0; r < W; r++) ok = ok && all[(size_t)r].ok;
    if (ok) {
        for (int r = 0; r < W && ok; r++) {
            auto &p = w.peers[r];
            if (r == me) { p.ptr = w.buf; p.gen = w.gen; continue; }
            if (p.ptr && p.gen == all[(size_t)r].gen) continue;
            if (p.ptr) cudaIpcCloseMemHandle(p.ptr);
            p.ptr = nullptr;
            if (cudaIpcOpenMemHandle(&p.ptr, all[(size_t)r].h, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { cudaGetLastError(); p.ptr = nullptr; ok = false; }
            p.gen = all[(size_t)r].gen;
        }
    }
    // second round: did every rank manage to map every window?
    mine.ok = ok ? 1 : 0;
    gather();
    for (int r = 0; r < W; r++) ok = ok && all[(size_t)r].ok;
    for (void *r : w.retired) cudaFree(r);  // every rank is past its remapping loop: nobody maps an outgrown buffer any more
    w.retired.clear();
    if (!ok) { w.state = -1; return false; }
    w.state = 1;
    for (int r = 0; r < W; r++) peer_ptrs[r] = w.peers[r].ptr;
    return true;
}
bool peer_window_prepare(Ctx *ctx, size_t need, void **peer_ptrs) {
    auto &w = ctx->win;
    const int W = ctx->opts.world;
    if (w.state < 0 || W > 64) return false;
    if (const char *e = getenv("DFGPU_EXCHANGE")) if (!strcmp(e, "nccl")) { w.state = -1; return false; }
    // Every window is large enough and mapped already (`need` is the same number on every rank, so this is decided alike
    // everywhere).  The caller's all-gather of the partition histograms, stream-ordered after whatever still read the
    // windows, already was the barrier that frees them for the next round of stores.
    if (w.state == 1 && need <= w.bytes) {
        for (int r = 0; r < W; r++) peer_ptrs[r] = w.peers[r].ptr;
        return true;
    }
    bool ok = true;
    if (need > w.bytes) {
        // grow: everything queued on the old window (peers' pushes of the previous exchange were fenced by its barrier)
        // the old window stays allocated until every peer has dropped its mapping of it (after the second gather)
        DF_CUDA(cudaStreamSynchronize(ctx->stream));
        if (w.buf) w.retired.push_back(w.buf);
        w.buf = nullptr;
        w.bytes = 0;
        size_t nb = (need + need / 2 + ((size_t)1 << 20)) & ~(size_t)255;
        if (cudaMalloc(&w.buf, nb) == cudaSuccess) { w.bytes = nb; w.gen++; }
        else { cudaGetLastError(); ok = false; }
    }
    return window_exchange_handles(ctx, w, ok, peer_ptrs);
}
bool arena_map(Ctx *ctx, const uint64_t *gens, void **peer_ptrs) {
    auto &w = ctx->swin;
    const int W = ctx->opts.world;
    if (w.state < 0 || W > 64) return false;
    // peers[r].gen (r == me included) is the generation of rank r's arena when handles were last exchanged: the same
    // numbers on every rank, so every rank decides alike whether handles have to travel again.
    bool stale = w.state != 1;
    for (int r = 0; r < W; r++) stale = stale || w.peers[r].gen != gens[r];
    if (!stale) {
        for (int r = 0; r < W; r++) peer_ptrs[r] = w.peers[r].ptr;
        return true;
    }
    bool ok = true;
    if (!w.buf) {   // a rank that placed nothing yet still needs something to export
        if (cudaMalloc(&w.buf, (size_t)1 << 20) == cudaSuccess) { w.bytes = (size_t)1 << 20; w.gen++; }
        else { cudaGetLastError(); ok = false; }
    }
    return window_exchange_handles(ctx, w, ok, peer_ptrs);
}

// ---- exchange helpers --------------------------------------------------------------------------------
__global__ void k_bitmap_to_bytes(const uint8_t *bm, int64_t bit0, int64_t n, uint8_t *out, int all_one) {
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    for (; i < n; i += (int64_t)gridDim.x * blockDim.x) out[i] = all_one ? 1 : (bit_get(bm, bit0 + i) ? 1 : 0);
}
__global__ void k_bytes_to_bits(const uint8_t *__restrict__ bytes, int64_t n, uint32_t *__restrict__ words) {
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    int64_t nround = (n + 31) / 32 * 32;
    for (; i < nround; i += (int64_t)gridDim.x * blockDim.x) {
        uint32_t m = __ballot_sync(0xffffffffu, i < n && bytes[i] != 0);
        if ((threadIdx.x & 31) == 0) words[i >> 5] = m;
    }
}
__global__ void k_lengths_of(ColView v, int64_t n, int64_t *out) {
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    for (; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const uint8_t *p; int64_t len;
        col_bytes(v, i, p, len);
        out[i] = len;
    }
}
__global__ void k_offsets_from_starts(const int64_t *starts, int64_t n, int64_t total, void *out, int large) {
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    for (; i <= n; i += (int64_t)gridDim.x * blockDim.x) {
        int64_t v = i < n ? starts[i] : total;
        if (large) ((int64_t *)out)[i] = v; else ((int32_t *)out)[i] = (int32_t)v;
    }
}
__global__ void k_iota_p(void *dst, int width, int64_t n) {
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    for (; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        if (width == 1) ((uint8_t *)dst)[i] = (uint8_t)i;
        else if (width == 2) ((uint16_t *)dst)[i] = (uint16_t)i;
        else if (width == 4) ((uint32_t *)dst)[i] = (uint32_t)i;
        else ((uint64_t *)dst)[i] = (uint64_t)i;
    }
}
static ColRef wrap_trivial_dictionary(Ctx *ctx, const ColRef &plain, const TypeInfo &dict_type) {
    int64_t n = plain->length;
    auto wrap = std::make_shared<DevCol>();
    wrap->type = dict_type;
    wrap->length = n;
    int iw = type_width(dict_type.index_id);
    TypeInfo it; it.id = dict_type.index_id;
    double maxv = iw == 8 ? 9.2e18 : (double)((1ULL << (8 * iw - (it.is_signed_int() ? 1 : 0))) - 1);
    if ((double)n > maxv) fail(DFGPU_EXECUTION, "dictionary index type too narrow");
    BufRef ib = make_buf(ctx, (size_t)n * iw + 16);
    if (n) { k_iota_p<<<grid_for(n, 256, ctx, 16), 256, 0, ctx->stream>>>(ib->ptr, iw, n); note_launch(ctx); }
    wrap->owned.push_back(ib);
    wrap->values = ib->ptr;
    wrap->validity = plain->validity;
    wrap->vbit = plain->vbit;
    wrap->null_count = plain->null_count;
    auto d = std::make_shared<DevCol>(*plain);
    d->validity = nullptr; d->null_count = 0;
    d->type = *dict_type.value;
    wrap->dict = d;
    wrap->keepalive = plain;
    return wrap;
}

// The all-to-all.  Per column we ship: values (rows x width), validity as one byte per row,
// var-width columns as (int64 length per row, bytes).  Counts go first (ncclAllGather).
Batch *exchange_all_to_all(Ctx *ctx, Batch *const *parts, const Schema &schema) {
    int W = ctx->opts.world, me = ctx->opts.rank;
    size_t nc = schema.fields.size();
    if (W <= 1) {
        if (!parts[0]) return nullptr;
        Batch *one[1] = {parts[0]};
        return batch_concat(ctx, one, 1);
    }
    comm_check(ctx);
    ncclComm_t comm = (ncclComm_t)ctx->nccl_comm;
    cudaStream_t st = ctx->stream;
    // decode dictionaries; count rows / var bytes per destination
    std::vector<std::vector<ColRef>> send((size_t)W, std::vector<ColRef>(nc));
    size_t rec = 1 + 2 * nc;  // rows, then per column: var bytes, has_validity
    std::vector<int64_t> mine((size_t)W * rec, 0);
    for (int r = 0; r < W; r++) {
        if (!parts[r]) continue;
        if (parts[r]->cols.size() != nc) fail(DFGPU_INTERNAL, "exchange: schema mismatch");
        mine[r * rec] = parts[r]->num_rows;
        for (size_t c = 0; c < nc; c++) {
            ColRef col = parts[r]->cols[c];
            if (col->type.id == T_DICT) col = col_decode_dictionary(ctx, *col);
            send[r][c] = col;
            if (col->type.is_var()) mine[r * rec + 1 + 2 * c] = col->data_bytes;
            mine[r * rec + 2 + 2 * c] = (col->validity && col->null_count != 0) ? 1 : 0;
        }
    }
    size_t row_bytes = (size_t)W * rec * 8;
    BufRef d_mine = make_buf(ctx, row_bytes), d_all = make_buf(ctx, row_bytes * W);
    DF_CUDA(cudaMemcpyAsync(d_mine->ptr, mine.data(), row_bytes, cudaMemcpyHostToDevice, st));
    DF_NCCL(nccl().AllGather(d_mine->ptr, d_all->ptr, row_bytes, ncclChar, comm, st));
    std::vector<int64_t> all((size_t)W * W * rec);
    DF_CUDA(cudaMemcpyAsync(all.data(), d_all->ptr, row_bytes * W, cudaMemcpyDeviceToHost, st));
    DF_CUDA(cudaStreamSynchronize(st));
    auto cnt = [&](int src, int dst, size_t k) { return all[((size_t)src * W + dst) * rec + k]; };
    std::vector<int64_t> row_off((size_t)W + 1, 0);
    for (int s = 0; s < W; s++) row_off[s + 1] = row_off[s] + cnt(s, me, 0);
    int64_t total = row_off[W];
    if (total == 0) {
        // still take part in the collective with zero-sized transfers below? nothing to send/recv to me,
        // but peers may expect data from me: fall through with empty receive buffers.
    }
    std::unique_ptr<Batch> out(new Batch());
    out->ctx = ctx;
    out->num_rows = total;
    // Three phases, so that the whole batch crosses NVLink in ONE grouped NCCL call (a group per
    // column buffer costs a launch and a rendezvous each; partial-state batches have dozens):
    //   1. per column: receive buffers, send staging (bitmap -> bytes, lengths of var-width rows);
    //   2. every ncclSend / ncclRecv of every column inside one ncclGroupStart / ncclGroupEnd;
    //   3. per column: rebuild validity bitmaps / offsets from what arrived.
    std::vector<std::function<void()>> comm_ops, post_ops;
    std::vector<std::shared_ptr<DevCol>> cols_out(nc);
    for (size_t c = 0; c < nc; c++) {
        TypeInfo ft = schema.fields[c];
        TypeInfo vt = ft.value_type();
        vt.name = ft.name; vt.nullable = ft.nullable;
        bool any_valid = false;
        for (int s = 0; s < W; s++)
            for (int d = 0; d < W; d++) any_valid = any_valid || cnt(s, d, 2 + 2 * c) != 0;
        auto col = std::make_shared<DevCol>();
        cols_out[c] = col;
        col->type = vt;
        col->length = total;
        // 1) validity bytes
        if (any_valid) {
            BufRef rv = make_buf(ctx, (size_t)total + 16);
            std::vector<BufRef> sv((size_t)W);
            for (int r = 0; r < W; r++) {
                int64_t n = mine[r * rec];
                sv[r] = make_buf(ctx, (size_t)n + 16);
                if (n) {
                    const ColRef &sc = send[r][c];
                    bool allv = !(sc->validity && sc->null_count != 0);
                    k_bitmap_to_bytes<<<grid_for(n, 256, ctx, 16), 256, 0, st>>>(sc->validity, sc->vbit, n, sv[r]->as<uint8_t>(), allv ? 1 : 0);
                    note_launch(ctx);
                }
            }
            comm_ops.push_back([=, &mine, &row_off, &cnt]() {
                for (int r = 0; r < W; r++) {
                    if (mine[r * rec]) DF_NCCL(nccl().Send(sv[r]->ptr, (size_t)mine[r * rec], ncclChar, r, comm, st));
                    if (cnt(r, me, 0)) DF_NCCL(nccl().Recv(rv->as<uint8_t>() + row_off[r], (size_t)cnt(r, me, 0), ncclChar, r, comm, st));
                    if (r != me) ctx->nvlink_bytes += mine[r * rec];
                }
            });
            post_ops.push_back([=]() {
                BufRef bits = make_buf(ctx, (size_t)((total + 31) / 32) * 4 + 16, true);
                if (total) { k_bytes_to_bits<<<grid_for(total, 256, ctx, 16), 256, 0, st>>>(rv->as<uint8_t>(), total, bits->as<uint32_t>()); note_launch(ctx); }
                col->owned.push_back(bits);
                col->validity = bits->as<uint8_t>();
                col->null_count = -1;
            });
        }
        // 2) payload
        int w = vt.width();
        if (vt.id == T_BOOL) {
            BufRef rb = make_buf(ctx, (size_t)total + 16);
            std::vector<BufRef> sb((size_t)W);
            for (int r = 0; r < W; r++) {
                int64_t n = mine[r * rec];
                sb[r] = make_buf(ctx, (size_t)n + 16);
                if (n) { k_bitmap_to_bytes<<<grid_for(n, 256, ctx, 16), 256, 0, st>>>((const uint8_t *)send[r][c]->values, send[r][c]->dbit, n, sb[r]->as<uint8_t>(), 0); note_launch(ctx); }
            }
            comm_ops.push_back([=, &mine, &row_off, &cnt]() {
                for (int r = 0; r < W; r++) {
                    if (mine[r * rec]) DF_NCCL(nccl().Send(sb[r]->ptr, (size_t)mine[r * rec], ncclChar, r, comm, st));
                    if (cnt(r, me, 0)) DF_NCCL(nccl().Recv(rb->as<uint8_t>() + row_off[r], (size_t)cnt(r, me, 0), ncclChar, r, comm, st));
                    if (r != me) ctx->nvlink_bytes += mine[r * rec];
                }
            });
            post_ops.push_back([=]() {
                BufRef bits = make_buf(ctx, (size_t)((total + 31) / 32) * 4 + 16, true);
                if (total) { k_bytes_to_bits<<<grid_for(total, 256, ctx, 16), 256, 0, st>>>(rb->as<uint8_t>(), total, bits->as<uint32_t>()); note_launch(ctx); }
                col->owned.push_back(bits);
                col->values = bits->ptr;
            });
        } else if (w > 0) {
            BufRef rb = make_buf(ctx, (size_t)total * w + 16);
            const std::vector<ColRef> *sendp = send.data();
            comm_ops.push_back([=, &mine, &row_off, &cnt]() {
                for (int r = 0; r < W; r++) {
                    if (mine[r * rec]) DF_NCCL(nccl().Send(sendp[r][c]->values, (size_t)mine[r * rec] * w, ncclChar, r, comm, st));
                    if (cnt(r, me, 0)) DF_NCCL(nccl().Recv(rb->as<uint8_t>() + row_off[r] * w, (size_t)cnt(r, me, 0) * w, ncclChar, r, comm, st));
                    if (r != me) ctx->nvlink_bytes += mine[r * rec] * w;
                }
            });
            col->owned.push_back(rb);
            col->values = rb->ptr;
        } else {
            // lengths, then bytes
            std::vector<int64_t> byte_off((size_t)W + 1, 0);
            for (int s = 0; s < W; s++) byte_off[s + 1] = byte_off[s] + cnt(s, me, 1 + 2 * c);
            int64_t tbytes = byte_off[W];
            BufRef rl = make_buf(ctx, (size_t)total * 8 + 16), rd = make_buf(ctx, (size_t)tbytes + 16);
            std::vector<BufRef> sl((size_t)W);
            std::vector<int64_t> firsts((size_t)W, 0);
            for (int r = 0; r < W; r++) {
                int64_t n = mine[r * rec];
                sl[r] = make_buf(ctx, (size_t)n * 8 + 16);
                if (n) {
                    k_lengths_of<<<grid_for(n, 256, ctx, 16), 256, 0, st>>>(send[r][c]->view(), n, sl[r]->as<int64_t>());
                    note_launch(ctx);
                    int ow = send[r][c]->type.is_large() ? 8 : 4;
                    uint8_t t0[8] = {0};
                    DF_CUDA(cudaMemcpyAsync(t0, send[r][c]->offsets, ow, cudaMemcpyDeviceToHost, st));
                    DF_CUDA(cudaStreamSynchronize(st));
                    firsts[r] = ow == 8 ? *(int64_t *)t0 : *(int32_t *)t0;
                }
            }
            const std::vector<ColRef> *sendp = send.data();
            comm_ops.push_back([=, &mine, &row_off, &cnt]() {
                for (int r = 0; r < W; r++) {
                    int64_t n = mine[r * rec], nb = mine[r * rec + 1 + 2 * c];
                    if (n) DF_NCCL(nccl().Send(sl[r]->ptr, (size_t)n * 8, ncclChar, r, comm, st));
                    if (nb) DF_NCCL(nccl().Send((const uint8_t *)sendp[r][c]->values + firsts[r], (size_t)nb, ncclChar, r, comm, st));
                    if (cnt(r, me, 0)) DF_NCCL(nccl().Recv(rl->as<int64_t>() + row_off[r], (size_t)cnt(r, me, 0) * 8, ncclChar, r, comm, st));
                    if (cnt(r, me, 1 + 2 * c)) DF_NCCL(nccl().Recv(rd->as<uint8_t>() + byte_off[r], (size_t)cnt(r, me, 1 + 2 * c), ncclChar, r, comm, st));
                    if (r != me) ctx->nvlink_bytes += mine[r * rec] * 8 + mine[r * rec + 1 + 2 * c];
                }
            });
            post_ops.push_back([=]() {
                BufRef starts = make_buf(ctx, (size_t)total * 8 + 16);
                int64_t tot2 = 0;
                exclusive_scan_i64(ctx, rl->as<int64_t>(), starts->as<int64_t>(), total, &tot2);
                int ow = vt.is_large() ? 8 : 4;
                if (ow == 4 && tbytes > 0x7fffffffLL) fail(DFGPU_EXECUTION, "offset overflow in exchanged Utf8/Binary column");
                BufRef offs = make_buf(ctx, (size_t)(total + 1) * ow + 16, true);
                k_offsets_from_starts<<<grid_for(total + 1, 256, ctx, 16), 256, 0, st>>>(starts->as<int64_t>(), total, tbytes, offs->ptr, ow == 8);
                note_launch(ctx);
                col->owned.push_back(offs);
                col->owned.push_back(rd);
                col->offsets = offs->ptr;
                col->values = rd->ptr;
                col->data_bytes = tbytes;
            });
        }
        DF_CUDA(cudaGetLastError());
    }
    {
        KTimer kt(ctx, "stage:exchange_nccl");
        DF_NCCL(nccl().GroupStart());
        for (auto &op : comm_ops) op();
        DF_NCCL(nccl().GroupEnd());
    }
    for (auto &op : post_ops) op();
    DF_CUDA(cudaGetLastError());
    for (size_t c = 0; c < nc; c++) {
        ColRef result = cols_out[c];
        if (schema.fields[c].id == T_DICT) result = wrap_trivial_dictionary(ctx, cols_out[c], schema.fields[c]);
        out->cols.push_back(result);
    }
    // send staging buffers are freed stream-ordered after the collective (the lambdas own them)
    if (total == 0) return nullptr;
    return out.release();
}


// ---- allocation-free partitioning (dfgpu_partition_i64_noalloc) ---------------------------------------------
// For callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that
// forbid it): the scratch partition_perm() takes from the stream-ordered pool (partition ids, per-block histogram,
// scanned write positions) lives in ONE file-scope device array instead.  Every GPU that loads the library holds
// these ~5 MiB until exit.  All calls on a device share the array and it keeps the last call's contents: the
// ticket is cleared at the start of every call, everything else is fully rewritten before it is read.  Calls that
// overlap on the GPU (different streams, different host threads) race on it.
static const int NA_ROWS_PER_BLOCK = 4096;
static const int NA_MAX_BLOCKS = DFGPU_NOALLOC_MAX_ROWS / NA_ROWS_PER_BLOCK;
struct NaScratch {
    unsigned int ticket;  // blocks of k_na_count that have published their histogram column; must start at 0
    unsigned int pad[63];
    uint32_t pid[DFGPU_NOALLOC_MAX_ROWS];                              // partition of every row
    int64_t hist[(size_t)DFGPU_NOALLOC_MAX_PARTS * NA_MAX_BLOCKS];     // [P][nblocks] rows per partition and block
    int64_t starts[(size_t)DFGPU_NOALLOC_MAX_PARTS * NA_MAX_BLOCKS];   // exclusive scan of hist
};
static __device__ __align__(256) NaScratch g_na_scratch;

// pass 1 of partition_perm() (k_part_count) with the hash computed in place, plus the scan: the block that
// publishes the last histogram column scans all of them (at most 65 536 counters) and writes the partition offsets.
__global__ void __launch_bounds__(PART_T) k_na_count(const int64_t *__restrict__ keys, const uint8_t *__restrict__ validity, int64_t n,
                                                     int64_t rows_per_block, int P, NaScratch *__restrict__ sc,
                                                     int64_t *__restrict__ offsets_out) {
    __shared__ uint32_t cnt[DFGPU_NOALLOC_MAX_PARTS];
    __shared__ int64_t tsum[PART_T];
    __shared__ bool last;
    for (int p = threadIdx.x; p < P; p += PART_T) cnt[p] = 0;
    __syncthreads();
    const int64_t lo = (int64_t)blockIdx.x * rows_per_block, hi = min(n, lo + rows_per_block);
    for (int64_t i = lo + threadIdx.x; i < hi; i += PART_T) {
        // create_hashes over one Int64 column, seeds (0,0,0,0): a NULL row keeps the initial hash 0 (hash_utils.rs:117-131)
        const uint64_t h = (validity && !bit_get(validity, i)) ? 0 : h_u64((uint64_t)keys[i], 0);
        const uint32_t p = (uint32_t)(h % (uint64_t)P);  // repartition/mod.rs:184-187
        sc->pid[i] = p;
        atomicAdd(&cnt[p], 1u);
    }
    __syncthreads();
    for (int p = threadIdx.x; p < P; p += PART_T) sc->hist[(int64_t)p * gridDim.x + blockIdx.x] = cnt[p];
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) last = atomicAdd(&sc->ticket, 1u) == gridDim.x - 1;
    __syncthreads();
    if (!last) return;
    __threadfence();
    const int64_t hn = (int64_t)P * gridDim.x, per = (hn + PART_T - 1) / PART_T;
    const int64_t a = min(hn, (int64_t)threadIdx.x * per), b = min(hn, a + per);
    int64_t s = 0;
    for (int64_t k = a; k < b; k++) s += __ldcg(&sc->hist[k]);
    tsum[threadIdx.x] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
        int64_t run = 0;
        for (int t = 0; t < PART_T; t++) { const int64_t v = tsum[t]; tsum[t] = run; run += v; }
    }
    __syncthreads();
    int64_t run = tsum[threadIdx.x];
    for (int64_t k = a; k < b; k++) {
        sc->starts[k] = run;
        if (k % gridDim.x == 0) offsets_out[k / gridDim.x] = run;  // partition p starts at starts[p][0]
        run += __ldcg(&sc->hist[k]);
    }
    if (threadIdx.x == 0) offsets_out[P] = n;
}

// the scratch array's device address, resolved once per device (the symbol's name is only a host-side stand-in)
static NaScratch *na_scratch_ptr() {
    static std::mutex mu;
    static NaScratch *ptrs[64] = {};
    int dev = 0;
    DF_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) fail(DFGPU_NOT_IMPLEMENTED, "dfgpu_partition_i64_noalloc: device ordinal above 63");
    std::lock_guard<std::mutex> lock(mu);
    if (!ptrs[dev]) {
        NaScratch *p = nullptr;
        cudaError_t e = cudaGetSymbolAddress((void **)&p, g_na_scratch);
        if (e != cudaSuccess || !p)
            fail(DFGPU_CUDA, std::string("dfgpu_partition_i64_noalloc: cudaGetSymbolAddress failed: ") + cudaGetErrorString(e));
        ptrs[dev] = p;
    }
    return ptrs[dev];
}

static void partition_i64_noalloc(cudaStream_t st, const int64_t *keys, const uint8_t *validity, int64_t n, int P, int64_t *perm_out,
                                  int64_t *offsets_out) {
    if (P <= 0 || P > DFGPU_NOALLOC_MAX_PARTS || n < 0 || n > DFGPU_NOALLOC_MAX_ROWS)
        fail(DFGPU_NOT_IMPLEMENTED, "dfgpu_partition_i64_noalloc: " + std::to_string(n) + " rows into " + std::to_string(P) +
                                        " partitions exceeds the static scratch (at most " + std::to_string(DFGPU_NOALLOC_MAX_ROWS) +
                                        " rows and " + std::to_string(DFGPU_NOALLOC_MAX_PARTS) + " partitions per call); use dfgpu_partition_hash");
    if (!offsets_out || (n > 0 && (!keys || !perm_out))) fail(DFGPU_INTERNAL, "dfgpu_partition_i64_noalloc: null pointer");
    if (n == 0) {
        DF_CUDA(cudaMemsetAsync(offsets_out, 0, (size_t)(P + 1) * 8, st));
        return;
    }
    NaScratch *sc = na_scratch_ptr();
    DF_CUDA(cudaMemsetAsync(&sc->ticket, 0, sizeof(unsigned int), st));  // the one field that must start at zero
    const int nblocks = (int)((n + NA_ROWS_PER_BLOCK - 1) / NA_ROWS_PER_BLOCK);
    k_na_count<<<nblocks, PART_T, 0, st>>>(keys, validity, n, NA_ROWS_PER_BLOCK, P, sc, offsets_out);
    const size_t smem = (size_t)P * 8 + (size_t)(PART_T / 32) * P * 4;  // <= 10 KiB at P = 256
    k_part_scatter<<<nblocks, PART_T, smem, st>>>(sc->pid, n, NA_ROWS_PER_BLOCK, P, sc->starts, perm_out);
    DF_CUDA(cudaGetLastError());
}

}  // namespace dfgpu

using namespace dfgpu;
extern "C" {
dfgpu_status dfgpu_nccl_get_unique_id(void *out128) {
    try {
        ncclUniqueId id;
        DF_NCCL(nccl().GetUniqueId(&id));
        memcpy(out128, &id, sizeof id);
    } catch (const dfgpu::Error &e) { dfgpu::set_last_error(e.msg); return e.code; }
    return DFGPU_OK;
}
dfgpu_status dfgpu_partition_hash(dfgpu_ctx *ctx, const dfgpu_batch *batch, const int32_t *key_cols, int32_t n_keys, int32_t n_parts,
                                  dfgpu_batch **out_parts) {
    try {
        DF_CUDA(cudaSetDevice(ctx->device));
        partition_hash(ctx, *batch, key_cols, n_keys, n_parts, out_parts);
    } catch (const dfgpu::Error &e) { dfgpu::set_last_error(e.msg); return e.code; }
    return DFGPU_OK;
}
dfgpu_status dfgpu_exchange_all_to_all(dfgpu_ctx *ctx, dfgpu_batch *const *parts, const ArrowSchema *schema, dfgpu_batch **out) {
    try {
        DF_CUDA(cudaSetDevice(ctx->device));
        Schema s = schema_from_arrow(schema);
        *out = exchange_all_to_all(ctx, parts, s);
    } catch (const dfgpu::Error &e) { dfgpu::set_last_error(e.msg); return e.code; }
    return DFGPU_OK;
}
dfgpu_status dfgpu_exchange_all_gather(dfgpu_ctx *ctx, const dfgpu_batch *local, const ArrowSchema *schema, dfgpu_batch **out) {
    try {
        DF_CUDA(cudaSetDevice(ctx->device));
        Schema s = schema_from_arrow(schema);
        int W = ctx->opts.world;
        // all-gather == all-to-all where every destination receives the same local batch
        std::vector<dfgpu_batch *> parts((size_t)std::max(1, W), const_cast<dfgpu_batch *>(local));
        *out = exchange_all_to_all(ctx, parts.data(), s);
    } catch (const dfgpu::Error &e) { dfgpu::set_last_error(e.msg); return e.code; }
    return DFGPU_OK;
}
dfgpu_status dfgpu_partition_i64_noalloc(void *stream, const int64_t *dev_keys, const uint8_t *dev_validity, int64_t n, int32_t n_parts,
                                         int64_t *dev_perm_out, int64_t *dev_offsets_out) {
    try {
        partition_i64_noalloc((cudaStream_t)stream, dev_keys, dev_validity, n, n_parts, dev_perm_out, dev_offsets_out);
    } catch (const dfgpu::Error &e) { dfgpu::set_last_error(e.msg); return e.code; }
    return DFGPU_OK;
}
}
