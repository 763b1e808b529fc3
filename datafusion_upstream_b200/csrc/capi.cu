// capi.cu — the extern "C" boundary (include/dfgpu.h): context, batches, create_hashes and the
// aggregation operator.  Join and repartition entry points live in join.cu / partition.cu.
#include <nccl.h>
#include <nvtx3/nvToolsExt.h>

#include "device.cuh"
#include "groups.h"

struct dfgpu_agg;
namespace dfgpu {
dfgpu_agg *agg_create(Ctx *ctx, int mode, const Schema &in, const int32_t *group_cols, int n_group, const dfgpu_agg_desc *descs, int n_aggs);
void agg_push(dfgpu_agg &a, const Batch &batch);
void agg_push_host(dfgpu_agg &a, const ArrowArray *batch);
void agg_push_host_many(dfgpu_agg &a, const ArrowArray *const *batches, int n);
void agg_flush_pending(dfgpu_agg &a);
Batch *agg_emit(dfgpu_agg &a);
Batch *agg_emit_first(dfgpu_agg &a, int64_t n);
const Schema &agg_out_schema(const dfgpu_agg &a);
const Schema &agg_in_schema(const dfgpu_agg &a);
Ctx *agg_ctx(const dfgpu_agg &a);
void agg_metrics(const dfgpu_agg &a, dfgpu_metrics *out);
int64_t agg_num_groups(dfgpu_agg &a);
void agg_exchange(dfgpu_agg &a);
void agg_exchange_virtual(dfgpu_agg *const *aggs, int n);
void agg_destroy(dfgpu_agg *a);
void ctx_init_nccl(Ctx *ctx);
void ctx_destroy_nccl(Ctx *ctx);

// create_hashes (common/src/hash_utils.rs:357-416): column 0 assigns, later columns combine;
// NULL rows leave the running hash untouched; dictionaries hash their value.
struct HashCols {
    ColView c[MAX_KEY_COLS];
    ColView d[MAX_KEY_COLS];  // dictionary values for T_DICT columns
    int32_t n;
};
__global__ void k_create_hashes(HashCols hc, uint64_t seed, int force_collisions, int64_t n, uint64_t *__restrict__ out) {
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    for (; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        uint64_t h = 0;
        if (!force_collisions) {
            for (int c = 0; c < hc.n; c++) {
                const ColView &v = hc.c[c];
                if (!col_valid(v, i)) continue;
                uint64_t hv;
                if (v.type == T_DICT) {
                    int64_t k = dict_index(v, i);
                    hv = col_valid(hc.d[c], k) ? col_hash_value(hc.d[c], k, seed) : 0;  // hash_utils.rs:188-208
                } else {
                    hv = col_hash_value(v, i, seed);
                }
                h = c == 0 ? hv : combine_hashes(hv, h);
            }
        }
        out[i] = h;
    }
}
NvtxRange::NvtxRange(const char *name) { nvtxRangePushA(name); }
NvtxRange::~NvtxRange() { nvtxRangePop(); }
void ctx_release(Ctx *c) {
    if (c->refs.fetch_sub(1) != 1) return;
    cudaSetDevice(c->device);
    cudaStreamSynchronize(c->stream);
    c->bufs.trim(c->stream);
    cudaStreamSynchronize(c->stream);
    if (c->guard) {  // no caller left to hand a status to: say it and stop the process, this is a debugging mode
        unsigned long long w[2] = {0, 0};
        cudaMemcpy(w, c->guard_words, sizeof w, cudaMemcpyDeviceToHost);
        cudaFree(c->guard_words);
        if (w[0]) {
            fprintf(stderr, "DFGPU_GUARD: %llu buffer(s) were written out of bounds; the first had %llu bytes\n", w[0], w[1]);
            abort();
        }
    }
    ctx_destroy_nccl(c);
    if (c->import_event) cudaEventDestroy(c->import_event);
    if (c->copy_stream) cudaStreamDestroy(c->copy_stream);
    cudaStreamDestroy(c->stream);
    delete c;
}
void create_hashes_device(Ctx *ctx, const std::vector<const DevCol *> &cols, int64_t row_off, int64_t n, uint64_t seed, uint64_t *out) {
    if ((int)cols.size() > MAX_KEY_COLS) fail(DFGPU_NOT_IMPLEMENTED, "more than 16 hash columns");
    HashCols hc;
    memset(&hc, 0, sizeof hc);
    hc.n = (int)cols.size();
    for (size_t c = 0; c < cols.size(); c++) {
        hc.c[c] = view_at(cols[c]->view(), row_off);
        if (cols[c]->type.id == T_DICT) hc.d[c] = cols[c]->dict->view();
    }
    if (n == 0) return;
    k_create_hashes<<<grid_for(n, 256, ctx, 32), 256, 0, ctx->stream>>>(hc, seed, ctx->opts.force_hash_collisions, n, out);
    note_launch(ctx);
    DF_CUDA(cudaGetLastError());
}
}  // namespace dfgpu

using namespace dfgpu;

#define DF_TRY try { dfgpu::NvtxRange _nvtx(__func__);
#define DF_CATCH                                                   \
    }                                                              \
    catch (const dfgpu::Error &e) {                                \
        dfgpu::set_last_error(e.msg);                              \
        return e.code;                                             \
    }                                                              \
    catch (const std::exception &e) {                              \
        dfgpu::set_last_error(std::string("Internal error: ") + e.what()); \
        return DFGPU_INTERNAL;                                     \
    }                                                              \
    return DFGPU_OK;

extern "C" {

const char *dfgpu_last_error(void) { return dfgpu::last_error().c_str(); }
const char *dfgpu_version(void) { return "dfgpu 0.2 (sm_100a)"; }
int32_t dfgpu_hash_contract(void) { return 1; }   // "dfhash v1" (device.cuh): mix64 = splitmix64 finaliser, combine_hashes as hash_utils.rs:38-41

void dfgpu_options_default(dfgpu_options *o) {
    memset(o, 0, sizeof *o);
    o->device = 0;
    o->batch_size = 8192;
    o->memory_limit = 0;
    o->load_factor = 0.5;
    o->force_hash_collisions = 0;
    o->rank = 0;
    o->world = 1;
    o->nccl_unique_id = nullptr;
    o->chunk_rows = 0;
}

dfgpu_status dfgpu_ctx_create(const dfgpu_options *opts, dfgpu_ctx **out) {
    DF_TRY
    if (!out) fail(DFGPU_INTERNAL, "null out pointer");
    dfgpu_options o;
    if (opts) o = *opts; else dfgpu_options_default(&o);
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0) {
        cudaGetLastError();
        // no CPU fallback: the product path fails loudly without a device
        fail(DFGPU_CUDA, std::string("no CUDA device available (") + (e != cudaSuccess ? cudaGetErrorString(e) : "device count 0") +
                             "); libdfgpu has no CPU fallback");
    }
    if (o.device < 0 || o.device >= count) fail(DFGPU_CUDA, "CUDA device ordinal out of range");
    DF_CUDA(cudaSetDevice(o.device));
    std::unique_ptr<dfgpu_ctx> c(new dfgpu_ctx());
    c->opts = o;
    c->device = o.device;
    DF_CUDA(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
    cudaDeviceProp prop;
    DF_CUDA(cudaGetDeviceProperties(&prop, o.device));
    c->sm_count = prop.multiProcessorCount;
    c->smem_optin = prop.sharedMemPerBlockOptin;
    // keep freed blocks in the stream-ordered pool (no trimming at synchronisation points)
    cudaMemPool_t pool;
    DF_CUDA(cudaDeviceGetDefaultMemPool(&pool, o.device));
    uint64_t thresh = UINT64_MAX;
    DF_CUDA(cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &thresh));
    {
        size_t free_b = 0, total_b = 0;
        if (cudaMemGetInfo(&free_b, &total_b) == cudaSuccess) c->bufs.limit = total_b / 3;
    }
    if (const char *g = getenv("DFGPU_GUARD")) c->guard = atoi(g);
    if (c->guard) {
        DF_CUDA(cudaMalloc((void **)&c->guard_words, 16));
        DF_CUDA(cudaMemset(c->guard_words, 0, 16));
    }
    if (o.world > 1) ctx_init_nccl(c.get());
    *out = c.release();
    DF_CATCH
}
void dfgpu_ctx_destroy(dfgpu_ctx *ctx) {
    if (ctx) dfgpu::ctx_release(ctx);  // torn down once the last batch / operator on it is gone
}
dfgpu_status dfgpu_ctx_synchronize(dfgpu_ctx *ctx) {
    DF_TRY
    DF_CUDA(cudaStreamSynchronize(ctx->stream));
    guard_poll(ctx);
    comm_poll(ctx);
    DF_CATCH
}
dfgpu_status dfgpu_ctx_set_profiling(dfgpu_ctx *ctx, int32_t enable) {
    DF_TRY
    DF_CUDA(cudaStreamSynchronize(ctx->stream));
    for (auto &r : ctx->prof) { cudaEventDestroy(r.start); cudaEventDestroy(r.stop); }
    ctx->prof.clear();
    ctx->profiling = enable != 0;
    DF_CATCH
}
dfgpu_status dfgpu_ctx_profile_json(dfgpu_ctx *ctx, char *buf, int64_t cap) {
    DF_TRY
    DF_CUDA(cudaStreamSynchronize(ctx->stream));
    struct Acc { int64_t launches = 0, bytes = 0; double ms = 0; };
    std::map<std::string, Acc> by;
    for (auto &r : ctx->prof) {
        float ms = 0;
        DF_CUDA(cudaEventElapsedTime(&ms, r.start, r.stop));
        Acc &a = by[r.name];
        a.launches++; a.ms += ms; a.bytes += r.algo_bytes;
        cudaEventDestroy(r.start); cudaEventDestroy(r.stop);
    }
    ctx->prof.clear();
    std::string s = "[";
    for (auto &kv : by) {
        if (s.size() > 1) s += ",";
        char tmp[256];
        snprintf(tmp, sizeof tmp, "{\"name\":\"%s\",\"launches\":%lld,\"ms\":%.6f,\"algo_bytes\":%lld}", kv.first.c_str(),
                 (long long)kv.second.launches, kv.second.ms, (long long)kv.second.bytes);
        s += tmp;
    }
    s += "]";
    if ((int64_t)s.size() + 1 > cap) fail(DFGPU_INTERNAL, "profile buffer too small");
    memcpy(buf, s.c_str(), s.size() + 1);
    DF_CATCH
}
void *dfgpu_ctx_stream(dfgpu_ctx *ctx) { return (void *)ctx->stream; }
int64_t dfgpu_ctx_kernel_launches(const dfgpu_ctx *ctx) { return ctx->launches; }
dfgpu_status dfgpu_host_alloc(dfgpu_ctx *ctx, int64_t bytes, void **out) {
    DF_TRY
    (void)ctx;
    *out = pinned_alloc((size_t)bytes);
    DF_CATCH
}
void dfgpu_host_free(dfgpu_ctx *ctx, void *p) {
    (void)ctx;
    pinned_free(p);
}

dfgpu_status dfgpu_batch_import(dfgpu_ctx *ctx, const ArrowArray *array, const ArrowSchema *schema, int32_t on_device, dfgpu_batch **out) {
    DF_TRY
    DF_CUDA(cudaSetDevice(ctx->device));
    Schema s = schema_from_arrow(schema);
    *out = batch_import(ctx, array, s, on_device != 0);
    DF_CATCH
}
dfgpu_status dfgpu_batch_export(dfgpu_ctx *ctx, const dfgpu_batch *batch, ArrowArray *out_array, ArrowSchema *out_schema) {
    DF_TRY
    DF_CUDA(cudaSetDevice(ctx->device));
    batch_export(ctx, batch, out_array, out_schema);
    guard_poll(ctx);
    DF_CATCH
}
int64_t dfgpu_batch_num_rows(const dfgpu_batch *b) { return b ? b->num_rows : 0; }
int32_t dfgpu_batch_num_columns(const dfgpu_batch *b) { return b ? (int32_t)b->cols.size() : 0; }
dfgpu_status dfgpu_batch_schema(const dfgpu_batch *b, ArrowSchema *out) {
    DF_TRY
    schema_to_arrow(b->schema(), out);
    DF_CATCH
}
dfgpu_status dfgpu_batch_concat(dfgpu_ctx *ctx, dfgpu_batch *const *batches, int32_t n, dfgpu_batch **out) {
    DF_TRY
    DF_CUDA(cudaSetDevice(ctx->device));
    *out = batch_concat(ctx, batches, n);
    DF_CATCH
}
dfgpu_status dfgpu_batch_slice(dfgpu_ctx *ctx, const dfgpu_batch *b, int64_t off, int64_t len, dfgpu_batch **out) {
    DF_TRY
    *out = batch_slice(ctx, b, off, len);
    DF_CATCH
}
dfgpu_status dfgpu_batch_project(dfgpu_ctx *ctx, const dfgpu_batch *b, const int32_t *cols, int32_t n_cols, dfgpu_batch **out) {
    DF_TRY
    std::unique_ptr<Batch> o(new Batch());
    o->ctx = ctx;
    o->num_rows = b->num_rows;
    for (int i = 0; i < n_cols; i++) {
        if (cols[i] < 0 || cols[i] >= (int)b->cols.size()) fail(DFGPU_PLAN, "project: column index out of range");
        o->cols.push_back(b->cols[cols[i]]);
    }
    *out = o.release();
    DF_CATCH
}
struct dfgpu_coalescer {
    dfgpu::CtxHold ctx;
    dfgpu::Schema schema;
    int64_t target = 0;
    std::unique_ptr<dfgpu::HostCoalescer> co;
};
dfgpu_status dfgpu_coalescer_create(dfgpu_ctx *ctx, const ArrowSchema *schema, int64_t target_rows, dfgpu_coalescer **out) {
    DF_TRY
    if (!ctx || !schema || !out) fail(DFGPU_INTERNAL, "null argument");
    if (target_rows <= 0) fail(DFGPU_PLAN, "CoalesceBatchesExec: target_batch_size must be positive");
    DF_CUDA(cudaSetDevice(ctx->device));
    std::unique_ptr<dfgpu_coalescer> c(new dfgpu_coalescer());
    c->ctx = ctx;
    c->schema = schema_from_arrow(schema);
    c->target = target_rows;
    if (!HostCoalescer::supports(c->schema))
        fail(DFGPU_NOT_IMPLEMENTED, "This feature is not implemented: host batch coalescing of dictionary / nested columns");
    c->co.reset(new HostCoalescer(ctx, c->schema, target_rows));
    *out = c.release();
    DF_CATCH
}
dfgpu_status dfgpu_coalescer_push(dfgpu_coalescer *c, const ArrowArray *batch, dfgpu_batch **out) {
    DF_TRY
    if (!c || !batch || !out) fail(DFGPU_INTERNAL, "null argument");
    *out = nullptr;
    Ctx *ctx = c->ctx;
    DF_CUDA(cudaSetDevice(ctx->device));
    if (batch->length >= c->target && c->co->rows() == 0) {  // coalesce_batches.rs:215-218
        *out = batch_import(ctx, batch, c->schema, false);
    } else if (batch->length > 0) {                          // :219-240
        if (!c->co->fits(batch)) *out = c->co->flush();      // i32 Utf8 offsets would overflow: hand over early
        c->co->append(batch);
        if (!*out && c->co->rows() >= c->target) *out = c->co->flush();
    }
    DF_CATCH
}
dfgpu_status dfgpu_coalescer_finish(dfgpu_coalescer *c, dfgpu_batch **out) {
    DF_TRY
    if (!c || !out) fail(DFGPU_INTERNAL, "null argument");
    DF_CUDA(cudaSetDevice(c->ctx->device));
    *out = c->co->flush();                                   // :242-262 (nullptr when the buffer is empty)
    DF_CATCH
}
void dfgpu_coalescer_destroy(dfgpu_coalescer *c) {
    if (!c) return;
    cudaSetDevice(c->ctx->device);
    delete c;
}
dfgpu_status dfgpu_batch_eval_decimal(dfgpu_ctx *ctx, const dfgpu_batch *b, const dfgpu_expr_node *nodes, int32_t n_nodes, const char *name,
                                      dfgpu_batch **out) {
    DF_TRY
    if (!ctx || !b || !out) fail(DFGPU_INTERNAL, "null argument");
    DF_CUDA(cudaSetDevice(ctx->device));
    *out = batch_eval_decimal(ctx, b, nodes, n_nodes, name);
    DF_CATCH
}
dfgpu_status dfgpu_batch_filter(dfgpu_ctx *ctx, const dfgpu_batch *b, int32_t predicate_col, dfgpu_batch **out) {
    DF_TRY
    if (!b || !out) fail(DFGPU_INTERNAL, "null argument");
    DF_CUDA(cudaSetDevice(ctx->device));
    *out = batch_filter(ctx, b, predicate_col);
    DF_CATCH
}
void dfgpu_batch_free(dfgpu_batch *b) {
    if (!b) return;
    cudaSetDevice(b->ctx->device);
    delete b;
}

dfgpu_status dfgpu_create_hashes(dfgpu_ctx *ctx, const dfgpu_batch *batch, const int32_t *cols, int32_t n_cols, uint64_t seed,
                                 uint64_t *host_out) {
    DF_TRY
    DF_CUDA(cudaSetDevice(ctx->device));
    std::vector<const DevCol *> cs;
    for (int i = 0; i < n_cols; i++) {
        if (cols[i] < 0 || cols[i] >= (int)batch->cols.size()) fail(DFGPU_PLAN, "hash column index out of range");
        cs.push_back(batch->cols[cols[i]].get());
    }
    int64_t n = batch->num_rows;
    DevBuf out(ctx, (size_t)n * 8 + 8);
    create_hashes_device(ctx, cs, 0, n, seed, out.as<uint64_t>());
    if (n) DF_CUDA(cudaMemcpyAsync(host_out, out.ptr, (size_t)n * 8, cudaMemcpyDeviceToHost, ctx->stream));
    DF_CUDA(cudaStreamSynchronize(ctx->stream));
    DF_CATCH
}

// ---- aggregation -----------------------------------------------------------------------------------
dfgpu_status dfgpu_agg_create(dfgpu_ctx *ctx, int32_t mode, const ArrowSchema *input_schema, const int32_t *group_cols,
                              int32_t n_group_cols, const dfgpu_agg_desc *aggs, int32_t n_aggs, dfgpu_agg **out) {
    DF_TRY
    DF_CUDA(cudaSetDevice(ctx->device));
    Schema s = schema_from_arrow(input_schema);
    *out = agg_create(ctx, mode, s, group_cols, n_group_cols, aggs, n_aggs);
    DF_CATCH
}
dfgpu_status dfgpu_agg_output_schema(const dfgpu_agg *agg, ArrowSchema *out) {
    DF_TRY
    schema_to_arrow(agg_out_schema(*agg), out);
    DF_CATCH
}
dfgpu_status dfgpu_agg_push(dfgpu_agg *agg, const dfgpu_batch *batch) {
    DF_TRY
    DF_CUDA(cudaSetDevice(agg_ctx(*agg)->device));
    agg_flush_pending(*agg);
    agg_push(*agg, *batch);
    DF_CATCH
}
dfgpu_status dfgpu_agg_push_arrow(dfgpu_agg *agg, const ArrowArray *batch) {
    DF_TRY
    Ctx *ctx = agg_ctx(*agg);
    DF_CUDA(cudaSetDevice(ctx->device));
    agg_push_host(*agg, batch);
    DF_CATCH
}
dfgpu_status dfgpu_agg_push_arrow_batches(dfgpu_agg *agg, const ArrowArray *const *batches, int32_t n) {
    DF_TRY
    if (n < 0 || (n > 0 && !batches)) fail(DFGPU_INTERNAL, "push_arrow_batches: null batch array");
    Ctx *ctx = agg_ctx(*agg);
    DF_CUDA(cudaSetDevice(ctx->device));
    agg_push_host_many(*agg, batches, n);
    DF_CATCH
}
dfgpu_status dfgpu_agg_num_groups(dfgpu_agg *agg, int64_t *out) {
    DF_TRY
    DF_CUDA(cudaSetDevice(agg_ctx(*agg)->device));
    *out = agg_num_groups(*agg);
    DF_CATCH
}
dfgpu_status dfgpu_agg_exchange(dfgpu_agg *agg) {
    DF_TRY
    DF_CUDA(cudaSetDevice(agg_ctx(*agg)->device));
    agg_exchange(*agg);
    DF_CATCH
}
dfgpu_status dfgpu_agg_exchange_virtual(dfgpu_agg *const *aggs, int32_t n) {
    DF_TRY
    if (n < 1 || !aggs) fail(DFGPU_PLAN, "virtual record exchange: no operators");
    DF_CUDA(cudaSetDevice(agg_ctx(*aggs[0])->device));
    agg_exchange_virtual(aggs, n);
    DF_CATCH
}
dfgpu_status dfgpu_agg_emit(dfgpu_agg *agg, dfgpu_batch **out) {
    DF_TRY
    DF_CUDA(cudaSetDevice(agg_ctx(*agg)->device));
    *out = agg_emit(*agg);
    DF_CATCH
}
dfgpu_status dfgpu_agg_emit_first(dfgpu_agg *agg, int64_t n, dfgpu_batch **out) {
    DF_TRY
    DF_CUDA(cudaSetDevice(agg_ctx(*agg)->device));
    *out = agg_emit_first(*agg, n);
    DF_CATCH
}
dfgpu_status dfgpu_agg_emit_arrow(dfgpu_agg *agg, ArrowArray *out_array, ArrowSchema *out_schema) {
    DF_TRY
    Ctx *ctx = agg_ctx(*agg);
    DF_CUDA(cudaSetDevice(ctx->device));
    std::unique_ptr<Batch> b(agg_emit(*agg));
    batch_export(ctx, b.get(), out_array, out_schema);
    DF_CATCH
}
dfgpu_status dfgpu_agg_metrics(const dfgpu_agg *agg, dfgpu_metrics *out) {
    DF_TRY
    agg_metrics(*agg, out);
    DF_CATCH
}
void dfgpu_agg_destroy(dfgpu_agg *agg) {
    if (!agg) return;
    cudaSetDevice(agg_ctx(*agg)->device);
    agg_destroy(agg);
}

}  // extern "C"
