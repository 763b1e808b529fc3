// batch.cu — context, memory, Arrow C Data import/export and the columnar utility kernels
// (take / concat / slice / bitmap realign / prefix sum) that stand in for arrow-rs
// `compute::{take, concat_batches}` on the device (joins/utils.rs:1180-1230, hash_join.rs:764).
#include <algorithm>
#include <array>

#include "common.h"
#include "device.cuh"

namespace dfgpu {

// ------------------------------------------------------------------------------------------------
// errors / types
// ------------------------------------------------------------------------------------------------
static thread_local std::string g_last_error;
void set_last_error(const std::string &m) { g_last_error = m; }
const std::string &last_error() { return g_last_error; }

int type_width(int32_t id) {
    switch (id) {
    case T_I8: case T_U8: return 1;
    case T_I16: case T_U16: return 2;
    case T_I32: case T_U32: case T_F32: case T_DATE32: case T_TIME32: return 4;
    case T_I64: case T_U64: case T_F64: case T_DATE64: case T_TS: case T_TIME64: return 8;
    case T_DEC128: return 16;
    default: return 0;
    }
}
int TypeInfo::width() const { return type_width(id); }
std::string TypeInfo::describe() const {
    if (id == T_DICT) return "Dictionary(" + std::to_string(index_id) + "," + value->describe() + ")";
    return format.empty() ? format_of(*this) : format;
}

std::string format_of(const TypeInfo &t) {
    if (!t.format.empty()) return t.format;
    switch (t.id) {
    case T_BOOL: return "b";
    case T_I8: return "c";
    case T_U8: return "C";
    case T_I16: return "s";
    case T_U16: return "S";
    case T_I32: return "i";
    case T_U32: return "I";
    case T_I64: return "l";
    case T_U64: return "L";
    case T_F32: return "f";
    case T_F64: return "g";
    case T_DATE32: return "tdD";
    case T_DATE64: return "tdm";
    case T_TS: return "tsu:";
    case T_TIME32: return "ttm";
    case T_TIME64: return "ttu";
    case T_DEC128: return "d:" + std::to_string(t.precision) + "," + std::to_string(t.scale);
    case T_UTF8: return "u";
    case T_LARGE_UTF8: return "U";
    case T_BINARY: return "z";
    case T_LARGE_BINARY: return "Z";
    }
    fail(DFGPU_INTERNAL, "format_of: unknown type");
}

static int32_t id_from_format(const std::string &f, int &prec, int &scale) {
    if (f == "b") return T_BOOL;
    if (f == "c") return T_I8;
    if (f == "C") return T_U8;
    if (f == "s") return T_I16;
    if (f == "S") return T_U16;
    if (f == "i") return T_I32;
    if (f == "I") return T_U32;
    if (f == "l") return T_I64;
    if (f == "L") return T_U64;
    if (f == "f") return T_F32;
    if (f == "g") return T_F64;
    if (f == "u") return T_UTF8;
    if (f == "U") return T_LARGE_UTF8;
    if (f == "z") return T_BINARY;
    if (f == "Z") return T_LARGE_BINARY;
    if (f == "tdD") return T_DATE32;
    if (f == "tdm") return T_DATE64;
    if (f == "tts" || f == "ttm") return T_TIME32;
    if (f == "ttu" || f == "ttn") return T_TIME64;
    if (f.rfind("ts", 0) == 0 || f.rfind("tD", 0) == 0) return T_TS;  // timestamps / durations: int64
    if (f.rfind("d:", 0) == 0) {
        int bits = 128;
        int n = sscanf(f.c_str(), "d:%d,%d,%d", &prec, &scale, &bits);
        if (n < 2) fail(DFGPU_NOT_IMPLEMENTED, "bad decimal format " + f);
        if (bits != 128) fail(DFGPU_NOT_IMPLEMENTED, "This feature is not implemented: Decimal" + std::to_string(bits));
        return T_DEC128;
    }
    fail(DFGPU_NOT_IMPLEMENTED, "This feature is not implemented: Arrow type with format '" + f + "'");
}

TypeInfo type_from_schema(const ArrowSchema *s) {
    TypeInfo t;
    t.name = s->name ? s->name : "";
    t.nullable = (s->flags & ARROW_FLAG_NULLABLE) != 0;
    std::string f = s->format;
    int p = 0, sc = 0;
    int32_t id = id_from_format(f, p, sc);
    if (s->dictionary) {
        t.id = T_DICT;
        t.index_id = id;
        t.dict_ordered = (s->flags & ARROW_FLAG_DICTIONARY_ORDERED) != 0;
        t.value = std::make_shared<TypeInfo>(type_from_schema(s->dictionary));
        t.format = f;  // index format
        return t;
    }
    t.id = id;
    t.precision = p;
    t.scale = sc;
    t.format = f;
    return t;
}
TypeInfo make_type(int32_t id, const std::string &name, bool nullable, int precision, int scale) {
    TypeInfo t;
    t.id = id; t.name = name; t.nullable = nullable; t.precision = precision; t.scale = scale;
    t.format = format_of(t);
    return t;
}
Schema schema_from_arrow(const ArrowSchema *s) {
    if (!s || !s->format || std::string(s->format) != "+s")
        fail(DFGPU_PLAN, "expected a struct schema ('+s') describing a RecordBatch");
    Schema out;
    for (int64_t i = 0; i < s->n_children; i++) out.fields.push_back(type_from_schema(s->children[i]));
    return out;
}

// ---- schema export -----------------------------------------------------------------------------
struct SchemaPriv {
    std::string format, name;
    std::vector<ArrowSchema> child_store;
    std::vector<ArrowSchema *> child_ptrs;
    ArrowSchema dict;
    bool has_dict = false;
};
static void release_schema(ArrowSchema *s) {
    if (!s || !s->release) return;
    auto *p = (SchemaPriv *)s->private_data;
    for (auto &c : p->child_store)
        if (c.release) c.release(&c);
    if (p->has_dict && p->dict.release) p->dict.release(&p->dict);
    delete p;
    s->release = nullptr;
}
static void field_to_arrow(const TypeInfo &t, ArrowSchema *out) {
    auto *p = new SchemaPriv();
    memset(out, 0, sizeof *out);
    p->name = t.name;
    if (t.id == T_DICT) {
        TypeInfo idx;
        idx.id = t.index_id;
        p->format = t.format.empty() ? format_of(idx) : t.format;
        p->has_dict = true;
        field_to_arrow(*t.value, &p->dict);
        out->dictionary = &p->dict;
        if (t.dict_ordered) out->flags |= ARROW_FLAG_DICTIONARY_ORDERED;
    } else {
        p->format = format_of(t);
    }
    out->format = p->format.c_str();
    out->name = p->name.c_str();
    if (t.nullable) out->flags |= ARROW_FLAG_NULLABLE;
    out->release = release_schema;
    out->private_data = p;
}
void schema_to_arrow(const Schema &s, ArrowSchema *out) {
    auto *p = new SchemaPriv();
    memset(out, 0, sizeof *out);
    p->format = "+s";
    p->child_store.resize(s.fields.size());
    for (size_t i = 0; i < s.fields.size(); i++) {
        field_to_arrow(s.fields[i], &p->child_store[i]);
        p->child_ptrs.push_back(&p->child_store[i]);
    }
    out->format = p->format.c_str();
    out->name = p->name.c_str();
    out->n_children = (int64_t)s.fields.size();
    out->children = p->child_ptrs.data();
    out->release = release_schema;
    out->private_data = p;
}

// ------------------------------------------------------------------------------------------------
// memory
// ------------------------------------------------------------------------------------------------
static PinnedPool &pinned_pool() {
    static PinnedPool *p = new PinnedPool();  // intentionally leaked: exported arrays may outlive any ctx
    return *p;
}
void *PinnedPool::alloc(size_t bytes) {
    size_t sz = 4096;
    while (sz < bytes) sz <<= 1;
    {
        std::lock_guard<std::mutex> g(mu);
        auto it = free_list.find(sz);
        if (it != free_list.end()) {
            void *p = it->second;
            free_list.erase(it);
            cached -= sz;
            return p;
        }
    }
    void *p = nullptr;
    cudaError_t e = cudaMallocHost(&p, sz);
    if (e != cudaSuccess) {
        trim();
        e = cudaMallocHost(&p, sz);
        if (e != cudaSuccess) fail(DFGPU_RESOURCES_EXHAUSTED, std::string("pinned host allocation failed: ") + cudaGetErrorString(e));
    }
    std::lock_guard<std::mutex> g(mu);
    sizes[p] = sz;
    return p;
}
void PinnedPool::release(void *p) {
    if (!p) return;
    std::lock_guard<std::mutex> g(mu);
    size_t sz = sizes[p];
    if (cached + sz > (size_t)16 << 30) {
        sizes.erase(p);
        cudaFreeHost(p);
        return;
    }
    free_list.emplace(sz, p);
    cached += sz;
}
void PinnedPool::trim() {
    std::lock_guard<std::mutex> g(mu);
    for (auto &kv : free_list) {
        sizes.erase(kv.second);
        cudaFreeHost(kv.second);
    }
    free_list.clear();
    cached = 0;
}
void *pinned_alloc(size_t n) { return pinned_pool().alloc(n); }
void pinned_free(void *p) { pinned_pool().release(p); }

size_t BufCache::size_class(size_t n) {
    if (n < ((size_t)256 << 10)) return 0;
    size_t p = (size_t)1 << (63 - __builtin_clzll((unsigned long long)n));
    size_t step = p >> 3;  // 8 classes per octave: <= 12.5% internal waste
    return (n + step - 1) / step * step;
}
void *BufCache::take(size_t cls) {
    std::lock_guard<std::mutex> g(mu);
    auto it = free_list.find(cls);
    if (it == free_list.end()) return nullptr;
    void *p = it->second;
    free_list.erase(it);
    cached -= cls;
    return p;
}
bool BufCache::give(size_t cls, void *p) {
    std::lock_guard<std::mutex> g(mu);
    if (cached + cls > limit) return false;
    free_list.emplace(cls, p);
    cached += cls;
    return true;
}
void BufCache::trim(cudaStream_t st) {
    std::lock_guard<std::mutex> g(mu);
    for (auto &kv : free_list) cudaFreeAsync(kv.second, st);
    free_list.clear();
    cached = 0;
}

// ---- DFGPU_GUARD=1: red zones ----------------------------------------------------------------------
// compute-sanitizer is not available on every box; this is the library's own out-of-bounds-write detector.  Each
// buffer is allocated exactly (no size classes), with GUARD_ZONE bytes of 0xA5 before and after; the zones are
// checked on the device when the buffer is freed and the verdict is read at dfgpu_ctx_synchronize,
// dfgpu_batch_export and context teardown.  tests: DFGPU_GUARD=1 python -m pytest -m gpu
static const size_t GUARD_ZONE = 256;
__global__ void k_guard_fill(uint8_t *front, uint8_t *back) {
    front[threadIdx.x] = 0xA5;
    back[threadIdx.x] = 0xA5;
}
__global__ void k_guard_check(const uint8_t *front, const uint8_t *back, unsigned long long bytes, unsigned long long *words) {
    bool bad = front[threadIdx.x] != 0xA5 || back[threadIdx.x] != 0xA5;
    if (__syncthreads_or(bad) && threadIdx.x == 0)
        if (atomicAdd(&words[0], 1ull) == 0) words[1] = bytes;
}
void guard_poll(Ctx *ctx) {
    if (!ctx->guard) return;
    unsigned long long w[2] = {0, 0};
    DF_CUDA(cudaMemcpyAsync(w, ctx->guard_words, sizeof w, cudaMemcpyDeviceToHost, ctx->stream));
    DF_CUDA(cudaStreamSynchronize(ctx->stream));
    if (w[0]) fail(DFGPU_INTERNAL, "DFGPU_GUARD: " + std::to_string(w[0]) + " buffer(s) were written out of bounds; the first had " +
                                       std::to_string(w[1]) + " bytes");
}

DevBuf::DevBuf(Ctx *c, size_t n, bool zero) : ctx(c), bytes(n) {
    if (n == 0) return;
    if (c->guard) {
        uint8_t *raw = nullptr;
        cudaError_t ge = cudaMallocAsync((void **)&raw, n + 2 * GUARD_ZONE, c->stream);
        if (ge != cudaSuccess) {
            cudaGetLastError();
            fail(DFGPU_RESOURCES_EXHAUSTED, "Resources exhausted: failed to allocate " + std::to_string(n) + " bytes of device memory: " + cudaGetErrorString(ge));
        }
        ptr = raw + GUARD_ZONE;
        k_guard_fill<<<1, (int)GUARD_ZONE, 0, c->stream>>>(raw, raw + GUARD_ZONE + n);
        if (zero) DF_CUDA(cudaMemsetAsync(ptr, 0, n, c->stream));
        return;
    }
    cls = BufCache::size_class(n);
    if (cls) ptr = c->bufs.take(cls);
    cudaError_t e = cudaSuccess;
    if (!ptr) {
        e = cudaMallocAsync(&ptr, cls ? cls : n, c->stream);
        if (e != cudaSuccess) {  // give the cached blocks back to the pool and retry once
            cudaGetLastError();
            c->bufs.trim(c->stream);
            e = cudaMallocAsync(&ptr, cls ? cls : n, c->stream);
        }
    }
    if (e != cudaSuccess) {
        ptr = nullptr;
        cudaGetLastError();
        // ResourcesExhausted, like MemoryReservation::try_grow failing (hash_join.rs:708,738)
        fail(DFGPU_RESOURCES_EXHAUSTED, "Resources exhausted: failed to allocate " + std::to_string(n) +
                                            " bytes of device memory: " + cudaGetErrorString(e));
    }
    if (zero) DF_CUDA(cudaMemsetAsync(ptr, 0, n, c->stream));
}
DevBuf::~DevBuf() {
    if (!ptr || parent) return;
    if (ctx->guard) {
        uint8_t *raw = (uint8_t *)ptr - GUARD_ZONE;
        k_guard_check<<<1, (int)GUARD_ZONE, 0, ctx->stream>>>(raw, raw + GUARD_ZONE + bytes, (unsigned long long)bytes, ctx->guard_words);
        cudaFreeAsync(raw, ctx->stream);
        return;
    }
    if (cls && ctx->bufs.give(cls, ptr)) return;
    cudaFreeAsync(ptr, ctx->stream);
}

ColView DevCol::view() const {
    ColView v;
    v.validity = validity;
    v.values = values;
    v.offsets = offsets;
    v.vbit = vbit;
    v.dbit = dbit;
    v.type = type.id;
    v.width = type.id == T_DICT ? type_width(type.index_id) : type.width();
    v.large = type.is_large() ? 1 : 0;
    TypeInfo it;
    it.id = type.index_id;
    v.index_signed = type.id == T_DICT ? (it.is_signed_int() ? 1 : 0) : 0;
    return v;
}

// ------------------------------------------------------------------------------------------------
// small kernels
// ------------------------------------------------------------------------------------------------
// dst bit i = src bit (sbit + i), i < n; dst is word-aligned, tail bits zero
__global__ void k_bitmap_realign(uint32_t *dst, const uint8_t *src, int64_t sbit, int64_t n) {
    int64_t w = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    int64_t nw = (n + 31) / 32;
    for (; w < nw; w += (int64_t)gridDim.x * blockDim.x) {
        int64_t b0 = sbit + 32 * w;
        int64_t byte = b0 >> 3;
        int sh = (int)(b0 & 7);
        int64_t last_byte = (sbit + n - 1) >> 3;
        uint64_t acc = 0;
        for (int k = 0; k < 5; k++)
            if (byte + k <= last_byte) acc |= (uint64_t)src[byte + k] << (8 * k);
        uint32_t v = (uint32_t)(acc >> sh);
        int64_t rem = n - 32 * w;
        if (rem < 32) v &= (rem <= 0) ? 0u : ((1u << rem) - 1u);
        dst[w] = v;
    }
}
// dst bit (dbit + i) = src bit (sbit + i) for i < n; dst pre-zeroed, concurrent writers use atomicOr
__global__ void k_bitmap_append(uint32_t *dst, int64_t dbit, const uint8_t *src, int64_t sbit, int64_t n, int all_valid) {
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    for (; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        bool v = all_valid ? true : bit_get(src, sbit + i);
        if (v) atomicOr(&dst[(dbit + i) >> 5], 1u << ((dbit + i) & 31));
    }
}
__global__ void k_count_unset(const uint8_t *bm, int64_t sbit, int64_t n, unsigned long long *out) {
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    unsigned long long c = 0;
    for (; i < n; i += (int64_t)gridDim.x * blockDim.x) c += bit_get(bm, sbit + i) ? 0 : 1;
    for (int o = 16; o; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
    if ((threadIdx.x & 31) == 0 && c) atomicAdd(out, c);
}
__global__ void k_rebase_offsets32(int32_t *dst, const int32_t *src, int64_t n1, int64_t base, int64_t add) {
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    for (; i < n1; i += (int64_t)gridDim.x * blockDim.x) dst[i] = (int32_t)((int64_t)src[i] - base + add);
}
__global__ void k_rebase_offsets64(int64_t *dst, const int64_t *src, int64_t n1, int64_t base, int64_t add) {
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    for (; i < n1; i += (int64_t)gridDim.x * blockDim.x) dst[i] = src[i] - base + add;
}
__global__ void k_iota(void *dst, int width, int64_t n) {
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    for (; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        switch (width) {
        case 1: ((uint8_t *)dst)[i] = (uint8_t)i; break;
        case 2: ((uint16_t *)dst)[i] = (uint16_t)i; break;
        case 4: ((uint32_t *)dst)[i] = (uint32_t)i; break;
        default: ((uint64_t *)dst)[i] = (uint64_t)i; break;
        }
    }
}

int64_t count_nulls(Ctx *ctx, const uint8_t *validity, int64_t vbit, int64_t n) {
    if (!validity || n == 0) return 0;
    DevBuf cnt(ctx, 8, true);
    k_count_unset<<<grid_for(n, 256, ctx, 8), 256, 0, ctx->stream>>>(validity, vbit, n, cnt.as<unsigned long long>());
    note_launch(ctx);
    unsigned long long h = 0;
    DF_CUDA(cudaMemcpyAsync(&h, cnt.ptr, 8, cudaMemcpyDeviceToHost, ctx->stream));
    DF_CUDA(cudaStreamSynchronize(ctx->stream));
    return (int64_t)h;
}

// ---- exclusive prefix sum: block scan -> recursive scan of block totals -> add back -------------
static const int SCAN_BLOCK = 256, SCAN_ITEMS = 8, SCAN_TILE = SCAN_BLOCK * SCAN_ITEMS;
__global__ void __launch_bounds__(SCAN_BLOCK) k_scan_tiles(const int64_t *in, int64_t *out, int64_t n, int64_t *tile_sums) {
    __shared__ int64_t warp_sums[SCAN_BLOCK / 32];
    int64_t base = (int64_t)blockIdx.x * SCAN_TILE + (int64_t)threadIdx.x * SCAN_ITEMS;
    int64_t v[SCAN_ITEMS];
    int64_t sum = 0;
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; k++) {
        v[k] = (base + k < n) ? in[base + k] : 0;
        sum += v[k];
    }
    int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    int64_t incl = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        int64_t t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += t;
    }
    if (lane == 31) warp_sums[wid] = incl;
    __syncthreads();
    if (wid == 0) {
        int64_t w = lane < SCAN_BLOCK / 32 ? warp_sums[lane] : 0;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            int64_t t = __shfl_up_sync(0xffffffffu, w, o);
            if (lane >= o) w += t;
        }
        if (lane < SCAN_BLOCK / 32) warp_sums[lane] = w;
    }
    __syncthreads();
    int64_t excl = incl - sum + (wid ? warp_sums[wid - 1] : 0);
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; k++) {
        if (base + k < n) out[base + k] = excl;
        excl += v[k];
    }
    if (threadIdx.x == SCAN_BLOCK - 1 && tile_sums) tile_sums[blockIdx.x] = excl;
}
__global__ void k_scan_add(int64_t *out, int64_t n, const int64_t *tile_offsets) {
    int64_t i = (int64_t)blockIdx.x * SCAN_TILE + threadIdx.x;
    int64_t add = tile_offsets[blockIdx.x];
    for (int k = 0; k < SCAN_ITEMS; k++, i += SCAN_BLOCK)
        if (i < n) out[i] += add;
}
static void scan_rec(Ctx *ctx, const int64_t *in, int64_t *out, int64_t n, int64_t *total_dev) {
    int64_t tiles = (n + SCAN_TILE - 1) / SCAN_TILE;
    if (tiles <= 1) {
        k_scan_tiles<<<1, SCAN_BLOCK, 0, ctx->stream>>>(in, out, n, total_dev);
        note_launch(ctx);
        return;
    }
    DevBuf sums(ctx, (size_t)tiles * 8), offs(ctx, (size_t)tiles * 8);
    k_scan_tiles<<<(int)tiles, SCAN_BLOCK, 0, ctx->stream>>>(in, out, n, sums.as<int64_t>());
    note_launch(ctx);
    scan_rec(ctx, sums.as<int64_t>(), offs.as<int64_t>(), tiles, total_dev);
    k_scan_add<<<(int)tiles, SCAN_BLOCK, 0, ctx->stream>>>(out, n, offs.as<int64_t>());
    note_launch(ctx);
}
void exclusive_scan_i64(Ctx *ctx, const int64_t *in, int64_t *out, int64_t n, int64_t *total_host) {
    if (n == 0) {
        if (total_host) *total_host = 0;
        return;
    }
    DevBuf total(ctx, 8, true);
    scan_rec(ctx, in, out, n, total.as<int64_t>());
    if (total_host) {
        DF_CUDA(cudaMemcpyAsync(total_host, total.ptr, 8, cudaMemcpyDeviceToHost, ctx->stream));
        DF_CUDA(cudaStreamSynchronize(ctx->stream));
    }
}

// ------------------------------------------------------------------------------------------------
// import (Arrow C Data Interface -> HBM)
// ------------------------------------------------------------------------------------------------
static const void *upload(Ctx *ctx, DevCol &c, const void *src, size_t bytes, bool on_device) {
    if (bytes == 0 || !src) return nullptr;
    if (on_device) return src;
    // 16 bytes of slack so vector loads at the tail stay in bounds
    BufRef b = make_buf(ctx, bytes + 16);
    DF_CUDA(cudaMemcpyAsync(b->ptr, src, bytes, cudaMemcpyHostToDevice, ctx->stream));
    // A pageable source is staged by the driver before cudaMemcpyAsync returns; a pinned (or registered) one is read by the DMA
    // engine later.  Inputs are borrowed for the duration of the call only, so batch_import waits for such copies before it returns.
    if (bytes >= (64 << 10) || !ctx->import_pinned_pending) {
        cudaPointerAttributes at;
        if (cudaPointerGetAttributes(&at, src) == cudaSuccess) {
            if (at.type == cudaMemoryTypeHost || at.type == cudaMemoryTypeManaged) ctx->import_pinned_pending = true;
        } else {
            cudaGetLastError();
        }
    }
    ctx->h2d_bytes += (int64_t)bytes;
    c.owned.push_back(b);
    return b->ptr;
}

static ColRef import_array(Ctx *ctx, const ArrowArray *a, const TypeInfo &t, bool on_device) {
    auto col = std::make_shared<DevCol>();
    DevCol &c = *col;
    c.type = t;
    c.length = a->length;
    int64_t off = a->offset, n = a->length;
    const uint8_t *val = a->n_buffers > 0 ? (const uint8_t *)a->buffers[0] : nullptr;
    c.null_count = a->null_count;
    if (val && a->null_count != 0 && n > 0) {
        int64_t b0 = off >> 3, b1 = (off + n + 7) >> 3;
        c.validity = (const uint8_t *)upload(ctx, c, val + b0, (size_t)(b1 - b0), on_device);
        c.vbit = (int32_t)(off & 7);
        if (c.null_count < 0) c.null_count = -1;
    } else {
        c.null_count = 0;
    }
    int32_t phys = t.id == T_DICT ? t.index_id : t.id;
    int w = type_width(phys);
    if (phys == T_BOOL) {
        if (a->n_buffers < 2) fail(DFGPU_INTERNAL, "boolean array without values buffer");
        if (n > 0) {
            int64_t b0 = off >> 3, b1 = (off + n + 7) >> 3;
            c.values = upload(ctx, c, (const uint8_t *)a->buffers[1] + b0, (size_t)(b1 - b0), on_device);
            c.dbit = (int32_t)(off & 7);
        }
    } else if (w > 0) {
        if (a->n_buffers < 2) fail(DFGPU_INTERNAL, "primitive array without values buffer");
        if (n > 0) c.values = upload(ctx, c, (const uint8_t *)a->buffers[1] + off * w, (size_t)(n * w), on_device);
    } else if (t.is_var()) {
        if (a->n_buffers < 3) fail(DFGPU_INTERNAL, "variable-width array needs 3 buffers");
        int ow = t.is_large() ? 8 : 4;
        const uint8_t *ho = (const uint8_t *)a->buffers[1] + off * ow;
        int64_t first = 0, last = 0;
        if (n > 0 || a->buffers[1]) {
            if (on_device) {
                uint8_t tmp[16];
                DF_CUDA(cudaMemcpy(tmp, ho, ow, cudaMemcpyDeviceToHost));
                DF_CUDA(cudaMemcpy(tmp + 8, ho + n * ow, ow, cudaMemcpyDeviceToHost));
                first = ow == 8 ? *(int64_t *)tmp : *(int32_t *)tmp;
                last = ow == 8 ? *(int64_t *)(tmp + 8) : *(int32_t *)(tmp + 8);
            } else {
                first = ow == 8 ? ((const int64_t *)ho)[0] : ((const int32_t *)ho)[0];
                last = ow == 8 ? ((const int64_t *)ho)[n] : ((const int32_t *)ho)[n];
            }
            c.offsets = upload(ctx, c, ho, (size_t)((n + 1) * ow), on_device);
        }
        c.data_bytes = last - first;
        const uint8_t *hd = (const uint8_t *)a->buffers[2];
        if (on_device) {
            c.values = hd;  // device data base already matches the offsets
        } else if (last > first) {
            const uint8_t *dd = (const uint8_t *)upload(ctx, c, hd + first, (size_t)(last - first), false);
            c.values = dd - first;
        }
        if (!c.values) {
            // zero data bytes: any non-null base works
            BufRef b = make_buf(ctx, 16, true);
            c.owned.push_back(b);
            c.values = (const uint8_t *)b->ptr - first;
        }
    } else {
        fail(DFGPU_NOT_IMPLEMENTED, "This feature is not implemented: import of type " + t.describe());
    }
    if (t.id == T_DICT) {
        if (!a->dictionary) fail(DFGPU_INTERNAL, "dictionary array without dictionary");
        c.dict = import_array(ctx, a->dictionary, *t.value, on_device);
    }
    return col;
}

Batch *batch_import(Ctx *ctx, const ArrowArray *arr, const Schema &schema, bool on_device) {
    if (!arr) fail(DFGPU_INTERNAL, "null ArrowArray");
    if (arr->n_children != (int64_t)schema.fields.size())
        fail(DFGPU_INTERNAL, "RecordBatch has " + std::to_string(arr->n_children) + " columns, schema has " +
                                 std::to_string(schema.fields.size()));
    if (arr->offset != 0) fail(DFGPU_NOT_IMPLEMENTED, "struct-level offset on an imported RecordBatch");
    std::unique_ptr<Batch> b(new Batch());
    b->ctx = ctx;
    b->num_rows = arr->length;
    for (int64_t i = 0; i < arr->n_children; i++) {
        const ArrowArray *ch = arr->children[i];
        if (ch->length != arr->length) fail(DFGPU_INTERNAL, "column length mismatch in RecordBatch");
        b->cols.push_back(import_array(ctx, ch, schema.fields[(size_t)i], on_device));
    }
    if (ctx->import_pinned_pending && !ctx->import_defer_wait) {
        // the caller may release or reuse its (pinned) buffers as soon as this call returns: the copies must have read them
        ctx->import_pinned_pending = false;
        if (!ctx->import_event) DF_CUDA(cudaEventCreateWithFlags(&ctx->import_event, cudaEventDisableTiming));
        DF_CUDA(cudaEventRecord(ctx->import_event, ctx->stream));
        DF_CUDA(cudaEventSynchronize(ctx->import_event));
    }
    return b.release();
}

// ------------------------------------------------------------------------------------------------
// export (HBM -> pinned host buffers, released by the consumer through ArrowArray.release)
// ------------------------------------------------------------------------------------------------
struct ArrayPriv {
    std::vector<void *> pinned;  // owned host buffers
    std::vector<const void *> buffers;
    std::vector<ArrowArray> child_store;
    std::vector<ArrowArray *> child_ptrs;
    ArrowArray dict;
    bool has_dict = false;
};
static void release_array(ArrowArray *a) {
    if (!a || !a->release) return;
    auto *p = (ArrayPriv *)a->private_data;
    for (auto &c : p->child_store)
        if (c.release) c.release(&c);
    if (p->has_dict && p->dict.release) p->dict.release(&p->dict);
    for (void *h : p->pinned) pinned_free(h);
    delete p;
    a->release = nullptr;
}
static const void *download(Ctx *ctx, ArrayPriv *p, const void *dev, size_t bytes) {
    size_t alloc = bytes ? bytes : 8;
    void *h = pinned_alloc(alloc);
    p->pinned.push_back(h);
    if (bytes) {
        DF_CUDA(cudaMemcpyAsync(h, dev, bytes, cudaMemcpyDeviceToHost, ctx->stream));
        ctx->d2h_bytes += (int64_t)bytes;
    } else {
        memset(h, 0, alloc);
    }
    return h;
}

static void export_array(Ctx *ctx, const DevCol &c, ArrowArray *out, std::vector<BufRef> &temps,
                         std::vector<std::pair<ArrowArray *, BufRef>> &pending_nulls) {
    auto *p = new ArrayPriv();
    memset(out, 0, sizeof *out);
    out->private_data = p;
    out->release = release_array;
    int64_t n = c.length;
    out->length = n;
    out->offset = 0;
    // validity (re-aligned to bit 0)
    const void *hval = nullptr;
    out->null_count = 0;
    if (c.validity && c.null_count != 0 && n > 0) {
        size_t words = (size_t)((n + 31) / 32);
        const uint8_t *src = c.validity;
        if (c.vbit != 0) {
            BufRef t = make_buf(ctx, words * 4);
            k_bitmap_realign<<<grid_for((int64_t)words, 256, ctx, 8), 256, 0, ctx->stream>>>(t->as<uint32_t>(), c.validity, c.vbit, n);
            note_launch(ctx);
            temps.push_back(t);
            src = t->as<uint8_t>();
        }
        hval = download(ctx, p, src, (size_t)((n + 7) / 8));
        if (c.null_count < 0) {
            BufRef cnt = make_buf(ctx, 8, true);
            k_count_unset<<<grid_for(n, 256, ctx, 8), 256, 0, ctx->stream>>>(c.validity, c.vbit, n, cnt->as<unsigned long long>());
            note_launch(ctx);
            pending_nulls.emplace_back(out, cnt);
        } else {
            out->null_count = c.null_count;
        }
    }
    p->buffers.push_back(hval);
    int32_t phys = c.type.id == T_DICT ? c.type.index_id : c.type.id;
    int w = type_width(phys);
    if (phys == T_BOOL) {
        const uint8_t *src = (const uint8_t *)c.values;
        size_t words = (size_t)((n + 31) / 32);
        if (c.dbit != 0 && n > 0) {
            BufRef t = make_buf(ctx, words * 4);
            k_bitmap_realign<<<grid_for((int64_t)words, 256, ctx, 8), 256, 0, ctx->stream>>>(t->as<uint32_t>(), src, c.dbit, n);
            note_launch(ctx);
            temps.push_back(t);
            src = t->as<uint8_t>();
        }
        p->buffers.push_back(download(ctx, p, src, (size_t)((n + 7) / 8)));
    } else if (w > 0) {
        p->buffers.push_back(download(ctx, p, c.values, (size_t)(n * w)));
    } else if (c.type.is_var()) {
        int ow = c.type.is_large() ? 8 : 4;
        // rebase offsets to start at zero
        int64_t first = 0;
        if (n > 0) {
            uint8_t tmp[8] = {0};
            DF_CUDA(cudaMemcpyAsync(tmp, c.offsets, ow, cudaMemcpyDeviceToHost, ctx->stream));
            DF_CUDA(cudaStreamSynchronize(ctx->stream));
            first = ow == 8 ? *(int64_t *)tmp : *(int32_t *)tmp;
        }
        if (n == 0) {
            void *h = pinned_alloc(16);
            memset(h, 0, 16);
            p->pinned.push_back(h);
            p->buffers.push_back(h);
        } else if (first != 0) {
            BufRef t = make_buf(ctx, (size_t)((n + 1) * ow));
            if (ow == 8)
                k_rebase_offsets64<<<grid_for(n + 1, 256, ctx, 8), 256, 0, ctx->stream>>>(t->as<int64_t>(), (const int64_t *)c.offsets, n + 1, first, 0);
            else
                k_rebase_offsets32<<<grid_for(n + 1, 256, ctx, 8), 256, 0, ctx->stream>>>(t->as<int32_t>(), (const int32_t *)c.offsets, n + 1, first, 0);
            note_launch(ctx);
            temps.push_back(t);
            p->buffers.push_back(download(ctx, p, t->ptr, (size_t)((n + 1) * ow)));
        } else {
            p->buffers.push_back(download(ctx, p, c.offsets, (size_t)((n + 1) * ow)));
        }
        p->buffers.push_back(download(ctx, p, (const uint8_t *)c.values + first, (size_t)c.data_bytes));
    } else {
        fail(DFGPU_NOT_IMPLEMENTED, "export of type " + c.type.describe());
    }
    if (c.type.id == T_DICT) {
        p->has_dict = true;
        export_array(ctx, *c.dict, &p->dict, temps, pending_nulls);
        out->dictionary = &p->dict;
    }
    out->n_buffers = (int64_t)p->buffers.size();
    out->buffers = p->buffers.data();
}

void batch_export(Ctx *ctx, const Batch *b, ArrowArray *out, ArrowSchema *out_schema) {
    auto *p = new ArrayPriv();
    memset(out, 0, sizeof *out);
    out->private_data = p;
    out->release = release_array;
    out->length = b->num_rows;
    p->buffers.push_back(nullptr);
    out->n_buffers = 1;
    out->buffers = p->buffers.data();
    p->child_store.resize(b->cols.size());
    std::vector<BufRef> temps;
    std::vector<std::pair<ArrowArray *, BufRef>> pending_nulls;
    try {
        for (size_t i = 0; i < b->cols.size(); i++) {
            export_array(ctx, *b->cols[i], &p->child_store[i], temps, pending_nulls);
        }
        std::vector<unsigned long long> counts(pending_nulls.size());
        for (size_t i = 0; i < pending_nulls.size(); i++)
            DF_CUDA(cudaMemcpyAsync(&counts[i], pending_nulls[i].second->ptr, 8, cudaMemcpyDeviceToHost, ctx->stream));
        DF_CUDA(cudaStreamSynchronize(ctx->stream));
        for (size_t i = 0; i < pending_nulls.size(); i++) pending_nulls[i].first->null_count = (int64_t)counts[i];
    } catch (...) {
        release_array(out);
        throw;
    }
    for (auto &c : p->child_store) p->child_ptrs.push_back(&c);
    out->n_children = (int64_t)p->child_ptrs.size();
    out->children = p->child_ptrs.data();
    if (out_schema) schema_to_arrow(b->schema(), out_schema);
}

// ------------------------------------------------------------------------------------------------
// nulls / slice / dictionary decode / take / concat
// ------------------------------------------------------------------------------------------------
ColRef col_nulls(Ctx *ctx, const TypeInfo &t, int64_t n) {  // new_null_array (joins/utils.rs:1209-1214)
    auto c = std::make_shared<DevCol>();
    c->type = t;
    c->length = n;
    c->null_count = n;
    size_t vb = (size_t)((n + 31) / 32) * 4;
    BufRef v = make_buf(ctx, vb + 16, true);
    c->owned.push_back(v);
    c->validity = v->as<uint8_t>();
    int32_t phys = t.id == T_DICT ? t.index_id : t.id;
    int w = type_width(phys);
    if (phys == T_BOOL) {
        BufRef b = make_buf(ctx, vb + 16, true);
        c->owned.push_back(b);
        c->values = b->ptr;
    } else if (w > 0) {
        BufRef b = make_buf(ctx, (size_t)(n * w) + 16, true);
        c->owned.push_back(b);
        c->values = b->ptr;
    } else {
        int ow = t.is_large() ? 8 : 4;
        BufRef o = make_buf(ctx, (size_t)((n + 1) * ow) + 16, true);
        BufRef d = make_buf(ctx, 16, true);
        c->owned.push_back(o);
        c->owned.push_back(d);
        c->offsets = o->ptr;
        c->values = d->ptr;
        c->data_bytes = 0;
    }
    if (t.id == T_DICT) {
        // empty dictionary; every index is NULL
        auto d = std::make_shared<DevCol>();
        d->type = *t.value;
        d->length = 0;
        if (d->type.is_var()) {
            BufRef o = make_buf(ctx, 32, true);
            d->owned.push_back(o);
            d->offsets = o->ptr;
            d->values = o->as<uint8_t>() + 16;
        }
        c->dict = d;
    }
    return c;
}

static ColRef col_slice(const ColRef &src, int64_t off, int64_t len, std::shared_ptr<void> keep) {
    auto c = std::make_shared<DevCol>(*src);
    c->length = len;
    c->keepalive = keep;
    if (src->validity) {
        int64_t bit = src->vbit + off;
        c->validity = src->validity + (bit >> 3);
        c->vbit = (int32_t)(bit & 7);
        c->null_count = (off == 0 && len == src->length) ? src->null_count : -1;
    }
    int32_t phys = src->type.id == T_DICT ? src->type.index_id : src->type.id;
    int w = type_width(phys);
    if (phys == T_BOOL) {
        int64_t bit = src->dbit + off;
        c->values = (const uint8_t *)src->values + (bit >> 3);
        c->dbit = (int32_t)(bit & 7);
    } else if (w > 0) {
        c->values = (const uint8_t *)src->values + off * w;
    } else {
        int ow = src->type.is_large() ? 8 : 4;
        c->offsets = (const uint8_t *)src->offsets + off * ow;
        // data_bytes is an upper bound for a slice (exact value needs the offsets)
    }
    return c;
}

Batch *batch_slice(Ctx *ctx, const Batch *b, int64_t off, int64_t len) {
    if (off < 0 || len < 0 || off + len > b->num_rows) fail(DFGPU_INTERNAL, "slice out of bounds");
    std::unique_ptr<Batch> o(new Batch());
    o->ctx = ctx;
    o->num_rows = len;
    for (auto &c : b->cols) {
        ColRef s = col_slice(c, off, len, nullptr);
        if (s->type.is_var() && len != c->length) {
            // exact byte count of the slice
            int ow = s->type.is_large() ? 8 : 4;
            uint8_t t0[8] = {0}, t1[8] = {0};
            DF_CUDA(cudaMemcpyAsync(t0, s->offsets, ow, cudaMemcpyDeviceToHost, ctx->stream));
            DF_CUDA(cudaMemcpyAsync(t1, (const uint8_t *)s->offsets + len * ow, ow, cudaMemcpyDeviceToHost, ctx->stream));
            DF_CUDA(cudaStreamSynchronize(ctx->stream));
            int64_t f = ow == 8 ? *(int64_t *)t0 : *(int32_t *)t0, l = ow == 8 ? *(int64_t *)t1 : *(int32_t *)t1;
            s->data_bytes = l - f;
        }
        o->cols.push_back(s);
    }
    return o.release();
}

// ---- take ------------------------------------------------------------------------------------------
template <int W>
__global__ void k_take_fixed(const uint8_t *__restrict__ src, ColView sv, const int64_t *__restrict__ idx, int64_t n,
                             uint8_t *__restrict__ dst, uint32_t *__restrict__ out_valid) {
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    int64_t stride = (int64_t)gridDim.x * blockDim.x;
    int64_t nround = (n + 31) / 32 * 32;
    for (; i < nround; i += stride) {
        bool valid = false;
        if (i < n) {
            int64_t j = idx[i];
            valid = j >= 0 && col_valid(sv, j);
            if (W == 16) {
                uint4 v = valid ? *(const uint4 *)(src + 16 * j) : make_uint4(0, 0, 0, 0);
                *(uint4 *)(dst + 16 * i) = v;
            } else if (W == 8) {
                ((uint64_t *)dst)[i] = valid ? ((const uint64_t *)src)[j] : 0;
            } else if (W == 4) {
                ((uint32_t *)dst)[i] = valid ? ((const uint32_t *)src)[j] : 0;
            } else if (W == 2) {
                ((uint16_t *)dst)[i] = valid ? ((const uint16_t *)src)[j] : 0;
            } else {
                dst[i] = valid ? src[j] : 0;
            }
        }
        uint32_t m = __ballot_sync(0xffffffffu, valid);
        if ((threadIdx.x & 31) == 0 && out_valid) out_valid[i >> 5] = m;
    }
}
__global__ void k_take_bool(ColView sv, const int64_t *__restrict__ idx, int64_t n, uint32_t *__restrict__ dst,
                            uint32_t *__restrict__ out_valid) {
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    int64_t stride = (int64_t)gridDim.x * blockDim.x;
    int64_t nround = (n + 31) / 32 * 32;
    for (; i < nround; i += stride) {
        bool valid = false, v = false;
        if (i < n) {
            int64_t j = idx[i];
            valid = j >= 0 && col_valid(sv, j);
            v = valid && bit_get((const uint8_t *)sv.values, sv.dbit + j);
        }
        uint32_t m = __ballot_sync(0xffffffffu, valid), b = __ballot_sync(0xffffffffu, v);
        if ((threadIdx.x & 31) == 0) {
            dst[i >> 5] = b;
            if (out_valid) out_valid[i >> 5] = m;
        }
    }
}
__global__ void k_take_lengths(ColView sv, const int64_t *__restrict__ idx, int64_t n, int64_t *__restrict__ lens,
                               uint32_t *__restrict__ out_valid) {
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    int64_t stride = (int64_t)gridDim.x * blockDim.x;
    int64_t nround = (n + 31) / 32 * 32;
    for (; i < nround; i += stride) {
        bool valid = false;
        if (i < n) {
            int64_t j = idx[i];
            valid = j >= 0 && col_valid(sv, j);
            int64_t len = 0;
            if (valid) {
                const uint8_t *p;
                col_bytes(sv, j, p, len);
            }
            lens[i] = len;
        }
        uint32_t m = __ballot_sync(0xffffffffu, valid);
        if ((threadIdx.x & 31) == 0 && out_valid) out_valid[i >> 5] = m;
    }
}
// one warp per row copies the bytes; offsets written in the requested width
__global__ void k_take_bytes(ColView sv, const int64_t *__restrict__ idx, int64_t n, const int64_t *__restrict__ starts,
                             uint8_t *__restrict__ data, void *__restrict__ out_offsets, int large, int64_t total) {
    int lane = threadIdx.x & 31;
    int64_t w = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
    int64_t nw = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t i = w; i < n; i += nw) {
        int64_t j = idx[i];
        int64_t st = starts[i];
        if (lane == 0) {
            if (large) ((int64_t *)out_offsets)[i] = st; else ((int32_t *)out_offsets)[i] = (int32_t)st;
            if (i == n - 1) {
                if (large) ((int64_t *)out_offsets)[n] = total; else ((int32_t *)out_offsets)[n] = (int32_t)total;
            }
        }
        if (j >= 0 && col_valid(sv, j)) {
            const uint8_t *p; int64_t len;
            col_bytes(sv, j, p, len);
            for (int64_t k = lane; k < len; k += 32) data[st + k] = p[k];
        }
    }
}

ColRef col_take(Ctx *ctx, const DevCol &src, const int64_t *idx, int64_t n, bool idx_may_be_null) {
    auto c = std::make_shared<DevCol>();
    c->type = src.type;
    c->length = n;
    bool need_valid = idx_may_be_null || (src.validity && src.null_count != 0);
    size_t vwords = (size_t)((n + 31) / 32);
    uint32_t *ov = nullptr;
    if (need_valid) {
        BufRef v = make_buf(ctx, vwords * 4 + 16, true);
        c->owned.push_back(v);
        ov = v->as<uint32_t>();
        c->validity = v->as<uint8_t>();
        c->null_count = -1;
    }
    ColView sv = src.view();
    int32_t phys = src.type.id == T_DICT ? src.type.index_id : src.type.id;
    int w = type_width(phys);
    int grid = grid_for(n, 256, ctx, 16);
    if (n == 0) {
        // nothing to gather; still need legal buffers
        if (src.type.is_var()) {
            BufRef o = make_buf(ctx, 32, true);
            c->owned.push_back(o);
            c->offsets = o->ptr;
            c->values = o->as<uint8_t>() + 16;
        }
        c->dict = src.dict;
        return c;
    }
    if (phys == T_BOOL) {
        BufRef b = make_buf(ctx, vwords * 4 + 16, true);
        c->owned.push_back(b);
        k_take_bool<<<grid, 256, 0, ctx->stream>>>(sv, idx, n, b->as<uint32_t>(), ov);
        c->values = b->ptr;
    } else if (w > 0) {
        BufRef b = make_buf(ctx, (size_t)(n * w) + 16);
        c->owned.push_back(b);
        const uint8_t *s = (const uint8_t *)src.values;
        uint8_t *d = b->as<uint8_t>();
        switch (w) {
        case 1: k_take_fixed<1><<<grid, 256, 0, ctx->stream>>>(s, sv, idx, n, d, ov); break;
        case 2: k_take_fixed<2><<<grid, 256, 0, ctx->stream>>>(s, sv, idx, n, d, ov); break;
        case 4: k_take_fixed<4><<<grid, 256, 0, ctx->stream>>>(s, sv, idx, n, d, ov); break;
        case 8: k_take_fixed<8><<<grid, 256, 0, ctx->stream>>>(s, sv, idx, n, d, ov); break;
        default: k_take_fixed<16><<<grid, 256, 0, ctx->stream>>>(s, sv, idx, n, d, ov); break;
        }
        c->values = b->ptr;
    } else {
        BufRef lens = make_buf(ctx, (size_t)n * 8), starts = make_buf(ctx, (size_t)n * 8);
        k_take_lengths<<<grid, 256, 0, ctx->stream>>>(sv, idx, n, lens->as<int64_t>(), ov);
        note_launch(ctx);
        int64_t total = 0;
        exclusive_scan_i64(ctx, lens->as<int64_t>(), starts->as<int64_t>(), n, &total);
        if (!src.type.is_large() && total > 0x7fffffffLL)
            fail(DFGPU_EXECUTION, "offset overflow: gathered Utf8/Binary data exceeds i32 offsets");  // binary_map.rs:464-471
        int ow = src.type.is_large() ? 8 : 4;
        BufRef o = make_buf(ctx, (size_t)((n + 1) * ow) + 16), d = make_buf(ctx, (size_t)total + 16);
        c->owned.push_back(o);
        c->owned.push_back(d);
        k_take_bytes<<<grid_for(n * 32, 256, ctx, 16), 256, 0, ctx->stream>>>(sv, idx, n, starts->as<int64_t>(), d->as<uint8_t>(),
                                                                             o->ptr, src.type.is_large() ? 1 : 0, total);
        c->offsets = o->ptr;
        c->values = d->ptr;
        c->data_bytes = total;
    }
    note_launch(ctx);
    c->dict = src.dict;
    return c;
}

// ---- take of several columns in one launch ----------------------------------------------------------
// build_batch_from_indices (joins/utils.rs:1180-1230) gathers every output column with the same one
// or two index arrays.  Fixed-width columns without NULLs (on either side of the gather) are taken
// together: the indices are read once per row and the row's gathers are all in flight at once.
static const int TAKE_MAXCOLS = 12;
struct TakeCol {
    const uint8_t *src;
    uint8_t *dst;
    const int64_t *idx;
    int32_t w, stride;   // element j of the source lies at src + stride * j (a plain column: stride = w; a field of packed rows: the row size)
};
struct TakePlan {
    TakeCol c[TAKE_MAXCOLS];
    int32_t n, pad;
};
DF_D uint4 take_load(const uint8_t *src, int w, int64_t j, int stride) {
    uint4 v = make_uint4(0, 0, 0, 0);
    src += (int64_t)stride * j;
    if (w == 16) v = *(const uint4 *)src;
    else if (w == 8) { const uint2 t = *(const uint2 *)src; v.x = t.x; v.y = t.y; }
    else if (w == 4) v.x = *(const uint32_t *)src;
    else if (w == 2) v.x = *(const uint16_t *)src;
    else v.x = *src;
    return v;
}
DF_D void take_store(uint8_t *dst, int w, int64_t i, const uint4 &v) {
    if (w == 16) *(uint4 *)(dst + 16 * i) = v;
    else if (w == 8) *(uint2 *)(dst + 8 * i) = make_uint2(v.x, v.y);
    else if (w == 4) *(uint32_t *)(dst + 4 * i) = v.x;
    else if (w == 2) *(uint16_t *)(dst + 2 * i) = (uint16_t)v.x;
    else dst[i] = (uint8_t)v.x;
}
// columns -> packed rows: TakeCol.src = the column, .dst = the field's place in row 0, .stride = the row size
__global__ void __launch_bounds__(256) k_pack_rows(const __grid_constant__ TakePlan p, int64_t n) {
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    for (; i < n; i += (int64_t)gridDim.x * blockDim.x)
        for (int c = 0; c < p.n; c++) {
            const uint4 v = take_load(p.c[c].src, p.c[c].w, i, p.c[c].w);
            take_store(p.c[c].dst + (int64_t)p.c[c].stride * i, p.c[c].w, 0, v);
        }
}
__global__ void __launch_bounds__(256) k_take_multi(const __grid_constant__ TakePlan p, int64_t n) {
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    for (; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        for (int c0 = 0; c0 < p.n; c0 += 4) {
            uint4 v[4];
#pragma unroll
            for (int q = 0; q < 4; q++)
                if (c0 + q < p.n) v[q] = take_load(p.c[c0 + q].src, p.c[c0 + q].w, p.c[c0 + q].idx[i], p.c[c0 + q].stride);
#pragma unroll
            for (int q = 0; q < 4; q++)
                if (c0 + q < p.n) take_store(p.c[c0 + q].dst, p.c[c0 + q].w, i, v[q]);
        }
    }
}

// Fixed-width columns without nulls packed row-wise: a gather of several of them by the same index then costs one DRAM line
// per row instead of one per column (a random 8-byte load pulls a whole 128-byte line on this part: ncu measured 117 B of
// DRAM read per gathered value).  Fields sit at offsets aligned to their width; rows are padded to 16 / 32 / 64 bytes so
// that no row straddles a line.  Returns an empty PackedRows when fewer than two columns qualify or a row would exceed 64 bytes.
PackedRows pack_rows(Ctx *ctx, const std::vector<ColRef> &cols, int64_t n) {
    PackedRows pr;
    pr.off.assign(cols.size(), -1);
    std::vector<int> take;
    for (int w : {16, 8, 4, 2, 1})   // widest first: every field is aligned to its width
        for (size_t c = 0; c < cols.size(); c++) {
            const DevCol &src = *cols[c];
            if (src.type.id == T_DICT || src.type.id == T_BOOL || type_width(src.type.id) != w || (src.validity && src.null_count != 0)) continue;
            take.push_back((int)c);
        }
    int bytes = 0;
    for (int c : take) bytes += type_width(cols[(size_t)c]->type.id);
    if (take.size() < 2 || bytes > 64 || n == 0 || (int)take.size() > TAKE_MAXCOLS) return pr;
    pr.row_bytes = bytes <= 16 ? 16 : bytes <= 32 ? 32 : 64;
    pr.rows = make_buf(ctx, (size_t)n * pr.row_bytes + 16);
    TakePlan p;
    memset(&p, 0, sizeof p);
    int at = 0;
    for (int c : take) {
        const int w = type_width(cols[(size_t)c]->type.id);
        pr.off[(size_t)c] = at;
        p.c[p.n++] = TakeCol{(const uint8_t *)cols[(size_t)c]->values, pr.rows->as<uint8_t>() + at, nullptr, w, pr.row_bytes};
        at += w;
    }
    k_pack_rows<<<grid_for(n, 256, ctx, 16), 256, 0, ctx->stream>>>(p, n);
    note_launch(ctx);
    DF_CUDA(cudaGetLastError());
    return pr;
}

std::vector<ColRef> cols_take(Ctx *ctx, const std::vector<const DevCol *> &srcs, const std::vector<const int64_t *> &idx, int64_t n,
                              const std::vector<bool> &idx_may_be_null, const std::vector<std::pair<const uint8_t *, int>> *packed) {
    std::vector<ColRef> out(srcs.size());
    TakePlan p;
    memset(&p, 0, sizeof p);
    auto flush = [&]() {
        if (!p.n) return;
        k_take_multi<<<grid_for(n, 256, ctx, 16), 256, 0, ctx->stream>>>(p, n);
        note_launch(ctx);
        p.n = 0;
    };
    for (size_t k = 0; k < srcs.size(); k++) {
        const DevCol &src = *srcs[k];
        const int32_t phys = src.type.id == T_DICT ? src.type.index_id : src.type.id;
        const int w = type_width(phys);
        const bool plain = n > 0 && w > 0 && phys != T_BOOL && !idx_may_be_null[k] && !(src.validity && src.null_count != 0);
        if (!plain) {
            out[k] = col_take(ctx, src, idx[k], n, idx_may_be_null[k]);
            continue;
        }
        auto c = std::make_shared<DevCol>();
        c->type = src.type;
        c->length = n;
        BufRef b = make_buf(ctx, (size_t)(n * w) + 16);
        c->owned.push_back(b);
        c->values = b->ptr;
        c->dict = src.dict;
        if (packed && (*packed)[k].first) p.c[p.n++] = TakeCol{(*packed)[k].first, b->as<uint8_t>(), idx[k], w, (*packed)[k].second};
        else p.c[p.n++] = TakeCol{(const uint8_t *)src.values, b->as<uint8_t>(), idx[k], w, w};
        out[k] = c;
        if (p.n == TAKE_MAXCOLS) flush();
    }
    flush();
    DF_CUDA(cudaGetLastError());
    return out;
}

__global__ void k_dict_to_idx(ColView sv, ColView dv, int64_t n, int64_t *idx) {
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    for (; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        int64_t j = -1;
        if (col_valid(sv, i)) {
            j = dict_index(sv, i);
            if (!col_valid(dv, j)) j = -1;
        }
        idx[i] = j;
    }
}
ColRef col_decode_dictionary(Ctx *ctx, const DevCol &src) {
    if (src.type.id != T_DICT) fail(DFGPU_INTERNAL, "col_decode_dictionary on a non-dictionary column");
    BufRef idx = make_buf(ctx, (size_t)src.length * 8 + 8);
    if (src.length) {
        k_dict_to_idx<<<grid_for(src.length, 256, ctx, 16), 256, 0, ctx->stream>>>(src.view(), src.dict->view(), src.length, idx->as<int64_t>());
        note_launch(ctx);
    }
    ColRef out = col_take(ctx, *src.dict, idx->as<int64_t>(), src.length, true);
    out->type.name = src.type.name;
    out->type.nullable = src.type.nullable;
    return out;
}

static ColRef col_concat(Ctx *ctx, const std::vector<ColRef> &parts, int64_t total) {
    const DevCol &first = *parts[0];
    if (first.type.id == T_DICT) {
        bool same = true;
        for (auto &p : parts) same = same && p->dict.get() == first.dict.get();
        std::vector<ColRef> plain;
        if (!same) {
            // decode, concatenate, and wrap in a trivial dictionary (indices = 0..n-1) to keep the type
            for (auto &p : parts) plain.push_back(col_decode_dictionary(ctx, *p));
            ColRef vals = col_concat(ctx, plain, total);
            auto c = std::make_shared<DevCol>();
            c->type = first.type;
            c->length = total;
            int iw = type_width(first.type.index_id);
            TypeInfo it; it.id = first.type.index_id;
            double maxv = iw == 8 ? 9.2e18 : (double)((1ULL << (8 * iw - (it.is_signed_int() ? 1 : 0))) - 1);
            if ((double)total > maxv) fail(DFGPU_NOT_IMPLEMENTED, "dictionary index type too narrow to concatenate dictionaries");
            BufRef ib = make_buf(ctx, (size_t)(total * iw) + 16);
            if (total) {
                k_iota<<<grid_for(total, 256, ctx, 16), 256, 0, ctx->stream>>>(ib->ptr, iw, total);
                note_launch(ctx);
            }
            c->owned.push_back(ib);
            c->values = ib->ptr;
            c->validity = vals->validity;
            c->vbit = vals->vbit;
            c->null_count = vals->null_count;
            auto d = std::make_shared<DevCol>(*vals);
            d->validity = nullptr; d->null_count = 0;  // nulls live on the indices
            d->type.name = first.type.value->name;
            c->dict = d;
            c->keepalive = vals;
            return c;
        }
    }
    auto c = std::make_shared<DevCol>();
    c->type = first.type;
    c->length = total;
    c->dict = first.dict;
    bool any_nulls = false;
    for (auto &p : parts) any_nulls = any_nulls || (p->validity && p->null_count != 0);
    size_t vwords = (size_t)((total + 31) / 32);
    if (any_nulls) {
        BufRef v = make_buf(ctx, vwords * 4 + 16, true);
        c->owned.push_back(v);
        c->validity = v->as<uint8_t>();
        int64_t pos = 0;
        c->null_count = 0;
        for (auto &p : parts) {
            if (p->length) {
                bool allv = !(p->validity && p->null_count != 0);
                k_bitmap_append<<<grid_for(p->length, 256, ctx, 16), 256, 0, ctx->stream>>>(v->as<uint32_t>(), pos, p->validity, p->vbit, p->length, allv ? 1 : 0);
                note_launch(ctx);
            }
            if (p->null_count < 0 || c->null_count < 0) c->null_count = -1; else c->null_count += p->null_count;
            pos += p->length;
        }
    }
    int32_t phys = first.type.id == T_DICT ? first.type.index_id : first.type.id;
    int w = type_width(phys);
    if (phys == T_BOOL) {
        BufRef b = make_buf(ctx, vwords * 4 + 16, true);
        c->owned.push_back(b);
        int64_t pos = 0;
        for (auto &p : parts) {
            if (p->length) {
                k_bitmap_append<<<grid_for(p->length, 256, ctx, 16), 256, 0, ctx->stream>>>(b->as<uint32_t>(), pos, (const uint8_t *)p->values, p->dbit, p->length, 0);
                note_launch(ctx);
            }
            pos += p->length;
        }
        c->values = b->ptr;
    } else if (w > 0) {
        BufRef b = make_buf(ctx, (size_t)(total * w) + 16);
        c->owned.push_back(b);
        int64_t pos = 0;
        for (auto &p : parts) {
            if (p->length) DF_CUDA(cudaMemcpyAsync(b->as<uint8_t>() + pos * w, p->values, (size_t)(p->length * w), cudaMemcpyDeviceToDevice, ctx->stream));
            pos += p->length;
        }
        c->values = b->ptr;
    } else {
        int ow = first.type.is_large() ? 8 : 4;
        // per-part first offsets (host) to rebase
        std::vector<int64_t> firsts(parts.size(), 0), lasts(parts.size(), 0);
        std::vector<uint8_t> tmp(parts.size() * 16, 0);
        for (size_t k = 0; k < parts.size(); k++) {
            if (!parts[k]->length) continue;
            DF_CUDA(cudaMemcpyAsync(&tmp[k * 16], parts[k]->offsets, ow, cudaMemcpyDeviceToHost, ctx->stream));
            DF_CUDA(cudaMemcpyAsync(&tmp[k * 16 + 8], (const uint8_t *)parts[k]->offsets + parts[k]->length * ow, ow, cudaMemcpyDeviceToHost, ctx->stream));
        }
        DF_CUDA(cudaStreamSynchronize(ctx->stream));
        int64_t bytes = 0;
        for (size_t k = 0; k < parts.size(); k++) {
            if (!parts[k]->length) continue;
            firsts[k] = ow == 8 ? *(int64_t *)&tmp[k * 16] : *(int32_t *)&tmp[k * 16];
            lasts[k] = ow == 8 ? *(int64_t *)&tmp[k * 16 + 8] : *(int32_t *)&tmp[k * 16 + 8];
            bytes += lasts[k] - firsts[k];
        }
        if (ow == 4 && bytes > 0x7fffffffLL) fail(DFGPU_EXECUTION, "offset overflow concatenating Utf8/Binary data");
        BufRef o = make_buf(ctx, (size_t)((total + 1) * ow) + 16, true), d = make_buf(ctx, (size_t)bytes + 16);
        c->owned.push_back(o);
        c->owned.push_back(d);
        int64_t pos = 0, bpos = 0;
        for (size_t k = 0; k < parts.size(); k++) {
            auto &p = parts[k];
            if (!p->length) continue;
            if (ow == 8)
                k_rebase_offsets64<<<grid_for(p->length + 1, 256, ctx, 16), 256, 0, ctx->stream>>>(o->as<int64_t>() + pos, (const int64_t *)p->offsets, p->length + 1, firsts[k], bpos);
            else
                k_rebase_offsets32<<<grid_for(p->length + 1, 256, ctx, 16), 256, 0, ctx->stream>>>(o->as<int32_t>() + pos, (const int32_t *)p->offsets, p->length + 1, firsts[k], bpos);
            note_launch(ctx);
            if (lasts[k] > firsts[k])
                DF_CUDA(cudaMemcpyAsync(d->as<uint8_t>() + bpos, (const uint8_t *)p->values + firsts[k], (size_t)(lasts[k] - firsts[k]), cudaMemcpyDeviceToDevice, ctx->stream));
            pos += p->length;
            bpos += lasts[k] - firsts[k];
        }
        c->offsets = o->ptr;
        c->values = d->ptr;
        c->data_bytes = bytes;
    }
    return c;
}

Batch *batch_concat(Ctx *ctx, Batch *const *batches, int n) {
    if (n <= 0) fail(DFGPU_INTERNAL, "concat of zero batches");
    std::unique_ptr<Batch> o(new Batch());
    o->ctx = ctx;
    size_t nc = batches[0]->cols.size();
    int64_t total = 0;
    for (int i = 0; i < n; i++) {
        if (batches[i]->cols.size() != nc) fail(DFGPU_INTERNAL, "concat: column count mismatch");
        total += batches[i]->num_rows;
    }
    o->num_rows = total;
    for (size_t c = 0; c < nc; c++) {
        std::vector<ColRef> parts;
        for (int i = 0; i < n; i++) parts.push_back(batches[i]->cols[c]);
        if (n == 1) o->cols.push_back(parts[0]);
        else o->cols.push_back(col_concat(ctx, parts, total));
    }
    return o.release();
}

}  // namespace dfgpu


// ------------------------------------------------------------------------------------------------
// FilterExec: batch_filter -> arrow `filter_record_batch` (physical-plan/src/filter.rs:315-327): keep the rows whose
// predicate value is true and not NULL, in input order.  Two passes over the predicate BITMAP (1 bit per row): block
// counts, a scan of the (n / 4096) block counts, then every block writes the row numbers of its survivors at its
// offset (warp ballots give the rank inside the block, so the order is the input order).  The survivors' columns are
// then gathered with the take kernels.
// ------------------------------------------------------------------------------------------------
namespace dfgpu {
static const int FLT_T = 256, FLT_R = 16;  // rows of one block = FLT_T * FLT_R
DF_D bool pred_true(const ColView &m, int64_t i) { return col_valid(m, i) && bit_get((const uint8_t *)m.values, m.dbit + i); }

__global__ void __launch_bounds__(FLT_T) k_filter_count(ColView m, int64_t n, int64_t *__restrict__ block_counts) {
    const int64_t base = (int64_t)blockIdx.x * FLT_T * FLT_R;
    int c = 0;
#pragma unroll 4
    for (int k = 0; k < FLT_R; k++) {
        int64_t row = base + (int64_t)k * FLT_T + threadIdx.x;
        c += (row < n && pred_true(m, row)) ? 1 : 0;
    }
    __shared__ int wsum[FLT_T / 32];
#pragma unroll
    for (int o = 16; o; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
    if ((threadIdx.x & 31) == 0) wsum[threadIdx.x >> 5] = c;
    __syncthreads();
    if (threadIdx.x == 0) {
        int t = 0;
        for (int w = 0; w < FLT_T / 32; w++) t += wsum[w];
        block_counts[blockIdx.x] = t;
    }
}
__global__ void __launch_bounds__(FLT_T) k_filter_indices(ColView m, int64_t n, const int64_t *__restrict__ block_offs, int64_t *__restrict__ idx) {
    const int64_t base = (int64_t)blockIdx.x * FLT_T * FLT_R;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    __shared__ int wcnt[FLT_R][FLT_T / 32];
    bool keep[FLT_R];
    unsigned bal[FLT_R];
#pragma unroll
    for (int k = 0; k < FLT_R; k++) {
        int64_t row = base + (int64_t)k * FLT_T + threadIdx.x;
        keep[k] = row < n && pred_true(m, row);
        bal[k] = __ballot_sync(0xffffffffu, keep[k]);
        if (lane == 0) wcnt[k][warp] = __popc(bal[k]);
    }
    __syncthreads();
    // rank of (k, warp) = survivors of all earlier (k', warp') in row order: k major, warp minor
    int64_t out = block_offs[blockIdx.x];
#pragma unroll
    for (int k = 0; k < FLT_R; k++) {
        int before = 0, total = 0;
#pragma unroll
        for (int w = 0; w < FLT_T / 32; w++) {
            int c = wcnt[k][w];
            before += w < warp ? c : 0;
            total += c;
        }
        if (keep[k]) idx[out + before + __popc(bal[k] & ((1u << lane) - 1))] = base + (int64_t)k * FLT_T + threadIdx.x;
        out += total;
    }
}

Batch *batch_filter(Ctx *ctx, const Batch *b, int predicate_col) {
    if (predicate_col < 0 || predicate_col >= (int)b->cols.size()) fail(DFGPU_PLAN, "filter: predicate column index out of range");
    const DevCol &p = *b->cols[(size_t)predicate_col];
    if (p.type.id != T_BOOL) fail(DFGPU_PLAN, "filter: predicate column must be Boolean");  // as_boolean_array (filter.rs:323)
    const int64_t n = b->num_rows;
    cudaStream_t st = ctx->stream;
    std::unique_ptr<Batch> out(new Batch());
    out->ctx = ctx;
    int64_t total = 0;
    BufRef idx;
    if (n > 0) {
        const int64_t nblk = (n + FLT_T * FLT_R - 1) / (FLT_T * FLT_R);
        BufRef counts = make_buf(ctx, (size_t)nblk * 8 + 16), offs = make_buf(ctx, (size_t)nblk * 8 + 16);
        {
            KTimer kt(ctx, "k_filter_count", n / 8);
            k_filter_count<<<(unsigned)nblk, FLT_T, 0, st>>>(p.view(), n, counts->as<int64_t>());
        }
        exclusive_scan_i64(ctx, counts->as<int64_t>(), offs->as<int64_t>(), nblk, &total);
        idx = make_buf(ctx, (size_t)total * 8 + 16);
        if (total) {
            KTimer kt(ctx, "k_filter_indices", n / 8 + total * 8);
            k_filter_indices<<<(unsigned)nblk, FLT_T, 0, st>>>(p.view(), n, offs->as<int64_t>(), idx->as<int64_t>());
        }
        note_launch(ctx, 2);
        DF_CUDA(cudaGetLastError());
    } else {
        idx = make_buf(ctx, 16);
    }
    out->num_rows = total;
    std::vector<const DevCol *> srcs;
    std::vector<const int64_t *> idxs;
    std::vector<bool> maybe_null;
    for (auto &c : b->cols) {
        srcs.push_back(c.get());
        idxs.push_back(idx->as<int64_t>());
        maybe_null.push_back(false);
    }
    KTimer kt(ctx, "stage:filter_take");
    out->cols = cols_take(ctx, srcs, idxs, total, maybe_null);
    for (size_t i = 0; i < out->cols.size(); i++) out->cols[i]->type = b->cols[i]->type;
    return out.release();
}
}  // namespace dfgpu


// ------------------------------------------------------------------------------------------------
// ProjectionExec arithmetic over Decimal128 (dfgpu_batch_eval_decimal, include/dfgpu.h): one thread per row walks the
// post-order node list; the result types and the 10^k rescale factors of + / - are resolved on the host.
// ------------------------------------------------------------------------------------------------
namespace dfgpu {
static const int EX_MAX_NODES = 16;
struct ExNode {
    int32_t kind, left, right, pad;
    ColView col;                 // COLUMN
    uint64_t lit_lo, lit_hi;     // LITERAL
    uint64_t lmul_lo, lmul_hi, rmul_lo, rmul_hi;  // ADD / SUB: 10^(scale - s_left), 10^(scale - s_right)
};
struct ExPlan {
    int32_t n, pad;
    ExNode nodes[EX_MAX_NODES];
};
DF_D bool mul128_checked(__int128 a, __int128 b, __int128 &r) {
    if (a == 0 || b == 0) { r = 0; return true; }
    const bool neg = (a < 0) != (b < 0);
    const unsigned __int128 ua = a < 0 ? (unsigned __int128)(-(a + 1)) + 1 : (unsigned __int128)a;
    const unsigned __int128 ub = b < 0 ? (unsigned __int128)(-(b + 1)) + 1 : (unsigned __int128)b;
    const unsigned __int128 lim = neg ? ((unsigned __int128)1 << 127) : (((unsigned __int128)1 << 127) - 1);
    unsigned __int128 m;
    if (((uint64_t)(ua >> 63) | (uint64_t)(ub >> 63)) == 0) {
        m = (unsigned __int128)(uint64_t)ua * (uint64_t)ub;  // both magnitudes below 2^63: the product is below 2^126 (the common case; no division)
    } else {
        if (ua > lim / ub) return false;
        m = ua * ub;
    }
    r = neg ? (__int128)(~m + 1) : (__int128)m;
    return true;
}
DF_D bool add128_checked(__int128 a, __int128 b, __int128 &r) {
    const unsigned __int128 s = (unsigned __int128)a + (unsigned __int128)b;
    r = (__int128)s;
    return !((a >= 0) == (b >= 0) && (r >= 0) != (a >= 0));
}
__global__ void __launch_bounds__(256) k_eval_decimal(ExPlan pl, int64_t n, uint64_t *__restrict__ out, uint32_t *__restrict__ valid_words,
                                                      uint32_t *__restrict__ err) {
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    const int64_t nround = (n + 31) / 32 * 32;
    for (; i < nround; i += (int64_t)gridDim.x * blockDim.x) {
        __int128 v[EX_MAX_NODES];
        bool ok[EX_MAX_NODES];
        bool valid = false;
        if (i < n) {
            for (int k = 0; k < pl.n; k++) {
                const ExNode &nd = pl.nodes[k];
                if (nd.kind == DFGPU_EXPR_COLUMN) {
                    ok[k] = col_valid(nd.col, i);
                    const uint64_t *p = (const uint64_t *)nd.col.values + 2 * i;
                    v[k] = ok[k] ? (__int128)(((unsigned __int128)p[1] << 64) | p[0]) : 0;
                } else if (nd.kind == DFGPU_EXPR_LITERAL) {
                    ok[k] = true;
                    v[k] = (__int128)(((unsigned __int128)nd.lit_hi << 64) | nd.lit_lo);
                } else {
                    ok[k] = ok[nd.left] && ok[nd.right];
                    v[k] = 0;
                    if (ok[k]) {
                        bool fine;
                        if (nd.kind == DFGPU_EXPR_MUL) {
                            fine = mul128_checked(v[nd.left], v[nd.right], v[k]);
                        } else {
                            __int128 a, b;
                            fine = mul128_checked(v[nd.left], (__int128)(((unsigned __int128)nd.lmul_hi << 64) | nd.lmul_lo), a) &&
                                   mul128_checked(v[nd.right], (__int128)(((unsigned __int128)nd.rmul_hi << 64) | nd.rmul_lo), b);
                            if (fine && nd.kind == DFGPU_EXPR_SUB) {
                                if (b == (__int128)((unsigned __int128)1 << 127)) fine = add128_checked(a, 1, a) && add128_checked(a, -(b + 1), v[k]);  // a - MIN
                                else fine = add128_checked(a, -b, v[k]);
                            } else if (fine) {
                                fine = add128_checked(a, b, v[k]);
                            }
                        }
                        if (!fine) { atomicExch(err, 1u); v[k] = 0; }
                    }
                }
            }
            valid = ok[pl.n - 1];
            const unsigned __int128 r = (unsigned __int128)v[pl.n - 1];
            out[2 * i] = (uint64_t)r;
            out[2 * i + 1] = (uint64_t)(r >> 64);
        }
        const uint32_t m = __ballot_sync(0xffffffffu, valid);
        if ((threadIdx.x & 31) == 0) valid_words[i >> 5] = m;
    }
}

Batch *batch_eval_decimal(Ctx *ctx, const Batch *b, const dfgpu_expr_node *nodes, int n_nodes, const char *name) {
    if (!nodes || n_nodes <= 0) fail(DFGPU_PLAN, "eval: empty expression");
    if (n_nodes > EX_MAX_NODES) fail(DFGPU_NOT_IMPLEMENTED, "This feature is not implemented: expressions with more than 16 nodes");
    ExPlan pl;
    memset(&pl, 0, sizeof pl);
    pl.n = n_nodes;
    int prec[EX_MAX_NODES], scale[EX_MAX_NODES];
    auto pow10 = [](int e) { unsigned __int128 r = 1; for (int k = 0; k < e; k++) r *= 10; return r; };
    bool any_nullable = false;
    for (int k = 0; k < n_nodes; k++) {
        const dfgpu_expr_node &s = nodes[k];
        ExNode &d = pl.nodes[k];
        d.kind = s.kind;
        if (s.kind == DFGPU_EXPR_COLUMN) {
            if (s.column < 0 || s.column >= (int)b->cols.size()) fail(DFGPU_PLAN, "eval: column index out of range");
            const DevCol &c = *b->cols[(size_t)s.column];
            if (c.type.id != T_DEC128) fail(DFGPU_NOT_IMPLEMENTED, "This feature is not implemented: decimal arithmetic over " + c.type.describe());
            d.col = c.view();
            prec[k] = c.type.precision; scale[k] = c.type.scale;
            any_nullable = any_nullable || c.validity != nullptr;
        } else if (s.kind == DFGPU_EXPR_LITERAL) {
            d.lit_lo = s.lit_lo; d.lit_hi = s.lit_hi;
            prec[k] = s.lit_precision; scale[k] = s.lit_scale;
            if (prec[k] < 1 || prec[k] > 38 || scale[k] < 0 || scale[k] > prec[k]) fail(DFGPU_PLAN, "eval: invalid Decimal128 literal type");
        } else if (s.kind == DFGPU_EXPR_ADD || s.kind == DFGPU_EXPR_SUB || s.kind == DFGPU_EXPR_MUL) {
            if (s.left < 0 || s.left >= k || s.right < 0 || s.right >= k) fail(DFGPU_PLAN, "eval: operands must precede their operator (post order)");
            d.left = s.left; d.right = s.right;
            const int p1 = prec[s.left], s1 = scale[s.left], p2 = prec[s.right], s2 = scale[s.right];
            if (s.kind == DFGPU_EXPR_MUL) {
                if (s1 + s2 > 38)
                    fail(DFGPU_EXECUTION, "Arrow error: Invalid argument error: Output scale of Decimal128(" + std::to_string(p1) + ", " + std::to_string(s1) +
                                              ") * Decimal128(" + std::to_string(p2) + ", " + std::to_string(s2) + ") would exceed max scale of 38");
                prec[k] = std::min(38, p1 + p2 + 1);
                scale[k] = s1 + s2;
            } else {
                scale[k] = std::max(s1, s2);
                prec[k] = std::min(38, scale[k] + std::max(p1 - s1, p2 - s2) + 1);
                unsigned __int128 lm = pow10(scale[k] - s1), rm = pow10(scale[k] - s2);
                d.lmul_lo = (uint64_t)lm; d.lmul_hi = (uint64_t)(lm >> 64);
                d.rmul_lo = (uint64_t)rm; d.rmul_hi = (uint64_t)(rm >> 64);
            }
        } else {
            fail(DFGPU_PLAN, "eval: unknown expression node kind");
        }
    }
    const int64_t n = b->num_rows;
    cudaStream_t st = ctx->stream;
    auto col = std::make_shared<DevCol>();
    col->type = make_type(T_DEC128, name ? name : "", true, prec[n_nodes - 1], scale[n_nodes - 1]);
    col->length = n;
    BufRef vals = make_buf(ctx, (size_t)n * 16 + 16), valid = make_buf(ctx, (size_t)((n + 31) / 32) * 4 + 16, true);
    BufRef err = make_buf(ctx, 4, true);
    col->owned.push_back(vals);
    col->owned.push_back(valid);
    col->values = vals->ptr;
    col->validity = valid->as<uint8_t>();
    col->null_count = any_nullable ? -1 : 0;
    if (n > 0) {
        int64_t abytes = n * 16;
        for (int k = 0; k < n_nodes; k++)
            if (nodes[k].kind == DFGPU_EXPR_COLUMN) abytes += n * 16;
        {
            KTimer kt(ctx, "k_eval_decimal", abytes);
            k_eval_decimal<<<grid_for(n, 256, ctx, 16), 256, 0, st>>>(pl, n, vals->as<uint64_t>(), valid->as<uint32_t>(), err->as<uint32_t>());
        }
        note_launch(ctx);
        DF_CUDA(cudaGetLastError());
        uint32_t e = 0;
        DF_CUDA(cudaMemcpyAsync(&e, err->ptr, 4, cudaMemcpyDeviceToHost, st));
        DF_CUDA(cudaStreamSynchronize(st));
        if (e) fail(DFGPU_EXECUTION, "Arrow error: Compute error: Overflow happened on a Decimal128 operation");
    }
    if (!any_nullable) { col->validity = nullptr; }
    std::unique_ptr<Batch> o(new Batch());
    o->ctx = ctx;
    o->num_rows = n;
    o->cols = b->cols;
    o->cols.push_back(col);
    return o.release();
}
}  // namespace dfgpu

#include "coalesce.inc"
