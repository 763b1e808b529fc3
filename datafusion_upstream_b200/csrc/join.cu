// join.cu — HashJoinExec on the device (physical-plan/src/joins/hash_join.rs, joins/utils.rs).
//   build  : collect_left_input (hash_join.rs:678-768) — concat the build batches, intern the
//            key of every build row to a dense key id (GroupTable), then lay the row ids of each
//            key out contiguously (count -> scan -> fill) instead of JoinHashMap's `next` chains
//            (utils.rs:121-229): span[key] = (start, count), rows[start .. start+count).
//   probe  : lookup_join_hashmap + equal_rows_arr (hash_join.rs:1024-1118) fused — the table is
//            keyed by the real key, so there are no false candidates to filter — as the two-pass
//            count -> exclusive scan -> scatter of (build_idx, probe_idx) pairs, with the
//            join-type index fix-ups of adjust_indices_by_join_type (utils.rs:1234-1279) folded
//            into the count/scatter, then build_batch_from_indices (utils.rs:1180-1230) = take.
//   final  : process_unmatched_build_batch / get_final_indices_from_bit_map
//            (hash_join.rs:1348-1388, utils.rs:1119-1141) from the visited flags.
#include <set>
#include <chrono>

#include "device.cuh"
#include "fastagg.h"
#include "groups.h"

struct dfgpu_join {
    dfgpu::CtxHold ctx;
    int join_type = 0, partition_mode = 0;
    bool null_equals_null = false;
    dfgpu::Schema build_schema, probe_schema, out_schema;
    std::vector<int> bkeys, pkeys;
    std::vector<std::unique_ptr<dfgpu_batch>> pending;  // build batches until build_finish
    std::unique_ptr<dfgpu_batch> build;
    std::unique_ptr<dfgpu::GroupTable> table;
    dfgpu::BufRef span, rows, visited, gid_scratch;
    dfgpu::BufRef jtags, jslots;  // probe-side index of the fused path (join_fast.inc)
    dfgpu::PackedRows brows;      // the build side's plain columns row-wise: one DRAM line per matched build row when the output is gathered
    uint64_t jgroups = 0;
    bool fast = false;
    int64_t build_rows = 0, num_keys = 0;
    bool finished = false;
    // residual JoinFilter (joins/utils.rs:1143-1176): the columns of the intermediate batch, and the candidate pairs of the
    // probe batch whose predicate the host is evaluating (between probe_candidates and probe_commit)
    bool has_filter = false;
    std::vector<std::pair<int, int>> filter_cols;  // (side: 0 build / 1 probe, column)
    struct Pending {
        bool valid = false;
        dfgpu::BufRef ob, op;
        int64_t total = 0, np = 0;
        std::vector<dfgpu::ColRef> probe_cols;
    } pend;
    dfgpu_metrics m;
};

namespace dfgpu {
typedef dfgpu_join JoinOp;

static bool need_final(int jt) {  // need_produce_result_in_final (utils.rs:1103-1108)
    return jt == DFGPU_JOIN_LEFT || jt == DFGPU_JOIN_LEFT_ANTI || jt == DFGPU_JOIN_LEFT_SEMI || jt == DFGPU_JOIN_FULL;
}

__global__ void k_build_count(TableView t, const uint32_t *__restrict__ gids, int64_t n, uint32_t *__restrict__ cnt) {
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    for (; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        uint32_t g = resolve_gid(t, gids[i]);
        if (g != GID_NONE) atomicAdd(&cnt[g], 1u);
    }
}
__global__ void k_fill_none(uint32_t *p, int64_t n) {
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    for (; i < n; i += (int64_t)gridDim.x * blockDim.x) p[i] = GID_NONE;
}
__global__ void k_u32_to_i64(const uint32_t *in, int64_t n, int64_t *out) {
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    for (; i < n; i += (int64_t)gridDim.x * blockDim.x) out[i] = in[i];
}
__global__ void k_make_span(const int64_t *__restrict__ starts, const uint32_t *__restrict__ cnt, int64_t G, uint64_t *__restrict__ span,
                            uint32_t *__restrict__ cursor) {
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    for (; i < G; i += (int64_t)gridDim.x * blockDim.x) {
        span[i] = ((uint64_t)starts[i] << 32) | cnt[i];
        cursor[i] = 0;
    }
}
__global__ void k_build_fill(TableView t, const uint32_t *__restrict__ gids, int64_t n, const uint64_t *__restrict__ span,
                             uint32_t *__restrict__ cursor, uint32_t *__restrict__ rows) {
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    for (; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        uint32_t g = resolve_gid(t, gids[i]);
        if (g == GID_NONE) continue;
        uint32_t pos = atomicAdd(&cursor[g], 1u);
        rows[(span[g] >> 32) + pos] = (uint32_t)i;
    }
}

// pass 1: output rows per probe row for this join type
__global__ void k_probe_count(const uint32_t *__restrict__ gids, int64_t n, const uint64_t *__restrict__ span, int jt,
                              int64_t *__restrict__ counts) {
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    for (; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        uint32_t g = gids[i];
        int64_t c = g == GID_NONE ? 0 : (int64_t)(uint32_t)span[g];
        int64_t o;
        switch (jt) {
        case DFGPU_JOIN_INNER: case DFGPU_JOIN_LEFT: o = c; break;
        case DFGPU_JOIN_RIGHT: case DFGPU_JOIN_FULL: o = c ? c : 1; break;  // + unmatched probe rows (utils.rs:1248-1254)
        case DFGPU_JOIN_RIGHT_SEMI: o = c ? 1 : 0; break;                   // get_semi_indices
        case DFGPU_JOIN_RIGHT_ANTI: o = c ? 0 : 1; break;                   // get_anti_indices
        default: o = 0; break;                                              // LeftSemi / LeftAnti: produced at the end
        }
        counts[i] = o;
    }
}
// pass 2: scatter (build_idx, probe_idx) pairs; mark visited build rows (hash_join.rs:1274-1278)
__global__ void k_probe_scatter(const uint32_t *__restrict__ gids, int64_t n, const uint64_t *__restrict__ span,
                                const uint32_t *__restrict__ rows, const int64_t *__restrict__ offs, int jt,
                                int64_t *__restrict__ out_build, int64_t *__restrict__ out_probe, uint8_t *__restrict__ visited) {
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    for (; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        uint32_t g = gids[i];
        uint64_t sp = g == GID_NONE ? 0 : span[g];
        uint32_t c = (uint32_t)sp;
        uint64_t st = sp >> 32;
        int64_t o = offs[i];
        switch (jt) {
        case DFGPU_JOIN_INNER: case DFGPU_JOIN_LEFT: case DFGPU_JOIN_RIGHT: case DFGPU_JOIN_FULL:
            for (uint32_t k = 0; k < c; k++) {
                uint32_t r = rows[st + k];
                out_build[o + k] = r;
                out_probe[o + k] = i;
                if (visited) visited[r] = 1;
            }
            if (c == 0 && (jt == DFGPU_JOIN_RIGHT || jt == DFGPU_JOIN_FULL)) { out_build[o] = -1; out_probe[o] = i; }
            break;
        case DFGPU_JOIN_RIGHT_SEMI: if (c) out_probe[o] = i; break;
        case DFGPU_JOIN_RIGHT_ANTI: if (!c) out_probe[o] = i; break;
        default:  // LeftSemi / LeftAnti: only the visited flags
            for (uint32_t k = 0; k < c; k++) visited[rows[st + k]] = 1;
            break;
        }
    }
}
__global__ void k_flags_to_counts(const uint8_t *__restrict__ flags, int64_t n, int want, int64_t *__restrict__ out) {
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    for (; i < n; i += (int64_t)gridDim.x * blockDim.x) out[i] = (flags[i] != 0) == (want != 0) ? 1 : 0;
}
__global__ void k_compact(const int64_t *__restrict__ counts, const int64_t *__restrict__ offs, int64_t n, int64_t *__restrict__ out) {
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    for (; i < n; i += (int64_t)gridDim.x * blockDim.x)
        if (counts[i]) out[offs[i]] = i;
}

#include "join_fast.inc"

static TypeInfo forced_nullable(TypeInfo t) {
    t.nullable = true;
    return t;
}

JoinOp *join_create(Ctx *ctx, int jt, int pmode, bool null_eq, const Schema &bs, const Schema &ps, const int32_t *bk, const int32_t *pk, int nk) {
    std::unique_ptr<JoinOp> j(new JoinOp());
    j->ctx = ctx;
    j->join_type = jt;
    j->partition_mode = pmode;
    j->null_equals_null = null_eq;
    j->build_schema = bs;
    j->probe_schema = ps;
    memset(&j->m, 0, sizeof j->m);
    if (jt < 0 || jt > DFGPU_JOIN_RIGHT_ANTI) fail(DFGPU_PLAN, "unknown JoinType");
    if (nk <= 0) fail(DFGPU_PLAN, "On constraints in HashJoinExec should be non-empty");  // hash_join.rs:320-322
    std::vector<TypeInfo> key_types;
    for (int i = 0; i < nk; i++) {
        if (bk[i] < 0 || bk[i] >= (int)bs.fields.size() || pk[i] < 0 || pk[i] >= (int)ps.fields.size())
            fail(DFGPU_PLAN, "join key column index out of range");
        const TypeInfo &l = bs.fields[bk[i]].value_type(), &r = ps.fields[pk[i]].value_type();
        bool both_str = (l.id == T_UTF8 || l.id == T_LARGE_UTF8) && (r.id == T_UTF8 || r.id == T_LARGE_UTF8);
        bool both_bin = (l.id == T_BINARY || l.id == T_LARGE_BINARY) && (r.id == T_BINARY || r.id == T_LARGE_BINARY);
        if (l.id != r.id && !both_str && !both_bin)
            fail(DFGPU_PLAN, "join key types differ: " + l.describe() + " vs " + r.describe());  // check_join_is_valid
        j->bkeys.push_back(bk[i]);
        j->pkeys.push_back(pk[i]);
        // the table is declared with plain value types: dictionary keys are interned by value
        TypeInfo kt = l;
        if (bs.fields[bk[i]].id == T_DICT || ps.fields[pk[i]].id == T_DICT) {
            // force the packed (multi-column style) path so both encodings meet in one string table
            kt = l;
        }
        key_types.push_back(kt);
    }
    j->table.reset(new GroupTable(ctx, key_types));
    // build_join_schema (utils.rs:657-729)
    bool left_null = jt == DFGPU_JOIN_RIGHT || jt == DFGPU_JOIN_FULL;
    bool right_null = jt == DFGPU_JOIN_LEFT || jt == DFGPU_JOIN_FULL;
    bool left_out = jt != DFGPU_JOIN_RIGHT_SEMI && jt != DFGPU_JOIN_RIGHT_ANTI;
    bool right_out = jt != DFGPU_JOIN_LEFT_SEMI && jt != DFGPU_JOIN_LEFT_ANTI;
    if (left_out) for (auto &f : bs.fields) j->out_schema.fields.push_back(left_null ? forced_nullable(f) : f);
    if (right_out) for (auto &f : ps.fields) j->out_schema.fields.push_back(right_null ? forced_nullable(f) : f);
    return j.release();
}

void join_build_push(JoinOp &j, const Batch &b) {
    if (j.finished) fail(DFGPU_INTERNAL, "build side already finished");
    if (b.cols.size() != j.build_schema.fields.size()) fail(DFGPU_INTERNAL, "build batch does not match the build schema");
    std::unique_ptr<Batch> copy(new Batch(b));  // shares column buffers
    j.m.build_input_batches += 1;
    j.m.build_input_rows += b.num_rows;
    j.pending.push_back(std::move(copy));
}

// key columns for interning: dictionary-encoded keys are decoded so both sides present plain values
static std::vector<ColRef> key_columns(Ctx *ctx, const Batch &b, const std::vector<int> &idx) {
    std::vector<ColRef> out;
    for (int c : idx) {
        ColRef col = b.cols[c];
        if (col->type.id == T_DICT) col = col_decode_dictionary(ctx, *col);
        out.push_back(col);
    }
    return out;
}

static void intern_chunks(JoinOp &j, const std::vector<ColRef> &keys, int64_t n, uint32_t *gids, bool find_only) {
    Ctx *ctx = j.ctx;
    std::vector<const DevCol *> kp;
    for (auto &k : keys) kp.push_back(k.get());
    int64_t chunk = ctx->opts.chunk_rows > 0 ? ctx->opts.chunk_rows : (4LL << 20);
    NullPolicy np = j.null_equals_null ? NULLS_ARE_KEYS : NULLS_EXCLUDED;  // eq vs not_distinct (hash_join.rs:1067-1076)
    for (int64_t off = 0; off < n; off += chunk) {
        int64_t m = std::min(chunk, n - off);
        j.table->intern(kp, off, m, gids + off, find_only, np);
    }
}

// The fused path applies when the key image fits 64 bits and no key column needs a pre-pass
// (per-column string ids).  DFGPU_DISABLE_JOIN_INDEX=1 keeps the general path.
static bool fast_eligible(const JoinOp &j, int64_t nb) {
    const bool off = getenv("DFGPU_DISABLE_JOIN_INDEX") && atoi(getenv("DFGPU_DISABLE_JOIN_INDEX")) != 0;
    if (off || j.ctx->opts.force_hash_collisions) return false;
    if (j.table->mode() != KM_U64_DIRECT && j.table->mode() != KM_U64_PACKED) return false;
    auto lay = j.table->key_layout();
    if ((int)lay.size() > FA_MAX_KEYCOLS || nb > (1LL << 30)) return false;
    for (size_t i = 0; i < lay.size(); i++)
        if (lay[i].kind == 2) return false;
    return true;
}

static JKeys key_plan(const JoinOp &j, const std::vector<ColRef> &keys) {
    JKeys jk;
    memset(&jk, 0, sizeof jk);
    auto lay = j.table->key_layout();
    jk.nkeys = (int)lay.size();
    jk.direct = j.table->mode() == KM_U64_DIRECT;
    jk.null_off = j.table->null_off();
    jk.nulls_excluded = j.null_equals_null ? 0 : 1;
    for (int k = 0; k < jk.nkeys; k++) {
        jk.keys[k].v = keys[k]->view();
        jk.keys[k].kind = lay[k].kind; jk.keys[k].off = lay[k].off; jk.keys[k].w = lay[k].w; jk.keys[k].null_bit = lay[k].null_bit;
    }
    return jk;
}

// direct build of the probe index from the build rows (join_fast.inc): partition by table block, one CTA per block.
// false: a block overflowed (keys that collide on the index hash) - the caller builds the general path instead.
static bool build_index(JoinOp &j) {
    Ctx *ctx = j.ctx;
    cudaStream_t st = ctx->stream;
    const int64_t nb = j.build_rows;
    const uint64_t want = std::max<uint64_t>(4, ((uint64_t)nb * 2 + JG - 1) / JG);  // load factor <= 1/2
    j.jgroups = (want + JB_G - 1) / JB_G * JB_G;
    const uint32_t nblocks = (uint32_t)(j.jgroups / JB_G), nbins = nblocks + 2;
    const uint64_t nslots = j.jgroups * JG + 2;
    j.jtags = make_buf(ctx, j.jgroups * JG + 16);
    j.jslots = make_buf(ctx, nslots * 16 + 16);
    j.rows = make_buf(ctx, (size_t)nb * 4 + 16);
    BufRef hist = make_buf(ctx, (size_t)nbins * 4 + 16, true), offs = make_buf(ctx, (size_t)(nbins + 1) * 4 + 16), cursors = make_buf(ctx, (size_t)nbins * 4 + 16);
    BufRef recs = make_buf(ctx, (size_t)nb * 16 + 16);
    DevBuf flags(ctx, 16, true);  // [0] = some key repeats, [1] = a block overflowed, [2..3] = rows[] cursor
    {
        static std::mutex mu;
        static std::set<int> done;
        std::lock_guard<std::mutex> lock(mu);
        if (done.insert(ctx->device).second) DF_CUDA(cudaFuncSetAttribute(k_jbp_build, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)jbp_build_smem()));
    }
    std::vector<ColRef> keys = key_columns(ctx, *j.build, j.bkeys);
    JKeys jk = key_plan(j, keys);
    const int64_t ntiles = (nb + JP_TILE - 1) / JP_TILE;
    const bool sm = nbins <= JB_SMEM_BINS;   // the bins fit shared memory: per-CTA counts, one global atomic per CTA and bin
    // (persistent CTAs: as many as shared memory lets an SM hold - the scatter is latency-bound, ncu: 5 % issue-slot use at 4 per SM)
    const int per_sm = sm ? std::max(1, std::min(8, (int)(((size_t)200 << 10) / ((size_t)nbins * 4 + 1024)))) : 8;
    const unsigned nt = (unsigned)std::min<int64_t>(ntiles, sm ? (int64_t)ctx->sm_count * per_sm : ntiles);
    const size_t smem = sm ? (size_t)nbins * 4 : 0;
    BufRef cta_cnt = make_buf(ctx, sm ? (size_t)nt * nbins * 4 + 16 : 16);
    {
        KTimer kt(ctx, "k_jb_partition", nb * 8);
        uint32_t *h = hist->as<uint32_t>(), *cc = cta_cnt->as<uint32_t>(), *of = offs->as<uint32_t>(), *cu = cursors->as<uint32_t>();
        const uint32_t ng = (uint32_t)j.jgroups;
        if (jk.direct && sm) k_jbp_hist<true, true><<<nt, JP_T, smem, st>>>(jk, nb, ng, nbins, h, cc);
        else if (jk.direct) k_jbp_hist<true, false><<<nt, JP_T, 0, st>>>(jk, nb, ng, nbins, h, cc);
        else if (sm) k_jbp_hist<false, true><<<nt, JP_T, smem, st>>>(jk, nb, ng, nbins, h, cc);
        else k_jbp_hist<false, false><<<nt, JP_T, 0, st>>>(jk, nb, ng, nbins, h, cc);
        k_jbp_scan<<<1, 1024, 0, st>>>(h, nbins, of, cu);
        if (jk.direct && sm) k_jbp_scatter<true, true><<<nt, JP_T, smem, st>>>(jk, nb, ng, nbins, of, cu, cc, recs->as<ulonglong2>());
        else if (jk.direct) k_jbp_scatter<true, false><<<nt, JP_T, 0, st>>>(jk, nb, ng, nbins, of, cu, cc, recs->as<ulonglong2>());
        else if (sm) k_jbp_scatter<false, true><<<nt, JP_T, smem, st>>>(jk, nb, ng, nbins, of, cu, cc, recs->as<ulonglong2>());
        else k_jbp_scatter<false, false><<<nt, JP_T, 0, st>>>(jk, nb, ng, nbins, of, cu, cc, recs->as<ulonglong2>());
    }
    {
        KTimer kt(ctx, "k_jb_build", nb * 16);
        k_jbp_build<<<nbins, JB_T, jbp_build_smem(), st>>>(recs->as<ulonglong2>(), offs->as<uint32_t>(), (uint32_t)j.jgroups, j.jtags->as<uint32_t>(), j.jslots->as<ulonglong2>(),
                                                          j.rows->as<uint32_t>(), flags.as<uint32_t>());
    }
    note_launch(ctx, 4);
    uint32_t hf[2] = {0, 0};
    DF_CUDA(cudaMemcpyAsync(hf, flags.ptr, 8, cudaMemcpyDeviceToHost, st));
    DF_CUDA(cudaStreamSynchronize(st));
    DF_CUDA(cudaGetLastError());
    if (hf[1]) {
        j.jtags.reset();
        j.jslots.reset();
        j.rows.reset();
        return false;
    }
    return true;
}

// fused probe: returns the (build_idx, probe_idx) pairs and their number
static int64_t probe_pairs_fast(JoinOp &j, const Batch &probe, BufRef &ob, BufRef &op, int jt_use, bool mark_visited) {
    Ctx *ctx = j.ctx;
    cudaStream_t st = ctx->stream;
    const int64_t np = probe.num_rows;
    std::vector<ColRef> keys = key_columns(ctx, probe, j.pkeys);  // dictionary keys are decoded to plain values
    JKeys jk = key_plan(j, keys);
    JIndexView ix{j.jtags->as<uint4>(), j.jslots->as<ulonglong2>(), j.jgroups};
    const int64_t nt = (np + JP_TILE - 1) / JP_TILE;
    // per-tile emit lists live at the tile's own row range, so they need no offsets of their own
    BufRef lsp = make_buf(ctx, (size_t)nt * JP_TILE * 8 + 16), lrow = make_buf(ctx, (size_t)nt * JP_TILE * 2 + 16);
    const int64_t nsup = (nt + JP_SUPER - 1) / JP_SUPER;
    BufRef tot = make_buf(ctx, (size_t)nt * 8 + 16), tne = make_buf(ctx, (size_t)nt * 4 + 16), sup = make_buf(ctx, (size_t)nsup * 8 + 16, true);
    {
        KTimer kt(ctx, "k_jp_count", np * 8);
        if (jk.direct)
            k_jp_count<true><<<(unsigned)nt, JP_T, 0, st>>>(jk, ix, np, jt_use, lsp->as<uint64_t>(), lrow->as<uint16_t>(), tne->as<uint32_t>(),
                                                            tot->as<unsigned long long>(), sup->as<unsigned long long>());
        else
            k_jp_count<false><<<(unsigned)nt, JP_T, 0, st>>>(jk, ix, np, jt_use, lsp->as<uint64_t>(), lrow->as<uint16_t>(), tne->as<uint32_t>(),
                                                             tot->as<unsigned long long>(), sup->as<unsigned long long>());
    }
    // the output size = sum of the super-tile sums (a few KB): summed on the host
    std::vector<unsigned long long> hsup((size_t)nsup);
    DF_CUDA(cudaMemcpyAsync(hsup.data(), sup->ptr, (size_t)nsup * 8, cudaMemcpyDeviceToHost, st));
    DF_CUDA(cudaStreamSynchronize(st));
    int64_t total = 0;
    for (auto v : hsup) total += (int64_t)v;
    ob = make_buf(ctx, (size_t)total * 8 + 16);
    op = make_buf(ctx, (size_t)total * 8 + 16);
    {
        KTimer kt(ctx, "k_jp_emit", total * 28);
        k_jp_emit<<<(unsigned)nt, JE_T, 0, st>>>(lsp->as<uint64_t>(), lrow->as<uint16_t>(), tne->as<uint32_t>(), tot->as<unsigned long long>(),
                                                 sup->as<unsigned long long>(), jt_use,
                                                 j.rows->as<uint32_t>(), ob->as<int64_t>(), op->as<int64_t>(),
                                                 (mark_visited && j.visited) ? j.visited->as<uint8_t>() : nullptr);
    }
    note_launch(ctx, 2);
    DF_CUDA(cudaGetLastError());
    return total;
}

void join_build_finish(JoinOp &j) {
    Ctx *ctx = j.ctx;
    cudaStream_t st = ctx->stream;
    auto t0 = std::chrono::steady_clock::now();
    if (j.finished) return;
    std::vector<Batch *> bs;
    for (auto &p : j.pending) bs.push_back(p.get());
    if (bs.empty()) {
        j.build.reset(new Batch());
        j.build->ctx = ctx;
        for (auto &f : j.build_schema.fields) j.build->cols.push_back(col_nulls(ctx, f, 0));
    } else {
        j.build.reset(batch_concat(ctx, bs.data(), (int)bs.size()));  // concat_batches (hash_join.rs:764)
    }
    j.pending.clear();
    int64_t nb = j.build->num_rows;
    if (nb >= (1LL << 32) - 1) fail(DFGPU_RESOURCES_EXHAUSTED, "Resources exhausted: build side larger than 2^32 rows");
    j.build_rows = nb;
    j.fast = nb > 0 && fast_eligible(j, nb) && build_index(j);
    if (j.fast) {
        j.num_keys = 1;
    } else if (nb) {
        BufRef gids = make_buf(ctx, (size_t)nb * 4 + 16);
        std::vector<ColRef> keys = key_columns(ctx, *j.build, j.bkeys);
        intern_chunks(j, keys, nb, gids->as<uint32_t>(), false);
        int64_t G = j.table->num_groups_sync();
        j.num_keys = G;
        TableView tv = j.table->view();
        BufRef cnt = make_buf(ctx, (size_t)G * 4 + 16, true), cnt64 = make_buf(ctx, (size_t)G * 8 + 16), starts = make_buf(ctx, (size_t)G * 8 + 16);
        j.span = make_buf(ctx, (size_t)G * 8 + 16);
        j.rows = make_buf(ctx, (size_t)nb * 4 + 16);
        int grid = grid_for(nb, 256, ctx, 32);
        k_build_count<<<grid, 256, 0, st>>>(tv, gids->as<uint32_t>(), nb, cnt->as<uint32_t>());
        if (G) {
            k_u32_to_i64<<<grid_for(G, 256, ctx, 32), 256, 0, st>>>(cnt->as<uint32_t>(), G, cnt64->as<int64_t>());
            exclusive_scan_i64(ctx, cnt64->as<int64_t>(), starts->as<int64_t>(), G, nullptr);
            // cnt64 is reused as the 32-bit fill cursor
            k_make_span<<<grid_for(G, 256, ctx, 32), 256, 0, st>>>(starts->as<int64_t>(), cnt->as<uint32_t>(), G, j.span->as<uint64_t>(), cnt64->as<uint32_t>());
            k_build_fill<<<grid, 256, 0, st>>>(tv, gids->as<uint32_t>(), nb, j.span->as<uint64_t>(), cnt64->as<uint32_t>(), j.rows->as<uint32_t>());
            note_launch(ctx, 3);
        }
        note_launch(ctx);
        DF_CUDA(cudaGetLastError());
    }
    if (need_final(j.join_type)) j.visited = make_buf(ctx, (size_t)nb + 16, true);
    // build columns that reach the output: row-wise copy for the gathers of the probe phase (not worth it for tiny build sides,
    // which stay in L2, nor for joins that never emit build columns)
    const int jt0 = j.join_type;
    if (nb >= (1 << 20) && jt0 != DFGPU_JOIN_RIGHT_SEMI && jt0 != DFGPU_JOIN_RIGHT_ANTI && jt0 != DFGPU_JOIN_LEFT_SEMI && jt0 != DFGPU_JOIN_LEFT_ANTI &&
        jt0 != DFGPU_JOIN_RIGHT && jt0 != DFGPU_JOIN_FULL)
        j.brows = pack_rows(ctx, j.build->cols, nb);
    j.finished = true;
    j.m.build_mem_used = (int64_t)(j.table->device_bytes() + (j.span ? j.span->bytes : 0) + (j.rows ? j.rows->bytes : 0) +
                                   (j.jslots ? j.jslots->bytes + j.jtags->bytes : 0) + (j.brows.rows ? j.brows.rows->bytes : 0));
    DF_CUDA(cudaStreamSynchronize(st));
    j.m.build_time_ns += std::chrono::duration_cast<std::chrono::nanoseconds>(std::chrono::steady_clock::now() - t0).count();
}

static Batch *empty_output(JoinOp &j) {
    std::unique_ptr<Batch> o(new Batch());
    o->ctx = j.ctx;
    for (auto &f : j.out_schema.fields) o->cols.push_back(col_nulls(j.ctx, f, 0));
    return o.release();
}

// general probe: intern the probe keys against the GroupTable (any key shape), then count -> scan -> scatter
static int64_t probe_pairs_general(JoinOp &j, const Batch &probe, BufRef &ob, BufRef &op, int jt_use, bool mark_visited) {
    Ctx *ctx = j.ctx;
    cudaStream_t st = ctx->stream;
    const int64_t np = probe.num_rows;
    const int jt = jt_use;
    BufRef gids = make_buf(ctx, (size_t)np * 4 + 16);
    int grid = grid_for(np, 256, ctx, 32);
    bool have_table = j.build_rows > 0 && j.num_keys > 0;
    if (have_table) {
        std::vector<ColRef> keys = key_columns(ctx, probe, j.pkeys);
        intern_chunks(j, keys, np, gids->as<uint32_t>(), true);
    } else {  // empty build side: every lookup misses (tests join_left_empty_right etc., hash_join.rs:2069-2140)
        k_fill_none<<<grid, 256, 0, st>>>(gids->as<uint32_t>(), np);
        note_launch(ctx);
    }
    BufRef counts = make_buf(ctx, (size_t)np * 8 + 16), offs = make_buf(ctx, (size_t)np * 8 + 16);
    const uint64_t *span = have_table ? j.span->as<uint64_t>() : nullptr;
    {
        KTimer kt(ctx, "k_probe_count", np * 12);
        k_probe_count<<<grid, 256, 0, st>>>(gids->as<uint32_t>(), np, span, jt, counts->as<int64_t>());
    }
    note_launch(ctx);
    int64_t total = 0;
    {
        KTimer kt(ctx, "probe_scan", np * 16);
        exclusive_scan_i64(ctx, counts->as<int64_t>(), offs->as<int64_t>(), np, &total);
    }
    ob = make_buf(ctx, (size_t)total * 8 + 16);
    op = make_buf(ctx, (size_t)total * 8 + 16);
    uint8_t *visited = (mark_visited && j.visited) ? j.visited->as<uint8_t>() : nullptr;
    {
        KTimer kt(ctx, "k_probe_scatter", np * 12 + total * 16);
        k_probe_scatter<<<grid, 256, 0, st>>>(gids->as<uint32_t>(), np, span, have_table ? j.rows->as<uint32_t>() : nullptr, offs->as<int64_t>(),
                                              jt, ob->as<int64_t>(), op->as<int64_t>(), visited);
    }
    note_launch(ctx);
    return total;
}

// build_batch_from_indices (joins/utils.rs:1180-1230): the output columns of `total` (build_idx, probe_idx) pairs, -1 = NULL side
static Batch *materialize(JoinOp &j, const std::vector<ColRef> &probe_cols, const BufRef &ob, const BufRef &op, int64_t total) {
    Ctx *ctx = j.ctx;
    const int jt = j.join_type;
    const bool left_out = jt != DFGPU_JOIN_RIGHT_SEMI && jt != DFGPU_JOIN_RIGHT_ANTI;
    const bool right_out = jt != DFGPU_JOIN_LEFT_SEMI && jt != DFGPU_JOIN_LEFT_ANTI;
    if (!right_out) return empty_output(j);  // LeftSemi / LeftAnti emit nothing per probe batch (utils.rs:1269-1277)
    std::unique_ptr<Batch> out(new Batch());
    out->ctx = ctx;
    out->num_rows = total;
    size_t fi = 0;
    const bool build_nullable = jt == DFGPU_JOIN_RIGHT || jt == DFGPU_JOIN_FULL;
    if (left_out && j.build_rows == 0)
        for (auto &c : j.build->cols) {
            ColRef t = col_nulls(ctx, c->type, total);
            t->type = j.out_schema.fields[fi++];
            out->cols.push_back(t);
        }
    std::vector<const DevCol *> srcs;
    std::vector<const int64_t *> idx;
    std::vector<bool> may_null;
    std::vector<std::pair<const uint8_t *, int>> packed;
    if (left_out && j.build_rows != 0)
        for (size_t c = 0; c < j.build->cols.size(); c++) {
            srcs.push_back(j.build->cols[c].get());
            idx.push_back(ob->as<int64_t>());
            may_null.push_back(build_nullable);
            const bool pk = j.brows.rows && j.brows.off[c] >= 0;
            packed.emplace_back(pk ? j.brows.rows->as<uint8_t>() + j.brows.off[c] : nullptr, j.brows.row_bytes);
        }
    for (auto &c : probe_cols) {
        srcs.push_back(c.get());
        idx.push_back(op->as<int64_t>());
        may_null.push_back(false);
        packed.emplace_back(nullptr, 0);
    }
    for (auto &t : cols_take(ctx, srcs, idx, total, may_null, &packed)) {
        t->type = j.out_schema.fields[fi++];
        out->cols.push_back(t);
    }
    return out.release();
}

Batch *join_probe(JoinOp &j, const Batch &probe) {
    Ctx *ctx = j.ctx;
    auto t0 = std::chrono::steady_clock::now();
    if (!j.finished) fail(DFGPU_INTERNAL, "probe before build_finish");
    if (j.has_filter) fail(DFGPU_INTERNAL, "this join has a residual JoinFilter: use dfgpu_join_probe_candidates / dfgpu_join_probe_commit");
    if (probe.cols.size() != j.probe_schema.fields.size()) fail(DFGPU_INTERNAL, "probe batch does not match the probe schema");
    int64_t np = probe.num_rows;
    j.m.input_batches += 1;
    j.m.input_rows += np;
    if (np == 0) return empty_output(j);
    BufRef ob, op;
    int64_t total = 0;
    if (j.fast) total = probe_pairs_fast(j, probe, ob, op, j.join_type, true);
    else total = probe_pairs_general(j, probe, ob, op, j.join_type, true);
    KTimer kt_gather(ctx, "probe_gather(col_take)", total * 16);
    DF_CUDA(cudaGetLastError());
    std::unique_ptr<Batch> out(materialize(j, probe.cols, ob, op, total));
    j.m.output_batches += 1;
    j.m.output_rows += out->num_rows;
    j.m.join_time_ns += std::chrono::duration_cast<std::chrono::nanoseconds>(std::chrono::steady_clock::now() - t0).count();
    return out.release();
}

// ---- residual JoinFilter ------------------------------------------------------------------------------------
// The reference applies the filter to the candidate pairs BEFORE it marks build rows visited and before the join-type
// fix-ups (hash_join.rs:1259-1311), so the predicate decides which rows count as matched.  PhysicalExpr::evaluate stays in
// the host engine, hence two calls: probe_candidates returns the intermediate batch of the filter's columns for every
// equi-join candidate pair (build_batch_from_indices over filter.column_indices, utils.rs:1154-1162), the host evaluates
// the predicate on it, and probe_commit takes the Boolean result and finishes the probe batch.
__global__ void k_mark_u8(const int64_t *__restrict__ idx, int64_t n, uint8_t *__restrict__ flags) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) flags[idx[i]] = 1;
}
__global__ void k_fill_i64(int64_t *p, int64_t v, int64_t n) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) p[i] = v;
}

void join_set_filter(JoinOp &j, const int32_t *sides, const int32_t *cols, int n) {
    if (j.finished || !j.pending.empty()) fail(DFGPU_INTERNAL, "set_filter after the build side was pushed");
    j.filter_cols.clear();
    for (int i = 0; i < n; i++) {
        const auto &sch = sides[i] == 0 ? j.build_schema : j.probe_schema;
        if ((sides[i] != 0 && sides[i] != 1) || cols[i] < 0 || cols[i] >= (int)sch.fields.size()) fail(DFGPU_PLAN, "JoinFilter column index out of range");
        j.filter_cols.push_back({sides[i], cols[i]});
    }
    j.has_filter = true;
}

Batch *join_probe_candidates(JoinOp &j, const Batch &probe) {
    Ctx *ctx = j.ctx;
    if (!j.finished) fail(DFGPU_INTERNAL, "probe before build_finish");
    if (!j.has_filter) fail(DFGPU_INTERNAL, "probe_candidates: the join has no JoinFilter (use dfgpu_join_probe)");
    if (j.pend.valid) fail(DFGPU_INTERNAL, "probe_candidates: the previous probe batch was not committed");
    if (probe.cols.size() != j.probe_schema.fields.size()) fail(DFGPU_INTERNAL, "probe batch does not match the probe schema");
    const int64_t np = probe.num_rows;
    j.m.input_batches += 1;
    j.m.input_rows += np;
    BufRef ob, op;
    int64_t total = 0;
    if (np > 0) {   // every equi-join match, nothing marked yet: the Inner pairs
        if (j.fast) total = probe_pairs_fast(j, probe, ob, op, DFGPU_JOIN_INNER, false);
        else total = probe_pairs_general(j, probe, ob, op, DFGPU_JOIN_INNER, false);
    }
    if (!ob) { ob = make_buf(ctx, 16); op = make_buf(ctx, 16); }
    j.pend.valid = true;
    j.pend.ob = ob;
    j.pend.op = op;
    j.pend.total = total;
    j.pend.np = np;
    j.pend.probe_cols = probe.cols;
    std::unique_ptr<Batch> out(new Batch());
    out->ctx = ctx;
    out->num_rows = total;
    for (auto &fc : j.filter_cols) {
        const DevCol &src = fc.first == 0 ? *j.build->cols[(size_t)fc.second] : *probe.cols[(size_t)fc.second];
        if (fc.first == 0 && j.build_rows == 0) { out->cols.push_back(col_nulls(ctx, src.type, 0)); continue; }
        out->cols.push_back(col_take(ctx, src, fc.first == 0 ? ob->as<int64_t>() : op->as<int64_t>(), total, false));
    }
    return out.release();
}

Batch *join_probe_commit(JoinOp &j, const Batch &mask, int mask_col) {
    Ctx *ctx = j.ctx;
    cudaStream_t st = ctx->stream;
    auto t0 = std::chrono::steady_clock::now();
    if (!j.pend.valid) fail(DFGPU_INTERNAL, "probe_commit without probe_candidates");
    if (mask_col < 0 || mask_col >= (int)mask.cols.size() || mask.cols[(size_t)mask_col]->type.id != T_BOOL)
        fail(DFGPU_PLAN, "probe_commit: the filter result must be a Boolean column");   // as_boolean_array (utils.rs:1168)
    if (mask.num_rows != j.pend.total) fail(DFGPU_INTERNAL, "probe_commit: the filter result has " + std::to_string(mask.num_rows) + " rows for " + std::to_string(j.pend.total) + " candidate pairs");
    JoinOp::Pending pd = std::move(j.pend);
    j.pend = JoinOp::Pending();
    const int jt = j.join_type;
    // compute::filter of both index arrays by the mask (utils.rs:1170-1171; a NULL predicate counts as false)
    auto i64col = [&](const BufRef &b, int64_t n) {
        auto c = std::make_shared<DevCol>();
        c->type = make_type(T_I64, "i", false);
        c->length = n;
        c->values = b->ptr;
        c->owned.push_back(b);
        return c;
    };
    Batch pairs;
    pairs.ctx = ctx;
    pairs.num_rows = pd.total;
    pairs.cols = {i64col(pd.ob, pd.total), i64col(pd.op, pd.total), mask.cols[(size_t)mask_col]};
    std::unique_ptr<Batch> kept(batch_filter(ctx, &pairs, 2));
    const int64_t K = kept->num_rows;
    const int64_t *fb = (const int64_t *)kept->cols[0]->values, *fp = (const int64_t *)kept->cols[1]->values;
    if (K && need_final(jt)) {   // mark joined build rows visited (hash_join.rs:1274-1278)
        k_mark_u8<<<grid_for(K, 256, ctx, 16), 256, 0, st>>>(fb, K, j.visited->as<uint8_t>());
        note_launch(ctx);
    }
    // adjust_indices_by_join_type (utils.rs:1234-1279) over the whole probe batch
    BufRef ob, op;
    int64_t total = 0;
    if (jt == DFGPU_JOIN_INNER || jt == DFGPU_JOIN_LEFT) {
        ob = K ? make_borrowed(ctx, (void *)fb, (size_t)K * 8) : make_buf(ctx, 16);   // (`kept` outlives the gather below)
        op = K ? make_borrowed(ctx, (void *)fp, (size_t)K * 8) : make_buf(ctx, 16);
        total = K;
    } else if (jt != DFGPU_JOIN_LEFT_SEMI && jt != DFGPU_JOIN_LEFT_ANTI) {
        const int64_t np = pd.np;
        BufRef hit = make_buf(ctx, (size_t)np + 16, true);
        if (K) k_mark_u8<<<grid_for(K, 256, ctx, 16), 256, 0, st>>>(fp, K, hit->as<uint8_t>());
        // probe rows with (RightSemi) / without (Right, Full, RightAnti) a surviving pair, in row order
        int64_t U = 0;
        BufRef uidx = make_buf(ctx, 16);
        if (np) {
            BufRef counts = make_buf(ctx, (size_t)np * 8 + 16), offs = make_buf(ctx, (size_t)np * 8 + 16);
            k_flags_to_counts<<<grid_for(np, 256, ctx, 32), 256, 0, st>>>(hit->as<uint8_t>(), np, jt == DFGPU_JOIN_RIGHT_SEMI ? 1 : 0, counts->as<int64_t>());
            exclusive_scan_i64(ctx, counts->as<int64_t>(), offs->as<int64_t>(), np, &U);
            uidx = make_buf(ctx, (size_t)U * 8 + 16);
            k_compact<<<grid_for(np, 256, ctx, 32), 256, 0, st>>>(counts->as<int64_t>(), offs->as<int64_t>(), np, uidx->as<int64_t>());
            note_launch(ctx, 3);
        }
        if (jt == DFGPU_JOIN_RIGHT || jt == DFGPU_JOIN_FULL) {   // pairs ++ (NULL, unmatched probe row) (utils.rs:1248-1254)
            total = K + U;
            ob = make_buf(ctx, (size_t)total * 8 + 16);
            op = make_buf(ctx, (size_t)total * 8 + 16);
            if (K) {
                DF_CUDA(cudaMemcpyAsync(ob->ptr, fb, (size_t)K * 8, cudaMemcpyDeviceToDevice, st));
                DF_CUDA(cudaMemcpyAsync(op->ptr, fp, (size_t)K * 8, cudaMemcpyDeviceToDevice, st));
            }
            if (U) {
                k_fill_i64<<<grid_for(U, 256, ctx, 16), 256, 0, st>>>(ob->as<int64_t>() + K, -1, U);
                DF_CUDA(cudaMemcpyAsync(op->as<int64_t>() + K, uidx->ptr, (size_t)U * 8, cudaMemcpyDeviceToDevice, st));
                note_launch(ctx);
            }
        } else {   // RightSemi / RightAnti: probe side only
            total = U;
            op = uidx;
            ob = make_buf(ctx, 16);
        }
    } else {
        ob = make_buf(ctx, 16);
        op = make_buf(ctx, 16);
    }
    DF_CUDA(cudaGetLastError());
    std::unique_ptr<Batch> out(materialize(j, pd.probe_cols, ob, op, total));
    j.m.output_batches += 1;
    j.m.output_rows += out->num_rows;
    j.m.join_time_ns += std::chrono::duration_cast<std::chrono::nanoseconds>(std::chrono::steady_clock::now() - t0).count();
    return out.release();
}

Batch *join_emit_unmatched(JoinOp &j) {
    Ctx *ctx = j.ctx;
    cudaStream_t st = ctx->stream;
    if (!j.finished) fail(DFGPU_INTERNAL, "emit_unmatched before build_finish");
    if (!need_final(j.join_type)) return nullptr;
    int64_t nb = j.build_rows;
    int jt = j.join_type;
    int64_t total = 0;
    BufRef idx = make_buf(ctx, 16);
    if (nb) {
        BufRef counts = make_buf(ctx, (size_t)nb * 8 + 16), offs = make_buf(ctx, (size_t)nb * 8 + 16);
        int grid = grid_for(nb, 256, ctx, 32);
        k_flags_to_counts<<<grid, 256, 0, st>>>(j.visited->as<uint8_t>(), nb, jt == DFGPU_JOIN_LEFT_SEMI ? 1 : 0, counts->as<int64_t>());
        exclusive_scan_i64(ctx, counts->as<int64_t>(), offs->as<int64_t>(), nb, &total);
        idx = make_buf(ctx, (size_t)total * 8 + 16);
        k_compact<<<grid, 256, 0, st>>>(counts->as<int64_t>(), offs->as<int64_t>(), nb, idx->as<int64_t>());
        note_launch(ctx, 2);
        DF_CUDA(cudaGetLastError());
    }
    std::unique_ptr<Batch> out(new Batch());
    out->ctx = ctx;
    out->num_rows = total;
    size_t fi = 0;
    for (auto &c : j.build->cols) {
        ColRef t = col_take(ctx, *c, idx->as<int64_t>(), total, false);
        t->type = j.out_schema.fields[fi++];
        out->cols.push_back(t);
    }
    if (jt == DFGPU_JOIN_LEFT || jt == DFGPU_JOIN_FULL)
        for (auto &f : j.probe_schema.fields) {
            (void)f;
            out->cols.push_back(col_nulls(ctx, j.out_schema.fields[fi], total));
            fi++;
        }
    j.m.output_batches += 1;
    j.m.output_rows += total;
    return out.release();
}

}  // namespace dfgpu

using namespace dfgpu;
#define J_TRY try { dfgpu::NvtxRange _nvtx(__func__);
#define J_CATCH } catch (const dfgpu::Error &e) { dfgpu::set_last_error(e.msg); return e.code; } \
    catch (const std::exception &e) { dfgpu::set_last_error(std::string("Internal error: ") + e.what()); return DFGPU_INTERNAL; } \
    return DFGPU_OK;

extern "C" {
dfgpu_status dfgpu_join_create(dfgpu_ctx *ctx, int32_t join_type, int32_t partition_mode, int32_t null_equals_null,
                               const ArrowSchema *build_schema, const ArrowSchema *probe_schema, const int32_t *build_keys,
                               const int32_t *probe_keys, int32_t n_keys, dfgpu_join **out) {
    J_TRY
    DF_CUDA(cudaSetDevice(ctx->device));
    *out = join_create(ctx, join_type, partition_mode, null_equals_null != 0, schema_from_arrow(build_schema), schema_from_arrow(probe_schema),
                       build_keys, probe_keys, n_keys);
    J_CATCH
}
dfgpu_status dfgpu_join_output_schema(const dfgpu_join *join, ArrowSchema *out) {
    J_TRY
    schema_to_arrow(join->out_schema, out);
    J_CATCH
}
dfgpu_status dfgpu_join_build_push(dfgpu_join *join, const dfgpu_batch *batch) {
    J_TRY
    join_build_push(*join, *batch);
    J_CATCH
}
dfgpu_status dfgpu_join_build_finish(dfgpu_join *join) {
    J_TRY
    DF_CUDA(cudaSetDevice(join->ctx->device));
    join_build_finish(*join);
    J_CATCH
}
dfgpu_status dfgpu_join_probe(dfgpu_join *join, const dfgpu_batch *probe, dfgpu_batch **out) {
    J_TRY
    DF_CUDA(cudaSetDevice(join->ctx->device));
    *out = join_probe(*join, *probe);
    J_CATCH
}
dfgpu_status dfgpu_join_set_filter(dfgpu_join *join, const int32_t *sides, const int32_t *columns, int32_t n) {
    J_TRY
    join_set_filter(*join, sides, columns, n);
    J_CATCH
}
dfgpu_status dfgpu_join_probe_candidates(dfgpu_join *join, const dfgpu_batch *probe, dfgpu_batch **out_intermediate) {
    J_TRY
    DF_CUDA(cudaSetDevice(join->ctx->device));
    *out_intermediate = join_probe_candidates(*join, *probe);
    J_CATCH
}
dfgpu_status dfgpu_join_probe_commit(dfgpu_join *join, const dfgpu_batch *filter_result, int32_t column, dfgpu_batch **out) {
    J_TRY
    DF_CUDA(cudaSetDevice(join->ctx->device));
    *out = join_probe_commit(*join, *filter_result, column);
    J_CATCH
}
dfgpu_status dfgpu_join_emit_unmatched(dfgpu_join *join, dfgpu_batch **out) {
    J_TRY
    DF_CUDA(cudaSetDevice(join->ctx->device));
    *out = join_emit_unmatched(*join);
    J_CATCH
}
dfgpu_status dfgpu_join_metrics(const dfgpu_join *join, dfgpu_metrics *out) {
    J_TRY
    *out = join->m;
    out->h2d_bytes = join->ctx->h2d_bytes;
    out->d2h_bytes = join->ctx->d2h_bytes;
    out->kernel_launches = join->ctx->launches;
    J_CATCH
}
void dfgpu_join_destroy(dfgpu_join *join) {
    if (!join) return;
    cudaSetDevice(join->ctx->device);
    delete join;
}
}
