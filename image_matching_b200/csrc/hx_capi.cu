// hx_capi.cu -- extern "C" boundary of libhydia_b200.so (declarations: include/hydia_b200.h).
#include <cstring>

#include "../../include/hydia_b200.h"
#include "hx_client.h"
#include "hx_engine.h"

using namespace hx;
namespace hx { void rescale_noalloc(Engine& e, const u64* in, int ncomp, int nl, u64* out); }   // hx_noalloc.cu

struct hx_ctx { std::shared_ptr<Engine> eng; };
struct hx_ct { Ct ct; };
struct hx_digits { std::shared_ptr<Engine::Digits> dg; };
struct hx_client { Client* cl; };
struct hx_rotset { std::shared_ptr<Engine::RotSet> rs; };

static thread_local std::string g_err;
static void set_err(const std::exception& e) { g_err = e.what(); }

#define HX_TRY_PTR(body)               \
    try { body }                       \
    catch (const std::exception& e) {  \
        set_err(e);                    \
        return nullptr;                \
    }
#define HX_TRY_INT(body)               \
    try { body return 0; }             \
    catch (const std::exception& e) {  \
        set_err(e);                    \
        return -1;                     \
    }

static hx_ct* wrap(const Ct& c) {
    hx_ct* h = new hx_ct;
    h->ct = c;
    return h;
}
static void need(const void* p, const char* what) {
    if (!p) throw std::runtime_error(std::string("null argument: ") + what);
}

extern "C" {

const char* hx_last_error(void) { return g_err.c_str(); }
int hx_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

hx_ctx* hx_ctx_create(int N, int L, int K, int dnum, const uint64_t* q, const uint64_t* p, const uint64_t* psi_q,
                      const uint64_t* psi_p, int log2_scale, int device) {
    HX_TRY_PTR(
        need(q, "q"); need(p, "p"); need(psi_q, "psi_q"); need(psi_p, "psi_p");
        int n = 0;
        if (cudaGetDeviceCount(&n) != cudaSuccess || n == 0)
            throw std::runtime_error("no CUDA device: libhydia_b200 has no CPU fallback");
        hx_ctx* c = new hx_ctx;
        c->eng = std::make_shared<Engine>(N, L, K, dnum, (const u64*)q, (const u64*)p, (const u64*)psi_q, (const u64*)psi_p, log2_scale, device);
        return c;)
}
void hx_ctx_destroy(hx_ctx* ctx) {
    if (!ctx) return;
    delete ctx;
}
int hx_ctx_sync(hx_ctx* ctx) { HX_TRY_INT(need(ctx, "ctx"); ctx->eng->sync();) }
uint64_t hx_ctx_launches(hx_ctx* ctx) { return ctx ? ctx->eng->launches : 0; }
#ifndef HX_KERNEL_STAMP
#define HX_KERNEL_STAMP "unstamped"
#endif
const char* hx_kernel_stamp(void) { return HX_KERNEL_STAMP; }
// diagnosis: cross-stream reuse policies of the stream-ordered pool (bit 0 follow-event, bit 1 opportunistic, bit 2 internal dependencies)
static void set_pool_reuse(int device, int value) {
    cudaMemPool_t pool;
    HX_CUDA(cudaDeviceGetDefaultMemPool(&pool, device));
    int a = value & 1;
    int b = (value >> 1) & 1;
    int c = (value >> 2) & 1;
    HX_CUDA(cudaMemPoolSetAttribute(pool, cudaMemPoolReuseFollowEventDependencies, &a));
    HX_CUDA(cudaMemPoolSetAttribute(pool, cudaMemPoolReuseAllowOpportunistic, &b));
    HX_CUDA(cudaMemPoolSetAttribute(pool, cudaMemPoolReuseAllowInternalDependencies, &c));
}
int hx_ctx_set_option(hx_ctx* ctx, const char* name, int value) {
    HX_TRY_INT(
        need(ctx, "ctx"); need(name, "name");
        Engine::Guard g(ctx->eng.get());
        std::string n(name);
        if (n == "force_unfused") ctx->eng->force_unfused = value != 0;
        else if (n == "rot_chunk") ctx->eng->rot_chunk = value;
        else if (n == "rot_streams") ctx->eng->rot_streams = std::max(1, std::min(4, value));
        else if (n == "rot_lane") ctx->eng->rot_lane = value != 0;
        else if (n == "batch_blocks") ctx->eng->batch_blocks = value != 0;
        else if (n == "pdl") g_pdl = value != 0;
        else if (n == "ks_moddown_w") ctx->eng->ks_moddown_w = value != 0;
        else if (n == "db_diag0" || n == "db_diag_count") {
            if (ctx->eng->db_blocks()) throw std::runtime_error("hx_ctx_set_option: the resident diagonal range must be set before any database block is registered");
            if (value < 0 || (n == "db_diag_count" && value < 1)) throw std::runtime_error("hx_ctx_set_option: bad diagonal range");
            (n == "db_diag0" ? ctx->eng->db_diag0 : ctx->eng->db_diag_count) = value;
        }
        else if (n == "pool_reuse") set_pool_reuse(ctx->eng->device, value);
        else throw std::runtime_error("hx_ctx_set_option: unknown option " + n);)
}
int hx_prof_kinds(void) { return PK_COUNT; }
const char* hx_prof_name(int kind) { return prof_kind_name(kind); }
int hx_prof_enable(hx_ctx* ctx, int mask) { HX_TRY_INT(need(ctx, "ctx"); ctx->eng->prof_enable((unsigned)mask);) }
int hx_prof_reset(hx_ctx* ctx) { HX_TRY_INT(need(ctx, "ctx"); ctx->eng->prof_reset();) }
int hx_prof_read(hx_ctx* ctx, int kind, uint64_t* launches, double* ms, double* bytes) {
    HX_TRY_INT(
        need(ctx, "ctx");
        if (kind < 0 || kind >= PK_COUNT) throw std::runtime_error("hx_prof_read: bad kind");
        ctx->eng->prof_collect();
        if (launches) *launches = ctx->eng->prof_launches[kind];
        if (ms) *ms = ctx->eng->prof_ms[kind];
        if (bytes) *bytes = ctx->eng->prof_bytes[kind];)
}
int hx_timer_start(hx_ctx* ctx) { HX_TRY_INT(need(ctx, "ctx"); ctx->eng->timer_start();) }
int hx_timer_stop(hx_ctx* ctx, double* ms) { HX_TRY_INT(need(ctx, "ctx"); need(ms, "ms"); *ms = ctx->eng->timer_stop_ms();) }
uint32_t hx_galois_for_rotation(hx_ctx* ctx, int rotation) { return ctx ? (uint32_t)ctx->eng->galois_for_rotation(rotation) : 0; }

int hx_key_load_mult(hx_ctx* ctx, const uint64_t* a, const uint64_t* b) {
    HX_TRY_INT(need(ctx, "ctx"); need(a, "a"); need(b, "b"); ctx->eng->load_mult_key((const u64*)a, (const u64*)b);)
}
int hx_key_load_auto(hx_ctx* ctx, uint32_t galois, const uint64_t* a, const uint64_t* b) {
    HX_TRY_INT(need(ctx, "ctx"); need(a, "a"); need(b, "b"); ctx->eng->load_rot_key(galois, (const u64*)a, (const u64*)b);)
}

hx_ct* hx_ct_upload(hx_ctx* ctx, int ncomp, int nlimbs, const uint64_t* data, int level, int noise_deg) {
    HX_TRY_PTR(
        need(ctx, "ctx"); need(data, "data");
        if (ncomp < 1 || ncomp > 3 || nlimbs < 1 || nlimbs > ctx->eng->L) throw std::runtime_error("hx_ct_upload: bad shape");
        return wrap(ctx->eng->upload((const u64*)data, ncomp, nlimbs, level, noise_deg));)
}
hx_ct* hx_ct_upload_dev(hx_ctx* ctx, int ncomp, int nlimbs, const void* dev_data, int level, int noise_deg) {
    HX_TRY_PTR(
        need(ctx, "ctx"); need(dev_data, "dev_data");
        if (ncomp < 1 || ncomp > 3 || nlimbs < 1 || nlimbs > ctx->eng->L) throw std::runtime_error("hx_ct_upload_dev: bad shape");
        Engine* e = ctx->eng.get();
        Engine::Guard g(e);
        Ct c = e->new_ct(ncomp, nlimbs, level, noise_deg);
        HX_CUDA(cudaMemcpyAsync(c->d, dev_data, sizeof(u64) * (size_t)ncomp * nlimbs * e->N, cudaMemcpyDeviceToDevice, e->st));
        e->sync();
        c->settled = true;
        return wrap(c);)
}
int hx_ct_download_dev(hx_ct* ct, void* dev_data) {
    HX_TRY_INT(
        need(ct, "ct"); need(dev_data, "dev_data");
        if (ct->ct->nb != 1) throw std::runtime_error("hx_ct_download_dev: batched ciphertext (hx_ct_unstack it first)");
        Engine* e = ct->ct->eng;
        Engine::Guard g(e);
        HX_CUDA(cudaMemcpyAsync(dev_data, ct->ct->d, sizeof(u64) * (size_t)ct->ct->ncomp * ct->ct->nl * e->N, cudaMemcpyDeviceToDevice, e->st));
        e->sync();)
}
// Stream-ordered exchange with a caller-owned stream (the transport stream of sharded.py): no host synchronisation.  Both directions
// are ordered: the copy runs after everything queued on `stream` so far, and `stream` waits for the copy (so the caller may reuse or
// free the buffer in its own stream order).
static void order_streams(cudaStream_t first, cudaStream_t then) {
    cudaEvent_t ev;
    HX_CUDA(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
    cudaError_t e1 = cudaEventRecord(ev, first), e2 = cudaStreamWaitEvent(then, ev, 0);
    cudaEventDestroy(ev);   // released once the recorded work has completed
    HX_CUDA(e1);
    HX_CUDA(e2);
}
int hx_ct_download_dev_async(hx_ct* ct, void* dev_data, void* stream) {
    HX_TRY_INT(
        need(ct, "ct"); need(dev_data, "dev_data");
        if (ct->ct->nb != 1) throw std::runtime_error("hx_ct_download_dev_async: batched ciphertext (hx_ct_unstack it first)");
        Engine* e = ct->ct->eng;
        Engine::Guard g(e);
        HX_CUDA(cudaSetDevice(e->device));
        order_streams((cudaStream_t)stream, e->st);
        HX_CUDA(cudaMemcpyAsync(dev_data, ct->ct->d, sizeof(u64) * (size_t)ct->ct->ncomp * ct->ct->nl * e->N, cudaMemcpyDeviceToDevice, e->st));
        order_streams(e->st, (cudaStream_t)stream);)
}
hx_ct* hx_ct_upload_dev_async(hx_ctx* ctx, int ncomp, int nlimbs, const void* dev_data, int level, int noise_deg, void* stream) {
    HX_TRY_PTR(
        need(ctx, "ctx"); need(dev_data, "dev_data");
        if (ncomp < 1 || ncomp > 3 || nlimbs < 1 || nlimbs > ctx->eng->L) throw std::runtime_error("hx_ct_upload_dev_async: bad shape");
        Engine* e = ctx->eng.get();
        Engine::Guard g(e);
        HX_CUDA(cudaSetDevice(e->device));
        Ct c = e->new_ct(ncomp, nlimbs, level, noise_deg);
        order_streams((cudaStream_t)stream, e->st);
        HX_CUDA(cudaMemcpyAsync(c->d, dev_data, sizeof(u64) * (size_t)ncomp * nlimbs * e->N, cudaMemcpyDeviceToDevice, e->st));
        order_streams(e->st, (cudaStream_t)stream);
        return wrap(c);)
}
int hx_ct_download(hx_ct* ct, uint64_t* data) {
    HX_TRY_INT(need(ct, "ct"); need(data, "data");
               if (ct->ct->nb != 1) throw std::runtime_error("hx_ct_download: batched ciphertext (hx_ct_unstack it first)");
               ct->ct->eng->download(ct->ct, (u64*)data);)
}
int hx_ct_download_many(hx_ct* const* cts, int n, uint64_t* data) {
    HX_TRY_INT(need(cts, "cts"); need(data, "data");
               if (n < 1) throw std::runtime_error("hx_ct_download_many: empty input");
               need(cts[0], "cts[0]");
               Engine* e = cts[0]->ct->eng;
               Engine::Guard g(e);
               size_t off = 0;
               for (int i = 0; i < n; i++) {
                   need(cts[i], "cts[i]");
                   const CtObj& c = *cts[i]->ct;
                   if (c.eng != e) throw std::runtime_error("hx_ct_download_many: ciphertexts of different contexts");
                   const size_t words = (size_t)c.nb * c.ncomp * c.nl * e->N;
                   HX_CUDA(cudaMemcpyAsync(data + off, c.d, sizeof(u64) * words, cudaMemcpyDeviceToHost, e->st));
                   off += words;
               }
               e->sync();)
}
int hx_ct_info(hx_ct* ct, int* ncomp, int* nlimbs, int* level, int* noise_deg) {
    HX_TRY_INT(
        need(ct, "ct");
        if (ncomp) *ncomp = ct->ct->ncomp;
        if (nlimbs) *nlimbs = ct->ct->nl;
        if (level) *level = ct->ct->level;
        if (noise_deg) *noise_deg = ct->ct->deg;)
}
hx_ct* hx_ct_clone(hx_ct* ct) { HX_TRY_PTR(need(ct, "ct"); return wrap(ct->ct->eng->clone(ct->ct));) }
void hx_ct_free(hx_ct* ct) { delete ct; }

#define E(x) ((x)->ct->eng)
hx_ct* hx_add(hx_ct* a, hx_ct* b) { HX_TRY_PTR(need(a, "a"); need(b, "b"); return wrap(E(a)->add(a->ct, b->ct));) }
hx_ct* hx_sub(hx_ct* a, hx_ct* b) { HX_TRY_PTR(need(a, "a"); need(b, "b"); return wrap(E(a)->sub(a->ct, b->ct));) }
hx_ct* hx_add_const(hx_ct* a, double c) { HX_TRY_PTR(need(a, "a"); return wrap(E(a)->add_const(a->ct, c));) }
hx_ct* hx_mult_const(hx_ct* a, double c) { HX_TRY_PTR(need(a, "a"); return wrap(E(a)->mult_const(a->ct, c));) }
hx_ct* hx_mult_norelin(hx_ct* a, hx_ct* b) { HX_TRY_PTR(need(a, "a"); need(b, "b"); return wrap(E(a)->mult_norelin(a->ct, b->ct));) }
hx_ct* hx_relin(hx_ct* a) { HX_TRY_PTR(need(a, "a"); return wrap(E(a)->relinearize(a->ct));) }
hx_ct* hx_mult(hx_ct* a, hx_ct* b) { HX_TRY_PTR(need(a, "a"); need(b, "b"); return wrap(E(a)->mult(a->ct, b->ct));) }
hx_ct* hx_square(hx_ct* a) { HX_TRY_PTR(need(a, "a"); return wrap(E(a)->square(a->ct));) }
hx_ct* hx_rescale(hx_ct* a) { HX_TRY_PTR(need(a, "a"); return wrap(E(a)->rescale(a->ct));) }
hx_ct* hx_level_reduce(hx_ct* a, int levels) { HX_TRY_PTR(need(a, "a"); return wrap(E(a)->level_reduce(a->ct, levels));) }
hx_ct* hx_rotate(hx_ct* a, int r) { HX_TRY_PTR(need(a, "a"); return wrap(E(a)->rotate(a->ct, r));) }
hx_ct* hx_eval_sum(hx_ct* a, int n) { HX_TRY_PTR(need(a, "a"); return wrap(E(a)->eval_sum(a->ct, n));) }
hx_ct* hx_add_many(hx_ct* const* cts, int n) {
    HX_TRY_PTR(
        need(cts, "cts");
        if (n < 1) throw std::runtime_error("hx_add_many: empty input");
        std::vector<Ct> v;
        for (int i = 0; i < n; i++) { need(cts[i], "cts[i]"); v.push_back(cts[i]->ct); }
        return wrap(v[0]->eng->add_many(v));)
}
hx_digits* hx_rot_precompute(hx_ct* a) {
    HX_TRY_PTR(need(a, "a"); hx_digits* d = new hx_digits; d->dg = E(a)->fast_rotation_precompute(a->ct); return d;)
}
hx_ct* hx_rot_fast(hx_ct* a, int r, hx_digits* d) { HX_TRY_PTR(need(a, "a"); need(d, "digits"); return wrap(E(a)->fast_rotation(a->ct, r, d->dg));) }
void hx_digits_free(hx_digits* d) { delete d; }
hx_ct* hx_cheby_series(hx_ct* a, const double* coeffs, int n, double lo, double hi) {
    HX_TRY_PTR(need(a, "a"); need(coeffs, "coeffs"); return wrap(E(a)->chebyshev_series(a->ct, std::vector<double>(coeffs, coeffs + n), lo, hi));)
}
hx_ct* hx_poly(hx_ct* a, const double* coeffs, int n) {
    HX_TRY_PTR(need(a, "a"); need(coeffs, "coeffs"); return wrap(E(a)->eval_poly(a->ct, std::vector<double>(coeffs, coeffs + n)));)
}
hx_ct* hx_compare(hx_ct* a, double delta, int sign_depth) { HX_TRY_PTR(need(a, "a"); return wrap(E(a)->chebyshev_compare(a->ct, delta, sign_depth));) }

int hx_db_register(hx_ctx* ctx, int block, int diag, const uint64_t* data) {
    HX_TRY_INT(need(ctx, "ctx"); need(data, "data"); ctx->eng->db_register(block, diag, (const u64*)data);)
}
int hx_db_blocks(hx_ctx* ctx) { return ctx ? ctx->eng->db_blocks() : 0; }
int hx_db_register_ct(hx_ctx* ctx, int block, int diag, hx_ct* ct) {
    HX_TRY_INT(need(ctx, "ctx"); need(ct, "ct");
               if (ct->ct->eng != ctx->eng.get() || ct->ct->ncomp != 2 || ct->ct->nl != ctx->eng->L || ct->ct->nb != 1)
                   throw std::runtime_error("hx_db_register_ct: a fresh 2-component ciphertext of this context is required");
               ctx->eng->db_store_device(block, diag, ct->ct->d);)
}
hx_ct* hx_db_get(hx_ctx* ctx, int block, int diag) { HX_TRY_PTR(need(ctx, "ctx"); return wrap(ctx->eng->db_entry(block, diag));) }
hx_rotset* hx_rot_fast_all(hx_ct* query, int ndiag) {
    HX_TRY_PTR(need(query, "query"); hx_rotset* r = new hx_rotset; r->rs = query->ct->eng->rotate_hoisted_all(query->ct, ndiag); return r;)
}
hx_ct* hx_rotset_get(hx_rotset* r, int i) { HX_TRY_PTR(need(r, "rotset"); return wrap(r->rs->eng->rotset_entry(*r->rs, i));) }
void hx_rotset_free(hx_rotset* r) { delete r; }
int hx_score_blocks(hx_rotset* r, const int* blocks, int nblocks, hx_ct** out) {
    HX_TRY_INT(need(r, "rotset"); need(blocks, "blocks"); need(out, "out");
               Engine* e = r->rs->eng;
               int k = 0;
               for (auto& part : e->score_rotset(*r->rs, std::vector<int>(blocks, blocks + nblocks)))
                   for (auto& c : e->unstack(part)) out[k++] = wrap(c);)
}
int hx_score_diag(hx_ctx* ctx, hx_ct* query, int ndiag, hx_ct** out, int max_out) {
    try {
        need(ctx, "ctx"); need(query, "query"); need(out, "out");
        std::vector<Ct> s = ctx->eng->score_diag(query->ct, ndiag);
        if ((int)s.size() > max_out) throw std::runtime_error("hx_score_diag: output array too small");
        for (size_t i = 0; i < s.size(); i++) out[i] = wrap(s[i]);
        return (int)s.size();
    } catch (const std::exception& e) { set_err(e); return -1; }
}
hx_ct* hx_diag_mac_partial(hx_ctx* ctx, hx_ct* query, int d0, int d1, const int* blocks, int nblocks) {
    HX_TRY_PTR(need(ctx, "ctx"); need(query, "query"); need(blocks, "blocks");
               return wrap(ctx->eng->diag_mac_partial(query->ct, d0, d1, std::vector<int>(blocks, blocks + nblocks)));)
}
int hx_ct_fold_mod(hx_ct* ct) { HX_TRY_INT(need(ct, "ct"); ct->ct->eng->fold_mod(ct->ct);) }
hx_ct* hx_ct_alloc(hx_ctx* ctx, int ncomp, int nlimbs, int level, int noise_deg, int nbatch) {
    HX_TRY_PTR(need(ctx, "ctx");
               if (ncomp < 1 || ncomp > 3 || nlimbs < 1 || nlimbs > ctx->eng->L || nbatch < 1) throw std::runtime_error("hx_ct_alloc: bad shape");
               return wrap(ctx->eng->alloc_like(ncomp, nlimbs, level, noise_deg, nbatch));)
}
int hx_ct_device_view(hx_ct* ct, void** dev_ptr, uint64_t* words, int* nbatch) {
    HX_TRY_INT(need(ct, "ct");
               if (dev_ptr) *dev_ptr = ct->ct->d;
               if (words) *words = (uint64_t)ct->ct->nb * ct->ct->batch_stride();
               if (nbatch) *nbatch = ct->ct->nb;)
}
int hx_stream_order(hx_ctx* ctx, void* stream, int engine_waits) {
    HX_TRY_INT(need(ctx, "ctx");
               Engine* e = ctx->eng.get();
               Engine::Guard g(e);
               if (engine_waits) order_streams((cudaStream_t)stream, e->st);
               else order_streams(e->st, (cudaStream_t)stream);)
}
hx_ct* hx_membership(hx_ctx* ctx, hx_ct* query, int ndiag, double delta, int sign_depth) {
    HX_TRY_PTR(need(ctx, "ctx"); need(query, "query"); return wrap(ctx->eng->membership(query->ct, ndiag, delta, sign_depth));)
}
int hx_score_hers_blocks(hx_ctx* ctx, hx_ct* const* query, int nquery, const int* matrices, int nmatrices, hx_ct** out) {
    HX_TRY_INT(need(ctx, "ctx"); need(query, "query"); need(matrices, "matrices"); need(out, "out");
               std::vector<Ct> q;
               for (int i = 0; i < nquery; i++) { need(query[i], "query[i]"); q.push_back(query[i]->ct); }
               std::vector<int> ids(matrices, matrices + nmatrices);
               std::vector<Ct> s = ctx->eng->score_hers(q, &ids);
               for (size_t i = 0; i < s.size(); i++) out[i] = wrap(s[i]);)
}
hx_ct* hx_ct_stack(hx_ct* const* cts, int n) {
    HX_TRY_PTR(need(cts, "cts"); if (n <= 0) throw std::runtime_error("hx_ct_stack: empty input");
               std::vector<Ct> v; for (int i = 0; i < n; i++) { need(cts[i], "cts[i]"); v.push_back(cts[i]->ct); }
               return wrap(v[0]->eng->concat(v));)
}
int hx_ct_unstack(hx_ct* ct, hx_ct** out, int max_out) {
    try {
        need(ct, "ct"); need(out, "out");
        if (ct->ct->nb > max_out) throw std::runtime_error("hx_ct_unstack: output array too small");
        std::vector<Ct> v = ct->ct->eng->split(ct->ct, ct->ct->nb);
        for (size_t i = 0; i < v.size(); i++) out[i] = wrap(v[i]);
        return (int)v.size();
    } catch (const std::exception& e) { set_err(e); return -1; }
}
hx_ct* hx_mult_plain(hx_ct* a, const uint64_t* pt, int nlimbs) {
    HX_TRY_PTR(need(a, "a"); need(pt, "pt"); return wrap(a->ct->eng->mult_plain(a->ct, (const u64*)pt, nlimbs));)
}
hx_ct* hx_inner_product(hx_ct* a, hx_ct* b, int n) {
    HX_TRY_PTR(need(a, "a"); need(b, "b"); return wrap(a->ct->eng->inner_product(a->ct, b->ct, n));)
}
int hx_score_hers(hx_ctx* ctx, hx_ct* const* query, int nquery, hx_ct** out, int max_out) {
    try {
        need(ctx, "ctx"); need(query, "query"); need(out, "out");
        std::vector<Ct> q;
        for (int i = 0; i < nquery; i++) { need(query[i], "query[i]"); q.push_back(query[i]->ct); }
        std::vector<Ct> s = ctx->eng->score_hers(q);
        if ((int)s.size() > max_out) throw std::runtime_error("hx_score_hers: output array too small");
        for (size_t i = 0; i < s.size(); i++) out[i] = wrap(s[i]);
        return (int)s.size();
    } catch (const std::exception& e) { set_err(e); return -1; }
}
int hx_compare_many(hx_ctx* ctx, hx_ct* const* in, int n, double delta, int sign_depth, hx_ct** out) {
    try {
        need(ctx, "ctx"); need(in, "in"); need(out, "out");
        std::vector<Ct> v;
        for (int i = 0; i < n; i++) { need(in[i], "in[i]"); v.push_back(in[i]->ct); }
        std::vector<Ct> s = ctx->eng->compare_many(v, delta, sign_depth);
        for (size_t i = 0; i < s.size(); i++) out[i] = wrap(s[i]);
        return 0;
    } catch (const std::exception& e) { set_err(e); return -1; }
}
int hx_index(hx_ctx* ctx, hx_ct* query, int ndiag, double delta, int sign_depth, hx_ct** out, int max_out) {
    try {
        need(ctx, "ctx"); need(query, "query"); need(out, "out");
        std::vector<Ct> s = ctx->eng->index(query->ct, ndiag, delta, sign_depth);
        if ((int)s.size() > max_out) throw std::runtime_error("hx_index: output array too small");
        for (size_t i = 0; i < s.size(); i++) out[i] = wrap(s[i]);
        return (int)s.size();
    } catch (const std::exception& e) { set_err(e); return -1; }
}

int hx_ntt(hx_ctx* ctx, uint64_t* data, int nbatch, const int* primes, int nprimes, int inverse) {
    HX_TRY_INT(
        need(ctx, "ctx"); need(data, "data"); need(primes, "primes");
        Engine* e = ctx->eng.get();
        Engine::Guard g(e);
        size_t words = (size_t)nbatch * e->N;
        u64* d = e->alloc_words(words);
        HX_CUDA(cudaMemcpyAsync(d, data, words * sizeof(u64), cudaMemcpyHostToDevice, e->st));
        e->ntt_limbs(d, nbatch, std::vector<int>(primes, primes + nprimes), inverse != 0);
        HX_CUDA(cudaMemcpyAsync(data, d, words * sizeof(u64), cudaMemcpyDeviceToHost, e->st));
        e->sync();
        e->free_words(d);)
}

int hx_params_generate(int depth, int scale_bits, int first_mod_bits, int ring_dim, int num_large_digits, int* N, int* L, int* K,
                       int* dnum, uint64_t* q, uint64_t* p, uint64_t* psi_q, uint64_t* psi_p) {
    HX_TRY_INT(
        ParamSet P = generate_params(depth, scale_bits, first_mod_bits, ring_dim, num_large_digits);
        *N = P.N; *L = P.L; *K = P.K; *dnum = P.dnum;
        for (int i = 0; i < P.L; i++) { q[i] = P.q[i]; psi_q[i] = P.psi_q[i]; }
        for (int i = 0; i < P.K; i++) { p[i] = P.p[i]; psi_p[i] = P.psi_p[i]; })
}

hx_client* hx_client_create(hx_ctx* ctx, uint64_t seed) {
    HX_TRY_PTR(need(ctx, "ctx"); hx_client* c = new hx_client; c->cl = new Client(ctx->eng, seed); return c;)
}
void hx_client_destroy(hx_client* c) {
    if (!c) return;
    delete c->cl;
    delete c;
}
int hx_client_gen_mult_key(hx_client* c) { HX_TRY_INT(need(c, "client"); c->cl->gen_mult_key();) }
int hx_client_gen_rot_keys(hx_client* c, const int* rotations, int n) { HX_TRY_INT(need(c, "client"); need(rotations, "rotations"); c->cl->gen_rot_keys(rotations, n);) }
int hx_client_gen_sum_keys(hx_client* c) { HX_TRY_INT(need(c, "client"); c->cl->gen_sum_keys();) }
hx_ct* hx_client_encrypt(hx_client* c, const double* values, int n) { HX_TRY_PTR(need(c, "client"); need(values, "values"); return wrap(c->cl->encrypt(values, n));) }
int hx_client_encode(hx_client* c, const double* values, int n, uint64_t* out) {
    HX_TRY_INT(need(c, "client"); need(values, "values"); need(out, "out"); c->cl->encode_rns(values, n, (u64*)out);)
}
int hx_client_enroll(hx_client* c, int block, int diag, const double* values, int n) {
    HX_TRY_INT(need(c, "client"); need(values, "values"); c->cl->enroll(block, diag, values, n);)
}
int hx_client_decrypt(hx_client* c, hx_ct* ct, double* out) { HX_TRY_INT(need(c, "client"); need(ct, "ct"); need(out, "out"); c->cl->decrypt(ct->ct, out);) }
// ---- (de)serialisation support (shim/hx_serial.h): parameters, evaluation keys and the key pair as plain residues
int hx_ctx_get_params(hx_ctx* ctx, int* N, int* L, int* K, int* dnum, int* log2_scale, uint64_t* q, uint64_t* p, uint64_t* psi_q, uint64_t* psi_p) {
    HX_TRY_INT(need(ctx, "ctx");
               Engine* e = ctx->eng.get();
               if (N) *N = e->N;
               if (L) *L = e->L;
               if (K) *K = e->K;
               if (dnum) *dnum = e->dnum;
               if (log2_scale) *log2_scale = e->scale_bits;
               for (int i = 0; i < e->L; i++) { if (q) q[i] = e->mods[i]; if (psi_q) psi_q[i] = e->psis[i]; }
               for (int i = 0; i < e->K; i++) { if (p) p[i] = e->mods[e->L + i]; if (psi_p) psi_p[i] = e->psis[e->L + i]; })
}
int hx_key_count(hx_ctx* ctx, int* has_mult) {
    if (!ctx) return -1;
    Engine::Guard g(ctx->eng.get());
    if (has_mult) *has_mult = ctx->eng->has_mult_key() ? 1 : 0;
    return (int)ctx->eng->rot_key_galois().size();
}
int hx_key_list(hx_ctx* ctx, uint32_t* galois, int max_out) {
    try {
        need(ctx, "ctx"); need(galois, "galois");
        Engine::Guard g(ctx->eng.get());
        std::vector<u64> v = ctx->eng->rot_key_galois();
        if ((int)v.size() > max_out) throw std::runtime_error("hx_key_list: output array too small");
        for (size_t i = 0; i < v.size(); i++) galois[i] = (uint32_t)v[i];
        return (int)v.size();
    } catch (const std::exception& e) { set_err(e); return -1; }
}
int hx_key_export(hx_ctx* ctx, uint32_t galois, uint64_t* a, uint64_t* b) {
    HX_TRY_INT(need(ctx, "ctx"); need(a, "a"); need(b, "b"); ctx->eng->export_key(galois, (u64*)a, (u64*)b);)
}
int hx_client_export_public(hx_client* c, uint64_t* pkb, uint64_t* pka) {
    HX_TRY_INT(need(c, "client"); need(pkb, "pkb"); need(pka, "pka"); c->cl->export_public((u64*)pkb, (u64*)pka);)
}
hx_client* hx_client_import(hx_ctx* ctx, uint64_t seed, const int8_t* sk, const uint64_t* pkb, const uint64_t* pka) {
    HX_TRY_PTR(need(ctx, "ctx"); need(sk, "sk"); need(pkb, "pkb"); need(pka, "pka");
               hx_client* c = new hx_client;
               c->cl = new Client(ctx->eng, seed, sk, (const u64*)pkb, (const u64*)pka);
               return c;)
}
int hx_client_secret(hx_client* c, int8_t* out) { HX_TRY_INT(need(c, "client"); need(out, "out"); c->cl->secret(out);) }

int hx_trace_start(hx_ctx* ctx) { HX_TRY_INT(need(ctx, "ctx"); ctx->eng->trace_start();) }
int hx_trace_mark(hx_ctx* ctx, const char* name, void* stream) {
    HX_TRY_INT(need(ctx, "ctx"); need(name, "name"); ctx->eng->trace_mark(name, (cudaStream_t)stream);)
}
int hx_trace_dump(hx_ctx* ctx, char* out, int cap) {
    HX_TRY_INT(need(ctx, "ctx"); need(out, "out");
               std::string s = ctx->eng->trace_dump();
               if (cap < 1) throw std::runtime_error("hx_trace_dump: empty buffer");
               std::strncpy(out, s.c_str(), (size_t)cap - 1);
               out[cap - 1] = 0;)
}
int hx_rescale_noalloc(hx_ctx* ctx, const uint64_t* d_in, int ncomp, int nlimbs, uint64_t* d_out) {
    HX_TRY_INT(need(ctx, "ctx"); rescale_noalloc(*ctx->eng, (const u64*)d_in, ncomp, nlimbs, (u64*)d_out);)
}
}  // extern "C"
