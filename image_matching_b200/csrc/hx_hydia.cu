// hx_hydia.cu -- the fused HyDia Sender path: what
// DiagonalSender::computeSimilarity / membershipScenario / indexScenario
// (/root/reference/src/sender/sender_diag.cpp:12-94) compute, with the database HBM-resident.
#include "hx_engine.h"

namespace hx {


// DB ciphertext `diag` of block `block` (serial/db_diagonal/index{block*512+diag}.bin in the
// reference, sender_diag.cpp:87): stored in Montgomery form for the lazy MAC.
static u64* db_slot(Engine& e, std::map<int, Engine::DbBlock>& db, int block, int diag) {
    if (diag < e.db_diag0 || diag >= e.db_diag0 + e.db_diag_count)
        throw std::runtime_error("db_register: diagonal index outside this context's resident range (options db_diag0 / db_diag_count)");
    Engine::DbBlock& b = db[block];
    size_t ct_words = (size_t)2 * e.L * e.N;
    if (!b.d) {
        HX_CUDA(cudaMalloc(&b.d, sizeof(u64) * ct_words * e.db_diag_count));
        HX_CUDA(cudaMemsetAsync(b.d, 0, sizeof(u64) * ct_words * e.db_diag_count, e.st));
        b.ndiag = e.db_diag_count;
        b.diag0 = e.db_diag0;
    }
    if (b.diag0 != e.db_diag0 || b.ndiag != e.db_diag_count) throw std::runtime_error("db_register: the resident diagonal range changed after this block was created");
    return b.d + ct_words * (diag - b.diag0);
}
static void db_finish(Engine& e, u64* dst) {
    std::vector<int> primes;
    for (int i = 0; i < e.L; i++) primes.push_back(i);
    Engine::Scope sc(&e, PK_MISC, 1, 32.0 * e.L * e.N);
    launch_to_mont(e.T, dst, 2 * e.L, e.limb_map(primes), e.st);
}
void Engine::db_register(int block, int diag, const u64* host_ct) {
    Guard g(this);
    u64* dst = db_slot(*this, db_, block, diag);
    HX_CUDA(cudaMemcpyAsync(dst, host_ct, sizeof(u64) * (size_t)2 * L * N, cudaMemcpyHostToDevice, st));
    db_finish(*this, dst);
    sync();
}
void Engine::db_store_device(int block, int diag, const u64* dev_ct) {
    Guard g(this);
    u64* dst = db_slot(*this, db_, block, diag);
    HX_CUDA(cudaMemcpyAsync(dst, dev_ct, sizeof(u64) * (size_t)2 * L * N, cudaMemcpyDeviceToDevice, st));
    db_finish(*this, dst);
}
void Engine::db_adopt(int block, u64* dev, int ndiag) {
    Guard g(this);
    DbBlock& b = db_[block];
    if (b.d) cudaFree(b.d);
    b.d = dev;
    b.ndiag = ndiag;
}

// computeSimilarity (sender_diag.cpp:12-33): hoisted rotations 1..ndiag-1 of the query
// (EvalFastRotationPrecompute + EvalFastRotation), then per block the 512 tensor products, their
// sum, one relinearisation and one rescale (computeSimilarityMatrix, :66-83).
Engine::RotSet::~RotSet() {
    if (slot >= 0 && eng) { Engine::Guard g(eng); eng->rot_slot_release(slot); }
}
int Engine::rot_slot_acquire(size_t words, cudaStream_t writer) {
    int pick = -1;
    for (size_t i = 0; i < rot_slots_.size(); i++)
        if (!rot_slots_[i].busy && rot_slots_[i].words >= words && (pick < 0 || rot_slots_[i].words < rot_slots_[pick].words)) pick = (int)i;
    if (pick < 0) {
        for (size_t i = 0; i < rot_slots_.size(); i++)   // an idle slot that is too small: grow it (its last reader first)
            if (!rot_slots_[i].busy) { pick = (int)i; break; }
        if (pick < 0) { rot_slots_.emplace_back(); pick = (int)rot_slots_.size() - 1; }
        RotSlot& r = rot_slots_[pick];
        if (r.d) {
            if (r.has_release) HX_CUDA(cudaEventSynchronize(r.released));
            HX_CUDA(cudaFree(r.d));
            r.d = nullptr; r.has_release = false;
        }
        HX_CUDA(cudaMalloc(&r.d, words * sizeof(u64)));
        r.words = words;
        if (!r.released) HX_CUDA(cudaEventCreateWithFlags(&r.released, cudaEventDisableTiming));
    }
    RotSlot& r = rot_slots_[pick];
    r.busy = true;
    if (r.has_release) HX_CUDA(cudaStreamWaitEvent(writer, r.released, 0));
    return pick;
}
void Engine::rot_slot_release(int slot) {
    RotSlot& r = rot_slots_[slot];
    HX_CUDA(cudaEventRecord(r.released, st_main));   // every reader of a rotation set runs on (or is joined into) the engine stream
    r.has_release = true;
    r.busy = false;
}
std::shared_ptr<Engine::RotSet> Engine::rotate_hoisted_range(const Ct& query, int d0, int d1) {
    Guard g(this);
    if (query->ncomp != 2 || query->nl != L || query->nb != 1) throw std::runtime_error("score_diag: query must be a fresh 2-component ciphertext");
    if (d0 < 0 || d1 <= d0) throw std::runtime_error("score_diag: empty or negative rotation range");
    const int ndiag = d1 - d0;
    const int nl = query->nl;
    const size_t ct_words = (size_t)2 * nl * N;
    auto rs = std::make_shared<RotSet>();
    rs->eng = this;
    rs->keep = weak_from_this().lock();
    rs->d0 = d0; rs->ndiag = ndiag; rs->nl = nl; rs->level = query->level; rs->deg = query->deg;
    const int chunk = std::max(1, std::min(rot_chunk, kMaxKsBatch));
    const int ns = rot_streams;
    cudaStream_t main_st = st;
    // a timing-enabled event record on the engine stream at the head of every rotation section and before every diagonal MAC
    // (mac_rotset): orders nothing new, but see the measurement quoted there
    if (!mac_gate_) HX_CUDA(cudaEventCreate(&mac_gate_));
    HX_CUDA(cudaEventRecord(mac_gate_, main_st));
    struct StRestore { Engine* e; cudaStream_t s; ~StRestore() { e->st = s; } } st_restore{this, main_st};   // also on exceptions
    const bool lane = rot_lane && query->settled && !force_unfused;
    if (lane) {
        // rotation lane: nothing below depends on what the engine stream still has queued (see hx_engine.h)
        for (int s = 0; s < ns; s++)
            if (!lane_streams[s]) HX_CUDA(cudaStreamCreateWithFlags(&lane_streams[s], cudaStreamNonBlocking));   // priority 0 = lowest
        for (int s = 0; s < 5; s++)
            if (!lane_events[s]) HX_CUDA(cudaEventCreateWithFlags(&lane_events[s], cudaEventDisableTiming));
        st = lane_streams[0];   // allocations, the copy of the query and its ModUp are ordered on the lane
    }
    // Everything from here to the join runs on the lane / auxiliary streams.  If anything throws (a missing rotation key after
    // some batches were queued, an allocation failure), those streams are drained before the exception leaves: the RotSet's
    // destructor and the digit buffer's release below order their frees on the engine stream only.
    u64* digits = nullptr;
    cudaStream_t first = st;
    cudaStream_t* others = lane ? lane_streams + 1 : aux_streams;
    cudaEvent_t* evs = lane ? lane_events : aux_events;
    std::unique_ptr<Scope> span;
    try {
        rs->slot = rot_slot_acquire(ct_words * ndiag, st);
        trace_mark("rot_begin", st);   // fires once the slot's last reader and the lane's earlier work are done
        u64* rotq = rs->rotq = rot_slots_[rs->slot].d;
        if (d0 == 0) HX_CUDA(cudaMemcpyAsync(rotq, query->d, sizeof(u64) * ct_words, cudaMemcpyDeviceToDevice, st));   // rotation 0 (sender_diag.cpp:21)
        int ndig;
        digits = ks_precompute(query->comp(1), nl, &ndig);
        ks_kind = PK_KS_HOISTED;
        // Independent batches may run on several streams so that the tail of one batch's kernels overlaps the next batch
        // (smaller batches keep their intermediates in L2).  Stream s > 0 waits for the digits, the first stream waits for all.
        // profiler: the whole section is timed as one span on its first stream (its batches overlap across streams);
        // the per-batch scopes inside only book launches and algorithmic bytes
        span.reset(new Scope(this, PK_KS_HOISTED, 0, 0.0));
        prof_no_events = true;
        if (ns > 1) {
            if (!lane) {
                for (int s = 0; s < ns - 1; s++)
                    if (!aux_streams[s]) {   // same priority as the engine stream: unequal priorities between the batch streams cost 30 %
                        int pr = 0;
                        HX_CUDA(cudaStreamGetPriority(main_st, &pr));
                        HX_CUDA(cudaStreamCreateWithPriority(&aux_streams[s], cudaStreamNonBlocking, pr));
                    }
                for (int s = 0; s < 4; s++)
                    if (!aux_events[s]) HX_CUDA(cudaEventCreateWithFlags(&aux_events[s], cudaEventDisableTiming));
            }
            HX_CUDA(cudaEventRecord(evs[3], first));
            for (int s = 0; s < ns - 1; s++) HX_CUDA(cudaStreamWaitEvent(others[s], evs[3], 0));
        }
        int turn = 0;
        for (int b0 = std::max(d0, 1); b0 < d1; b0 += chunk, turn++) {
            int b1 = std::min(d1, b0 + chunk);
            st = (ns > 1 && turn % ns) ? others[turn % ns - 1] : first;
            std::vector<const EvalKey*> keys;
            std::vector<u64> gal;
            for (int d = b0; d < b1; d++) {
                u64 k = galois_for_rotation(d);
                keys.push_back(&rot_key(k));
                gal.push_back(k);
            }
            ks_core_batch(digits, ndig, nl, keys, query->comp(0), 1, gal, rotq + ct_words * (b0 - d0), ct_words);
        }
    } catch (...) {
        prof_no_events = false;
        ks_kind = PK_KS_FUSED;
        cudaStreamSynchronize(first);
        for (int s = 0; s < ns - 1; s++)
            if (others[s]) cudaStreamSynchronize(others[s]);
        st = first;
        if (digits) free_words(digits);
        st = main_st;
        throw;
    }
    st = first;
    if (ns > 1)
        for (int s = 0; s < ns - 1; s++) {
            HX_CUDA(cudaEventRecord(evs[s], others[s]));
            HX_CUDA(cudaStreamWaitEvent(first, evs[s], 0));
        }
    prof_no_events = false;
    span.reset();
    free_words(digits);   // on `first`, behind every reader: scratch allocated on a stream is released on that stream
    trace_mark("rot_end", first);
    if (lane) {   // the engine stream joins: everything after this call (diagonal MAC, slot release) is ordered behind the lane
        HX_CUDA(cudaEventRecord(lane_events[4], first));
        HX_CUDA(cudaStreamWaitEvent(main_st, lane_events[4], 0));
    }
    st = main_st;
    ks_kind = PK_KS_FUSED;
    return rs;
}
Ct Engine::rotset_entry(const RotSet& rs, int i) {
    Guard g(this);
    if (i < 0 || i >= rs.ndiag) throw std::runtime_error("rotset_entry: index out of range");
    Ct c = new_ct(2, rs.nl, rs.level, rs.deg);
    const size_t ct_words = (size_t)2 * rs.nl * N;
    HX_CUDA(cudaMemcpyAsync(c->d, rs.rotq + ct_words * i, sizeof(u64) * ct_words, cudaMemcpyDeviceToDevice, st));
    return c;
}
Ct Engine::db_entry(int block, int diag) {
    Guard g(this);
    auto it = db_.find(block);
    if (it == db_.end() || diag < 0 || diag >= it->second.ndiag) throw std::runtime_error("db_entry: no such database ciphertext");
    Ct c = new_ct(2, L, 0, 1);
    const size_t ct_words = (size_t)2 * L * N;
    HX_CUDA(cudaMemcpyAsync(c->d, it->second.d + ct_words * diag, sizeof(u64) * ct_words, cudaMemcpyDeviceToDevice, st));
    std::vector<int> primes;
    for (int i = 0; i < L; i++) primes.push_back(i);
    Scope sc(this, PK_MISC, 1, 32.0 * L * N);
    launch_to_mont(T, c->d, 2 * L, limb_map(primes), st, true);
    return c;
}
// database pointers of the named blocks at the rotation set's first diagonal
static std::vector<const u64*> db_ptrs_for(const Engine& e, const std::map<int, Engine::DbBlock>& db, const Engine::RotSet& rs, const std::vector<int>& block_ids) {
    std::vector<const u64*> blocks;
    const size_t ct_words = (size_t)2 * e.L * e.N;
    for (int id : block_ids) {
        auto it = db.find(id);
        if (it == db.end()) throw std::runtime_error("score_diag: database block not registered");
        const Engine::DbBlock& b = it->second;
        if (rs.d0 < b.diag0 || rs.d0 + rs.ndiag > b.diag0 + b.ndiag) throw std::runtime_error("score_diag: database block does not hold every diagonal of the rotation set");
        blocks.push_back(b.d + ct_words * (rs.d0 - b.diag0));
    }
    return blocks;
}
Ct Engine::mac_rotset(const RotSet& rs, const std::vector<int>& block_ids) {
    Guard g(this);
    const int nl = rs.nl, ndiag = rs.ndiag;
    std::vector<const u64*> blocks = db_ptrs_for(*this, db_, rs, block_ids);
    if (blocks.empty()) throw std::runtime_error("score_diag: no database block named");
    Ct acc = new_ct(3, nl, rs.level, rs.deg + 1, (int)blocks.size());
    trace_mark("mac_begin", st);
    // An event record between the lane join (cudaStreamWaitEvent in rotate_hoisted_range) and the MAC launch.  It orders nothing
    // new; measured at 2 GPUs (32 blocks per rank) it removes a 20 ms stall of rank 0's second MAC in steps that end with a
    // device-to-host download (e2e 152 -> 134 ms per step; CUDA_DEVICE_MAX_CONNECTIONS=1 removes it too, 32 does not): DESIGN.md 7.
    if (!mac_gate_) HX_CUDA(cudaEventCreate(&mac_gate_));   // timing enabled: the record that was measured
    HX_CUDA(cudaEventRecord(mac_gate_, st));
    for (size_t g0 = 0; g0 < blocks.size(); g0 += kMaxDbBlocks) {   // the rotated query is fetched from HBM once per launch
        const int n = (int)std::min((size_t)kMaxDbBlocks, blocks.size() - g0);
        DbBlockPtrs ptrs;
        for (int i = 0; i < n; i++) ptrs.p[i] = blocks[g0 + i];
        Scope sc(this, PK_DIAG_MAC, 1, 8.0 * N * nl * 2 * ndiag * n);   // algorithmic bytes: the database diagonals, read once
        launch_diag_mac(T, rs.rotq, ptrs, n, acc->d + g0 * acc->batch_stride(), ndiag, nl, st);
    }
    trace_mark("mac_end", st);
    return acc;
}
Ct Engine::diag_mac_partial(const Ct& query, int d0, int d1, const std::vector<int>& block_ids) {
    Guard g(this);
    std::shared_ptr<RotSet> rs = rotate_hoisted_range(query, d0, d1);
    return mac_rotset(*rs, block_ids);
}
void Engine::fold_mod(const Ct& a) {
    Guard g(this);
    Scope sc(this, PK_EW, 1, 16.0 * N * a->nl * a->comps_total());
    launch_fold_mod(T, a->d, a->nl, a->comps_total(), st);
}
std::vector<Ct> Engine::score_rotset(const RotSet& rs, const std::vector<int>& block_ids) {
    Guard g(this);
    // all blocks of this GPU in one launch each: diagonal MAC (the rotated query is fetched from HBM once
    // for all of them), relinearisation and rescale on the batch [G][3][nl][N]
    std::vector<Ct> parts;
    const int group = (batch_blocks && !force_unfused && fused_ks_supported(logN)) ? std::min(kMaxKsBatch, kMaxDbBlocks) : 1;
    for (size_t g0 = 0; g0 < block_ids.size(); g0 += group) {
        const size_t n = std::min((size_t)group, block_ids.size() - g0);
        Ct acc = mac_rotset(rs, std::vector<int>(block_ids.begin() + g0, block_ids.begin() + g0 + n));
        parts.push_back(rescale(relinearize(acc)));
    }
    return parts;
}
std::vector<Ct> Engine::score_diag_batched(const Ct& query, int ndiag) {
    Guard g(this);
    if (db_.empty()) throw std::runtime_error("score_diag: no database block registered");
    std::shared_ptr<RotSet> rs = rotate_hoisted_all(query, ndiag);
    std::vector<int> ids;
    for (auto& kv : db_) ids.push_back(kv.first);
    return score_rotset(*rs, ids);
}
std::vector<Ct> Engine::score_diag(const Ct& query, int ndiag) {
    std::vector<Ct> out;
    for (auto& p : score_diag_batched(query, ndiag))
        for (auto& c : unstack(p)) out.push_back(c);
    return out;
}

// indexScenario (sender_diag.cpp:52-63): the compare chain runs once per batch of blocks.
std::vector<Ct> Engine::index(const Ct& query, int ndiag, double delta, int depth) {
    std::vector<Ct> out;
    for (auto& p : score_diag_batched(query, ndiag))
        for (auto& c : unstack(chebyshev_compare(p, delta, depth))) out.push_back(c);
    return out;
}
// membershipScenario (sender_diag.cpp:35-50).
Ct Engine::membership(const Ct& query, int ndiag, double delta, int depth) {
    std::vector<Ct> s = index(query, ndiag, delta, depth);
    Ct m = add_many(s);
    return eval_sum(m, N / 2);
}

// HersSender::computeSimilarity (sender_hers.cpp:13-26, 61-98): per matrix, sum over the coordinates i of
// Rescale(Relinearize(EvalMultNoRelin(query[i], db[matrix][i]))).  The coordinates run in batches of 32: one
// tensor-product launch against the HBM-resident database, one batched relinearisation (same EvalMultKey, 32
// digit sets), one batched rescale, one modular sum over the batch.  Modular addition is associative and
// commutative, so the result equals the reference's sequential EvalAddInPlace chain bit for bit.
std::vector<Ct> Engine::score_hers(const std::vector<Ct>& query, const std::vector<int>* matrices) {
    Guard g(this);
    if (query.empty()) throw std::runtime_error("score_hers: empty query");
    if (db_.empty()) throw std::runtime_error("score_hers: no database matrix registered");
    const int dim = (int)query.size(), nl = query[0]->nl;
    for (auto& q : query)
        if (q->ncomp != 2 || q->nl != nl || q->nb != 1 || q->deg != query[0]->deg) throw std::runtime_error("score_hers: query ciphertexts differ in shape");
    if (nl != L) throw std::runtime_error("score_hers: query must be fresh (database ciphertexts are stored at full level)");
    if (!(fused_ks_supported(logN) && !force_unfused)) throw std::runtime_error("score_hers needs the fused pipeline");
    const size_t cs = (size_t)nl * N;
    std::vector<Ct> out;
    std::vector<std::pair<int, DbBlock>> todo;
    if (matrices) {
        for (int id : *matrices) {
            auto it = db_.find(id);
            if (it == db_.end()) throw std::runtime_error("score_hers: database matrix not registered");
            todo.emplace_back(it->first, it->second);
        }
    } else {
        for (auto& kv : db_) todo.emplace_back(kv.first, kv.second);
    }
    for (auto& kv : todo) {
        if (kv.second.ndiag < dim) throw std::runtime_error("score_hers: database matrix has fewer columns than the query");
        Ct acc = new_ct(2, nl - 1, query[0]->level + 1, query[0]->deg + 1 - 1);
        for (int i0 = 0; i0 < dim; i0 += kMaxKsBatch) {
            const int n = std::min(kMaxKsBatch, dim - i0);
            CtPtrs ap;
            for (int i = 0; i < n; i++) ap.p[i] = query[i0 + i]->d;
            Ct t = new_ct(3, nl, query[0]->level, query[0]->deg + 1, n);
            {   // algorithmic bytes: database ciphertexts once, query ciphertexts once, products once
                Scope sc(this, PK_TENSOR, 1, 8.0 * N * nl * n * (2 + 2 + 3));
                launch_tensor_db(T, t->d, ap, cs, kv.second.d + (size_t)i0 * 2 * cs, cs, nl, n, st);
            }
            Ct r = rescale(relinearize(t));
            {
                Scope sc(this, PK_EW, 1, 8.0 * N * r->nl * 2 * (n + 1));
                launch_batch_sum(T, acc->d, r->d, r->nl, 2, n, i0 != 0, st);
            }
        }
        out.push_back(acc);
    }
    return out;
}

std::vector<Ct> Engine::compare_many(const std::vector<Ct>& cts, double delta, int depth) {
    Guard g(this);
    std::vector<Ct> out;
    const bool batch = batch_blocks && !force_unfused && fused_ks_supported(logN);
    for (size_t i0 = 0; i0 < cts.size(); i0 += kMaxKsBatch) {
        const size_t n = std::min((size_t)kMaxKsBatch, cts.size() - i0);
        std::vector<Ct> part(cts.begin() + i0, cts.begin() + i0 + n);
        if (batch && n > 1) {
            for (auto& c : unstack(chebyshev_compare(stack(part), delta, depth))) out.push_back(c);
        } else {
            for (auto& c : part) out.push_back(chebyshev_compare(c, delta, depth));
        }
    }
    return out;
}

}  // namespace hx
