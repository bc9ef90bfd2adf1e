/*
 * hydia_b200.h -- C ABI of libhydia_b200.so: the B200-native CKKS evaluation engine behind the
 * HyDia Sender hot path (n7koirala/image_matching).  Plain pointers and sizes only; opaque handles.
 *
 * Every entry point names the reference interface it replaces (paths under /root/reference).
 * The reference reaches all of them through lbcrypto::CryptoContext<DCRTPoly> (OpenFHE v1.2.3);
 * INTEGRATION.md shows the C++ shim a maintainer adds so src/sender/ *.cpp and
 * src/openFHE_wrapper.cpp call these instead.
 *
 * Conventions
 *  - Polynomials are uint64 arrays, limb-major [component][limb][N], EVALUATION (NTT) form,
 *    canonical residues, OpenFHE limb order (q_0..q_{L-1}, then p_0..p_{K-1} for keys).
 *  - Functions returning a handle return NULL on error; functions returning int return 0 on
 *    success and non-zero on error; hx_last_error() gives the message (thread-local).  OpenFHE
 *    throws OpenFHEException in the same situations (missing key, level exhausted).
 *  - Inputs are never consumed or modified ("InPlace" reference calls become: call, then rebind).
 *  - All calls are thread-safe per context (the reference enters from up to 48 OpenMP threads,
 *    include/config.h:21, e.g. src/sender/sender_diag.cpp:23-26).
 */
#ifndef HYDIA_B200_H
#define HYDIA_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct hx_ctx hx_ctx;        /* CryptoContext<DCRTPoly>            (include/sender.h:39) */
typedef struct hx_ct hx_ct;          /* Ciphertext<DCRTPoly>                                      */
typedef struct hx_digits hx_digits;  /* shared_ptr<vector<DCRTPoly>> from EvalFastRotationPrecompute */

const char* hx_last_error(void);
int hx_device_count(void);

/* ---- context ----------------------------------------------------------------------------- */
/* GenCryptoContext result (src/main.cpp:169-175): ring dimension, RNS chain and roots as OpenFHE
 * generated them; dnum = number of hybrid key-switching digits. */
hx_ctx* hx_ctx_create(int N, int L, int K, int dnum, const uint64_t* q, const uint64_t* p, const uint64_t* psi_q,
                      const uint64_t* psi_p, int log2_scale, int device);
void hx_ctx_destroy(hx_ctx* ctx);
int hx_ctx_sync(hx_ctx* ctx);                       /* wait for all queued device work */
uint64_t hx_ctx_launches(hx_ctx* ctx);              /* kernels launched so far (bench accounting) */
/* Hash of the kernel sources this library was built from (csrc Makefile): committed ncu summaries under profiles/ carry the stamp
 * they were captured at, and bench.py ignores them when it differs. */
const char* hx_kernel_stamp(void);
/* Engine tuning / test switches: "rot_chunk" (hoisted rotations per key-switch batch), "rot_streams" (streams the
 * independent batches are issued on, default 2: the tail of one batch overlaps the head of the next),
 * "force_unfused" (1 = run the unfused reference kernels instead of the fused key-switch pipeline),
 * "ks_moddown_w" (default 1: on rings of 2^15 / 2^16 coefficients ModDown's coefficient-domain middle runs as warp tasks in one
 * kernel, hx_moddown.cu; 0 = the two-launch tile sequence every other ring uses; same bits either way),
 * "pdl" (default 1: dependent kernels are launched with programmatic stream serialization; 0 = plain stream order),
 * "rot_lane" (default 1: the hoisted rotations of an uploaded query are issued on low-priority side streams that do not wait for
 * work still queued on the engine stream, so the previous call's compare chain overlaps them; 0 = one ordered stream),
 * "batch_blocks" (default 1: the fused HyDia entries relinearise / rescale / compare all database blocks of
 * this GPU as one batch; 0 = one chain per block). */
int hx_ctx_set_option(hx_ctx* ctx, const char* name, int value);
/* Launch accounting and the optional CUDA-event profiler (events on the engine's launching stream).
 * The reference has only wall-clock phase timers (src/main.cpp:342-347); these feed bench.py's
 * roofline object.  kind in [0, hx_prof_kinds()). bytes = algorithmic bytes of the launches so far. */
int hx_prof_kinds(void);
const char* hx_prof_name(int kind);
int hx_prof_enable(hx_ctx* ctx, int mask);   /* bit k times kernel family k; -1 = all; 0 = off */
int hx_prof_reset(hx_ctx* ctx);
int hx_prof_read(hx_ctx* ctx, int kind, uint64_t* launches, double* ms, double* bytes);
/* device timer on the engine's stream: start records an event, stop records + waits, returns ms */
int hx_timer_start(hx_ctx* ctx);
int hx_timer_stop(hx_ctx* ctx, double* ms);
/* FindAutomorphismIndex2nComplex: Galois element of a slot rotation (5^r or 5^-|r| mod 2N) */
uint32_t hx_galois_for_rotation(hx_ctx* ctx, int rotation);

/* ---- evaluation keys (src/main.cpp:188-206) ------------------------------------------------ */
/* EvalKey a/b vectors, [dnum][L+K][N] each, plain residues; converted to the engine's layout. */
int hx_key_load_mult(hx_ctx* ctx, const uint64_t* a, const uint64_t* b);                      /* EvalMultKeyGen */
int hx_key_load_auto(hx_ctx* ctx, uint32_t galois, const uint64_t* a, const uint64_t* b);     /* EvalRotateKeyGen / EvalSumKeyGen */

/* ---- ciphertexts --------------------------------------------------------------------------- */
hx_ct* hx_ct_upload(hx_ctx* ctx, int ncomp, int nlimbs, const uint64_t* data, int level, int noise_deg);
int hx_ct_download(hx_ct* ct, uint64_t* data);      /* data: [ncomp][nlimbs][N] */
/* same, with DEVICE pointers on the context's GPU (NCCL staging buffers of the sharded sender) */
hx_ct* hx_ct_upload_dev(hx_ctx* ctx, int ncomp, int nlimbs, const void* dev_data, int level, int noise_deg);
int hx_ct_download_dev(hx_ct* ct, void* dev_data);
/* The same two, stream-ordered instead of synchronising the host: `stream` is the caller's cudaStream_t (NULL = the legacy default
 * stream), e.g. the stream its NCCL collectives are issued from.  The copy runs after the work queued on `stream` so far and `stream`
 * waits for the copy.  Used by the multi-GPU sender so that a scenario's result combine does not stall the host and the
 * next scenario's hoisted rotations overlap the previous compare chain ("rot_lane"); the uploaded ciphertext is not a lane candidate. */
int hx_ct_download_dev_async(hx_ct* ct, void* dev_data, void* stream);
hx_ct* hx_ct_upload_dev_async(hx_ctx* ctx, int ncomp, int nlimbs, const void* dev_data, int level, int noise_deg, void* stream);
int hx_ct_info(hx_ct* ct, int* ncomp, int* nlimbs, int* level, int* noise_deg);
hx_ct* hx_ct_clone(hx_ct* ct);
/* n ciphertexts back to back into one host buffer, one synchronisation (the G result ciphertexts of indexScenario) */
int hx_ct_download_many(hx_ct* const* cts, int n, uint64_t* data);
void hx_ct_free(hx_ct* ct);

/* ---- primitive evaluation ops (call sites: SURVEY.md section 8(b)) ---------------------------- */
hx_ct* hx_add(hx_ct* a, hx_ct* b);                  /* EvalAdd / EvalAddInPlace   sender_diag.cpp:76 */
hx_ct* hx_sub(hx_ct* a, hx_ct* b);                  /* EvalSub                                       */
hx_ct* hx_add_const(hx_ct* a, double c);            /* EvalAddInPlace(ct,double)  openFHE_wrapper.cpp:182 */
hx_ct* hx_mult_const(hx_ct* a, double c);           /* EvalMult(ct,double)                           */
hx_ct* hx_mult_norelin(hx_ct* a, hx_ct* b);         /* EvalMultNoRelin            sender_diag.cpp:93 */
hx_ct* hx_relin(hx_ct* a);                          /* RelinearizeInPlace         sender_diag.cpp:79 */
hx_ct* hx_mult(hx_ct* a, hx_ct* b);                 /* EvalMult(ct,ct)            sender_hers.cpp:156 */
hx_ct* hx_square(hx_ct* a);                         /* EvalSquare(InPlace)        sender_hers.cpp:124 */
/* EvalMult(ct, Plaintext) (openFHE_wrapper.cpp:235,244,289; sender_hers.cpp:110): pt = the encoded plaintext of
 * MakeCKKSPackedPlaintext, [nlimbs][N] evaluation form, nlimbs >= limbs of a (the first ones are used). */
hx_ct* hx_mult_plain(hx_ct* a, const uint64_t* pt, int nlimbs);
hx_ct* hx_inner_product(hx_ct* a, hx_ct* b, int n); /* EvalInnerProduct           sender_base.cpp:93  */
hx_ct* hx_rescale(hx_ct* a);                        /* RescaleInPlace             sender_diag.cpp:80 */
hx_ct* hx_level_reduce(hx_ct* a, int levels);       /* LevelReduceInPlace                            */
hx_ct* hx_rotate(hx_ct* a, int rotation);           /* EvalRotate                 openFHE_wrapper.cpp:124 */
hx_ct* hx_eval_sum(hx_ct* a, int n);                /* EvalSum                    sender_diag.cpp:47 */
hx_ct* hx_add_many(hx_ct* const* cts, int n);       /* EvalAddManyInPlace         sender_diag.cpp:46 */
hx_digits* hx_rot_precompute(hx_ct* a);             /* EvalFastRotationPrecompute sender_diag.cpp:22 */
hx_ct* hx_rot_fast(hx_ct* a, int rotation, hx_digits* digits);   /* EvalFastRotation sender_diag.cpp:25 */
void hx_digits_free(hx_digits* d);
/* EvalChebyshevSeries on precomputed coefficients / EvalPoly (openFHE_wrapper.cpp:174,179) */
hx_ct* hx_cheby_series(hx_ct* a, const double* coeffs, int n, double lo, double hi);
hx_ct* hx_poly(hx_ct* a, const double* coeffs, int n);
/* OpenFHEWrapper::chebyshevCompare (openFHE_wrapper.cpp:143-185) */
hx_ct* hx_compare(hx_ct* a, double delta, int sign_depth);

/* ---- fused HyDia Sender path (src/sender/sender_diag.cpp) ---------------------------------- */
/* Registers database ciphertext serial/db_diagonal/index{block*512+diag}.bin (sender_diag.cpp:87)
 * as HBM resident; data: [2][L][N]. */
int hx_db_register(hx_ctx* ctx, int block, int diag, const uint64_t* data);
int hx_db_blocks(hx_ctx* ctx);
/* DiagonalSender::computeSimilarity (sender_diag.cpp:12): one score ciphertext per block. */
int hx_score_diag(hx_ctx* ctx, hx_ct* query, int ndiag, hx_ct** out, int max_out);
/* DiagonalSender::membershipScenario (sender_diag.cpp:35) */
hx_ct* hx_membership(hx_ctx* ctx, hx_ct* query, int ndiag, double delta, int sign_depth);
/* DiagonalSender::indexScenario (sender_diag.cpp:52) */
int hx_index(hx_ctx* ctx, hx_ct* query, int ndiag, double delta, int sign_depth, hx_ct** out, int max_out);

/* Multi-GPU building blocks (image_matching_b200/sharded.py).  The 512 diagonals of EVERY block are sharded over
 * the ranks (context options "db_diag0" / "db_diag_count", set before registering): a rank computes only its own hoisted rotations
 * (sender_diag.cpp:20-26 restricted to d0 <= i < d1) and the partial sum over its diagonals of rotQ[i] (x) DB[g][i] for all blocks
 * (sender_diag.cpp:70-77), one batched 3-component ciphertext [nblocks][3][L][N].  The ranks' partial sums are added by a
 * reduce-scatter on the raw 64-bit words (world * q < 2^63), hx_ct_fold_mod brings the words back to residues, and
 * hx_relin / hx_rescale / hx_compare run on each rank's share of the blocks.  Modular addition is associative, so the result is
 * bit-identical to the single-GPU sum.  hx_ct_alloc / hx_ct_device_view / hx_stream_order let the collective read and write
 * ciphertext storage in place on a caller-owned stream (engine_waits = 0: `stream` waits for the engine stream; 1: the reverse). */
hx_ct* hx_diag_mac_partial(hx_ctx* ctx, hx_ct* query, int d0, int d1, const int* blocks, int nblocks);
int hx_ct_fold_mod(hx_ct* ct);
hx_ct* hx_ct_alloc(hx_ctx* ctx, int ncomp, int nlimbs, int level, int noise_deg, int nbatch);
int hx_ct_device_view(hx_ct* ct, void** dev_ptr, uint64_t* words, int* nbatch);
int hx_stream_order(hx_ctx* ctx, void* stream, int engine_waits);

/* The same path in two steps, for callers that hold the database as ciphertext handles and score blocks one at a time
 * (the deferred-execution layer of shim/openfhe.h): hx_rot_fast_all = sender_diag.cpp:20-26 (rotations 0..ndiag-1 of the
 * query, hoisted, batched; entry 0 is the query), hx_score_blocks = computeSimilarityMatrix (sender_diag.cpp:66-83) for the
 * listed blocks against that set (out: one ciphertext per block). hx_db_register_ct stores a device-resident ciphertext as
 * serial/db_diagonal/index{block*512+diag}.bin; hx_db_get returns a plain copy of a stored one. */
typedef struct hx_rotset hx_rotset;
int hx_db_register_ct(hx_ctx* ctx, int block, int diag, hx_ct* ct);
hx_ct* hx_db_get(hx_ctx* ctx, int block, int diag);
hx_rotset* hx_rot_fast_all(hx_ct* query, int ndiag);
hx_ct* hx_rotset_get(hx_rotset* r, int i);
void hx_rotset_free(hx_rotset* r);
int hx_score_blocks(hx_rotset* r, const int* blocks, int nblocks, hx_ct** out);

/* ---- batched ciphertexts ---------------------------------------------------------------------------------- */
/* A batch is n ciphertexts of one shape and level behind one handle. Every evaluation entry above accepts a batch and
 * applies the operation to each entry with the same constants / plaintext / keys in one launch per kernel
 * (a single ciphertext multiplied with a batch is broadcast). This is how the "#pragma omp parallel for" loops over
 * ciphertexts in src/sender/ *.cpp and src/openFHE_wrapper.cpp:199-202, 287-292 map onto the GPU. */
hx_ct* hx_ct_stack(hx_ct* const* cts, int n);            /* copies; inputs stay valid */
int hx_ct_unstack(hx_ct* ct, hx_ct** out, int max_out);  /* returns the number of entries */

/* ---- other senders ------------------------------------------------- */
/* HersSender::computeSimilarity (src/sender/sender_hers.cpp:13-26, 61-98): query = nquery ciphertexts, one per
 * coordinate (HersReceiver::encryptQuery, src/receiver/receiver_hers.cpp:13-24); database matrix m, column j =
 * serial/db_hers/matrix{m}/index{j}.bin, registered with hx_db_register(ctx, m, j, data).  Per matrix:
 * sum_j Rescale(Relinearize(EvalMultNoRelin(query[j], db[m][j]))).  Returns the number of matrices. */
int hx_score_hers(hx_ctx* ctx, hx_ct* const* query, int nquery, hx_ct** out, int max_out);
/* same for the listed matrices only (computeSimilarityHelper, sender_hers.cpp:61-86); out: nmatrices handles */
int hx_score_hers_blocks(hx_ctx* ctx, hx_ct* const* query, int nquery, const int* matrices, int nmatrices, hx_ct** out);
/* OpenFHEWrapper::chebyshevCompare over a vector of score ciphertexts (the "#pragma omp parallel for" loops of
 * sender_hers.cpp:33-36, sender_diag.cpp:40-43, ...): all n inputs go through one batched chain. out: n handles. */
int hx_compare_many(hx_ctx* ctx, hx_ct* const* in, int n, double delta, int sign_depth, hx_ct** out);

/* ---- building blocks exposed for parity tests ---------------------------------------------- */
/* DCRTPoly::SetFormat on a host buffer [nbatch][N]; primes[b % nprimes] = prime index of entry b */
int hx_ntt(hx_ctx* ctx, uint64_t* data, int nbatch, const int* primes, int nprimes, int inverse);

/* ---- synthetic client side (keys / inputs for smoke, bench and tests; not on the Sender path) */
/* The parameter recipe of GenCryptoContext(HEStd_128_classic, depth, scale 2^45, FIXEDMANUAL)
 * (src/main.cpp:169-175). ring_dim = 0 selects the smallest standard ring. Arrays sized >= 40. */
int hx_params_generate(int depth, int scale_bits, int first_mod_bits, int ring_dim, int num_large_digits, int* N, int* L, int* K,
                       int* dnum, uint64_t* q, uint64_t* p, uint64_t* psi_q, uint64_t* psi_p);
typedef struct hx_client hx_client;  /* KeyGen + Enroller/Receiver crypto (src/main.cpp:184-206) */
hx_client* hx_client_create(hx_ctx* ctx, uint64_t seed);
void hx_client_destroy(hx_client* c);
int hx_client_gen_mult_key(hx_client* c);                                   /* EvalMultKeyGen */
int hx_client_gen_rot_keys(hx_client* c, const int* rotations, int n);      /* EvalRotateKeyGen */
int hx_client_gen_sum_keys(hx_client* c);                                   /* EvalSumKeyGen  */
/* MakeCKKSPackedPlaintext + Encrypt (openFHE_wrapper.cpp:74-77); values: up to N/2 reals */
hx_ct* hx_client_encrypt(hx_client* c, const double* values, int n);
/* MakeCKKSPackedPlaintext (openFHE_wrapper.cpp:75, :257-271 masks): out = [L][N] residues, evaluation form */
int hx_client_encode(hx_client* c, const double* values, int n, uint64_t* out);
/* Encrypt + register as database ciphertext (DiagonalEnroller::serializeDBThread, enroller_diag.cpp:158) */
int hx_client_enroll(hx_client* c, int block, int diag, const double* values, int n);
/* Decrypt + GetRealPackedValue (openFHE_wrapper.cpp:81-85); out: N/2 doubles */
int hx_client_decrypt(hx_client* c, hx_ct* ct, double* out);
/* secret key (ternary, N int8) for cross-checks against the CPU oracle in tests */
int hx_client_secret(hx_client* c, int8_t* out);
/* (De)serialisation support for shim/hx_serial.h (Serial::SerializeToFile / DeserializeFromFile and cc->Serialize...Key of
 * /root/reference/src/main.cpp:124-165, 251-285): the context's parameters, the loaded evaluation keys as plain residues
 * [dnum][L+K][N] (galois 0 = the EvalMult key; the reverse direction is hx_key_load_mult / hx_key_load_auto) and the key pair
 * (secret as coefficients in {-1,0,1}, public key (b, a) as residues [L][N]). */
int hx_ctx_get_params(hx_ctx* ctx, int* N, int* L, int* K, int* dnum, int* log2_scale, uint64_t* q, uint64_t* p, uint64_t* psi_q, uint64_t* psi_p);
int hx_key_count(hx_ctx* ctx, int* has_mult);                        /* number of automorphism keys loaded */
int hx_key_list(hx_ctx* ctx, uint32_t* galois, int max_out);         /* their Galois elements */
int hx_key_export(hx_ctx* ctx, uint32_t galois, uint64_t* a, uint64_t* b);
int hx_client_export_public(hx_client* c, uint64_t* pkb, uint64_t* pka);
hx_client* hx_client_import(hx_ctx* ctx, uint64_t seed, const int8_t* sk, const uint64_t* pkb, const uint64_t* pka);

/* ---- timeline marks (diagnostic, no counterpart in the reference): named CUDA events on a stream (NULL = the engine stream),
 * dumped as "name:ms;name:ms;..." relative to hx_trace_start; hx_trace_dump waits for the marks' events (no device-wide synchronisation) and ends the trace.  The engine
 * marks its own rotation lane (rot_begin / rot_end) and diagonal MAC (mac_begin / mac_end) while a trace is on. */
int hx_trace_start(hx_ctx* ctx);
int hx_trace_mark(hx_ctx* ctx, const char* name, void* stream);
int hx_trace_dump(hx_ctx* ctx, char* out, int cap);

/* ---- allocation-free variant ----------------------------------------------------------------------------------------
 * No counterpart in the reference: added for callers that cannot call cudaMalloc (code
 * under CUDA graph capture, tight serving loops, harnesses that forbid it).  hx_rescale_noalloc is hx_rescale (RescaleInPlace,
 * the same four kernels) on device buffers the caller owns: d_in = [ncomp][nlimbs][N] residues in evaluation form, d_out =
 * [ncomp][nlimbs-1][N], both on the context's GPU and not overlapping.  It calls no cudaMalloc / cudaFree / allocating library
 * routine and creates no handle: its scratch (the dropped limb's coefficients, the converted limbs, the 1-limb conversion table)
 * is a file-scope __device__ array of HX_NOALLOC_MAX_WORDS 64-bit words (+ 4 KiB) that every GPU loading the library holds until
 * the process exits.  ncomp * nlimbs * N above that limit is rejected with an error (use hx_rescale).  The work is queued on
 * the context's engine stream like every other call (hx_ctx_sync / hx_stream_order to consume d_out elsewhere).  The array is
 * shared by all calls and contexts on a device and keeps the last call's contents; nothing in it needs clearing, but calls that
 * overlap on the GPU (two contexts on one device) race on it. */
#define HX_NOALLOC_MAX_WORDS (1u << 21) /* 16 MiB: 2 components x 16 limbs at N = 2^16, 2 x 32 limbs at N = 2^15 */
int hx_rescale_noalloc(hx_ctx* ctx, const uint64_t* d_in, int ncomp, int nlimbs, uint64_t* d_out);

#ifdef __cplusplus
}
#endif
#endif /* HYDIA_B200_H */
