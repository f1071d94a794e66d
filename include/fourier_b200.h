/* fourier_b200.h - C ABI of the B200-native KZG / Pianist hot path.
 *
 * Drop-in boundary for apollozkp/fourier's engine: every entry point below
 * replaces one call the reference makes from src/engine/piano.rs into the
 * rust-kzg-blst / blst crates (the file:line after "replaces").  A Rust `-sys`
 * crate, cgo, JNI or ctypes binds these symbols directly: plain pointers and
 * sizes, caller-allocated outputs, `int` status (0 = ok) with the message
 * available from fb_last_error().
 *
 * Encodings at the boundary are the reference's wire encodings:
 *   Fr  : 32 bytes big-endian, canonical (< r) else FB_ERR_SCALAR ("Invalid scalar")
 *   G1  : 48 bytes ZCash-compressed (FsG1::to_bytes / from_bytes), or 96 bytes
 *         uncompressed (FsG1::serialize / deserialize) where stated
 * Montgomery limbs, affine/XYZZ points and window tables never leave the library.
 *
 * A context is bound to ONE CUDA device (one process per GPU); it owns the SRS,
 * the window tables, the twiddle tables and all scratch memory in HBM.  The SRS
 * and its tables are read-only after setup; every call runs on one of a small
 * pool of lanes (a CUDA stream with its own scratch, error flags and pinned
 * buffers), so concurrent callers (the reference keeps the backend in an Arc and
 * runs one tokio task per connection, src/rpc.rs:165, 414-416, 502) are safe AND
 * overlap on the GPU.  Setup calls are exclusive.
 * There is no CPU fallback: without a usable CUDA device fb_ctx_create fails.
 */
#ifndef FOURIER_B200_H
#define FOURIER_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct fb_ctx fb_ctx;

enum {
    FB_OK = 0,
    FB_ERR_CUDA = 1,     /* CUDA runtime / launch failure */
    FB_ERR_ARG = 2,      /* bad argument (row index, length not a power of two, ...) */
    FB_ERR_SCALAR = 3,   /* "Invalid scalar": an Fr encoding >= r */
    FB_ERR_POINT = 4,    /* a G1 encoding that does not decode to a curve point */
    FB_ERR_STATE = 5,    /* SRS not loaded / wrong shard */
    FB_ERR_IO = 6        /* file errors */
};

/* ---- context -------------------------------------------------------------
 * replaces PianoBackend::new + PianoFFTSettings::new (piano.rs:82-85, 1061-1077):
 * scale n (total 2^n coefficients), machines_scale m (2^m rows of 2^(n-m)).
 * `device` is the CUDA ordinal.  Fails with FB_ERR_ARG if m > n. */
int fb_ctx_create(uint32_t n, uint32_t m, int device, fb_ctx** out);
void fb_ctx_destroy(fb_ctx* ctx);
/* message of the CALLING THREAD's last failing call (thread-local: another thread's
 * failure cannot overwrite it); valid until that thread's next failing call */
const char* fb_last_error(const fb_ctx* ctx);
const char* fb_version(void);
/* the caller's CUDA stream S (cudaStream_t passed as void*), e.g. torch's current
 * stream; NULL = a stream of the context.  The device-resident calls (fb_dev_*)
 * are ordered against S: they see everything enqueued on S before the call, and
 * with the default fb_ctx_set_lanes(1) they simply run on S. */
int fb_ctx_set_stream(fb_ctx* ctx, void* cuda_stream);
/* block the host until every lane and S are idle.  The device-resident decoders (fb_dev_fr_from_be,
 * fb_dev_combine_commit_open, the fb_dev_g1_* codecs) cannot fail synchronously: a scalar >= r or a malformed point sets
 * a sticky flag, and this call returns FB_ERR_SCALAR / FB_ERR_POINT for it (once; the flag is cleared). */
int fb_ctx_sync(fb_ctx* ctx);
/* lanes >= 2: the heavy device-resident calls (fb_dev_worker_commit / _open /
 * _commit_open, fb_dev_commit_rows / _open_rows, fb_dev_msm*, and the _partial
 * forms) rotate over `lanes` internal streams forked from S and return at once,
 * so that e.g. the commit and the open of one step overlap on the GPU.  Their
 * outputs are ordered on S only after fb_ctx_join (a device-side wait, the host
 * does not block) or fb_ctx_sync.  lanes = 1 (default): plain stream order on S.
 * Host calls always take a free lane of their own (1..4 concurrent callers). */
int fb_ctx_set_lanes(fb_ctx* ctx, int lanes);
int fb_ctx_join(fb_ctx* ctx);
/* window bits for the resident tables built by the next SRS load/generate
 * (0 = pick from the row length; -1 = no tables: plain Pippenger per call) */
int fb_ctx_set_window(fb_ctx* ctx, int window_bits);
/* window bits / number of windows of the resident row tables as built (0, 0: no tables) */
int fb_ctx_get_window(fb_ctx* ctx, int* window_bits, int* num_windows);
/* how fb_worker_open forms the quotient's evaluations: 0 (default) straight from the
 * evaluations (no transform; identical field elements), 1 the reference's literal
 * sequence iNTT -> eval -> divide -> NTT (piano.rs:193-236).  Same bytes either way. */
int fb_ctx_set_open_mode(fb_ctx* ctx, int mode);
/* parts >= 2: every large MSM runs as `parts` independent sub-MSMs (point ranges, or row ranges of a batch) on
 * `parts` streams of its lane, so that one sub-MSM's latency-bound phases run under another's multiplier-bound
 * ones; the partial sums are added on the device.  Same bytes as parts = 1 (default). */
int fb_ctx_set_msm_split(fb_ctx* ctx, int parts);
/* how the two MSMs of a commit + open of one row run (fb_*_commit_open*): 0 (default) by size - rows of >= 2^18 points as
 * two MSMs on two streams of the call's lane, smaller ones as one batch-of-two pass; 1 always the batch; 2 always two
 * streams.  Same bytes. */
int fb_ctx_set_pair_mode(fb_ctx* ctx, int mode);
/* uploads from PAGEABLE host memory (a Rust Vec<u8>, Python bytes): 0 (default) copy
 * through the lane's pinned double buffer, 1 hand the pointer to cudaMemcpyAsync.
 * Pinned / registered sources are always copied directly. */
int fb_ctx_set_staging(fb_ctx* ctx, int mode);

/* Column sharding (one MSM split over GPUs by point ranges): keep only columns
 * [col0, col0+ncols) of every resident row.  Takes effect at the next SRS
 * generate / load; ncols = 0 restores whole rows.  A column-sharded context
 * serves the *_partial calls below; the 96-byte partial sums of all shards are
 * added with fb_g1_sum (format 96). */
int fb_ctx_set_col_shard(fb_ctx* ctx, uint64_t col0, uint64_t ncols);

/* ---- SRS -------------------------------------------------------------------
 * Rows [row0, row0+nrows) of U are kept resident on this context's device
 * (nrows = 0 means all 2^m rows).  Sharding rows over GPUs = one context per GPU
 * with disjoint row ranges.
 *
 * fb_srs_generate replaces generate_trusted_setup (piano.rs:883-1037): tau_X,
 * tau_Y = the 32-byte secrets read big-endian mod r; U[i][j] = g^(R_i(tau_Y) *
 * L_j(tau_X)), g_tau_y[k] = g^(tau_Y^k).  Lagrange values use the closed form
 * (identical field elements to the reference's iFFT + Horner). */
int fb_srs_generate(fb_ctx* ctx, const uint8_t secret_x[32], const uint8_t secret_y[32], uint64_t row0,
                    uint64_t nrows);
/* replaces PianoSettings::load_setup_from_file (piano.rs:711-846); G1 points are
 * decoded on the GPU.  compressed != 0: 48/96-byte elements, else 96/192. */
int fb_srs_load_file(fb_ctx* ctx, const char* path, int compressed, uint64_t row0, uint64_t nrows);
int fb_srs_load_bytes(fb_ctx* ctx, const uint8_t* buf, size_t len, int compressed, uint64_t row0, uint64_t nrows);
/* replaces PianoSettings::save_setup_to_file (piano.rs:649-706); needs all rows resident */
int fb_srs_save_file(fb_ctx* ctx, const char* path, int compressed);
/* size of / copy of the serialised setup (same layout as the file) */
int fb_srs_to_bytes(fb_ctx* ctx, int compressed, uint8_t* out, size_t cap, size_t* written);

/* ---- engine: worker side -----------------------------------------------------
 * fb_worker_commit replaces PianoBackend::worker_commit (piano.rs:143-161):
 * out = sum_j coeffs[j] * U[i][j].  len <= 2^(n-m). */
int fb_worker_commit(fb_ctx* ctx, uint64_t i, const uint8_t* coeffs_be, uint64_t len, uint8_t out48[48]);
/* replaces PianoBackend::worker_open (piano.rs:187-238): p = iNTT(coeffs),
 * y = p(alpha), q = (p - y)/(X - alpha), proof = sum_j NTT(q||0)[j] * U[i][j]. */
int fb_worker_open(fb_ctx* ctx, uint64_t i, const uint8_t* coeffs_be, uint64_t len, const uint8_t alpha_be[32],
                   uint8_t out_y_be[32], uint8_t out_proof48[48]);
/* batched rows [row0, row0+nrows): coeffs is nrows*len Fr, outputs nrows entries */
int fb_commit_rows(fb_ctx* ctx, uint64_t row0, uint64_t nrows, const uint8_t* coeffs_be, uint64_t len,
                   uint8_t* out48s);
int fb_open_rows(fb_ctx* ctx, uint64_t row0, uint64_t nrows, const uint8_t* coeffs_be, uint64_t len,
                 const uint8_t alpha_be[32], uint8_t* out_ys_be, uint8_t* out_proof48s);

/* commit + open of rows [row0, row0+nrows) in one call (extension): the matrix goes up once, the commit MSMs overlap the
 * field half and the proof MSMs.  Byte-identical to fb_commit_rows + fb_open_rows.  len must be a power of two. */
int fb_commit_open_rows(fb_ctx* ctx, uint64_t row0, uint64_t nrows, const uint8_t* coeffs_be, uint64_t len,
                        const uint8_t alpha_be[32], uint8_t* out_commit48s, uint8_t* out_ys_be, uint8_t* out_proof48s);

/* partial results over the resident column range (uncompressed, 96 bytes); the
 * coefficient vector is always the FULL row */
int fb_worker_commit_partial(fb_ctx* ctx, uint64_t i, const uint8_t* coeffs_be, uint64_t len, uint8_t out96[96]);
int fb_worker_open_partial(fb_ctx* ctx, uint64_t i, const uint8_t* coeffs_be, uint64_t len, const uint8_t alpha_be[32],
                           uint8_t out_y_be[32], uint8_t out96[96]);

/* commit + open of one row in one pass (extension; the two calls above remain): the
 * coefficients go up once and both MSMs - same resident row - run as one batch of two
 * scalar vectors.  Byte-identical to fb_worker_commit followed by fb_worker_open.
 * len must be a power of two (as for fb_worker_open). */
int fb_worker_commit_open(fb_ctx* ctx, uint64_t i, const uint8_t* coeffs_be, uint64_t len, const uint8_t alpha_be[32],
                          uint8_t out_commit48[48], uint8_t out_y_be[32], uint8_t out_proof48[48]);
int fb_worker_commit_open_partial(fb_ctx* ctx, uint64_t i, const uint8_t* coeffs_be, uint64_t len,
                                  const uint8_t alpha_be[32], uint8_t out_commit96[96], uint8_t out_y_be[32],
                                  uint8_t out_proof96[96]);

/* Column shards with the field half sharded too (extension): a shard uploads and works on ITS slice of the row only.
 * begin: coeffs_slice_be = the slice_len coefficients of columns [col0, col0+ncols) cut to the row length len (a power
 * of two); writes 128 opaque bytes of state (this shard's partial sums).  The caller gathers the states of all shards in
 * any order (e.g. one all-gather) and gives them to finish, which returns the partial commitment and partial proof (96
 * bytes each, to be added over the shards: fb_g1_sum / fb_dev_combine_commit_open) and y.  The slice stays resident in
 * the context between the two calls: one sharded open in flight per context.  Same bytes as the unsharded calls. */
int fb_worker_commit_open_shard_begin(fb_ctx* ctx, uint64_t i, const uint8_t* coeffs_slice_be, uint64_t slice_len,
                                      uint64_t len, const uint8_t alpha_be[32], uint8_t out_state128[128]);
int fb_worker_commit_open_shard_finish(fb_ctx* ctx, uint64_t i, const uint8_t* states128, uint64_t nparts,
                                       const uint8_t alpha_be[32], uint8_t out_commit96[96], uint8_t out_y_be[32],
                                       uint8_t out_proof96[96]);

/* ---- engine: master side ------------------------------------------------------
 * replaces PianoBackend::master_commit (piano.rs:164-175): sum of the commitments */
int fb_master_commit(fb_ctx* ctx, const uint8_t* commitments48, uint64_t count, uint8_t out48[48]);
/* replaces PianoBackend::master_open (piano.rs:250-298) */
int fb_master_open(fb_ctx* ctx, const uint8_t* evals_be, const uint8_t* proofs48, uint64_t count,
                   const uint8_t beta_be[32], uint8_t out_z_be[32], uint8_t out_pi0_48[48], uint8_t out_pi1_48[48]);

/* ---- verification -----------------------------------------------------------------
 * replaces PianoBackend::worker_verify (piano.rs:309-356):
 *   e(C - y * g^{R_i(tau_Y)}, g2) == e(pi, g2^{tau_X} - alpha * g2)
 * The G1 side runs on the GPU, the pairing product (pairings_verify_single,
 * piano.rs:358-388) on the host.  *out_valid = 1 / 0; a malformed input is an error
 * (the reference rejects it while parsing the request, rpc.rs:330-331). */
int fb_worker_verify(fb_ctx* ctx, uint64_t i, const uint8_t commitment48[48], const uint8_t alpha_be[32],
                     const uint8_t y_be[32], const uint8_t proof48[48], int* out_valid);
/* replaces PianoBackend::master_verify (piano.rs:393-420, pairings_verify :422-464):
 *   e(C - z * g, g2) == e(pi0, g2^{tau_X} - alpha g2) * e(pi1, g2^{tau_Y} - beta g2) */
int fb_master_verify(fb_ctx* ctx, const uint8_t commitment48[48], const uint8_t beta_be[32], const uint8_t alpha_be[32],
                     const uint8_t z_be[32], const uint8_t pi0_48[48], const uint8_t pi1_48[48], int* out_valid);
/* the pairing product by itself, on serialized points (g1: n x 96 bytes as
 * blst_p1_serialize, g2: n x 192 bytes as blst_p2_serialize); needs no context and
 * no GPU.  *out_is_one = (prod_k e(g1[k], g2[k]) == 1). */
int fb_pairing_check(const uint8_t* g1_96, const uint8_t* g2_192, uint64_t n, int* out_is_one);
/* the compared value, (prod_k e(g1[k], g2[k]))^3 in GT: 12 Fp coefficients, 48-byte
 * big-endian, of w^0..w^5 (real part first), Fp12 = Fp2[w]/(w^6 - (1+u)). */
int fb_pairing_value(const uint8_t* g1_96, const uint8_t* g2_192, uint64_t n, uint8_t out576[576]);

/* ---- field-side RPC helpers -----------------------------------------------------
 * replaces PianoFFTSettings::fft_left / fft_right (piano.rs:1096-1103) as exposed
 * by the `fft` RPC (src/rpc.rs:66-70): left != 0 uses the 2^(n-m) domain. */
int fb_fft(fb_ctx* ctx, int left, int inverse, const uint8_t* in_be, uint64_t len, uint8_t* out_be);
/* replaces FsPoly::eval as exposed by the `evaluate` RPC (src/rpc.rs:60-64) */
int fb_evaluate(fb_ctx* ctx, const uint8_t* poly_be, uint64_t len, const uint8_t x_be[32], uint8_t out_y_be[32]);

/* ---- raw kernels (benches, sharded MSM) ------------------------------------------
 * A base set is an array of G1 points resident on the device, optionally with
 * window tables.  format: 48 = compressed, 96 = uncompressed. */
typedef struct fb_bases fb_bases;
int fb_bases_from_points(fb_ctx* ctx, const uint8_t* points, uint64_t count, int format, fb_bases** out);
/* P_k = s_k * G for the given scalars (synthetic bases with known discrete logs) */
int fb_bases_from_scalars(fb_ctx* ctx, const uint8_t* scalars_be, uint64_t count, fb_bases** out);
int fb_bases_precompute(fb_ctx* ctx, fb_bases* b, int window_bits);
int fb_bases_get(fb_ctx* ctx, const fb_bases* b, uint64_t index, uint8_t out48[48]);
/* window bits / number of windows of the set's table as built (0, 0: none) */
int fb_bases_get_window(fb_ctx* ctx, const fb_bases* b, int* window_bits, int* num_windows);
void fb_bases_free(fb_ctx* ctx, fb_bases* b);
/* replaces g1_linear_combination on arbitrary points: out = sum_k s_k * P_k */
int fb_msm(fb_ctx* ctx, const fb_bases* b, const uint8_t* scalars_be, uint64_t count, uint8_t out48[48]);
/* same, result left uncompressed (96 bytes) so partial sums of a sharded MSM can
 * be combined with fb_g1_sum */
int fb_msm_partial(fb_ctx* ctx, const fb_bases* b, const uint8_t* scalars_be, uint64_t count, uint8_t out96[96]);
int fb_g1_sum(fb_ctx* ctx, const uint8_t* points, uint64_t count, int format, uint8_t out48[48]);
/* NTT of any power-of-two length (tables cached per size) */
int fb_ntt(fb_ctx* ctx, int inverse, const uint8_t* in_be, uint64_t len, uint8_t* out_be);

/* ---- device-resident entry points ----------------------------------------------------
 * Same kernels, operands already in HBM (device pointers, e.g. torch tensors):
 * Fr vectors are arrays of 32-byte Montgomery elements as produced by
 * fb_dev_fr_from_be.  Asynchronous on the context's stream. */
int fb_dev_fr_from_be(fb_ctx* ctx, const void* d_in_be, void* d_out_fr, uint64_t count);
int fb_dev_fr_to_be(fb_ctx* ctx, const void* d_in_fr, void* d_out_be, uint64_t count);
int fb_dev_ntt(fb_ctx* ctx, int inverse, const void* d_in_fr, void* d_out_fr, uint64_t len);
/* d_out: 48 bytes (device) */
int fb_dev_msm(fb_ctx* ctx, const fb_bases* b, const void* d_scalars_fr, uint64_t count, void* d_out48);
int fb_dev_worker_commit(fb_ctx* ctx, uint64_t i, const void* d_coeffs_fr, uint64_t len, void* d_out48);
/* d_out: 32 bytes y (big-endian) followed by 48 bytes proof */
int fb_dev_worker_open(fb_ctx* ctx, uint64_t i, const void* d_coeffs_fr, uint64_t len, const void* d_alpha_fr,
                       void* d_out80);

int fb_dev_worker_commit_partial(fb_ctx* ctx, uint64_t i, const void* d_coeffs_fr, uint64_t len, void* d_out96);
/* d_out: 32 bytes y followed by 96 bytes partial proof */
int fb_dev_worker_open_partial(fb_ctx* ctx, uint64_t i, const void* d_coeffs_fr, uint64_t len, const void* d_alpha_fr,
                               void* d_out128);

/* fused commit + open: d_out = 48 bytes commitment | 32 bytes y | 48 bytes proof */
int fb_dev_worker_commit_open(fb_ctx* ctx, uint64_t i, const void* d_coeffs_fr, uint64_t len, const void* d_alpha_fr,
                              void* d_out128);
/* column shard: d_out = 96 bytes partial commitment | 32 bytes y | 96 bytes partial proof */
int fb_dev_worker_commit_open_partial(fb_ctx* ctx, uint64_t i, const void* d_coeffs_fr, uint64_t len,
                                      const void* d_alpha_fr, void* d_out224);

/* sharded field half, operands in HBM: d_coeffs_slice_fr = the resident columns' coefficients (Montgomery), d_states =
 * nparts x 128 bytes as gathered; d_out224 = 96 | 32 | 96.  Both run on the caller's stream (fb_ctx_set_stream). */
int fb_dev_commit_open_shard_begin(fb_ctx* ctx, uint64_t i, const void* d_coeffs_slice_fr, uint64_t len,
                                   const void* d_alpha_fr, void* d_out_state128);
int fb_dev_commit_open_shard_finish(fb_ctx* ctx, uint64_t i, const void* d_coeffs_slice_fr, uint64_t len,
                                    const void* d_alpha_fr, const void* d_states, uint64_t nparts, void* d_out224);
/* adds the shards' partials on the device: d_parts224 = nparts x (96-byte partial commitment | 32-byte y | 96-byte partial
 * proof) -> d_out128 = 48-byte commitment | 32-byte y | 48-byte proof.  On the caller's stream. */
int fb_dev_combine_commit_open(fb_ctx* ctx, const void* d_parts224, uint64_t nparts, void* d_out128);

/* uncompressed (96-byte) result, for the partial sums of an MSM sharded by point ranges */
int fb_dev_msm_partial(fb_ctx* ctx, const fb_bases* b, const void* d_scalars_fr, uint64_t count, void* d_out96);
/* batched rows (the bivariate row loop of src/bipoly.rs / benches/piano.rs:41-63), operands
 * in HBM: d_coeffs_fr = nrows x len Montgomery Fr; outputs nrows x 48 / nrows x 32 bytes */
int fb_dev_commit_rows(fb_ctx* ctx, uint64_t row0, uint64_t nrows, const void* d_coeffs_fr, uint64_t len,
                       void* d_out48s);
int fb_dev_open_rows(fb_ctx* ctx, uint64_t row0, uint64_t nrows, const void* d_coeffs_fr, uint64_t len,
                     const void* d_alpha_fr, void* d_out_ys_be, void* d_out_proof48s);

/* ---- allocation-free evaluation ---------------------------------------------------------
 * No counterpart in the reference: added for callers that
 * cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid
 * it).  y = sum_k poly[k] x^k (FsPoly::eval) on operands already in HBM: d_poly_fr = len Montgomery
 * Fr (as written by fb_dev_fr_from_be), d_x_fr and d_out_y_fr = one Montgomery Fr each.  Needs no
 * context; asynchronous on `cuda_stream` (a cudaStream_t) of device `device`.  Calls no cudaMalloc /
 * cudaFree: the scratch of the blocked Horner scan is a static __device__ array of 1.04 MiB that
 * every GPU loading the library holds until exit, sized for len <= 2^20 (FB_ERR_ARG beyond that,
 * message from fb_last_error(NULL)).  All calls share that array: calls that overlap on the GPU
 * race, so serialise them (one stream, or events).  The first call on a device looks the array's
 * address up (cudaGetSymbolAddress): make one call before beginning a graph capture. */
int fb_dev_evaluate_noalloc(int device, const void* d_poly_fr, uint64_t len, const void* d_x_fr, void* d_out_y_fr,
                            void* cuda_stream);

int fb_dev_commit_open_rows(fb_ctx* ctx, uint64_t row0, uint64_t nrows, const void* d_coeffs_fr, uint64_t len,
                            const void* d_alpha_fr, void* d_out_commit48s, void* d_out_ys_be, void* d_out_proof48s);

/* number of kernel launches issued by this library so far */
uint64_t fb_launch_count(const fb_ctx* ctx);
/* per-phase device timing (CUDA events on the launching stream); fb_prof_read
 * writes a JSON object {"phase": [total_ms, count], ...} and resets the log */
int fb_prof_enable(fb_ctx* ctx, int on);
int fb_prof_read(fb_ctx* ctx, char* buf, size_t cap);
/* work counters of the MSMs run while profiling was on (summed over lanes, reset by the read): out[0] MSMs,
 * [1] sorted (point, window) entries, [2] output slots of the batched-affine rounds, [3] survivors added by the
 * XYZZ kernel, [4] buckets.  bench.py turns them into the number of Fp multiplications actually issued. */
int fb_msm_stats(fb_ctx* ctx, uint64_t out[8]);

#ifdef __cplusplus
}
#endif
#endif /* FOURIER_B200_H */
