/*
 * shelfi_b200.h — C ABI of the B200-native CKKS aggregation engine (libshelfi_b200.so).
 *
 * Drop-in boundary for the reference's SHELFI_FHE pybind module.  Every entry point names the
 * reference interface it replaces; paths are relative to the reference repository
 * (wzjin2017/fhe-fed), directory palisade_pybind/SHELFI_FHE/.
 *
 *   reference (pybind, src/binding.cpp:18-31)            this library
 *   ------------------------------------------------    ------------------------------------------
 *   CKKS(scheme, batchSize, scaleFactorBits, cryptodir)  shelfi_ckks_create          (src/ckks.cpp:5-9)
 *   CKKS::loadCryptoParams()                             shelfi_load_crypto_params   (src/ckks.cpp:11-23)
 *   CKKS::genCryptoContextAndKeyGen()                    shelfi_gen_crypto_context_and_keygen (src/ckks.cpp:25-59)
 *   CKKS::encrypt(ndarray) -> bytes                      shelfi_encrypt[_into]       (src/ckks.cpp:61-104)
 *   CKKS::computeWeightedAverage(list, list) -> bytes    shelfi_compute_weighted_average[_into] (src/ckks.cpp:264-320)
 *   CKKS::decrypt(bytes, n) -> ndarray                   shelfi_decrypt              (src/ckks.cpp:170-213)
 *   encrypt_cpp / decrypt_cpp / computeWeightedAverage_cpp  same three calls (binding.cpp:27,29,31 bind the
 *                                                        *_cpp names to the same methods; include/ckks.h:46-52)
 *
 * All hot-path work runs in hand-written sm_100a CUDA kernels; there is no CPU fallback: every call
 * fails with SHELFI_ERR_CUDA when no usable device is present.
 *
 * Conventions: functions return 0 on success or a negative SHELFI_ERR_* code (exceptions noted);
 * shelfi_last_error(h) describes the last failure on that handle.  Inputs are borrowed for the
 * duration of the call; outputs returned through `uint8_t **` are owned by the library and released
 * with shelfi_free.  A handle serialises its calls internally (the reference object is not
 * thread-safe either, binding.cpp holds the GIL throughout).
 */
#ifndef SHELFI_B200_H
#define SHELFI_B200_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct shelfi_ckks shelfi_ckks;

enum {
    SHELFI_OK = 0,
    SHELFI_ERR_ARG = -1,        /* bad argument (null, zero length, mismatched sizes) */
    SHELFI_ERR_NOT_LOADED = -2, /* context / key needed by the call is not loaded */
    SHELFI_ERR_FORMAT = -3,     /* malformed or unsupported serialized object */
    SHELFI_ERR_PARAMS = -4,     /* object belongs to different crypto parameters */
    SHELFI_ERR_CUDA = -5,       /* CUDA runtime failure (including: no device) */
    SHELFI_ERR_IO = -6,         /* file could not be read / written */
    SHELFI_ERR_UNSUPPORTED = -7, /* outside what the engine implements (e.g. ring dimension != 8192) */
    SHELFI_ERR_MEMORY = -8       /* device memory for resident ciphertexts exhausted (limit or the device itself) */
};

/* ---- lifecycle: CKKS::CKKS (src/ckks.cpp:5-9, defaults src/binding.cpp:19-23) ---- */
/* cryptodir is concatenated raw with the three file names, so it must end in '/', as in the reference.
 * device = CUDA ordinal (one process per GPU: pass LOCAL_RANK). */
int shelfi_ckks_create(const char *scheme, uint32_t batch_size, uint32_t scale_factor_bits, const char *cryptodir,
                       int device, shelfi_ckks **out);
void shelfi_ckks_destroy(shelfi_ckks *h);
const char *shelfi_last_error(const shelfi_ckks *h);

/* CKKS::loadCryptoParams (src/ckks.cpp:11-23).  Returns a bit mask of what could NOT be loaded
 * (1 context, 2 public key, 4 private key; 0 = all three) — the reference prints and carries on; the
 * Python shim reproduces that.  A negative value is a SHELFI_ERR_* (e.g. no CUDA device). */
int shelfi_load_crypto_params(shelfi_ckks *h);
/* same, from memory images of cryptocontext.txt / key-public.txt / key-private.txt (any may be NULL) */
int shelfi_load_crypto_params_from_memory(shelfi_ckks *h, const uint8_t *cc, size_t cc_len, const uint8_t *pk,
                                          size_t pk_len, const uint8_t *sk, size_t sk_len);
/* CKKS::genCryptoContextAndKeyGen (src/ckks.cpp:25-59): multDepth 1, writes the three files into
 * cryptodir, does not load them.  Returns 1 on success, 0 on failure — the reference's convention. */
int shelfi_gen_crypto_context_and_keygen(shelfi_ckks *h);
/* Extension: the same with PALISADE's multDepth and ring dimension free (the reference hard-codes multDepth = 1,
 * src/ckks.cpp:26, and lets PALISADE pick the ring dimension).  mult_depth in [1, 7] gives mult_depth + 1 limbs
 * (60-bit q_0, then the EXACTRESCALE chain around 2^scaleFactorBits); ring_dim 0 = PALISADE's choice from the
 * HEStd_128_classic table, else 8192 / 16384 / 32768.  (mult_depth 3, 52 bits) -> N = 16384, L = 4 and
 * (7, 52) -> N = 32768, L = 8 are BASELINE configs 3 and 4.  Returns 1 / 0 like the reference call. */
int shelfi_gen_crypto_context_and_keygen_ex(shelfi_ckks *h, uint32_t mult_depth, uint32_t ring_dim);

/* Randomness.  The reference draws v, e0, e1 (Encrypt, src/ckks.cpp:81,95) and s, e, a (KeyGen, src/ckks.cpp:46) from
 * PALISADE's Blake2-based PRNG, seeded from OS entropy.  Here every draw is ChaCha20 in counter mode (the RFC 8439
 * block function, 20 rounds) under a 256-bit key taken from getrandom(2): one fresh key per encrypt call and learner;
 * key generation uses two independent keys, one for the secret material (s, e) and one for what is published (a, the
 * key tag).  A fixed seed replaces those keys by ones laid out from the seed so that the device sampler is
 * reproducible (tests, parity against the CPU oracle); never use it for real data. */
int shelfi_set_sampler_seed(shelfi_ckks *h, uint64_t seed);
int shelfi_clear_sampler_seed(shelfi_ckks *h);
/* Semantics of PALISADE that no fixture of the reference pins (DESIGN.md 5 lists them), switchable so
 * that a deployment checked against a real PALISADE can be corrected without a rebuild.  The CPU oracle has the same
 * switches under the same ids.
 *   SHELFI_OPT_EVALMULT_ROUND  EvalMult(ct, double)'s constant (src/ckks.cpp:288): 0 = (int64)(c * Delta + 0.5), the
 *                              default (PALISADE v1.11 source); 1 = llround(c * Delta) (differs for negative weights)
 *   SHELFI_OPT_DECODE_NOISE    cc->Decrypt (src/ckks.cpp:189): 1 = add PALISADE >= 1.10.6's Gaussian decode noise (the
 *                              Li-Micciancio key-recovery countermeasure: standard deviation estimated from the
 *                              conjugate-asymmetric part of the decrypted polynomial, floored at sqrt(N)/8, times
 *                              sqrt(2)) before FFTSpecial; 0 = off (default: deterministic, which the bit-exact parity
 *                              tier needs).  Also SHELFI_DECODE_NOISE=1 in the environment.
 *   SHELFI_OPT_WIRE_VARIANT    RLWE parameter block of GENERATED contexts: 9 words (default, newer PALISADE) or 8;
 *                              loaded contexts are re-emitted in the variant they came in. */
enum { SHELFI_OPT_EVALMULT_ROUND = 1, SHELFI_OPT_DECODE_NOISE = 2, SHELFI_OPT_WIRE_VARIANT = 3 };
int shelfi_set_option(shelfi_ckks *h, int option, int64_t value);
int64_t shelfi_get_option(const shelfi_ckks *h, int option);
/* known-answer tap: one ChaCha20 block computed by the device sampler code (key 8 words, block counter, nonce 3
 * words -> 16 words); tests hold the RFC 8439 vectors */
int shelfi_chacha20_block_device(shelfi_ckks *h, const uint32_t key[8], uint32_t counter, const uint32_t nonce[3],
                                 uint32_t out[16]);

/* ---- introspection ---- */
uint32_t shelfi_ring_dim(const shelfi_ckks *h);
uint32_t shelfi_num_limbs(const shelfi_ckks *h);
uint32_t shelfi_slots(const shelfi_ckks *h);
uint32_t shelfi_batch_size(const shelfi_ckks *h);
uint64_t shelfi_modulus(const shelfi_ckks *h, uint32_t limb);
uint64_t shelfi_root_of_unity(const shelfi_ckks *h, uint32_t limb);
double shelfi_scaling_factor(const shelfi_ckks *h);               /* q_last as double: level-0 scaling factor */
int shelfi_key_tag(const shelfi_ckks *h, char out[33]);
uint64_t shelfi_num_ciphertexts(const shelfi_ckks *h, uint64_t n); /* ceil(n / batchSize), src/ckks.cpp:65 */
uint64_t shelfi_ciphertext_words(const shelfi_ckks *h);            /* 2 * L * N u64 per ciphertext (dense form) */
uint64_t shelfi_blob_size(const shelfi_ckks *h, uint64_t n_ct);    /* bytes of a serialized vector of n_ct ciphertexts */
uint64_t shelfi_blob_size_level(const shelfi_ckks *h, uint64_t n_ct, uint32_t limbs); /* ... that kept `limbs` towers */
/* EvalMult constant for a learner weight: (int64)((double)(float)w * scalingFactor + 0.5)  (src/ckks.cpp:287-288) */
int64_t shelfi_weight_constant(const shelfi_ckks *h, float w);

/* ---- reference-facing calls: host buffers in, host buffers out ---- */
/* CKKS::encrypt (src/ckks.cpp:61-104): n doubles -> serialized vector<Ciphertext<DCRTPoly>>. */
int shelfi_encrypt(shelfi_ckks *h, const double *x, uint64_t n, uint8_t **blob, uint64_t *blob_len);
/* same, into a caller buffer of at least shelfi_blob_size(h, shelfi_num_ciphertexts(h, n)) bytes
 * (pinned memory from shelfi_host_alloc avoids the staging copy) */
int shelfi_encrypt_into(shelfi_ckks *h, const double *x, uint64_t n, uint8_t *dst, uint64_t cap, uint64_t *blob_len);
/* CKKS::computeWeightedAverage (src/ckks.cpp:264-320): k blobs, k weights (float on purpose: the
 * reference narrows each Python float to float32, src/ckks.cpp:287) -> one blob, depth 2. */
int shelfi_compute_weighted_average(shelfi_ckks *h, const uint8_t *const *blobs, const uint64_t *blob_lens,
                                    const float *weights, uint32_t k, uint8_t **blob, uint64_t *blob_len);
int shelfi_compute_weighted_average_into(shelfi_ckks *h, const uint8_t *const *blobs, const uint64_t *blob_lens,
                                         const float *weights, uint32_t k, uint8_t *dst, uint64_t cap,
                                         uint64_t *blob_len);
/* Rescale-before-transport (opt-in extension; the reference leaves the average at depth 2, src/ckks.cpp:287-303):
 * the same weighted average followed by PALISADE's ModReduce (EXACTRESCALE: DCRTPoly::DropLastElementAndScale) on
 * every ciphertext.  The result blob has one tower fewer (half the bytes at two towers), depth 1, level 1,
 * scalingFactor / q_last; shelfi_decrypt and a PALISADE Decrypt read it like any other ciphertext vector.
 * shelfi_rescale does the ModReduce alone on an already aggregated (depth >= 2) blob. */
int shelfi_compute_weighted_average_rescaled(shelfi_ckks *h, const uint8_t *const *blobs, const uint64_t *blob_lens,
                                             const float *weights, uint32_t k, uint8_t **blob, uint64_t *blob_len);
int shelfi_compute_weighted_average_rescaled_into(shelfi_ckks *h, const uint8_t *const *blobs, const uint64_t *blob_lens,
                                                  const float *weights, uint32_t k, uint8_t *dst, uint64_t cap,
                                                  uint64_t *blob_len);
int shelfi_rescale(shelfi_ckks *h, const uint8_t *blob, uint64_t blob_len, uint8_t **out, uint64_t *out_len);
int shelfi_rescale_into(shelfi_ckks *h, const uint8_t *blob, uint64_t blob_len, uint8_t *dst, uint64_t cap, uint64_t *out_len);
/* CKKS::decrypt (src/ckks.cpp:170-213): blob -> n doubles (n = data_dimensions).  Accepts level-0 and rescaled blobs. */
int shelfi_decrypt(shelfi_ckks *h, const uint8_t *blob, uint64_t blob_len, uint64_t n, double *out);

/* ---- many independent vectors in one call ----
 * The reference's callers encrypt, aggregate and decrypt a model layer by layer (code/benchmark.py:490-493,
 * code/benchmark_selection.py:152-193: one CKKS::encrypt / computeWeightedAverage / decrypt per state_dict entry).
 * These entry points do the same work for n_layers vectors with one host->device copy, the kernels of every layer on
 * one stream and one device->host copy.  Every layer is its own serialized ciphertext vector, byte-identical to
 * what the single-vector call returns for it (under a fixed sampler seed).
 *   encrypt: x[s] = layer s, layer_len[s] values.  *blobs receives ONE buffer
 *            (shelfi_free) holding blob s at offsets[s] (8-byte aligned), lens[s] bytes; *total_len its size.
 *   weighted average: blobs[i * n_layers + s] / blob_lens[...] = learner i, layer s; result laid out as above.
 *   decrypt: out = data_dimensions[0] values of layer 0, then layer 1, ... */
int shelfi_encrypt_layers(shelfi_ckks *h, const double *const *x, const uint64_t *layer_len, uint32_t n_layers, uint8_t **blobs,
                          uint64_t *offsets, uint64_t *lens, uint64_t *total_len);
int shelfi_compute_weighted_average_layers(shelfi_ckks *h, const uint8_t *const *blobs, const uint64_t *blob_lens,
                                           const float *weights, uint32_t k, uint32_t n_layers, uint8_t **out,
                                           uint64_t *offsets, uint64_t *lens, uint64_t *total_len);
/* the same into a caller buffer (e.g. page-locked memory from shelfi_host_alloc): layer s of n_ct_s ciphertexts needs
 * shelfi_blob_size(h, n_ct_s) bytes rounded up to a multiple of 8; the blobs are laid out back to back in that order */
int shelfi_encrypt_layers_into(shelfi_ckks *h, const double *const *x, const uint64_t *layer_len, uint32_t n_layers, uint8_t *dst,
                               uint64_t cap, uint64_t *offsets, uint64_t *lens, uint64_t *total_len);
int shelfi_compute_weighted_average_layers_into(shelfi_ckks *h, const uint8_t *const *blobs, const uint64_t *blob_lens,
                                                const float *weights, uint32_t k, uint32_t n_layers, uint8_t *dst,
                                                uint64_t cap, uint64_t *offsets, uint64_t *lens, uint64_t *total_len);
/* the same followed by ModReduce on every ciphertext of every layer (rescale-before-transport, opt-in; see
 * shelfi_compute_weighted_average_rescaled): blob s is byte-identical to that call on layer s alone */
int shelfi_compute_weighted_average_layers_rescaled(shelfi_ckks *h, const uint8_t *const *blobs, const uint64_t *blob_lens,
                                                    const float *weights, uint32_t k, uint32_t n_layers, uint8_t **out,
                                                    uint64_t *offsets, uint64_t *lens, uint64_t *total_len);
int shelfi_compute_weighted_average_layers_rescaled_into(shelfi_ckks *h, const uint8_t *const *blobs,
                                                         const uint64_t *blob_lens, const float *weights, uint32_t k,
                                                         uint32_t n_layers, uint8_t *dst, uint64_t cap, uint64_t *offsets,
                                                         uint64_t *lens, uint64_t *total_len);
int shelfi_decrypt_layers(shelfi_ckks *h, const uint8_t *const *blobs, const uint64_t *blob_lens,
                          const uint64_t *data_dimensions, uint32_t n_layers, double *out);

/* ---- resident ciphertext vectors: the same three calls with the ciphertexts left in HBM ----
 * The reference's callers keep every serialized vector inside one process and only hand it to the next call
 * (code/benchmark.py:459-460 encrypt, :490-493 the per-layer loop, :507-513 computeWeightedAverage, :527-528 decrypt), so
 * nothing forces 64 bytes per parameter and learner across PCIe twice.  A shelfi_ct is a vector<Ciphertext<DCRTPoly>>
 * held in device memory by the handle that made it: the dense payload plus what the wire format needs to write it
 * out again.  shelfi_ct_to_blob yields exactly the bytes shelfi_encrypt / shelfi_compute_weighted_average would have
 * returned (Serial::Serialize, src/ckks.cpp:98-101,308-310); shelfi_ct_from_blob is the inverse (Deserialize,
 * src/ckks.cpp:172-176,278-281), so vectors enter and leave the device only when a caller needs the bytes.
 *   shelfi_ct_encrypt          CKKS::encrypt (src/ckks.cpp:61-104).  x: n values, dtype 0 = float64, 1 = float32 (widened
 *                              exactly on the device — what pybind's array_t<double> caster does on the host for the
 *                              float32 state_dict arrays the reference's callers pass), in pageable or page-locked host
 *                              memory or in device memory.  Returns when x has been read (the caller may reuse it); the
 *                              encryption itself continues on the device, so the upload of the next learner's vector
 *                              overlaps it.
 *   shelfi_ct_weighted_average CKKS::computeWeightedAverage (src/ckks.cpp:264-320) over k resident vectors -> a new one;
 *                              rescale != 0 appends PALISADE's ModReduce (see shelfi_compute_weighted_average_rescaled).
 *   shelfi_ct_decrypt          CKKS::decrypt (src/ckks.cpp:170-213) -> n doubles in host memory; returns when they are there.
 * Device memory comes from a per-handle pool (freed vectors are reused, cudaFree never runs on the hot path) with a
 * byte limit: 80 % of the free device memory at first use, SHELFI_RESIDENT_LIMIT_MB, or shelfi_resident_set_limit.
 * A call that would exceed it fails with SHELFI_ERR_MEMORY and says how much is held.  Free every vector before
 * destroying its handle (shelfi_ckks_destroy releases the device memory of those still alive). */
typedef struct shelfi_ct shelfi_ct;
int shelfi_ct_encrypt(shelfi_ckks *h, const void *x, uint64_t n, int dtype, shelfi_ct **out);
int shelfi_ct_from_blob(shelfi_ckks *h, const uint8_t *blob, uint64_t blob_len, shelfi_ct **out);
int shelfi_ct_weighted_average(shelfi_ckks *h, shelfi_ct *const *cts, const float *weights, uint32_t k, int rescale,
                               shelfi_ct **out);
int shelfi_ct_decrypt(shelfi_ckks *h, const shelfi_ct *ct, uint64_t n, double *out);
uint64_t shelfi_ct_blob_size(const shelfi_ct *ct);   /* bytes shelfi_ct_to_blob writes */
int shelfi_ct_to_blob(shelfi_ckks *h, const shelfi_ct *ct, uint8_t *dst, uint64_t cap, uint64_t *blob_len);
/* number of ciphertexts, towers per polynomial, depth / level / scaling factor of ciphertext 0; any pointer may be NULL */
int shelfi_ct_info(const shelfi_ct *ct, uint64_t *n_ct, uint32_t *limbs, uint64_t *depth, uint64_t *level,
                   double *scaling_factor);
/* the dense payload u64 [n_ct][2][limbs][N] in device memory (for the shelfi_*_device calls); valid until the free */
const uint64_t *shelfi_ct_device_ptr(const shelfi_ct *ct);
void shelfi_ct_free(shelfi_ckks *h, shelfi_ct *ct);
/* bytes held by live vectors, bytes parked on the free list, the limit */
int shelfi_resident_stats(shelfi_ckks *h, uint64_t *in_use, uint64_t *pooled, uint64_t *limit);
int shelfi_resident_set_limit(shelfi_ckks *h, uint64_t bytes);   /* also returns the parked buffers to the driver */
/* waits for everything the handle has queued on the device (resident calls return before their kernels finish) */
int shelfi_synchronize(shelfi_ckks *h);

void shelfi_free(void *p);
/* page-locked host memory for the *_into calls and for inputs */
void *shelfi_host_alloc(size_t bytes);
void shelfi_host_free(void *p);

/* ---- device-resident calls: the same three stages with inputs and outputs in HBM ----
 * Ciphertext batches are dense limb-major u64 [n_ct][2][L][N] (the coefficient runs of the wire
 * format, in wire order, without the bytes between them).  `stream` is a cudaStream_t (NULL =
 * default stream); calls are asynchronous on it.  ct_index0 is the global index of the first
 * ciphertext so that a shard draws the same sampler stream as the unsharded call. */
int shelfi_encrypt_device(shelfi_ckks *h, const double *d_x, uint64_t n, uint64_t ct_index0, uint64_t *d_ct,
                          void *stream);
/* k learners' vectors of equal length n in ONE launch (a server simulating its clients, as code/benchmark.py:490-493
 * does on the CPU): same result as k shelfi_encrypt_device calls, better machine fill for small models. */
int shelfi_encrypt_many_device(shelfi_ckks *h, const double *const *d_x, uint32_t k, uint64_t n, uint64_t ct_index0,
                               uint64_t *const *d_ct, void *stream);
int shelfi_weighted_sum_device(shelfi_ckks *h, const uint64_t *const *d_cts, const float *weights, uint32_t k,
                               uint64_t n_ct, uint64_t *d_out, void *stream);
/* the same folded into an accumulator that already holds a (reduced) partial sum: learners streamed through a resident
 * accumulator in groups when K x ciphertext bytes exceed HBM (BASELINE config 4: 100 learners x 6.5 GB) */
int shelfi_weighted_sum_accumulate_device(shelfi_ckks *h, const uint64_t *const *d_cts, const float *weights, uint32_t k,
                                          uint64_t n_ct, uint64_t *d_acc, void *stream);
int shelfi_decrypt_device(shelfi_ckks *h, const uint64_t *d_ct, uint64_t n_ct, double scaling_factor, double *d_out,
                          uint64_t n, void *stream);
/* shelfi_decrypt_device without any allocation.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA stream capture, tight serving loops, harnesses
 * that forbid it).  Same arguments and the same results as shelfi_decrypt_device; the scratch between the inverse
 * transforms and the decode (and, for rings above 8192 or more than two towers, the decode's own scratch) is a
 * file-scope device array of SHELFI_STATIC_SCRATCH_BYTES that every GPU which loads the library holds until the
 * process exits.  Per ciphertext a call needs 8 * (L*N [+ N + 2*slots if slots > 4096, in the general decode]) bytes:
 * 64 ciphertexts (262,144 values) per call for N = 8192, L = 2; more returns SHELFI_ERR_ARG with the limit in
 * shelfi_last_error.  The call makes no cudaMalloc/cudaFree and creates no stream or event; the handle must already
 * hold its context and secret key (else SHELFI_ERR_NOT_LOADED), and SHELFI_OPT_DECODE_NOISE must be off.
 * Every call on a device shares the one array, which keeps the previous call's contents (each call writes every
 * word before reading it, nothing has to be cleared): calls that OVERLAP on the GPU - two streams, handles or
 * threads - race on it; serialise them on one stream or with events. */
#define SHELFI_STATIC_SCRATCH_BYTES (8u << 20)
int shelfi_decrypt_static_device(shelfi_ckks *h, const uint64_t *d_ct, uint64_t n_ct, double scaling_factor, double *d_out,
                                 uint64_t n, void *stream);
/* Decrypt-gather: the optional reassembly of the decoded model across the GPUs of one box (the reference decrypts every
 * ciphertext into one float64 array, ckks.cpp:181-205; sharded by ciphertext index that array is the concatenation of
 * the ranks' slices) folded into the decode kernel: decoded value i, i < n, is stored to d_outs[p][offset + i] for every
 * p < n_outs <= 8.  d_outs[0] is normally this GPU's copy of the full vector and the others are the peers' copies,
 * mapped with shelfi_peer_buffer_open (or any other peer-accessible device pointers); offset = first parameter index
 * of this rank's slice.  The stores cross NVLink while the kernel computes; no collective re-reads the slices.  The
 * caller orders the ranks afterwards (stream synchronise + a process-group barrier) before anyone reads its copy. */
int shelfi_decrypt_gather_device(shelfi_ckks *h, const uint64_t *d_ct, uint64_t n_ct, double scaling_factor,
                                 double *const *d_outs, uint32_t n_outs, uint64_t offset, uint64_t n, void *stream);
/* Device buffers that the other processes of the box can map (CUDA IPC, one process per GPU): create on the owning
 * handle's device and pass the 64-byte handle to the peers by any byte transport; each peer opens it once and closes it
 * before the owner destroys the buffer. */
#define SHELFI_PEER_HANDLE_BYTES 64
int shelfi_peer_buffer_create(shelfi_ckks *h, uint64_t bytes, void **d_ptr, uint8_t handle[SHELFI_PEER_HANDLE_BYTES]);
int shelfi_peer_buffer_open(shelfi_ckks *h, const uint8_t handle[SHELFI_PEER_HANDLE_BYTES], void **d_ptr);
int shelfi_peer_buffer_close(shelfi_ckks *h, void *d_ptr);
int shelfi_peer_buffer_destroy(shelfi_ckks *h, void *d_ptr);
/* Barrier between the ranks of one box, queued on `stream` (no host synchronisation): rank my_rank stores `epoch` into
 * slot my_rank of every d_flags[r] and waits until all n_ranks slots of d_flags[my_rank] show it.  d_flags[r] points
 * at 9 u64 inside rank r's peer buffer (zero after shelfi_peer_buffer_create); every rank passes the same epoch,
 * larger on every call.  One before shelfi_decrypt_gather_device (the peers are done reading the previous result) and
 * one after it (every slice has landed) replace the collective's synchronisation.  A rank that waits longer than
 * SHELFI_PEER_TIMEOUT_MS (default 20000) writes the epoch to slot 8 of its own array and lets the stream continue. */
int shelfi_peer_barrier_device(shelfi_ckks *h, uint64_t *const *d_flags, uint32_t n_ranks, uint32_t my_rank, uint64_t epoch,
                               void *stream);
/* The barrier AFTER a decrypt-gather, guarding its result: if the wait for rank r times out, d_values[r * block] — the
 * first value of the slice rank r owed this GPU, in this GPU's copy of the model — is set to NaN besides the status
 * word, so that a consumer which never reads the status cannot take a half-gathered model for a whole one. */
int shelfi_peer_barrier_guard_device(shelfi_ckks *h, uint64_t *const *d_flags, uint32_t n_ranks, uint32_t my_rank,
                                     uint64_t epoch, double *d_values, uint64_t block, void *stream);
/* ModReduce on a dense batch: d_ct[n_ct][2][limbs_in][N] -> d_out[n_ct][2][limbs_in - 1][N] (limbs_in 0 = the context's
 * chain); the decrypt calls below take the towers the ciphertexts kept.  After one rescale of a depth-2 average the
 * scaling factor to pass is scalingFactor^2 / q_last. */
int shelfi_rescale_device(shelfi_ckks *h, const uint64_t *d_ct, uint64_t n_ct, uint32_t limbs_in, uint64_t *d_out, void *stream);
int shelfi_decrypt_level_device(shelfi_ckks *h, const uint64_t *d_ct, uint64_t n_ct, uint32_t limbs, double scaling_factor,
                                double *d_out, uint64_t n, void *stream);
int shelfi_decrypt_words_level_device(shelfi_ckks *h, const uint64_t *d_ct, uint64_t n_ct, uint32_t limbs, uint64_t *d_words,
                                      void *stream);
/* Per-slot plaintext weights: d_out = sum_l d_cts[l] (*) Encode(d_w[l]) — PALISADE EvalMult(ct, plaintext) + EvalAdd,
 * the ct x plaintext-vector aggregation of the TenSEAL/FLARE-style workload (code/benchmark_nvidia.py:439,461-465);
 * the reference's own computeWeightedAverage (ckks.cpp:287-288) only takes one scalar per learner.  d_w[l] holds n
 * doubles (one per model parameter, chunked like encrypt); learners may share one vector (equal pointers are encoded
 * once).  Result: depth 2, scale Delta^2, no rescale — decrypt with scaling_factor^2 like the scalar form. */
int shelfi_weighted_sum_vector_device(shelfi_ckks *h, const uint64_t *const *d_cts, const double *const *d_w, uint32_t k,
                                      uint64_t n, uint64_t *d_out, void *stream);
/* ---- selective encryption: only a chosen subset of the parameters goes through CKKS, the rest is averaged in
 * plaintext (the callers code/benchmark_selection.py:152-193 pick whole layers, benchmark_selection_rate.py:133-141 a
 * leading fraction of every layer; a sensitivity mask is the general case).  d_sel[n_sel] holds the indices of the
 * encrypted parameters (u32, any order, each < the length of the learners' vectors); it is shared by all learners.
 *   encrypt_selected_many : ciphertexts of the gathered vectors y_l[i] = d_x[l][d_sel[i]] — bit-identical to
 *                           shelfi_encrypt_many_device on the gathered copies; the gather is fused into the encode load
 *   decrypt_scatter       : d_out[d_sel[i]] = decoded[i]; entries of d_out not named by d_sel are left untouched
 *   plain_weighted_sum    : d_out[i] = sum_l (double)weights[l] * d_x[l][i] for all i < n, evaluated in learner order
 *                           with separate multiply and add (== numpy's sum of w_l * x_l); run it first, then scatter the
 *                           decrypted part over it.  Needs no keys. */
int shelfi_encrypt_selected_many_device(shelfi_ckks *h, const double *const *d_x, uint32_t k, const uint32_t *d_sel,
                                        uint64_t n_sel, uint64_t ct_index0, uint64_t *const *d_ct, void *stream);
int shelfi_decrypt_scatter_device(shelfi_ckks *h, const uint64_t *d_ct, uint64_t n_ct, double scaling_factor,
                                  double *d_out, const uint32_t *d_sel, uint64_t n_sel, void *stream);
int shelfi_plain_weighted_sum_device(shelfi_ckks *h, const double *const *d_x, const float *weights, uint32_t k,
                                     uint64_t n, double *d_out, void *stream);
/* dense <-> serialized image on the device (image buffers must be 8-byte aligned and hold
 * shelfi_blob_size bytes rounded up to 8, plus 8) */
int shelfi_pack_device(shelfi_ckks *h, const uint64_t *d_dense, uint64_t n_ct, uint8_t *d_image, void *stream);
int shelfi_unpack_device(shelfi_ckks *h, const uint8_t *d_image, uint64_t n_ct, uint64_t *d_dense, void *stream);

/* Host-only self-test of the staging copy threads (runs without a device): `rounds` verified copies of `bytes` bytes
 * split over `threads` workers; 0 = all correct. */
int shelfi_selftest_parallel_copy(uint64_t bytes, uint32_t threads, uint32_t rounds);

/* ---- primitives, exposed for parity tests and the NTT bandwidth metric ---- */
/* rows limb polynomials [rows][N], limb = row % L; PALISADE ForwardTransformToBitReverse /
 * InverseTransformFromBitReverse semantics, in place */
int shelfi_ntt_forward_device(shelfi_ckks *h, uint64_t *d_rows, uint32_t rows, void *stream);
int shelfi_ntt_inverse_device(shelfi_ckks *h, uint64_t *d_rows, uint32_t rows, void *stream);
/* measurement tap (not used by the product): the same forward transform with its last stage done by a warp-shuffle lane
 * exchange instead of a shared-memory pass; ring dimension 8192; bit-identical results */
int shelfi_ntt_forward_shfl_device(shelfi_ckks *h, uint64_t *d_rows, uint32_t rows, void *stream);
/* MakeCKKSPackedPlaintext up to the integer coefficients: n doubles -> int64 [n_ct][N] */
int shelfi_encode_device(shelfi_ckks *h, const double *d_x, uint64_t n, int64_t *d_coeff, void *stream);
/* Decrypt up to the centred CRT integer of every coefficient (two's complement hi/lo words) */
int shelfi_decrypt_integer_device(shelfi_ckks *h, const uint64_t *d_ct, uint64_t n_ct, int64_t *d_hi, uint64_t *d_lo,
                                  void *stream);
/* the same tap for any limb count: d_words[n_ct][N][8], little-endian two's complement (512 bits) of the centred
 * CRT integer of every coefficient */
int shelfi_decrypt_words_device(shelfi_ckks *h, const uint64_t *d_ct, uint64_t n_ct, uint64_t *d_words, void *stream);

/* ---- per-kernel timing (benchmarks): CUDA events on the launching stream around every launch ----
 * kernel ids: 0 encode+sample, 1 encrypt, 2 weighted sum, 3 decrypt inverse NTT, 4 decode, 5 pack, 6 unpack,
 * 7 unpack-scale-accumulate, 8 forward NTT, 9 inverse NTT, 10 rescale, 11 / 12 the host->device / device->host copies
 * of the resident calls.  shelfi_profile_get waits for recorded launches.  With SHELFI_TIMELINE=1 in the environment it
 * also prints, per recorded launch, one JSON line {"timeline": id, "t0_ms", "t1_ms"} to stderr: begin and end on the
 * device clock relative to the first launch recorded since the last call (a poor man's trace; the boxes have no nsys). */
int shelfi_profile_enable(shelfi_ckks *h, int on);
int shelfi_profile_reset(shelfi_ckks *h);
int shelfi_profile_get(shelfi_ckks *h, int kernel, double *ms_total, uint64_t *launches);

const char *shelfi_version(void);

#ifdef __cplusplus
}
#endif
#endif
