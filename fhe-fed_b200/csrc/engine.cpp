// engine.cpp — the C ABI of include/shelfi_b200.h: handle, parameter/table setup, and the three
// reference-facing calls built from the kernels in kernels.cu.  Host code only orchestrates; every
// coefficient is touched on the GPU.  No CPU fallback: without a CUDA device every call fails.
#include "../../include/shelfi_b200.h"

#include <algorithm>
#include <atomic>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cerrno>
#include <cstring>
#include <condition_variable>
#include <functional>
#include <map>
#include <mutex>
#include <thread>
#include <string>
#include <vector>

#include <sys/random.h>

#include <nvtx3/nvToolsExt.h>   // header-only: ranges around the reference-facing calls show up in Nsight timelines

#include "kernels.h"
#include "numtheory.h"
#include "tables.h"
#include "wire.h"

using namespace shelfi;

namespace {

struct NvtxRange {
    explicit NvtxRange(const char *name) { nvtxRangePushA(name); }
    ~NvtxRange() { nvtxRangePop(); }
};

struct DeviceBuffer {  // grow-only device scratch
    void *p = nullptr;
    size_t cap = 0;
    cudaError_t reserve(size_t bytes) {
        if (bytes <= cap) return cudaSuccess;
        if (p) cudaFree(p);
        p = nullptr; cap = 0;
        size_t want = bytes + bytes / 8 + 256;
        cudaError_t e = cudaMalloc(&p, want);
        if (e == cudaSuccess) cap = want;
        return e;
    }
    void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
    template <class T> T *as() const { return static_cast<T *>(p); }
};

// Optional per-kernel timing: CUDA events recorded on the launching stream around each launch.
enum KernelId { kKEncodeSample, kKEncrypt, kKWeightedSum, kKDecryptIntt, kKDecode, kKPack, kKUnpack,
                kKUnpackScaleAcc, kKNttForward, kKNttInverse, kKRescale, kKCopyIn, kKCopyOut, kKernelCount };
struct Profiler {
    bool on = false;
    struct Rec { int id; cudaEvent_t a, b; };
    std::vector<Rec> pending;
    std::vector<cudaEvent_t> spare;
    double total_ms[kKernelCount] = {0};
    uint64_t launches[kKernelCount] = {0};
    cudaEvent_t take() {
        if (!spare.empty()) { cudaEvent_t e = spare.back(); spare.pop_back(); return e; }
        cudaEvent_t e = nullptr;
        cudaEventCreate(&e);
        return e;
    }
    // SHELFI_TIMELINE=1: every drain also prints where each recorded launch / copy began and ended on the device
    // clock, relative to the first record of the batch (there is no nsys on the GPU boxes): one JSON line per record
    bool timeline = std::getenv("SHELFI_TIMELINE") != nullptr;
    void drain() {
        for (auto &r : pending) {
            float ms = 0;
            if (cudaEventSynchronize(r.b) == cudaSuccess && cudaEventElapsedTime(&ms, r.a, r.b) == cudaSuccess) {
                total_ms[r.id] += ms;
                launches[r.id]++;
                float t0 = 0;
                if (timeline && cudaEventElapsedTime(&t0, pending[0].a, r.a) == cudaSuccess)
                    std::fprintf(stderr, "{\"timeline\": %d, \"t0_ms\": %.4f, \"t1_ms\": %.4f}\n", r.id, t0, t0 + ms);
            }
            spare.push_back(r.a); spare.push_back(r.b);
        }
        pending.clear();
    }
    void release() {
        drain();
        for (cudaEvent_t e : spare) cudaEventDestroy(e);
        spare.clear();
    }
};

// Host-side memcpy split over a few persistent threads.  Pageable caller buffers (Python bytes, numpy arrays) are moved
// through page-locked staging chunks; a single thread copying into freshly allocated memory is bound by page faults
// (~3 GB/s), several threads first-touching disjoint ranges are not.
class ParallelCopier {
  public:
    explicit ParallelCopier(unsigned n) {
        parts_ = n ? n : 1;
        for (unsigned i = 1; i < parts_; i++) workers_.emplace_back([this, i] { loop(i); });
    }
    ~ParallelCopier() {
        { std::lock_guard<std::mutex> l(mu_); stop_.store(true); gen_.fetch_add(1, std::memory_order_release); }
        cv_.notify_all();
        for (auto &t : workers_) t.join();
    }
    // hot: the workers have just served a copy and are still polling (the later pieces of one transfer)
    void copy(void *dst, const void *src, size_t bytes, bool hot = false) {
        // waking sleeping workers costs a few hundred microseconds on a virtualised host (measured: 2.2 MB takes 0.5 ms
        // with eight threads), so only copies that take longer than that alone are split
        if (bytes < (hot ? (512u << 10) : (4u << 20)) || parts_ == 1) { std::memcpy(dst, src, bytes); return; }
        dst_ = static_cast<char *>(dst); src_ = static_cast<const char *>(src); bytes_ = bytes;
        pending_.store(parts_ - 1, std::memory_order_relaxed);
        { std::lock_guard<std::mutex> l(mu_); gen_.fetch_add(1, std::memory_order_release); }
        cv_.notify_all();
        piece(0);
        if (!spin_until([this] { return pending_.load(std::memory_order_acquire) == 0; }, 5000)) {
            std::unique_lock<std::mutex> l(mu_);
            done_.wait(l, [this] { return pending_.load(std::memory_order_acquire) == 0; });
        }
    }

  private:
    static void relax() {
#if defined(__x86_64__) || defined(__i386__)
        __builtin_ia32_pause();
#else
        std::this_thread::yield();
#endif
    }
    // poll `ready` for at most `us` microseconds; chunked copies follow each other within that time, so the workers
    // of a multi-chunk transfer never go back to sleep between chunks
    template <class F> static bool spin_until(F ready, unsigned us) {
        const auto t0 = std::chrono::steady_clock::now();
        for (unsigned it = 0;; it++) {
            if (ready()) return true;
            if ((it & 63) == 63 && std::chrono::steady_clock::now() - t0 > std::chrono::microseconds(us)) return false;
            relax();
        }
    }
    void piece(unsigned i) {
        size_t per = ((bytes_ / parts_) + 4095) & ~(size_t)4095;
        size_t a = std::min(bytes_, per * i), b = (i + 1 == parts_) ? bytes_ : std::min(bytes_, per * (i + 1));
        if (b > a) std::memcpy(dst_ + a, src_ + a, b - a);
    }
    void loop(unsigned i) {
        uint64_t seen = 0;
        for (;;) {
            if (!spin_until([&] { return gen_.load(std::memory_order_acquire) != seen; }, 500)) {
                std::unique_lock<std::mutex> l(mu_);
                cv_.wait(l, [&] { return gen_.load(std::memory_order_acquire) != seen; });
            }
            seen = gen_.load(std::memory_order_acquire);
            if (stop_.load()) return;
            piece(i);
            if (pending_.fetch_sub(1, std::memory_order_acq_rel) == 1) {
                std::lock_guard<std::mutex> l(mu_);
                done_.notify_one();
            }
        }
    }
    std::vector<std::thread> workers_;
    std::mutex mu_;
    std::condition_variable cv_, done_;
    const char *src_ = nullptr;
    char *dst_ = nullptr;
    size_t bytes_ = 0;
    unsigned parts_ = 1;
    std::atomic<unsigned> pending_{0};
    std::atomic<uint64_t> gen_{0};
    std::atomic<bool> stop_{false};
};

constexpr size_t kStageBytes = 16u << 20;   // two page-locked staging chunks of this size per handle
constexpr size_t kPieceBytes = 16u << 20;   // h2d moves pageable memory through them in pieces of this size (one buffer each) ...
constexpr int kStagePieces = (int)(2 * kStageBytes / kPieceBytes);   // ... a ring over both buffers

// Device memory of the resident ciphertext vectors (shelfi_ct): freed buffers are kept on a size-keyed free list,
// because cudaFree synchronises the whole device and a federation round makes and drops K + 1 vectors of the same size.
// A parked buffer carries the event that was recorded on the compute stream when it was given back (after everything
// that read or wrote it had been queued there): whoever takes it on ANOTHER stream waits for that event first.
struct PoolBuf { void *p = nullptr; cudaEvent_t ev = nullptr; };
struct DevicePool {
    static constexpr size_t kGrain = 2u << 20;
    std::multimap<size_t, PoolBuf> free_list;
    size_t in_use = 0, pooled = 0, limit = 0;   // bytes; limit 0 = not yet sized
    static size_t round_up(size_t b) { return (b + kGrain - 1) / kGrain * kGrain; }
    void trim() {
        for (auto &kv : free_list) {
            cudaFree(kv.second.p);
            if (kv.second.ev) cudaEventDestroy(kv.second.ev);
        }
        free_list.clear();
        pooled = 0;
    }
    // returns {nullptr} when the limit or the device is exhausted (after giving the pooled buffers back)
    PoolBuf take(size_t bytes, size_t *cap) {
        const size_t want = round_up(bytes);
        auto it = free_list.lower_bound(want);
        if (it != free_list.end() && it->first <= want + want / 4) {
            PoolBuf b = it->second;
            *cap = it->first;
            pooled -= it->first; in_use += it->first;
            free_list.erase(it);
            return b;
        }
        if (limit && in_use + pooled + want > limit) trim();
        if (limit && in_use + want > limit) return PoolBuf{};
        void *p = nullptr;
        if (cudaMalloc(&p, want) != cudaSuccess) {
            cudaGetLastError();
            trim();
            if (cudaMalloc(&p, want) != cudaSuccess) { cudaGetLastError(); return PoolBuf{}; }
        }
        *cap = want; in_use += want;
        return PoolBuf{p, nullptr};
    }
    void give(PoolBuf b, size_t cap) {
        in_use -= cap; pooled += cap;
        free_list.emplace(cap, b);
    }
};

constexpr int kEncStreams = 8;   // resident encrypt calls rotate over this many streams (and input slots)

}  // namespace

// A vector<Ciphertext<DCRTPoly>> that lives in HBM: the dense payload [n_ct][2][limbs][N] plus what the wire format
// needs to write it out again byte for byte (context variant, per-ciphertext metadata).
struct shelfi_ct {
    shelfi_ckks *owner = nullptr;
    u64 *d = nullptr;
    size_t cap = 0;
    cudaEvent_t buf_ev = nullptr;   // travels with the buffer (DevicePool)
    // a vector made by encrypt is written on one of the encrypt streams: `ready` is recorded there, and the first use
    // on the compute stream (or the free) makes that stream wait for it
    cudaEvent_t ready = nullptr;
    bool ready_pending = false;
    // a large weighted average is written on the compute stream in kCtParts groups of ciphertexts (group g =
    // [n_ct g / kCtParts, n_ct (g + 1) / kCtParts)); part_ev[g] is recorded after group g, so that decrypt can start on
    // the first group (and its values cross PCIe) while the later ones are still being summed
    cudaEvent_t part_ev[8] = {nullptr};
    uint32_t parts = 0;
    // a vector encrypted in kUploadChunks groups on as many streams keeps an event per group (group g =
    // [n_ct g / chunks, n_ct (g + 1) / chunks)): a weighted sum in parts starts on the first ciphertexts while the
    // last group of the last learner is still being encrypted
    cudaEvent_t chunk_ev[4] = {nullptr, nullptr, nullptr, nullptr};
    uint32_t chunks = 0;
    uint64_t n_ct = 0;
    uint32_t limbs = 0;
    wire::ContextParams ctx;
    std::vector<wire::CiphertextMeta> meta;   // one entry (every ciphertext alike) or n_ct entries
};

struct shelfi_ckks {
    std::string scheme, cryptodir, err;
    uint32_t batch = 0, scale_bits = 0;
    int device = 0;
    bool cuda_ready = false;
    bool have_cc = false, have_pk = false, have_sk = false;
    wire::ContextParams ctx;
    std::string key_tag;
    RingConst rc{};
    EncodeConst ec{};
    Crt2Const crt{};
    CrtNConst crtn{};
    uint64_t gauss_cdf[32] = {0};
    // device-resident parameters
    DeviceBuffer d_tw, d_itw, d_dtw, d_pk_b, d_pk_a, d_sk, d_fft_inv, d_fft_fwd;
    // scratch
    DeviceBuffer s_x, s_m, s_ve, s_coeff, s_dense, s_image, s_image2, s_image3, s_out, s_vals, s_zs, s_rows, s_pt, s_resc, s_dense2;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev_stage_out = nullptr;
    // rings above one tile: learner i+1 is encoded and sampled on this stream while learner i's transforms run
    cudaStream_t enc_stream = nullptr;
    cudaEvent_t ev_enc_fork = nullptr, ev_enc_ready[2] = {nullptr, nullptr}, ev_enc_free[2] = {nullptr, nullptr};
    cudaStream_t d2h_stream = nullptr;   // created on first use: results leave while the next inputs arrive (full duplex)
    cudaEvent_t ev_copy[2] = {nullptr, nullptr}, ev_done[2] = {nullptr, nullptr};
    cudaStream_t copy_stream = nullptr;
    bool seeded = false;
    uint64_t seed = 0;
    Profiler prof;
    std::mutex mu;
    // pageable host buffers: staging chunks + copy threads (created on first use)
    void *pin_x = nullptr;       // page-locked copy of a pageable encrypt input (grow-only)
    size_t pin_x_cap = 0;
    void *stage[2] = {nullptr, nullptr};
    cudaEvent_t ev_stage[2] = {nullptr, nullptr};
    // host -> device staging of pageable memory: the two buffers as a ring of kStagePieces pieces, each with the event
    // of the copy that last read it (h2d)
    cudaEvent_t ev_piece[kStagePieces] = {nullptr};
    bool piece_busy[kStagePieces] = {false};
    unsigned piece_turn = 0;
    ParallelCopier *copier = nullptr;
    // encrypt writes the same bookkeeping bytes around the coefficient runs every time a vector of n_ct ciphertexts is
    // serialized under the loaded keys: the plan (2035 small segments at 407 ciphertexts, 0.24 ms to build and 0.09 ms
    // to copy in on the host) is built once per n_ct and its segments live on the device, where a kernel drops them
    // into the image before it leaves.
    struct EncPlan {
        wire::BlobPlan plan;
        DeviceBuffer d_desc, d_bytes;
        uint32_t n_seg = 0;
    };
    std::map<std::string, EncPlan> enc_plans;   // key: context bytes + n_ct + the (uniform) ciphertext metadata
    // resident ciphertext vectors (shelfi_ct_*): pool, two input slots so that the upload of one vector overlaps the
    // encryption of the one before, and the events that order the copy streams against the compute stream
    DevicePool pool;
    std::vector<shelfi_ct *> live;
    // Consecutive encrypt calls are independent of one another: they rotate over kEncStreams streams (each with its own
    // input slot), so the next learner's kernel fills the SMs the current one leaves idle (407 CTAs on 444 slots at
    // config 2) and small vectors (3 CTAs per call in the pythonApi example) run side by side.  Everything else that
    // touches a vector runs on the compute stream `stream`.
    cudaStream_t enc_streams[kEncStreams] = {nullptr};
    bool opt_no_chunk_wait = std::getenv("SHELFI_NO_CHUNK_WAIT") != nullptr;   // A/B switch: the weighted sum waits for whole vectors
    bool opt_unfused_decrypt = std::getenv("SHELFI_DECRYPT_UNFUSED") != nullptr;   // A/B switch: two launches as in round 1
    cudaEvent_t ev_fork = nullptr, ev_part_done[kEncStreams] = {nullptr};   // ct_decrypt: groups on the side streams
    DeviceBuffer r_x[kEncStreams], r_out;
    cudaEvent_t ev_x_in[kEncStreams] = {nullptr}, ev_x_free[kEncStreams] = {nullptr}, ev_r_out = nullptr;
    bool x_busy[kEncStreams] = {false};
    unsigned r_turn = 0;
    std::vector<cudaEvent_t> spare_events;
    // The device-resident calls run on the CALLER's stream but share this handle's scratch (encode / sample output,
    // decrypt coefficients, CRT doubles, rescale tower).  Two calls on different streams would otherwise be free to
    // overlap on the GPU and corrupt each other's intermediates: every such call first makes its stream wait for the
    // event the previous one recorded after its last use of the scratch.
    cudaEvent_t ev_scratch = nullptr;
    cudaStream_t scratch_stream = nullptr;
    bool scratch_used = false;
    // the semantics the reference's fixtures do not pin, switchable (shelfi_set_option; same ids in the oracle)
    int opt_evalmult_round = 0;   // 0: (int64)(c * Delta + 0.5);  1: llround(c * Delta)
    int opt_decode_noise = 0;     // 1: PALISADE >= 1.10.6 Decode noise flooding
    int opt_wire_variant = 0;     // generated contexts: 0 = default (9 RLWE words), 8 or 9
    SamplerKey noise_key{};       // key of the decode noise of the current reference-facing call
};

namespace {

int fail(shelfi_ckks *h, int code, const std::string &msg) {
    if (h) h->err = msg;
    return code;
}
int cuda_fail(shelfi_ckks *h, cudaError_t e, const char *where) {
    return fail(h, SHELFI_ERR_CUDA, std::string(where) + ": " + cudaGetErrorString(e));
}
#define CU(call)                                                   \
    do {                                                           \
        cudaError_t e__ = (call);                                  \
        if (e__ != cudaSuccess) return cuda_fail(h, e__, #call);   \
    } while (0)

template <class F>
void timed(shelfi_ckks *h, int id, cudaStream_t s, F &&launch) {
    if (!h->prof.on || h->prof.pending.size() >= 65536) { launch(); return; }
    Profiler::Rec r{id, h->prof.take(), h->prof.take()};
    cudaEventRecord(r.a, s);
    launch();
    cudaEventRecord(r.b, s);
    h->prof.pending.push_back(r);
}

int ensure_cuda(shelfi_ckks *h) {
    if (h->cuda_ready) {
        CU(cudaSetDevice(h->device));
        return 0;
    }
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0)
        return fail(h, SHELFI_ERR_CUDA,
                    std::string("no CUDA device: this engine has no CPU path (") + cudaGetErrorString(e) + ")");
    if (h->device < 0 || h->device >= count) return fail(h, SHELFI_ERR_ARG, "device ordinal out of range");
    CU(cudaSetDevice(h->device));
    cudaDeviceProp prop;
    CU(cudaGetDeviceProperties(&prop, h->device));
    if (prop.major < 10)
        return fail(h, SHELFI_ERR_CUDA, std::string("device ") + prop.name + " is not sm_100-class; kernels are built for sm_100a only");
    CU(kernels_init());
    CU(cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking));
    CU(cudaStreamCreateWithFlags(&h->copy_stream, cudaStreamNonBlocking));
    for (int i = 0; i < 2; i++) {
        CU(cudaEventCreateWithFlags(&h->ev_copy[i], cudaEventDisableTiming));
        CU(cudaEventCreateWithFlags(&h->ev_done[i], cudaEventDisableTiming));
    }
    h->cuda_ready = true;
    return 0;
}

bool read_file(const std::string &path, std::vector<uint8_t> *out) {
    FILE *f = std::fopen(path.c_str(), "rb");
    if (!f) return false;
    std::fseek(f, 0, SEEK_END);
    long n = std::ftell(f);
    std::fseek(f, 0, SEEK_SET);
    out->resize(n > 0 ? (size_t)n : 0);
    bool ok = n >= 0 && std::fread(out->data(), 1, out->size(), f) == out->size();
    std::fclose(f);
    return ok;
}
bool write_file(const std::string &path, const std::vector<uint8_t> &data) {
    FILE *f = std::fopen(path.c_str(), "wb");
    if (!f) return false;
    bool ok = std::fwrite(data.data(), 1, data.size(), f) == data.size();
    return std::fclose(f) == 0 && ok;
}

void drop_enc_plans(shelfi_ckks *h) {
    for (auto &kv : h->enc_plans) { kv.second.d_desc.release(); kv.second.d_bytes.release(); }
    h->enc_plans.clear();
}
void free_params(shelfi_ckks *h) {
    drop_enc_plans(h);
    h->d_tw.release(); h->d_itw.release(); h->d_dtw.release(); h->d_pk_b.release(); h->d_pk_a.release(); h->d_sk.release();
    h->d_fft_inv.release(); h->d_fft_fwd.release();
}

int upload(shelfi_ckks *h, DeviceBuffer *buf, const void *src, size_t bytes) {
    CU(buf->reserve(bytes));
    CU(cudaMemcpy(buf->p, src, bytes, cudaMemcpyHostToDevice));
    return 0;
}

bool is_pinned(const void *p) {
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return false; }
    return a.type == cudaMemoryTypeHost || a.type == cudaMemoryTypeManaged;
}
int ensure_staging(shelfi_ckks *h) {
    if (h->copier) return 0;
    for (int i = 0; i < 2; i++) {
        CU(cudaHostAlloc(&h->stage[i], kStageBytes, cudaHostAllocDefault));
        CU(cudaEventCreateWithFlags(&h->ev_stage[i], cudaEventDisableTiming));
    }
    for (int i = 0; i < kStagePieces; i++) CU(cudaEventCreateWithFlags(&h->ev_piece[i], cudaEventDisableTiming));
    // (one process per GPU: the box's cores are shared by LOCAL_WORLD_SIZE processes, and polling copy threads that
    // outnumber them slow every rank down)
    unsigned cores = std::max(1u, std::thread::hardware_concurrency());
    if (const char *e = std::getenv("LOCAL_WORLD_SIZE")) cores = std::max(1u, cores / (unsigned)std::max(1, std::atoi(e)));
    unsigned n = std::min(8u, cores);
    if (const char *e = std::getenv("SHELFI_COPY_THREADS")) n = (unsigned)std::max(1, std::atoi(e));
    h->copier = new ParallelCopier(n);
    return 0;
}
// Host -> device, queued on `s`.  Page-locked sources go straight to the copy engine (asynchronous); pageable ones are
// copied chunk by chunk into page-locked staging by the copy threads (the call returns when the last chunk is queued).
int h2d(shelfi_ckks *h, void *dev, const void *host, size_t bytes, cudaStream_t s) {
    // (from 1 MiB: the per-group inputs of a pipelined encrypt are 2 MB of pageable numpy memory, which the driver's
    // own pageable path moves at a few GB/s and synchronously)
    if (bytes < (1u << 20) || is_pinned(host)) {
        CU(cudaMemcpyAsync(dev, host, bytes, cudaMemcpyHostToDevice, s));
        return 0;
    }
    int rcode = ensure_staging(h);
    if (rcode) return rcode;
    // piece by piece through the ring: the copy threads fill piece i + 1 while the DMA engine reads piece i; a piece is
    // reused (by this call or a later one) only after the copy that read it has finished, so the call can return as
    // soon as the last piece is queued — the caller's memory has been read by then
    for (size_t off = 0; off < bytes; off += kPieceBytes) {
        const unsigned i = h->piece_turn++ % kStagePieces;
        const size_t len = std::min(kPieceBytes, bytes - off);
        char *piece = static_cast<char *>(h->stage[i / (kStagePieces / 2)]) + (size_t)(i % (kStagePieces / 2)) * kPieceBytes;
        if (h->piece_busy[i]) CU(cudaEventSynchronize(h->ev_piece[i]));
        h->copier->copy(piece, static_cast<const char *>(host) + off, len, off > 0);
        CU(cudaMemcpyAsync(static_cast<char *>(dev) + off, piece, len, cudaMemcpyHostToDevice, s));
        CU(cudaEventRecord(h->ev_piece[i], s));
        h->piece_busy[i] = true;
    }
    return 0;
}
// Device -> host after everything queued on `s`; returns when the bytes are in `host`.
int d2h_sync(shelfi_ckks *h, void *host, const void *dev, size_t bytes, cudaStream_t s) {
    if (bytes < (4u << 20) || is_pinned(host)) {
        CU(cudaMemcpyAsync(host, dev, bytes, cudaMemcpyDeviceToHost, s));
        CU(cudaStreamSynchronize(s));
        return 0;
    }
    int rcode = ensure_staging(h);
    if (rcode) return rcode;
    for (int i = 0; i < kStagePieces; i++)   // uploads still reading the staging buffers (h2d's ring)
        if (h->piece_busy[i]) { CU(cudaEventSynchronize(h->ev_piece[i])); h->piece_busy[i] = false; }
    const size_t chunks = (bytes + kStageBytes - 1) / kStageBytes;
    auto issue = [&](size_t c) -> cudaError_t {
        const size_t off = c * kStageBytes, len = std::min(kStageBytes, bytes - off);
        cudaError_t e = cudaMemcpyAsync(h->stage[c & 1], static_cast<const char *>(dev) + off, len, cudaMemcpyDeviceToHost, s);
        return e != cudaSuccess ? e : cudaEventRecord(h->ev_stage[c & 1], s);
    };
    CU(issue(0));
    if (chunks > 1) CU(issue(1));
    for (size_t c = 0; c < chunks; c++) {
        const size_t off = c * kStageBytes, len = std::min(kStageBytes, bytes - off);
        CU(cudaEventSynchronize(h->ev_stage[c & 1]));
        h->copier->copy(static_cast<char *>(host) + off, h->stage[c & 1], len);
        if (c + 2 < chunks) CU(issue(c + 2));
    }
    return 0;
}

int install_context(shelfi_ckks *h, const wire::ContextParams &p) {
    uint32_t logn = 0;
    while ((1u << logn) < p.ring_dim) logn++;
    if (p.ring_dim != (1u << logn) || logn < (uint32_t)kLogN || logn > (uint32_t)(kLogN + kMaxLogBlocks) ||
        p.cyclotomic_order != 2u * p.ring_dim)
        return fail(h, SHELFI_ERR_UNSUPPORTED, "ring dimension " + std::to_string(p.ring_dim) +
                                                   " not supported: the kernels cover N = 8192, 16384, 32768");
    const uint32_t N = p.ring_dim;
    if (p.batch_size == 0 || (p.batch_size & (p.batch_size - 1)) || p.batch_size > N / 2)
        return fail(h, SHELFI_ERR_UNSUPPORTED, "context batch size must be a power of two <= N/2");
    if (p.limbs < 1 || p.limbs > (uint32_t)kMaxLimbs) return fail(h, SHELFI_ERR_UNSUPPORTED, "unsupported tower count");
    RingConst rc{};
    rc.n = N; rc.logn = logn; rc.log_blocks = logn - kLogN; rc.limbs = p.limbs; rc.slots = p.batch_size;
    std::vector<tw_t> tw_all((size_t)p.limbs * N), itw_all((size_t)p.limbs * N), tw, itw;
    std::vector<double> dtw_all((size_t)p.limbs * 2 * N, 0.0), dtw;   // per limb: forward twiddles, then inverse
    bool has_dtw[kMaxLimbs] = {false};
    // SHELFI_FP64_NTT=0 keeps every limb on the integer pipe (A/B measurements; the results are the same bits)
    const char *fp64_env = std::getenv("SHELFI_FP64_NTT");
    const bool want_fp64 = !(fp64_env && std::atoi(fp64_env) == 0);
    for (uint32_t l = 0; l < p.limbs; l++) {
        std::string e = make_limb_const(p.modulus[l], p.root[l], logn, &rc.limb[l], &tw, &itw, &dtw);
        if (!e.empty()) return fail(h, SHELFI_ERR_UNSUPPORTED, "limb " + std::to_string(l) + ": " + e);
        std::memcpy(&tw_all[(size_t)l * N], tw.data(), sizeof(tw_t) * N);
        std::memcpy(&itw_all[(size_t)l * N], itw.data(), sizeof(tw_t) * N);
        has_dtw[l] = want_fp64 && !dtw.empty();
        if (has_dtw[l]) std::memcpy(&dtw_all[(size_t)l * 2 * N], dtw.data(), sizeof(double) * 2 * N);
    }
    int rcode;
    if ((rcode = upload(h, &h->d_dtw, dtw_all.data(), dtw_all.size() * sizeof(double)))) return rcode;
    if ((rcode = upload(h, &h->d_tw, tw_all.data(), tw_all.size() * sizeof(tw_t)))) return rcode;
    if ((rcode = upload(h, &h->d_itw, itw_all.data(), itw_all.size() * sizeof(tw_t)))) return rcode;
    for (uint32_t l = 0; l < p.limbs; l++) {
        rc.limb[l].tw = h->d_tw.as<tw_t>() + (size_t)l * N;
        rc.limb[l].itw = h->d_itw.as<tw_t>() + (size_t)l * N;
        rc.limb[l].dtw = has_dtw[l] ? h->d_dtw.as<double>() + (size_t)l * 2 * N : nullptr;
        rc.limb[l].ditw = has_dtw[l] ? rc.limb[l].dtw + N : nullptr;
    }
    const uint32_t slots = p.batch_size;
    std::vector<cplx> t_inv, t_fwd;
    make_fft_tables(p.cyclotomic_order, slots, &t_inv, &t_fwd);
    if ((rcode = upload(h, &h->d_fft_inv, t_inv.data(), slots * sizeof(cplx)))) return rcode;
    if ((rcode = upload(h, &h->d_fft_fwd, t_fwd.data(), slots * sizeof(cplx)))) return rcode;
    EncodeConst ec{};
    ec.slots = slots; ec.gap = (N / 2) / slots; ec.batch = h->batch;
    ec.log_slots = 0;
    while ((1u << ec.log_slots) < slots) ec.log_slots++;
    ec.pow_p = (double)p.modulus[p.limbs - 1];
    ec.inv_size = 1.0 / (double)slots;
    ec.tw_inv = h->d_fft_inv.as<cplx>();
    ec.tw_fwd = h->d_fft_fwd.as<cplx>();
    Crt2Const crt = make_crt2(p.modulus[0], p.limbs >= 2 ? p.modulus[1] : 0);
    build_gauss_cdf((double)p.sigma, h->gauss_cdf);
    h->ctx = p; h->rc = rc; h->ec = ec; h->crt = crt;
    h->crtn = make_crtn(p.modulus, p.limbs);
    h->have_cc = true;
    return 0;
}

// key polynomial rows -> {value, Shoup companion} pairs on the device
int upload_key_poly(shelfi_ckks *h, DeviceBuffer *buf, const uint64_t *rows) {
    const uint32_t L = h->ctx.limbs, N = h->ctx.ring_dim;
    std::vector<tw_t> packed((size_t)L * N);
    for (uint32_t l = 0; l < L; l++)
        for (uint32_t j = 0; j < N; j++) {
            uint64_t v = rows[(size_t)l * N + j];
            if (v >= h->ctx.modulus[l]) return fail(h, SHELFI_ERR_FORMAT, "key coefficient not reduced");
            packed[(size_t)l * N + j] = tw_t{v, nt::shoup(v, h->ctx.modulus[l])};
        }
    return upload(h, buf, packed.data(), packed.size() * sizeof(tw_t));
}

bool same_ring(const wire::ContextParams &a, const wire::ContextParams &b) {
    return a.ring_dim == b.ring_dim && a.limbs == b.limbs &&
           !std::memcmp(a.modulus, b.modulus, sizeof(uint64_t) * a.limbs) &&
           !std::memcmp(a.root, b.root, sizeof(uint64_t) * a.limbs);
}

BlobRuns to_runs(const shelfi_ckks *h, const uint64_t *run0, const uint64_t *run1, uint64_t stride, uint32_t limbs = 0) {
    const uint32_t L = limbs ? limbs : h->ctx.limbs;
    BlobRuns br{};
    br.runs = 2 * L;
    br.n = h->ctx.ring_dim;
    for (uint32_t r = 0; r < 2 * L; r++) { br.off0[r] = run0[r]; br.off1[r] = run1[r]; }
    br.stride = stride;
    return br;
}
size_t image_bytes(uint64_t blob_size) { return ((size_t)blob_size + 7) / 8 * 8 + 8; }

int64_t weight_constant(const shelfi_ckks *h, float w) {
    double sc = (double)w;  // ckks.cpp:287 narrows to float32; PALISADE's EvalMult(ct, double) widens it back
    if (h->opt_evalmult_round == 1) return (int64_t)std::llround(sc * (double)h->ctx.modulus[h->ctx.limbs - 1]);
    return (int64_t)(sc * (double)h->ctx.modulus[h->ctx.limbs - 1] + 0.5);
}

// 256 bits from the kernel's CSPRNG (getrandom(2), blocking until the pool is initialised): one ChaCha20 key per
// encrypt call and learner, two independent ones per key generation (secret material / public material).  The
// reference gets the same from PALISADE's PRNG, which seeds itself from OS entropy.
bool fresh_key(SamplerKey *k) {
    unsigned char *p = reinterpret_cast<unsigned char *>(k->k);
    size_t got = 0;
    while (got < sizeof k->k) {
        ssize_t r = getrandom(p + got, sizeof k->k - got, 0);
        if (r < 0) {
            if (errno == EINTR) continue;
            return false;
        }
        got += (size_t)r;
    }
    return true;
}
// the key of one call: from the OS, or (tests) laid out from the fixed seed for the given domain
int call_key(shelfi_ckks *h, uint32_t domain, SamplerKey *k) {
    if (h->seeded) { *k = seed_key(h->seed, domain); return 0; }
    if (!fresh_key(k)) return fail(h, SHELFI_ERR_IO, std::string("getrandom failed: ") + std::strerror(errno));
    return 0;
}

// scratch hand-over between caller streams (see shelfi_ckks::ev_scratch)
int scratch_begin(shelfi_ckks *h, cudaStream_t s) {
    if (!h->ev_scratch) CU(cudaEventCreateWithFlags(&h->ev_scratch, cudaEventDisableTiming));
    if (h->scratch_used && s != h->scratch_stream) CU(cudaStreamWaitEvent(s, h->ev_scratch, 0));
    return 0;
}
int scratch_end(shelfi_ckks *h, cudaStream_t s) {
    CU(cudaEventRecord(h->ev_scratch, s));
    h->scratch_stream = s;
    h->scratch_used = true;
    return 0;
}

// --- stage drivers (all asynchronous on `s`) ---
// k learners' vectors (all of length n) in one launch: the grid covers k * n_ct ciphertexts, so small models still
// fill the machine and co-resident CTAs are out of phase with one another.
// dtype 1: the d_x[i] hold float32 values (widened on load)
int encrypt_dense_many(shelfi_ckks *h, const double *const *d_x, uint32_t k, uint64_t n, uint64_t ct_index0,
                       uint64_t *const *d_ct, cudaStream_t s, const uint32_t *d_sel = nullptr, int dtype = 0) {
    uint64_t n_ct = (n + h->batch - 1) / h->batch;
    const uint64_t N = h->ctx.ring_dim;
    if (n_ct * k > (0x7fffffffull / (2 * kMaxLimbs)) >> h->rc.log_blocks)
        return fail(h, SHELFI_ERR_ARG, "too many ciphertexts for one call");
    SamplerConst sc{};
    sc.ct_index0 = ct_index0;
    std::memcpy(sc.gauss_cdf, h->gauss_cdf, sizeof sc.gauss_cdf);
    if (h->rc.log_blocks > 0) {
        int src = scratch_begin(h, s);
        if (src) return src;
        // ring dimension above one tile: encode per FFT block, sampler pass, then one CTA per (ciphertext, limb, tile);
        // a single learner already gives thousands of CTAs, so learners go one after another
        // The encode/sample kernels are latency-bound and light on the arithmetic pipes; on a second stream they run
        // under the previous learner's transforms (two scratch slots, events both ways).
        const bool overlap = k > 1;
        const size_t m_words = (size_t)n_ct * N;
        CU(h->s_m.reserve((overlap ? 2 : 1) * m_words * sizeof(int64_t)));
        CU(h->s_ve.reserve((overlap ? 2 : 1) * m_words * sizeof(uint8_t)));
        cudaStream_t es = s;
        if (overlap) {
            if (!h->enc_stream) {
                CU(cudaStreamCreateWithFlags(&h->enc_stream, cudaStreamNonBlocking));
                CU(cudaEventCreateWithFlags(&h->ev_enc_fork, cudaEventDisableTiming));
                for (int i = 0; i < 2; i++) {
                    CU(cudaEventCreateWithFlags(&h->ev_enc_ready[i], cudaEventDisableTiming));
                    CU(cudaEventCreateWithFlags(&h->ev_enc_free[i], cudaEventDisableTiming));
                }
            }
            es = h->enc_stream;
            CU(cudaEventRecord(h->ev_enc_fork, s));          // inputs and scratch are ordered after the caller's work
            CU(cudaStreamWaitEvent(es, h->ev_enc_fork, 0));
        }
        for (uint32_t i = 0; i < k; i++) {
            const int b = overlap ? (int)(i & 1) : 0;
            i64 *m_slot = h->s_m.as<i64>() + (size_t)b * m_words;
            uint8_t *ve_slot = h->s_ve.as<uint8_t>() + (size_t)b * m_words;
            int krc = call_key(h, kKeyDomainEncrypt, &sc.key);
            if (krc) return krc;
            if (overlap && i >= 2) CU(cudaStreamWaitEvent(es, h->ev_enc_free[b], 0));   // slot b was read by learner i-2
            timed(h, kKEncodeSample, es, [&] {
                launch_encode_sample(h->rc, h->ec, sc, d_x[i], n, (u32)n_ct, m_slot, ve_slot, es, d_sel, dtype == 1);
            });
            if (overlap) {
                CU(cudaEventRecord(h->ev_enc_ready[b], es));
                CU(cudaStreamWaitEvent(s, h->ev_enc_ready[b], 0));
            }
            timed(h, kKEncrypt, s, [&] {
                launch_encrypt(h->rc, m_slot, ve_slot, h->d_pk_b.as<tw_t>(), h->d_pk_a.as<tw_t>(),
                               reinterpret_cast<u64 *>(d_ct[i]), (u32)n_ct, s);
            });
            if (overlap) CU(cudaEventRecord(h->ev_enc_free[b], s));
        }
        CU(cudaGetLastError());
        return scratch_end(h, s);
    }
    for (uint32_t base = 0; base < k; base += kMaxEncryptBatch) {
        EncryptBatch eb{};
        eb.k = std::min<uint32_t>(kMaxEncryptBatch, k - base);
        eb.n_ct = (u32)n_ct;
        eb.sel = d_sel;
        eb.f32 = dtype == 1 ? 1u : 0u;
        for (uint32_t i = 0; i < eb.k; i++) {
            eb.x[i] = d_x[base + i];
            eb.ct[i] = reinterpret_cast<u64 *>(d_ct[base + i]);
            int krc = call_key(h, kKeyDomainEncrypt, &eb.key[i]);
            if (krc) return krc;
        }
        timed(h, kKEncrypt, s, [&] {
            launch_encrypt_fused(h->rc, h->ec, sc, eb, n, h->d_pk_b.as<tw_t>(), h->d_pk_a.as<tw_t>(), s);
        });
    }
    CU(cudaGetLastError());
    return 0;
}
int encrypt_dense(shelfi_ckks *h, const double *d_x, uint64_t n, uint64_t ct_index0, uint64_t *d_ct, cudaStream_t s) {
    return encrypt_dense_many(h, &d_x, 1, n, ct_index0, &d_ct, s);
}

// EvalMult(ct, (float)w) + EvalAdd over k dense batches (ckks.cpp:286-297), 16 learners per launch
int weighted_sum_dense(shelfi_ckks *h, const uint64_t *const *d_cts, const float *weights, uint32_t k, uint64_t n_ct,
                       uint64_t *d_out, bool accumulate, cudaStream_t s) {
    const uint32_t L = h->ctx.limbs;
    for (uint32_t base = 0; base < k; base += kMaxLearnersPerLaunch) {
        WavgArgs wa{};
        wa.k = std::min<uint32_t>(kMaxLearnersPerLaunch, k - base);
        wa.accumulate = (base || accumulate) ? 1u : 0u;
        for (uint32_t i = 0; i < wa.k; i++) {
            wa.in[i] = reinterpret_cast<const u64 *>(d_cts[base + i]);
            int64_t sc = weight_constant(h, weights[base + i]);
            for (uint32_t l = 0; l < L; l++) wa.c[i][l] = nt::signed_mod(sc, h->ctx.modulus[l]);
        }
        timed(h, kKWeightedSum, s, [&] { launch_weighted_sum(h->rc, wa, reinterpret_cast<u64 *>(d_out), n_ct, s); });
    }
    CU(cudaGetLastError());
    return 0;
}

bool general_decode(const shelfi_ckks *h, uint32_t limbs = 0) {
    return h->rc.log_blocks > 0 || (limbs ? limbs : h->ctx.limbs) > 2;
}

// A rescaled ciphertext keeps the first `limbs` towers of the chain: the same ring and tables, a shorter CRT.
// ([PAL] Decrypt drops the secret key's extra towers to match; here the kernels simply stop at rc.limbs.)
struct LevelView {
    RingConst rc;
    Crt2Const crt;
    CrtNConst crtn;
};
LevelView level_view(const shelfi_ckks *h, uint32_t limbs) {
    LevelView v{h->rc, h->crt, h->crtn};
    if (limbs && limbs != h->ctx.limbs) {
        v.rc.limbs = limbs;
        v.crt = make_crt2(h->ctx.modulus[0], limbs >= 2 ? h->ctx.modulus[1] : 0);
        v.crtn = make_crtn(h->ctx.modulus, limbs);
    }
    return v;
}

// limbs = towers per polynomial of d_ct (0 = the context's chain, i.e. level 0)
// peers[0..n_peer): further destinations of every decoded value (decrypt-gather), already offset to this rank's slice
// ct_index0: index of d_ct[0] within the vector the caller decrypts (names the decode-noise stream of each ciphertext)
// slice: null = the handle's scratch from its start, handed over between streams here; else the caller has reserved
//        the three buffers for `slice->of` ciphertexts, done the hand-over, and this launch works at ciphertext
//        `slice->at` of them (several launches on different streams then share the scratch without touching)
struct ScratchSlice { uint64_t at, of; };
constexpr uint64_t kFusedDecryptMinCts = 256;
int decrypt_scratch_reserve(shelfi_ckks *h, uint64_t n_ct, uint32_t limbs) {
    const uint64_t N = h->ctx.ring_dim;
    CU(h->s_coeff.reserve(n_ct * limbs * N * sizeof(uint64_t)));
    if (general_decode(h, limbs) || h->opt_decode_noise) {
        CU(h->s_vals.reserve(n_ct * N * sizeof(double)));
        if (h->ec.slots > kFftFast) CU(h->s_zs.reserve(n_ct * (size_t)h->ec.slots * sizeof(cplx)));
    }
    return 0;
}
int decrypt_dense(shelfi_ckks *h, const uint64_t *d_ct, uint64_t n_ct, double sf, double *d_out, uint64_t n,
                  cudaStream_t s, const uint32_t *d_sel = nullptr, uint32_t limbs = 0, double *const *peers = nullptr,
                  uint32_t n_peer = 0, uint64_t ct_index0 = 0, const ScratchSlice *slice = nullptr) {
    const uint64_t N = h->ctx.ring_dim;
    const LevelView lv = level_view(h, limbs);
    int src;
    if (!slice) {
        if ((src = scratch_begin(h, s))) return src;
        if ((src = decrypt_scratch_reserve(h, n_ct, lv.rc.limbs))) return src;
    }
    const uint64_t at = slice ? slice->at : 0;
    u64 *coeff = h->s_coeff.as<u64>() + at * lv.rc.limbs * N;
    // one launch for both stages from kFusedDecryptMinCts ciphertexts; below that (the groups of a pipelined call,
    // small vectors) a launch is less than a wave and two kernels with one CTA per limb / ciphertext finish sooner
    const bool fused = !general_decode(h, lv.rc.limbs) && !h->opt_decode_noise && !h->opt_unfused_decrypt &&
                       n_ct >= kFusedDecryptMinCts;
    if (!fused)
        timed(h, kKDecryptIntt, s, [&] { launch_decrypt_intt(lv.rc, d_ct, h->d_sk.as<tw_t>(), coeff, (u32)n_ct, s); });
    DecodeConst dc{};
    dc.crt = lv.crt;
    double p = (double)h->ctx.plaintext_modulus;
    dc.pre = std::pow(sf, -1) * std::pow(2, p);   // [PAL] Decode: scale to 2^p before the FFT, back after it
    dc.pow_p = std::pow(2, -p);
    for (uint32_t i = 0; i < n_peer; i++) dc.peer[i] = peers[i];
    dc.n_peer = n_peer;
    if (fused) {   // ring dimension 8192, at most two limbs: one launch (timed as the inverse-transform stage)
        timed(h, kKDecryptIntt, s, [&] {
            launch_decrypt_fused(lv.rc, h->ec, dc, d_ct, h->d_sk.as<tw_t>(), coeff, (u32)n_ct, d_out, n, s, d_sel);
        });
    } else if (!general_decode(h, lv.rc.limbs) && !h->opt_decode_noise) {
        timed(h, kKDecode, s, [&] { launch_decode(lv.rc, h->ec, dc, coeff, (u32)n_ct, d_out, n, s, d_sel); });
    } else {
        double *vals = h->s_vals.as<double>() + at * N;
        cplx *zs = h->s_zs.as<cplx>() ? h->s_zs.as<cplx>() + at * (size_t)h->ec.slots : nullptr;
        timed(h, kKDecode, s, [&] {
            launch_crt(lv.rc, dc, lv.crtn, coeff, (u32)n_ct, vals, nullptr, nullptr, nullptr, s);
            if (h->opt_decode_noise)   // [PAL] Decode adds its noise at scale 2^p, before FFTSpecial
                launch_decode_noise(lv.rc, h->ec, h->noise_key, ct_index0, vals, (u32)n_ct, s);
            launch_decode_fft(lv.rc, h->ec, dc, vals, zs, (u32)n_ct, d_out, n, s, d_sel);
        });
    }
    CU(cudaGetLastError());
    return slice ? 0 : scratch_end(h, s);
}

// [PAL] ModReduce under EXACTRESCALE (DCRTPoly::DropLastElementAndScale) on a dense batch:
// d_ct[n_ct][2][limbs_in][N] -> d_out[n_ct][2][limbs_in - 1][N].  (c - [c]_{q_l}) / q_l with the centred remainder.
int rescale_dense(shelfi_ckks *h, const uint64_t *d_ct, uint64_t n_ct, uint32_t limbs_in, uint64_t *d_out, cudaStream_t s) {
    if (limbs_in < 2 || limbs_in > h->ctx.limbs) return fail(h, SHELFI_ERR_ARG, "no tower left to drop");
    if (n_ct > (0x7fffffffull / (2 * kMaxLimbs)) >> h->rc.log_blocks) return fail(h, SHELFI_ERR_ARG, "too many ciphertexts for one call");
    const LevelView lv = level_view(h, limbs_in);
    RescaleConst rs{};
    rs.q_last = h->ctx.modulus[limbs_in - 1];
    for (uint32_t i = 0; i + 1 < limbs_in; i++) {
        const uint64_t q = h->ctx.modulus[i];
        rs.inv[i] = nt::inv_mod(rs.q_last % q, q);
        rs.inv_s[i] = nt::shoup(rs.inv[i], q);
    }
    int src = scratch_begin(h, s);
    if (src) return src;
    CU(h->s_resc.reserve(n_ct * 2 * h->ctx.ring_dim * sizeof(uint64_t)));
    timed(h, kKRescale, s, [&] { launch_rescale(lv.rc, rs, d_ct, h->s_resc.as<u64>(), d_out, (u32)n_ct, s); });
    CU(cudaGetLastError());
    return scratch_end(h, s);
}

int check_stage(shelfi_ckks *h, bool need_pk, bool need_sk) {
    if (!h) return SHELFI_ERR_ARG;
    int rcode = ensure_cuda(h);
    if (rcode) return rcode;
    if (!h->have_cc) return fail(h, SHELFI_ERR_NOT_LOADED, "crypto context not loaded (call loadCryptoParams)");
    if (need_pk && !h->have_pk) return fail(h, SHELFI_ERR_NOT_LOADED, "public key not loaded");
    if (need_sk && !h->have_sk) return fail(h, SHELFI_ERR_NOT_LOADED, "private key not loaded");
    if (need_sk && h->opt_decode_noise && (rcode = call_key(h, kKeyDomainDecode, &h->noise_key))) return rcode;   // one key per decrypt call
    if (h->batch == 0 || h->batch > h->ctx.batch_size)
        return fail(h, SHELFI_ERR_PARAMS, "batchSize exceeds the slots of the loaded context");
    return 0;
}

// copy the header segments of a planned blob around the coefficient runs already in `dst`
void write_segments(const wire::BlobPlan &plan, uint8_t *dst) {
    for (const auto &seg : plan.segments) std::memcpy(dst + seg.offset, seg.bytes.data(), seg.bytes.size());
}

// Where layer s of a multi-layer call lives: its ciphertexts inside the dense batch, its blob inside the image.
struct LayerSlot {
    uint64_t n = 0, n_ct = 0, ct0 = 0;      // values, ciphertexts, first ciphertext in the concatenated dense batch
    uint64_t x0 = 0;                        // first value in the concatenated plaintext vector
    uint64_t off = 0;                       // byte offset of the blob in the image (8-byte aligned)
    wire::BlobPlan plan;
    shelfi_ckks::EncPlan *ep = nullptr;     // encrypt: the cached plan (then `plan` stays empty)
};
constexpr uint64_t kDeviceHeaderMinCts = 32;   // from here on the bookkeeping bytes are written by a kernel, not the host
const wire::BlobPlan &plan_of(const LayerSlot &l) { return l.ep ? l.ep->plan : l.plan; }
uint64_t align8(uint64_t v) { return (v + 7) & ~(uint64_t)7; }
constexpr uint64_t kPipelineMinCts = 128;   // below this a call is launch latency, not copy time
constexpr uint32_t kCtParts = 8;            // most parts of a shelfi_ct (part_ev); at most kEncStreams
constexpr uint32_t kUploadChunks = 4;       // ct_encrypt: a long page-locked input goes up (and is encrypted) in this many groups
static_assert(kCtParts <= (uint32_t)kEncStreams, "one side stream per part");
// parts of a weighted average / groups of a pipelined decrypt for a vector of n_ct ciphertexts (both sides must agree)
uint32_t part_count(uint64_t n_ct) {
    static const int forced = std::getenv("SHELFI_PARTS") ? std::atoi(std::getenv("SHELFI_PARTS")) : 0;
    if (forced == 4 || forced == 8) return (uint32_t)forced;
    return n_ct >= 2 * kPipelineMinCts ? 8 : 4;
}

// Called where a reference-facing call starts (everything queued by earlier calls has completed by then).
void trim_enc_plans(shelfi_ckks *h) {
    if (h->enc_plans.size() > 64) drop_enc_plans(h);
}
// The serialization plan of a vector of n_ct ciphertexts that all carry metadata m under context ctx (what encrypt
// and, with every learner-0 ciphertext alike, the weighted average write), cached with its segments on the device.
bool same_meta(const wire::CiphertextMeta &a, const wire::CiphertextMeta &b) {
    return a.depth == b.depth && a.level == b.level && std::memcmp(&a.scaling_factor, &b.scaling_factor, 8) == 0 &&
           a.encoding_type == b.encoding_type && a.limbs == b.limbs && std::memcmp(a.tag, b.tag, 32) == 0;
}
int enc_plan(shelfi_ckks *h, const wire::ContextParams &ctx, uint64_t n_ct, const wire::CiphertextMeta &m,
             shelfi_ckks::EncPlan **out) {
    const std::vector<uint8_t> cb = wire::emit_context(ctx);
    std::string key(cb.begin(), cb.end());
    key.append(reinterpret_cast<const char *>(&n_ct), 8).append(reinterpret_cast<const char *>(&m.depth), 8);
    key.append(reinterpret_cast<const char *>(&m.level), 8).append(reinterpret_cast<const char *>(&m.scaling_factor), 8);
    key.append(reinterpret_cast<const char *>(&m.encoding_type), 4).append(reinterpret_cast<const char *>(&m.limbs), 4);
    key.append(m.tag, 32);
    auto it = h->enc_plans.find(key);
    if (it == h->enc_plans.end()) {   // (eviction happens between calls, trim_enc_plans: a call keeps pointers into the map)
        shelfi_ckks::EncPlan ep;
        ep.plan = wire::plan_ciphertext_blob(ctx, n_ct, std::vector<wire::CiphertextMeta>(1, m));
        std::vector<SegDesc> desc;
        std::vector<uint8_t> bytes;
        for (const auto &seg : ep.plan.segments) {
            desc.push_back(SegDesc{seg.offset, (uint32_t)bytes.size(), (uint32_t)seg.bytes.size()});
            bytes.insert(bytes.end(), seg.bytes.begin(), seg.bytes.end());
        }
        ep.n_seg = (uint32_t)desc.size();
        if (n_ct >= kDeviceHeaderMinCts) {   // smaller vectors get their few segments from the host
            int rcode;
            if ((rcode = upload(h, &ep.d_desc, desc.data(), desc.size() * sizeof(SegDesc)))) return rcode;
            if ((rcode = upload(h, &ep.d_bytes, bytes.data(), bytes.size()))) { ep.d_desc.release(); return rcode; }
        }
        it = h->enc_plans.emplace(std::move(key), std::move(ep)).first;
    }
    *out = &it->second;
    return 0;
}

// ckks.cpp:61-104 for n_layers independent vectors xs[s]: copies and encrypt kernels of all layers queued on one
// stream, one D2H.  n_layers == 1 is CKKS::encrypt.  Blob s lands at dst + offsets[s], lens[s] bytes.
int encrypt_host(shelfi_ckks *h, const double *const *xs, const uint64_t *layer_len, uint32_t n_layers, uint8_t *dst,
                 uint64_t cap, uint8_t **alloc_out, uint64_t *offsets, uint64_t *lens, uint64_t *total_len) {
    NvtxRange nvtx_range("shelfi.encrypt");
    int rcode = check_stage(h, true, false);
    if (rcode) return rcode;
    trim_enc_plans(h);
    if (!xs || !layer_len || !lens || n_layers == 0) return fail(h, SHELFI_ERR_ARG, "null argument");
    for (uint32_t s = 0; s < n_layers; s++)
        if (!xs[s]) return fail(h, SHELFI_ERR_ARG, "null argument");

    std::vector<wire::CiphertextMeta> meta(1);
    meta[0].depth = 1; meta[0].level = 0; meta[0].encoding_type = 4;
    meta[0].scaling_factor = (double)h->ctx.modulus[h->ctx.limbs - 1];
    std::memcpy(meta[0].tag, h->key_tag.c_str(), 32);
    std::vector<LayerSlot> slot(n_layers);
    uint64_t n = 0, n_ct = 0, image = 0;
    for (uint32_t s = 0; s < n_layers; s++) {
        if (layer_len[s] == 0) return fail(h, SHELFI_ERR_ARG, "empty input (undefined behaviour in the reference, ckks.cpp:66,95)");
        slot[s].n = layer_len[s]; slot[s].n_ct = (layer_len[s] + h->batch - 1) / h->batch;
        slot[s].x0 = n; slot[s].ct0 = n_ct; slot[s].off = image;
        if ((rcode = enc_plan(h, h->ctx, slot[s].n_ct, meta[0], &slot[s].ep))) return rcode;
        n += slot[s].n; n_ct += slot[s].n_ct; image = align8(image + slot[s].ep->plan.size);
    }
    const uint64_t total = n_layers == 1 ? slot[0].ep->plan.size : image;
    if (alloc_out) {
        dst = static_cast<uint8_t *>(std::malloc(total));
        if (!dst) return fail(h, SHELFI_ERR_ARG, "out of host memory");
    } else if (!dst || cap < total) {
        return fail(h, SHELFI_ERR_ARG, "destination buffer too small: need " + std::to_string(total) + " bytes");
    }
    cudaStream_t s = h->stream;
    const uint64_t ct_words = 2ull * h->ctx.limbs * h->ctx.ring_dim;
    // an error return must not leave copies in flight to or from the caller's buffers (they are only borrowed for the call)
    auto bail = [&](int code) {
        cudaStreamSynchronize(h->stream);
        cudaStreamSynchronize(h->copy_stream);
        if (h->d2h_stream) cudaStreamSynchronize(h->d2h_stream);
        if (alloc_out) std::free(dst);
        return code;
    };
    // bookkeeping bytes of a large vector: dropped into the image by a kernel (disjoint from the runs the pack kernel
    // writes, byte stores on both sides), so the blob arrives complete and the host touches none of it
    auto device_headers = [&](const LayerSlot &l) { return l.n_ct >= kDeviceHeaderMinCts; };
    auto queue_headers = [&](const LayerSlot &l) {
        if (device_headers(l))
            launch_write_segments(l.ep->d_desc.as<SegDesc>(), l.ep->d_bytes.as<uint8_t>(), h->s_image.as<uint8_t>() + l.off,
                                  l.ep->n_seg, s);
    };
#define CUB(call)                                                              \
    do {                                                                       \
        cudaError_t e__ = (call);                                              \
        if (e__ != cudaSuccess) return bail(cuda_fail(h, e__, #call));         \
    } while (0)
    CUB(h->s_x.reserve(n * sizeof(double)));
    CUB(h->s_dense.reserve(n_ct * ct_words * 8));
    CUB(h->s_image.reserve(image_bytes(image)));
    if (n_layers == 1 && n_ct >= kPipelineMinCts && is_pinned(dst)) {
        // One large vector into a page-locked destination: the blob is 8x the input, so the call is the D2H copy plus
        // whatever is not hidden under it.  Ciphertext groups go through copy-in / encrypt / pack on the compute stream
        // while the previous group's bytes are already crossing PCIe on the copy stream.
        const LayerSlot &ls = slot[0];
        const wire::BlobPlan &pl = ls.ep->plan;
        // measured on the config-2 vector (407 ciphertexts, 2.20 ms of D2H): 1 group 2.82 ms per call, 2: 2.72, 4: 2.62,
        // 6: 2.59, 16: 3.53 (partial waves of the encrypt kernel and launch gaps take over)
        uint32_t groups = n_ct >= 3 * kPipelineMinCts ? 6 : 3;
        if (const char *e = std::getenv("SHELFI_ENC_GROUPS")) groups = (uint32_t)std::min<uint64_t>(n_ct, std::max(1, std::atoi(e)));
        cudaStream_t cs = h->copy_stream;
        queue_headers(ls);
        // A pageable input (a numpy array) is moved into a page-locked copy by the copy threads, group by group, each
        // while the GPU works on the group before; nothing waits on anything.  (Group-sized cudaMemcpyAsync calls on
        // pageable memory are synchronous and slow: 4.0 ms per call on the config-2 vector.)
        const double *src = xs[0];
        const bool stage_x = !is_pinned(src) && ls.n * sizeof(double) <= ((size_t)512 << 20);
        if (stage_x) {
            if ((rcode = ensure_staging(h))) return bail(rcode);
            if (h->pin_x_cap < ls.n * sizeof(double)) {
                if (h->pin_x) cudaFreeHost(h->pin_x);
                h->pin_x = nullptr; h->pin_x_cap = 0;
                const size_t want = ls.n * sizeof(double) + (ls.n * sizeof(double)) / 8;
                CUB(cudaHostAlloc(&h->pin_x, want, cudaHostAllocDefault));
                h->pin_x_cap = want;
            }
        }
        auto run_start = [&](uint64_t j) { return j == 0 ? pl.run0[0] : pl.run1[0] + (j - 1) * pl.stride; };
        uint64_t sent = 0;
        for (uint32_t g = 0; g < groups; g++) {
            const uint64_t c0 = n_ct * g / groups, c1 = n_ct * (g + 1) / groups;
            const uint64_t first = c0 * h->batch, count = std::min(ls.n, c1 * h->batch) - first;
            if (stage_x) {
                h->copier->copy(static_cast<double *>(h->pin_x) + first, src + first, count * sizeof(double));
                CUB(cudaMemcpyAsync(h->s_x.as<double>() + first, static_cast<const double *>(h->pin_x) + first,
                                    count * sizeof(double), cudaMemcpyHostToDevice, s));
            } else if ((rcode = h2d(h, h->s_x.as<double>() + first, src + first, count * sizeof(double), s))) {
                return bail(rcode);
            }
            if ((rcode = encrypt_dense(h, h->s_x.as<double>() + first, count, c0, h->s_dense.as<u64>() + c0 * ct_words, s)))
                return bail(rcode);
            uint64_t r0[2 * kMaxLimbs], r1[2 * kMaxLimbs];   // run offsets of ciphertexts c0 and c0 + 1
            for (uint32_t r = 0; r < 2 * h->ctx.limbs; r++) {
                r0[r] = c0 == 0 ? pl.run0[r] : pl.run1[r] + (c0 - 1) * pl.stride;
                r1[r] = pl.run1[r] + c0 * pl.stride;
            }
            timed(h, kKPack, s, [&] {
                launch_pack(h->rc, to_runs(h, r0, r1, pl.stride), h->s_dense.as<u64>() + c0 * ct_words, h->s_image.as<u64>(),
                            c1 - c0, s);
            });
            CUB(cudaGetLastError());
            CUB(cudaEventRecord(h->ev_done[g & 1], s));
            CUB(cudaStreamWaitEvent(cs, h->ev_done[g & 1], 0));
            // cut on a 4 KiB boundary inside the finished part: a copy that starts aligned runs at 55.6 GB/s, one that
            // starts at an odd byte at 52.5 (measured, 107 MB, page-locked destination)
            const uint64_t upto = g + 1 == groups ? total : std::max(sent, run_start(c1) & ~(uint64_t)4095);
            if (upto > sent)
                CUB(cudaMemcpyAsync(dst + sent, h->s_image.as<uint8_t>() + sent, upto - sent, cudaMemcpyDeviceToHost, cs));
            sent = upto;
        }
        CUB(cudaStreamSynchronize(cs));
        if (!device_headers(ls)) write_segments(pl, dst);
        if (offsets) offsets[0] = 0;
        lens[0] = pl.size;
        if (total_len) *total_len = total;
        return 0;
    }
    for (uint32_t i = 0; i < n_layers; i++)
        if ((rcode = h2d(h, h->s_x.as<double>() + slot[i].x0, xs[i], slot[i].n * sizeof(double), s))) return bail(rcode);
    for (const LayerSlot &ls : slot) {   // every layer is its own ciphertext vector: sampler index restarts at 0
        if ((rcode = encrypt_dense(h, h->s_x.as<double>() + ls.x0, ls.n, 0, h->s_dense.as<u64>() + ls.ct0 * ct_words, s)))
            return bail(rcode);
        const wire::BlobPlan &pl = ls.ep->plan;
        timed(h, kKPack, s, [&] {
            launch_pack(h->rc, to_runs(h, pl.run0, pl.run1, pl.stride), h->s_dense.as<u64>() + ls.ct0 * ct_words,
                        h->s_image.as<u64>() + ls.off / 8, ls.n_ct, s);
        });
        queue_headers(ls);
    }
    CUB(cudaGetLastError());
    if ((rcode = d2h_sync(h, dst, h->s_image.p, total, s))) return bail(rcode);
    for (uint32_t i = 0; i < n_layers; i++) {
        if (!device_headers(slot[i])) write_segments(slot[i].ep->plan, dst + slot[i].off);
        if (offsets) offsets[i] = slot[i].off;
        lens[i] = slot[i].ep->plan.size;
    }
    if (alloc_out) *alloc_out = dst;
    if (total_len) *total_len = total;
    return 0;
}

// ckks.cpp:264-320 for n_layers ciphertext vectors per learner (blobs[i * n_layers + s] = learner i, layer s).
// n_layers == 1 is CKKS::computeWeightedAverage.  Learner i+1 streams in over PCIe while learner i is folded into the
// resident accumulator; one D2H returns all result blobs (blob s at dst + offsets[s]).
// rescale: the average is brought back to depth 1 / one tower fewer before it is serialized (opt-in,
// the reference leaves it at depth 2), which halves the bytes the server sends at two towers.
int wavg_host(shelfi_ckks *h, const uint8_t *const *blobs, const uint64_t *blob_lens, const float *w, uint32_t k,
              uint32_t n_layers, uint8_t *dst, uint64_t cap, uint8_t **alloc_out, uint64_t *offsets, uint64_t *lens,
              uint64_t *total_len, bool rescale = false) {
    NvtxRange nvtx_range("shelfi.computeWeightedAverage");
    int rcode = check_stage(h, false, false);
    if (rcode) return rcode;
    trim_enc_plans(h);
    if (!blobs || !blob_lens || !w || !lens || n_layers == 0) return fail(h, SHELFI_ERR_ARG, "null argument");
    if (k == 0) return fail(h, SHELFI_ERR_ARG, "no learners");
    const uint32_t L = h->ctx.limbs;
    const double sf0 = (double)h->ctx.modulus[L - 1];
    if (rescale && L < 2) return fail(h, SHELFI_ERR_UNSUPPORTED, "rescale needs a chain of at least two towers");
    const uint32_t Lo = rescale ? L - 1 : L;   // towers of the result
    std::vector<wire::BlobMap> maps((size_t)k * n_layers);
    std::vector<uint64_t> in_off((size_t)k * n_layers);   // where each input blob sits in its learner's image
    size_t max_image = 0;
    for (uint32_t i = 0; i < k; i++) {  // ckks.cpp:278-281 Deserialize, minus the payload copy
        uint64_t img = 0;
        for (uint32_t s = 0; s < n_layers; s++) {
            const size_t id = (size_t)i * n_layers + s;
            const std::string who = "learner " + std::to_string(i) + (n_layers > 1 ? " layer " + std::to_string(s) : "");
            if (!blobs[id]) return fail(h, SHELFI_ERR_ARG, "null blob");
            std::string e = wire::map_ciphertext_blob(blobs[id], blob_lens[id], false, &maps[id]);
            if (!e.empty()) return fail(h, SHELFI_ERR_FORMAT, who + ": " + e);
            if (!same_ring(maps[id].ctx, h->ctx))
                return fail(h, SHELFI_ERR_PARAMS, who + ": ciphertext parameters differ from the loaded context");
            if (maps[id].n_ct != maps[s].n_ct) return fail(h, SHELFI_ERR_PARAMS, who + ": ciphertext count differs");
            if (maps[id].ct_limbs != L)
                return fail(h, SHELFI_ERR_UNSUPPORTED, who + ": only fresh depth-1 level-0 ciphertexts are supported");
            for (const auto &m : maps[id].meta)
                if (m.depth != 1 || m.level != 0)
                    return fail(h, SHELFI_ERR_UNSUPPORTED, who + ": only fresh depth-1 level-0 ciphertexts are supported");
            in_off[id] = img;
            img = align8(img + blob_lens[id]);
        }
        max_image = std::max(max_image, image_bytes(img));
    }
    std::vector<LayerSlot> slot(n_layers);
    uint64_t n_ct = 0, image = 0;
    for (uint32_t s = 0; s < n_layers; s++) {
        std::vector<wire::CiphertextMeta> meta = maps[s].meta;  // ckks.cpp:291-292: result starts as learner 0's vector
        for (auto &m : meta) { m.depth += 1; m.scaling_factor *= sf0; }  // [PAL] EvalMult(ct, double), no rescale
        if (rescale)   // [PAL] ModReduce: depth - 1, level + 1, scalingFactor / q_l (double), last tower dropped
            for (auto &m : meta) { m.depth -= 1; m.level += 1; m.scaling_factor /= sf0; m.limbs = Lo; }
        slot[s].n_ct = maps[s].n_ct; slot[s].ct0 = n_ct; slot[s].off = image;
        bool uniform = slot[s].n_ct >= kDeviceHeaderMinCts;   // large vector of like ciphertexts: cached plan, headers by kernel
        for (size_t j = 1; uniform && j < meta.size(); j++) uniform = same_meta(meta[j], meta[0]);
        if (uniform) {
            if ((rcode = enc_plan(h, maps[s].ctx, slot[s].n_ct, meta[0], &slot[s].ep))) return rcode;
        } else {
            slot[s].plan = wire::plan_ciphertext_blob(maps[s].ctx, slot[s].n_ct, meta);
        }
        n_ct += slot[s].n_ct; image = align8(image + plan_of(slot[s]).size);
    }
    const uint64_t total = n_layers == 1 ? plan_of(slot[0]).size : image;
    if (alloc_out) {
        dst = static_cast<uint8_t *>(std::malloc(total));
        if (!dst) return fail(h, SHELFI_ERR_ARG, "out of host memory");
    } else if (!dst || cap < total) {
        return fail(h, SHELFI_ERR_ARG, "destination buffer too small: need " + std::to_string(total) + " bytes");
    }
    // an error return must not leave copies in flight to or from the caller's buffers (they are only borrowed for the call)
    auto bail = [&](int code) {
        cudaStreamSynchronize(h->stream);
        cudaStreamSynchronize(h->copy_stream);
        if (h->d2h_stream) cudaStreamSynchronize(h->d2h_stream);
        if (alloc_out) std::free(dst);
        return code;
    };
    cudaStream_t st = h->stream, cs = h->copy_stream;
    const uint64_t ct_words = 2ull * L * h->ctx.ring_dim, out_words = 2ull * Lo * h->ctx.ring_dim;
    CUB(h->s_dense.reserve(n_ct * ct_words * 8));
    if (rescale) CUB(h->s_dense2.reserve(n_ct * out_words * 8));
    CUB(h->s_image.reserve(std::max(max_image, image_bytes(image))));
    CUB(h->s_image2.reserve(max_image));
    DeviceBuffer *img[2] = {&h->s_image, &h->s_image2};
    // One large vector into a page-locked destination: the ciphertext range is cut in two so that the first part's
    // result crosses PCIe device->host while the second part's inputs are still arriving host->device.
    const bool split = n_layers == 1 && n_ct >= 2 * kPipelineMinCts && is_pinned(dst);
    if (split) {
        if (!h->d2h_stream) {
            CUB(cudaStreamCreateWithFlags(&h->d2h_stream, cudaStreamNonBlocking));
            CUB(cudaEventCreateWithFlags(&h->ev_stage_out, cudaEventDisableTiming));
        }
        CUB(h->s_image3.reserve(image_bytes(image)));
    }
    const uint32_t groups = split ? 2 : 1;
    u64 *out_image = split ? h->s_image3.as<u64>() : h->s_image.as<u64>();
    auto run_start = [](const uint64_t *run0, const uint64_t *run1, uint64_t stride, uint64_t j) {
        return j == 0 ? run0[0] : run1[0] + (j - 1) * stride;
    };
    auto sub_runs = [&](const uint64_t *run0, const uint64_t *run1, uint64_t stride, uint64_t c0, uint32_t limbs) {
        uint64_t r0[2 * kMaxLimbs], r1[2 * kMaxLimbs];   // run offsets of ciphertexts c0 and c0 + 1
        for (uint32_t r = 0; r < 2 * limbs; r++) {
            r0[r] = c0 == 0 ? run0[r] : run1[r] + (c0 - 1) * stride;
            r1[r] = run1[r] + c0 * stride;
        }
        return to_runs(h, r0, r1, stride, limbs);
    };
    uint64_t sent = 0;
    for (uint32_t g = 0, seq = 0; g < groups; g++) {
        // ciphertexts [c0, c1) of every layer's vector (split == false: everything; split: 3/4 then 1/4, so the second
        // part's inputs take about as long as the first part's result)
        const uint64_t cut = n_ct - n_ct / 4;
        const uint64_t c0 = split && g == 1 ? cut : 0, c1 = split && g == 0 ? cut : 0;   // c1 == 0: to the end
        for (uint32_t i = 0; i < k; i++, seq++) {
            int b = seq & 1;
            if (seq >= 2) CUB(cudaStreamWaitEvent(cs, h->ev_done[b], 0));
            for (uint32_t s = 0; s < n_layers; s++) {
                const size_t id = (size_t)i * n_layers + s;
                const wire::BlobMap &m = maps[id];
                // the second part starts on a 4 KiB boundary a little before its first run (aligned copies are 6 % faster)
                const uint64_t lo = c0 ? run_start(m.run0, m.run1, m.stride, c0) & ~(uint64_t)4095 : 0;
                const uint64_t hi = c1 ? run_start(m.run0, m.run1, m.stride, c1) : blob_lens[id];
                if ((rcode = h2d(h, img[b]->as<uint8_t>() + in_off[id] + lo, blobs[id] + lo, hi - lo, cs))) return bail(rcode);
            }
            CUB(cudaEventRecord(h->ev_copy[b], cs));
            CUB(cudaStreamWaitEvent(st, h->ev_copy[b], 0));
            int64_t sc = weight_constant(h, w[i]);
            LimbScalars cs_l{};
            for (uint32_t l = 0; l < L; l++) cs_l.c[l] = nt::signed_mod(sc, h->ctx.modulus[l]);
            for (uint32_t s = 0; s < n_layers; s++) {
                const size_t id = (size_t)i * n_layers + s;
                const uint64_t cnt = (c1 ? c1 : slot[s].n_ct) - c0;
                timed(h, kKUnpackScaleAcc, st, [&] {
                    launch_unpack_scale_accumulate(h->rc, sub_runs(maps[id].run0, maps[id].run1, maps[id].stride, c0, L),
                                                   img[b]->as<u64>() + in_off[id] / 8, cs_l, i == 0 ? 1u : 0u,
                                                   h->s_dense.as<u64>() + (slot[s].ct0 + c0) * ct_words, cnt, st);
                });
            }
            CUB(cudaGetLastError());
            CUB(cudaEventRecord(h->ev_done[b], st));
        }
        for (const LayerSlot &ls : slot) {
            const uint64_t cnt = (c1 ? c1 : ls.n_ct) - c0;
            const u64 *result = h->s_dense.as<u64>() + (ls.ct0 + c0) * ct_words;
            if (rescale) {
                u64 *reduced = h->s_dense2.as<u64>() + (ls.ct0 + c0) * out_words;
                if ((rcode = rescale_dense(h, result, cnt, L, reduced, st))) return bail(rcode);
                result = reduced;
            }
            const wire::BlobPlan &pl = plan_of(ls);
            timed(h, kKPack, st, [&] {
                launch_pack(h->rc, sub_runs(pl.run0, pl.run1, pl.stride, c0, Lo), result, out_image + ls.off / 8, cnt, st);
            });
            if (ls.ep && g == 0)   // bookkeeping bytes of the whole vector, once (the input images have been consumed)
                launch_write_segments(ls.ep->d_desc.as<SegDesc>(), ls.ep->d_bytes.as<uint8_t>(),
                                      reinterpret_cast<uint8_t *>(out_image) + ls.off, ls.ep->n_seg, st);
        }
        CUB(cudaGetLastError());
        if (split) {   // this part's bytes start their way back while the next part's inputs are still coming in
            const wire::BlobPlan &pl = plan_of(slot[0]);
            const uint64_t upto = c1 ? run_start(pl.run0, pl.run1, pl.stride, c1) & ~(uint64_t)4095 : total;
            CUB(cudaEventRecord(h->ev_stage_out, st));
            CUB(cudaStreamWaitEvent(h->d2h_stream, h->ev_stage_out, 0));
            CUB(cudaMemcpyAsync(dst + sent, reinterpret_cast<const uint8_t *>(out_image) + sent, upto - sent,
                                cudaMemcpyDeviceToHost, h->d2h_stream));
            sent = upto;
        }
    }
    if (split) {
        CUB(cudaStreamSynchronize(h->d2h_stream));
        CUB(cudaStreamSynchronize(st));
    } else if ((rcode = d2h_sync(h, dst, h->s_image.p, total, st))) {
        return bail(rcode);
    }
    for (uint32_t s = 0; s < n_layers; s++) {
        if (!slot[s].ep) write_segments(slot[s].plan, dst + slot[s].off);
        if (offsets) offsets[s] = slot[s].off;
        lens[s] = plan_of(slot[s]).size;
    }
    if (alloc_out) *alloc_out = dst;
    if (total_len) *total_len = total;
    return 0;
}

// ckks.cpp:170-213 for n_layers ciphertext vectors: out = concatenation of the layers' data_dimensions[s] values.
// n_layers == 1 is CKKS::decrypt.
int decrypt_host(shelfi_ckks *h, const uint8_t *const *blobs, const uint64_t *blob_lens, const uint64_t *dims,
                 uint32_t n_layers, double *out) {
    NvtxRange nvtx_range("shelfi.decrypt");
    int rcode = check_stage(h, false, true);
    if (rcode) return rcode;
    if (!blobs || !blob_lens || !dims || !out || n_layers == 0) return fail(h, SHELFI_ERR_ARG, "null argument");
    const uint64_t batch = h->batch;
    std::vector<wire::BlobMap> maps(n_layers);
    std::vector<LayerSlot> slot(n_layers);
    uint64_t n = 0, n_ct = 0, image = 0, max_ct = 0;
    for (uint32_t s = 0; s < n_layers; s++) {
        if (!blobs[s]) return fail(h, SHELFI_ERR_ARG, "null blob");
        std::string e = wire::map_ciphertext_blob(blobs[s], blob_lens[s], false, &maps[s]);
        if (!e.empty()) return fail(h, SHELFI_ERR_FORMAT, e);
        if (!same_ring(maps[s].ctx, h->ctx)) return fail(h, SHELFI_ERR_PARAMS, "ciphertext parameters differ from the loaded context");
        const uint64_t c = maps[s].n_ct;
        if (maps[s].ct_limbs < 1 || maps[s].ct_limbs > h->ctx.limbs) return fail(h, SHELFI_ERR_FORMAT, "unexpected tower count");
        if (dims[s] == 0 || dims[s] > c * batch || dims[s] + batch <= c * batch)
            return fail(h, SHELFI_ERR_ARG, "data_dimensions inconsistent with the number of ciphertexts");
        slot[s].n = dims[s]; slot[s].n_ct = c; slot[s].x0 = n; slot[s].ct0 = n_ct; slot[s].off = image;
        n += dims[s]; n_ct += c; image = align8(image + blob_lens[s]);
        max_ct = std::max(max_ct, c);
    }
    const uint32_t L = h->ctx.limbs;
    const uint64_t N = h->ctx.ring_dim, ct_words = 2ull * L * N;
    cudaStream_t st = h->stream;
    // an error return must not leave copies in flight to or from the caller's buffers (they are only borrowed for the call)
    auto bail = [&](int code) {
        cudaStreamSynchronize(h->stream);
        cudaStreamSynchronize(h->copy_stream);
        if (h->d2h_stream) cudaStreamSynchronize(h->d2h_stream);
        return code;
    };
    CUB(h->s_image.reserve(image_bytes(image)));
    CUB(h->s_dense.reserve(n_ct * ct_words * 8));
    CUB(h->s_out.reserve(n * sizeof(double)));
    // decrypt_dense's scratch, sized once so that no launch of this call waits on a reallocation
    CUB(h->s_coeff.reserve(max_ct * L * N * sizeof(uint64_t)));
    if (general_decode(h)) {
        CUB(h->s_vals.reserve(max_ct * N * sizeof(double)));
        if (h->ec.slots > kFftFast) CUB(h->s_zs.reserve(max_ct * (size_t)h->ec.slots * sizeof(cplx)));
    }
    bool one_scale = true;
    for (const auto &m : maps[0].meta) one_scale = one_scale && m.scaling_factor == maps[0].meta[0].scaling_factor;
    if (n_layers == 1 && n_ct >= kPipelineMinCts && one_scale && is_pinned(blobs[0]) && is_pinned(out)) {
        // One large vector between page-locked buffers: the call is the H2D copy of the blob plus whatever is not hidden
        // under it.  Ciphertext groups arrive on the copy stream; each is unpacked and decrypted while the next one is
        // still crossing PCIe, and its values leave on a third stream (the link is full duplex).
        const wire::BlobMap &m = maps[0];
        const uint32_t groups = 4;
        cudaStream_t cs = h->copy_stream;
        if (!h->d2h_stream) {
            CUB(cudaStreamCreateWithFlags(&h->d2h_stream, cudaStreamNonBlocking));
            CUB(cudaEventCreateWithFlags(&h->ev_stage_out, cudaEventDisableTiming));
        }
        CUB(cudaEventRecord(h->ev_done[0], st));          // the image and scratch are free once earlier work on st is done
        CUB(cudaStreamWaitEvent(cs, h->ev_done[0], 0));
        const uint64_t layer_words = 2ull * m.ct_limbs * N;
        auto run_start = [&](uint64_t j) { return j == 0 ? m.run0[0] : m.run1[0] + (j - 1) * m.stride; };
        uint64_t got = 0;
        for (uint32_t g = 0; g < groups; g++) {
            const uint64_t c0 = n_ct * g / groups, c1 = n_ct * (g + 1) / groups;
            // bytes up to the first run of ciphertext c1, cut on a 4 KiB boundary at or after it (aligned copies are faster)
            const uint64_t upto = g + 1 == groups ? blob_lens[0] : std::min<uint64_t>(blob_lens[0], (run_start(c1) + 4095) & ~(uint64_t)4095);
            if (upto > got) CUB(cudaMemcpyAsync(h->s_image.as<uint8_t>() + got, blobs[0] + got, upto - got, cudaMemcpyHostToDevice, cs));
            got = std::max(got, upto);
            CUB(cudaEventRecord(h->ev_copy[g & 1], cs));
            CUB(cudaStreamWaitEvent(st, h->ev_copy[g & 1], 0));
            uint64_t r0[2 * kMaxLimbs], r1[2 * kMaxLimbs];   // run offsets of ciphertexts c0 and c0 + 1
            for (uint32_t r = 0; r < 2 * m.ct_limbs; r++) {
                r0[r] = c0 == 0 ? m.run0[r] : m.run1[r] + (c0 - 1) * m.stride;
                r1[r] = m.run1[r] + c0 * m.stride;
            }
            u64 *dense = h->s_dense.as<u64>() + c0 * layer_words;
            timed(h, kKUnpack, st, [&] {
                launch_unpack(h->rc, to_runs(h, r0, r1, m.stride, m.ct_limbs), h->s_image.as<u64>(), dense, c1 - c0, st);
            });
            CUB(cudaGetLastError());
            const uint64_t first = c0 * batch, last = std::min(n, c1 * batch);
            if ((rcode = decrypt_dense(h, dense, c1 - c0, m.meta[0].scaling_factor, h->s_out.as<double>() + first, last - first,
                                       st, nullptr, m.ct_limbs, nullptr, 0, c0))) {
                return bail(rcode);
            }
            CUB(cudaEventRecord(h->ev_stage_out, st));
            CUB(cudaStreamWaitEvent(h->d2h_stream, h->ev_stage_out, 0));
            CUB(cudaMemcpyAsync(out + first, h->s_out.as<double>() + first, (last - first) * sizeof(double), cudaMemcpyDeviceToHost,
                               h->d2h_stream));
        }
        CUB(cudaStreamSynchronize(h->d2h_stream));
        CUB(cudaStreamSynchronize(st));
        return 0;
    }
    for (uint32_t s = 0; s < n_layers; s++) {
        const LayerSlot &ls = slot[s];
        if ((rcode = h2d(h, h->s_image.as<uint8_t>() + ls.off, blobs[s], blob_lens[s], st))) return bail(rcode);
        u64 *dense = h->s_dense.as<u64>() + ls.ct0 * ct_words;
        timed(h, kKUnpack, st, [&] {
            launch_unpack(h->rc, to_runs(h, maps[s].run0, maps[s].run1, maps[s].stride, maps[s].ct_limbs),
                          h->s_image.as<u64>() + ls.off / 8, dense, ls.n_ct, st);
        });
        CUB(cudaGetLastError());
        // ciphertexts usually share one scaling factor; split the launch where they do not
        for (uint64_t j0 = 0; j0 < ls.n_ct;) {
            uint64_t j1 = j0 + 1;
            while (j1 < ls.n_ct && maps[s].meta[j1].scaling_factor == maps[s].meta[j0].scaling_factor) j1++;
            uint64_t first = j0 * batch, last = std::min(ls.n, j1 * batch);
            const uint64_t layer_words = 2ull * maps[s].ct_limbs * N;   // a rescaled layer is packed tighter
            if ((rcode = decrypt_dense(h, dense + j0 * layer_words, j1 - j0, maps[s].meta[j0].scaling_factor,
                                       h->s_out.as<double>() + ls.x0 + first, last - first, st, nullptr, maps[s].ct_limbs,
                                       nullptr, 0, ls.ct0 + j0)))
                return bail(rcode);
            j0 = j1;
        }
    }
    if ((rcode = d2h_sync(h, out, h->s_out.p, n * sizeof(double), st))) return bail(rcode);
    return 0;
}

// [PAL] cc->ModReduce on every ciphertext of a serialized vector (rescale-before-transport): the blob
// comes back one tower shorter, depth - 1, level + 1, scalingFactor / q_l.
int rescale_host(shelfi_ckks *h, const uint8_t *blob, uint64_t blob_len, uint8_t *dst, uint64_t cap, uint8_t **alloc_out,
                 uint64_t *out_len) {
    int rcode = check_stage(h, false, false);
    if (rcode) return rcode;
    if (!blob || !out_len) return fail(h, SHELFI_ERR_ARG, "null argument");
    wire::BlobMap map;
    std::string e = wire::map_ciphertext_blob(blob, blob_len, false, &map);
    if (!e.empty()) return fail(h, SHELFI_ERR_FORMAT, e);
    if (!same_ring(map.ctx, h->ctx)) return fail(h, SHELFI_ERR_PARAMS, "ciphertext parameters differ from the loaded context");
    const uint32_t Lin = map.ct_limbs;
    if (Lin < 2) return fail(h, SHELFI_ERR_UNSUPPORTED, "no tower left to drop");
    std::vector<wire::CiphertextMeta> meta = map.meta;
    const double ql = (double)h->ctx.modulus[Lin - 1];
    for (auto &m : meta) {
        if (m.depth < 2) return fail(h, SHELFI_ERR_UNSUPPORTED, "ciphertext is already at depth 1");
        m.depth -= 1; m.level += 1; m.scaling_factor /= ql; m.limbs = Lin - 1;
    }
    wire::BlobPlan plan = wire::plan_ciphertext_blob(map.ctx, map.n_ct, meta);
    if (alloc_out) {
        dst = static_cast<uint8_t *>(std::malloc(plan.size));
        if (!dst) return fail(h, SHELFI_ERR_ARG, "out of host memory");
    } else if (!dst || cap < plan.size) {
        return fail(h, SHELFI_ERR_ARG, "destination buffer too small: need " + std::to_string(plan.size) + " bytes");
    }
    // an error return must not leave copies in flight to or from the caller's buffers (they are only borrowed for the call)
    auto bail = [&](int code) {
        cudaStreamSynchronize(h->stream);
        cudaStreamSynchronize(h->copy_stream);
        if (h->d2h_stream) cudaStreamSynchronize(h->d2h_stream);
        if (alloc_out) std::free(dst);
        return code;
    };
    const uint64_t N = h->ctx.ring_dim;
    cudaStream_t st = h->stream;
    CUB(h->s_image.reserve(image_bytes(blob_len)));
    CUB(h->s_image2.reserve(image_bytes(plan.size)));
    CUB(h->s_dense.reserve(map.n_ct * 2 * Lin * N * 8));
    CUB(h->s_dense2.reserve(map.n_ct * 2 * (Lin - 1) * N * 8));
    if ((rcode = h2d(h, h->s_image.p, blob, blob_len, st))) return bail(rcode);
    timed(h, kKUnpack, st, [&] {
        launch_unpack(h->rc, to_runs(h, map.run0, map.run1, map.stride, Lin), h->s_image.as<u64>(), h->s_dense.as<u64>(),
                      map.n_ct, st);
    });
    if ((rcode = rescale_dense(h, h->s_dense.as<u64>(), map.n_ct, Lin, h->s_dense2.as<u64>(), st))) return bail(rcode);
    timed(h, kKPack, st, [&] {
        launch_pack(h->rc, to_runs(h, plan.run0, plan.run1, plan.stride, Lin - 1), h->s_dense2.as<u64>(),
                    h->s_image2.as<u64>(), map.n_ct, st);
    });
    CUB(cudaGetLastError());
    if ((rcode = d2h_sync(h, dst, h->s_image2.p, plan.size, st))) return bail(rcode);
    write_segments(plan, dst);
    if (alloc_out) *alloc_out = dst;
    *out_len = plan.size;
    return 0;
}


// ---------------------------------------------------------------------------------------------
// Resident ciphertext vectors (shelfi_ct_*): the three reference calls with the ciphertexts left in HBM.
// The reference's callers keep every blob inside one process and only hand it to the next call
// (code/benchmark.py:459-460,490-493,507-513,527-528), so nothing forces 64 bytes per parameter and learner across
// PCIe twice; what must cross it is the plaintext in (8 or 4 bytes per parameter) and the average out (8 bytes).
// ---------------------------------------------------------------------------------------------
int ensure_resident(shelfi_ckks *h) {
    if (h->ev_r_out) return 0;
    for (int i = 0; i < kEncStreams; i++) {
        CU(cudaStreamCreateWithFlags(&h->enc_streams[i], cudaStreamNonBlocking));
        CU(cudaEventCreateWithFlags(&h->ev_x_in[i], cudaEventDisableTiming));
        CU(cudaEventCreateWithFlags(&h->ev_x_free[i], cudaEventDisableTiming));
    }
    if (!h->d2h_stream) {
        CU(cudaStreamCreateWithFlags(&h->d2h_stream, cudaStreamNonBlocking));
        CU(cudaEventCreateWithFlags(&h->ev_stage_out, cudaEventDisableTiming));
    }
    if (!h->pool.limit) {
        size_t free_b = 0, total_b = 0;
        CU(cudaMemGetInfo(&free_b, &total_b));
        h->pool.limit = free_b - free_b / 5;   // 80 % of what is free now; shelfi_resident_set_limit changes it
        if (const char *e = std::getenv("SHELFI_RESIDENT_LIMIT_MB")) h->pool.limit = (size_t)std::max(1, std::atoi(e)) << 20;
    }
    CU(cudaEventCreateWithFlags(&h->ev_fork, cudaEventDisableTiming));
    for (int i = 0; i < kEncStreams; i++) CU(cudaEventCreateWithFlags(&h->ev_part_done[i], cudaEventDisableTiming));
    CU(cudaEventCreateWithFlags(&h->ev_r_out, cudaEventDisableTiming));
    return 0;
}
cudaEvent_t take_event(shelfi_ckks *h) {
    if (!h->spare_events.empty()) { cudaEvent_t e = h->spare_events.back(); h->spare_events.pop_back(); return e; }
    cudaEvent_t e = nullptr;
    if (cudaEventCreateWithFlags(&e, cudaEventDisableTiming) != cudaSuccess) { cudaGetLastError(); return nullptr; }
    return e;
}
// `on`: the stream that will write the new vector first; a recycled buffer's last uses were ordered on h->stream
int new_ct(shelfi_ckks *h, uint64_t n_ct, uint32_t limbs, shelfi_ct **out, cudaStream_t on = nullptr) {
    const size_t bytes = (size_t)n_ct * 2 * limbs * h->ctx.ring_dim * sizeof(uint64_t);
    shelfi_ct *c = new shelfi_ct();
    PoolBuf b = h->pool.take(bytes, &c->cap);
    if (!b.p) {
        delete c;
        return fail(h, SHELFI_ERR_MEMORY, "device pool exhausted: " + std::to_string(bytes >> 20) + " MiB requested, " +
                                              std::to_string(h->pool.in_use >> 20) + " MiB in resident ciphertexts, limit " +
                                              std::to_string(h->pool.limit >> 20) + " MiB (free vectors you no longer need, or raise SHELFI_RESIDENT_LIMIT_MB)");
    }
    c->d = static_cast<u64 *>(b.p);
    c->buf_ev = b.ev;
    if (b.ev && on && on != h->stream && cudaStreamWaitEvent(on, b.ev, 0) != cudaSuccess) {
        h->pool.give(b, c->cap);
        delete c;
        return cuda_fail(h, cudaGetLastError(), "cudaStreamWaitEvent");
    }
    c->owner = h; c->n_ct = n_ct; c->limbs = limbs; c->ctx = h->ctx;
    h->live.push_back(c);
    *out = c;
    return 0;
}
// before the compute stream touches a vector that an encrypt stream may still be writing
void ct_ready(shelfi_ckks *h, shelfi_ct *c) {
    if (c->ready_pending) {
        cudaStreamWaitEvent(h->stream, c->ready, 0);
        c->ready_pending = false;
    }
}
void drop_ct(shelfi_ckks *h, shelfi_ct *c) {
    auto it = std::find(h->live.begin(), h->live.end(), c);
    if (it == h->live.end()) return;
    h->live.erase(it);
    if (c->d) {
        ct_ready(h, c);                                  // an unused encryption may still be running
        if (!c->buf_ev) c->buf_ev = take_event(h);
        if (c->buf_ev) cudaEventRecord(c->buf_ev, h->stream);   // after every use queued so far
        else cudaStreamSynchronize(h->stream);
        h->pool.give(PoolBuf{c->d, c->buf_ev}, c->cap);
    }
    if (c->ready) h->spare_events.push_back(c->ready);
    for (cudaEvent_t e : c->part_ev)
        if (e) h->spare_events.push_back(e);
    for (cudaEvent_t e : c->chunk_ev)
        if (e) h->spare_events.push_back(e);
    delete c;
}
bool is_device_ptr(const void *p) {
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return false; }
    return a.type == cudaMemoryTypeDevice;
}
// the serialization plan of a resident vector (cached for vectors of like ciphertexts)
int ct_plan(shelfi_ckks *h, const shelfi_ct *c, shelfi_ckks::EncPlan **ep, wire::BlobPlan *own) {
    *ep = nullptr;
    if (c->meta.size() == 1) return enc_plan(h, c->ctx, c->n_ct, c->meta[0], ep);
    *own = wire::plan_ciphertext_blob(c->ctx, c->n_ct, c->meta);
    return 0;
}

// ckks.cpp:61-104 with the result left on the device.  x: n values, float64 (dtype 0) or float32 (dtype 1), in host
// memory (page-locked: copied by the DMA engine straight from the caller's pages; pageable: staged through the copy
// threads) or in device memory.  Returns once x has been read; the encryption itself continues on the compute stream.
int ct_encrypt(shelfi_ckks *h, const void *x, uint64_t n, int dtype, shelfi_ct **out) {
    NvtxRange nvtx_range("shelfi.encrypt[resident]");
    int rcode = check_stage(h, true, false);
    if (rcode) return rcode;
    if (!x || !out || n == 0) return fail(h, SHELFI_ERR_ARG, "null or empty argument (an empty input is undefined behaviour in the reference, ckks.cpp:66,95)");
    if (dtype != 0 && dtype != 1) return fail(h, SHELFI_ERR_ARG, "dtype: 0 = float64, 1 = float32");
    if ((rcode = ensure_resident(h))) return rcode;
    trim_enc_plans(h);
    const uint64_t n_ct = (n + h->batch - 1) / h->batch;
    const size_t in_bytes = n * (dtype ? 4 : 8);
    const bool on_device = is_device_ptr(x);
    // a long vector from page-locked memory goes up and is encrypted in kUploadChunks groups (see below)
    const bool borrowed = !on_device && (in_bytes < (1u << 20) || is_pinned(x));   // the DMA engine reads the caller's pages (h2d)
    // (pageable memory goes through the copy threads, which only pay off on whole inputs: it is staged whole into one
    // piece of the ring when it fits and goes up from there in groups; larger pageable inputs: one upload, one launch)
    const bool staged = !on_device && !borrowed && in_bytes <= kPieceBytes;
    const bool chunked = (borrowed || staged) && h->rc.log_blocks == 0 && n_ct >= 2 * kPipelineMinCts;
    const unsigned turn = h->r_turn++;
    const unsigned b = turn % kEncStreams;                                       // input slot
    const unsigned s0 = chunked ? (turn * kUploadChunks) % kEncStreams : b;      // a chunked call takes kUploadChunks streams
    cudaStream_t st = h->enc_streams[s0], cs = h->copy_stream;
    shelfi_ct *c = nullptr;
    if ((rcode = new_ct(h, n_ct, h->ctx.limbs, &c, st))) return rcode;
    wire::CiphertextMeta m;
    m.depth = 1; m.level = 0; m.encoding_type = 4;
    m.scaling_factor = (double)h->ctx.modulus[h->ctx.limbs - 1];
    std::memcpy(m.tag, h->key_tag.c_str(), 32);
    c->meta.assign(1, m);
    auto bail = [&](int code) {
        cudaStreamSynchronize(cs);
        for (cudaStream_t e : h->enc_streams) cudaStreamSynchronize(e);   // a chunked call spreads over several
        drop_ct(h, c);
        return code;
    };
    if (!(c->ready = take_event(h))) return bail(fail(h, SHELFI_ERR_CUDA, "cudaEventCreate failed"));
    const void *d_x = x;
    if (!on_device) {
        // slot b was last read by the encryption kEncStreams calls ago
        if (h->x_busy[b]) CUB(cudaStreamWaitEvent(cs, h->ev_x_free[b], 0));
        if (h->r_x[b].cap < in_bytes) {   // growing reallocates: wait until nothing reads the old buffer
            CUB(cudaStreamSynchronize(st));
            CUB(h->r_x[b].reserve(in_bytes));
        }
        if (!chunked) {
            timed(h, kKCopyIn, cs, [&] { rcode = h2d(h, h->r_x[b].p, x, in_bytes, cs); });
            if (rcode) return bail(rcode);
            CUB(cudaEventRecord(h->ev_x_in[b], cs));
            CUB(cudaStreamWaitEvent(st, h->ev_x_in[b], 0));
        }
        d_x = h->r_x[b].p;
    }
    const double *dx64 = static_cast<const double *>(d_x);
    u64 *ct_ptr = c->d;
    if (chunked) {
        // A long vector from page-locked memory: uploaded and encrypted in kUploadChunks groups of ciphertexts, group g
        // on encrypt stream s0 + g, so that the first kernel starts after a quarter of the upload (it matters for the
        // first vector of a round, which nothing else hides).  The draws are addressed by ciphertext index, so the
        // ciphertexts are the ones of a single launch.  Stream s0 joins the others before `ready` is recorded.
        const size_t esize = dtype ? 4 : 8;
        const uint64_t words = 2ull * h->ctx.limbs * h->ctx.ring_dim;
        const char *stage = nullptr;
        unsigned piece = 0;
        if (staged) {   // the caller's pageable memory, whole, into a page-locked piece (the copy threads at work)
            if ((rcode = ensure_staging(h))) return bail(rcode);
            piece = h->piece_turn++ % kStagePieces;
            char *p = static_cast<char *>(h->stage[piece / (kStagePieces / 2)]) + (size_t)(piece % (kStagePieces / 2)) * kPieceBytes;
            if (h->piece_busy[piece]) CUB(cudaEventSynchronize(h->ev_piece[piece]));
            h->copier->copy(p, x, in_bytes);
            stage = p;
        }
        for (uint32_t g = 0; g < kUploadChunks; g++) {
            const uint64_t c0 = n_ct * g / kUploadChunks, c1 = n_ct * (g + 1) / kUploadChunks;
            const uint64_t first = c0 * h->batch, last = std::min<uint64_t>(n, c1 * h->batch);
            cudaStream_t gs = h->enc_streams[(s0 + g) % kEncStreams];
            if (g && c->buf_ev) CUB(cudaStreamWaitEvent(gs, c->buf_ev, 0));   // new_ct ordered stream s0 only
            char *dst = static_cast<char *>(h->r_x[b].p) + first * esize;
            timed(h, kKCopyIn, cs, [&] {
                if (stage) {
                    cudaError_t ce = cudaMemcpyAsync(dst, stage + first * esize, (last - first) * esize, cudaMemcpyHostToDevice, cs);
                    rcode = ce == cudaSuccess ? 0 : cuda_fail(h, ce, "cudaMemcpyAsync");
                } else {
                    rcode = h2d(h, dst, static_cast<const char *>(x) + first * esize, (last - first) * esize, cs);
                }
            });
            if (rcode) return bail(rcode);
            CUB(cudaEventRecord(h->ev_x_in[b], cs));
            CUB(cudaStreamWaitEvent(gs, h->ev_x_in[b], 0));
            const double *dxg = reinterpret_cast<const double *>(dst);
            u64 *ctg = c->d + c0 * words;
            if ((rcode = encrypt_dense_many(h, &dxg, 1, last - first, c0, reinterpret_cast<uint64_t *const *>(&ctg), gs, nullptr, dtype)))
                return bail(rcode);
            if ((c->chunk_ev[g] = take_event(h))) CUB(cudaEventRecord(c->chunk_ev[g], gs));
            if (g) {
                CUB(cudaEventRecord(h->ev_part_done[g], gs));
                CUB(cudaStreamWaitEvent(st, h->ev_part_done[g], 0));
            }
        }
        if (stage) {   // the piece is free again when the last group has gone up
            CUB(cudaEventRecord(h->ev_piece[piece], cs));
            h->piece_busy[piece] = true;
        }
        c->chunks = kUploadChunks;
        for (uint32_t g = 0; g < kUploadChunks; g++)
            if (!c->chunk_ev[g]) c->chunks = 0;   // out of events: consumers fall back to `ready`
    } else if ((rcode = encrypt_dense_many(h, &dx64, 1, n, 0, reinterpret_cast<uint64_t *const *>(&ct_ptr), st, nullptr, dtype))) {
        return bail(rcode);
    }
    CUB(cudaEventRecord(c->ready, st));
    c->ready_pending = true;
    if (!on_device) {
        CUB(cudaEventRecord(h->ev_x_free[b], st));
        h->x_busy[b] = true;
        if (borrowed) CUB(cudaEventSynchronize(h->ev_x_in[b]));   // the caller's buffer is only borrowed for the call
        // (pageable memory has been copied into the staging ring by the time h2d returns)
    } else {
        CUB(cudaStreamSynchronize(st));             // likewise: the kernel read the caller's device buffer
    }
    *out = c;
    return 0;
}

// a serialized vector (bytes from encrypt / computeWeightedAverage, here or in a PALISADE peer) -> resident
int ct_from_blob(shelfi_ckks *h, const uint8_t *blob, uint64_t len, shelfi_ct **out) {
    int rcode = check_stage(h, false, false);
    if (rcode) return rcode;
    if (!blob || !out) return fail(h, SHELFI_ERR_ARG, "null argument");
    if ((rcode = ensure_resident(h))) return rcode;
    wire::BlobMap map;
    std::string e = wire::map_ciphertext_blob(blob, len, false, &map);
    if (!e.empty()) return fail(h, SHELFI_ERR_FORMAT, e);
    if (!same_ring(map.ctx, h->ctx)) return fail(h, SHELFI_ERR_PARAMS, "ciphertext parameters differ from the loaded context");
    if (map.ct_limbs < 1 || map.ct_limbs > h->ctx.limbs) return fail(h, SHELFI_ERR_FORMAT, "unexpected tower count");
    shelfi_ct *c = nullptr;
    if ((rcode = new_ct(h, map.n_ct, map.ct_limbs, &c))) return rcode;
    c->ctx = map.ctx;
    bool uniform = true;
    for (size_t j = 1; uniform && j < map.meta.size(); j++) uniform = same_meta(map.meta[j], map.meta[0]);
    for (auto &m : map.meta) m.limbs = map.ct_limbs == h->ctx.limbs ? 0 : map.ct_limbs;
    if (uniform) c->meta.assign(1, map.meta[0]); else c->meta = map.meta;
    cudaStream_t st = h->stream;
    auto bail = [&](int code) {
        cudaStreamSynchronize(st);
        drop_ct(h, c);
        return code;
    };
    CUB(h->s_image.reserve(image_bytes(len)));
    if ((rcode = h2d(h, h->s_image.p, blob, len, st))) return bail(rcode);
    timed(h, kKUnpack, st, [&] {
        launch_unpack(h->rc, to_runs(h, map.run0, map.run1, map.stride, map.ct_limbs), h->s_image.as<u64>(), c->d, map.n_ct, st);
    });
    CUB(cudaGetLastError());
    CUB(cudaStreamSynchronize(st));   // the blob is only borrowed for the call (and s_image is shared scratch)
    *out = c;
    return 0;
}

// ckks.cpp:264-320 over resident vectors; the result is a new resident vector
int ct_weighted_average(shelfi_ckks *h, shelfi_ct *const *cts, const float *w, uint32_t k, int rescale, shelfi_ct **out) {
    NvtxRange nvtx_range("shelfi.computeWeightedAverage[resident]");
    int rcode = check_stage(h, false, false);
    if (rcode) return rcode;
    if (!cts || !w || !out) return fail(h, SHELFI_ERR_ARG, "null argument");
    if (k == 0) return fail(h, SHELFI_ERR_ARG, "no learners");
    const uint32_t L = h->ctx.limbs;
    const double sf0 = (double)h->ctx.modulus[L - 1];
    if (rescale && L < 2) return fail(h, SHELFI_ERR_UNSUPPORTED, "rescale needs a chain of at least two towers");
    for (uint32_t i = 0; i < k; i++) {
        const std::string who = "learner " + std::to_string(i);
        if (!cts[i] || cts[i]->owner != h) return fail(h, SHELFI_ERR_ARG, who + ": not a resident vector of this handle");
        if (cts[i]->n_ct != cts[0]->n_ct) return fail(h, SHELFI_ERR_PARAMS, who + ": ciphertext count differs");
        if (cts[i]->limbs != L) return fail(h, SHELFI_ERR_UNSUPPORTED, who + ": only fresh depth-1 level-0 ciphertexts are supported");
        for (const auto &m : cts[i]->meta)
            if (m.depth != 1 || m.level != 0)
                return fail(h, SHELFI_ERR_UNSUPPORTED, who + ": only fresh depth-1 level-0 ciphertexts are supported");
    }
    const uint64_t n_ct = cts[0]->n_ct;
    shelfi_ct *acc = nullptr;
    if ((rcode = new_ct(h, n_ct, L, &acc))) return rcode;
    acc->ctx = cts[0]->ctx;
    acc->meta = cts[0]->meta;                     // ckks.cpp:291-292: the result starts as learner 0's vector
    for (auto &m : acc->meta) { m.depth += 1; m.scaling_factor *= sf0; }   // [PAL] EvalMult(ct, double), no rescale
    cudaStream_t st = h->stream;
    std::vector<const uint64_t *> ptrs(k);
    // a large result that is not rescaled: in kCtParts launches with an event after each (see shelfi_ct::part_ev)
    bool in_parts = !rescale && n_ct >= kPipelineMinCts;
    const uint32_t n_parts = part_count(n_ct);
    for (uint32_t g = 0; in_parts && g < n_parts; g++)
        if (!(acc->part_ev[g] = take_event(h))) in_parts = false;
    // learners encrypted in groups (shelfi_ct::chunk_ev) are waited for group by group, as the parts reach them
    const bool by_chunk = in_parts && n_parts % kUploadChunks == 0 && !h->opt_no_chunk_wait;
    for (uint32_t i = 0; i < k; i++) {
        if (!(by_chunk && cts[i]->ready_pending && cts[i]->chunks == kUploadChunks)) ct_ready(h, cts[i]);
        ptrs[i] = cts[i]->d;
    }
    if (in_parts) {
        const uint64_t words = 2ull * L * h->ctx.ring_dim;
        std::vector<const uint64_t *> part(k);
        for (uint32_t g = 0; g < n_parts; g++) {
            const uint64_t c0 = n_ct * g / n_parts, c1 = n_ct * (g + 1) / n_parts;
            if (by_chunk && g % (n_parts / kUploadChunks) == 0) {   // this part is the first of encrypt group j
                const uint32_t j = g / (n_parts / kUploadChunks);
                for (uint32_t i = 0; i < k; i++)
                    if (cts[i]->ready_pending && cts[i]->chunks == kUploadChunks) {
                        if (cudaStreamWaitEvent(st, cts[i]->chunk_ev[j], 0) != cudaSuccess) { drop_ct(h, acc); return cuda_fail(h, cudaGetLastError(), "cudaStreamWaitEvent"); }
                        if (j + 1 == kUploadChunks) cts[i]->ready_pending = false;   // every group has been waited for
                    }
            }
            for (uint32_t i = 0; i < k; i++) part[i] = ptrs[i] + c0 * words;
            if ((rcode = weighted_sum_dense(h, part.data(), w, k, c1 - c0, acc->d + c0 * words, false, st))) { drop_ct(h, acc); return rcode; }
            if (cudaEventRecord(acc->part_ev[g], st) != cudaSuccess) { drop_ct(h, acc); return cuda_fail(h, cudaGetLastError(), "cudaEventRecord"); }
        }
        acc->parts = n_parts;
        *out = acc;
        return 0;
    }
    if ((rcode = weighted_sum_dense(h, ptrs.data(), w, k, n_ct, acc->d, false, st))) { drop_ct(h, acc); return rcode; }
    if (rescale) {
        shelfi_ct *red = nullptr;
        if ((rcode = new_ct(h, n_ct, L - 1, &red))) { drop_ct(h, acc); return rcode; }
        red->ctx = acc->ctx; red->meta = acc->meta;
        for (auto &m : red->meta) { m.depth -= 1; m.level += 1; m.scaling_factor /= sf0; m.limbs = L - 1; }
        if ((rcode = rescale_dense(h, acc->d, n_ct, L, red->d, st))) { drop_ct(h, red); drop_ct(h, acc); return rcode; }
        drop_ct(h, acc);   // stream order: whoever gets this buffer next writes it after the rescale has read it
        acc = red;
    }
    *out = acc;
    return 0;
}

// ckks.cpp:170-213 from a resident vector into host memory (any kind); returns when `out` holds the values
int ct_decrypt(shelfi_ckks *h, const shelfi_ct *c, uint64_t n, double *out) {
    NvtxRange nvtx_range("shelfi.decrypt[resident]");
    int rcode = check_stage(h, false, true);
    if (rcode) return rcode;
    if (!c || c->owner != h || !out) return fail(h, SHELFI_ERR_ARG, "null argument or foreign vector");
    if ((rcode = ensure_resident(h))) return rcode;
    ct_ready(h, const_cast<shelfi_ct *>(c));
    const uint64_t batch = h->batch, n_ct = c->n_ct;
    if (n == 0 || n > n_ct * batch || n + batch <= n_ct * batch)
        return fail(h, SHELFI_ERR_ARG, "data_dimensions inconsistent with the number of ciphertexts");
    cudaStream_t st = h->stream, ds = h->d2h_stream;
    const uint64_t words = 2ull * c->limbs * h->ctx.ring_dim;
    auto bail = [&](int code) {
        cudaStreamSynchronize(st);
        for (cudaStream_t e : h->enc_streams) cudaStreamSynchronize(e);
        cudaStreamSynchronize(ds);
        return code;
    };
    CUB(cudaStreamWaitEvent(st, h->ev_r_out, 0));   // an earlier call's copy out of r_out (already finished: calls return after it)
    if (h->r_out.cap < n * sizeof(double)) {
        CUB(cudaStreamSynchronize(st));
        CUB(h->r_out.reserve(n * sizeof(double)));
    }
    double *d_out = h->r_out.as<double>();
    auto sf_of = [&](uint64_t j) { return c->meta[c->meta.size() == 1 ? 0 : j].scaling_factor; };
    const bool one_scale = c->meta.size() == 1;
    const bool pinned = is_pinned(out);
    // a large vector into page-locked memory: groups are decrypted while the group before crosses PCIe
    const uint32_t groups = (one_scale && pinned && n_ct >= kPipelineMinCts) ? part_count(n_ct) : 1;
    if (groups > 1) {
        // The groups run on side streams, side by side, each in its own slice of the scratch: a quarter of the vector
        // is less than one wave of CTAs, so four launches in a row would take four kernel latencies.  A vector written
        // in parts (the weighted average just queued) releases group g as soon as ITS part is there, under the sums of
        // the later parts.  The call ends with every side stream drained (the copy stream waits for each and is
        // synchronised) and the compute stream ordered after them, so nothing outlives the call.
        const bool by_part = c->parts == groups;
        if ((rcode = decrypt_scratch_reserve(h, n_ct, c->limbs))) return bail(rcode);
        if (!h->ev_scratch) CUB(cudaEventCreateWithFlags(&h->ev_scratch, cudaEventDisableTiming));
        if (!by_part) CUB(cudaEventRecord(h->ev_fork, st));
        for (uint32_t g = 0; g < groups; g++) {
            cudaStream_t gs = h->enc_streams[g];
            CUB(cudaStreamWaitEvent(gs, by_part ? c->part_ev[g] : h->ev_fork, 0));
            CUB(cudaStreamWaitEvent(gs, h->ev_r_out, 0));
            if (h->scratch_used && h->scratch_stream != gs) CUB(cudaStreamWaitEvent(gs, h->ev_scratch, 0));
        }
        for (uint32_t g = 0; g < groups; g++) {
            const uint64_t c0 = n_ct * g / groups, c1 = n_ct * (g + 1) / groups;
            const uint64_t first = c0 * batch, last = std::min(n, c1 * batch);
            cudaStream_t gs = h->enc_streams[g];
            const ScratchSlice slice{c0, n_ct};
            if ((rcode = decrypt_dense(h, c->d + c0 * words, c1 - c0, sf_of(0), d_out + first, last - first, gs, nullptr, c->limbs,
                                       nullptr, 0, c0, &slice)))
                return bail(rcode);
            CUB(cudaEventRecord(h->ev_part_done[g], gs));
            CUB(cudaStreamWaitEvent(ds, h->ev_part_done[g], 0));
            CUB(cudaStreamWaitEvent(st, h->ev_part_done[g], 0));
            cudaError_t ce = cudaSuccess;
            timed(h, kKCopyOut, ds, [&] {
                ce = cudaMemcpyAsync(out + first, d_out + first, (last - first) * sizeof(double), cudaMemcpyDeviceToHost, ds);
            });
            CUB(ce);
        }
        if ((rcode = scratch_end(h, st))) return bail(rcode);   // the compute stream now follows every user of the scratch
        CUB(cudaEventRecord(h->ev_r_out, ds));
        CUB(cudaStreamSynchronize(ds));
        return 0;
    }
    for (uint64_t j0 = 0; j0 < n_ct;) {   // ciphertexts usually share one scaling factor; split the launch where they do not
        uint64_t j1 = one_scale ? n_ct : j0 + 1;
        while (j1 < n_ct && sf_of(j1) == sf_of(j0)) j1++;
        const uint64_t first = j0 * batch, last = std::min(n, j1 * batch);
        if ((rcode = decrypt_dense(h, c->d + j0 * words, j1 - j0, sf_of(j0), d_out + first, last - first, st, nullptr, c->limbs,
                                   nullptr, 0, j0)))
            return bail(rcode);
        j0 = j1;
    }
    if ((rcode = d2h_sync(h, out, d_out, n * sizeof(double), st))) return bail(rcode);
    return 0;
}

uint64_t ct_blob_size(const shelfi_ct *c) {
    return wire::ciphertext_blob_size(c->ctx, c->n_ct, c->limbs == c->ctx.limbs ? 0 : c->limbs);
}
// Serial::Serialize of a resident vector (ckks.cpp:98-101,308-310): the exact bytes the host-buffer calls return
int ct_to_blob(shelfi_ckks *h, const shelfi_ct *c, uint8_t *dst, uint64_t cap, uint64_t *len) {
    int rcode = check_stage(h, false, false);
    if (rcode) return rcode;
    if (!c || c->owner != h || !dst || !len) return fail(h, SHELFI_ERR_ARG, "null argument or foreign vector");
    trim_enc_plans(h);
    ct_ready(h, const_cast<shelfi_ct *>(c));
    shelfi_ckks::EncPlan *ep = nullptr;
    wire::BlobPlan own;
    if ((rcode = ct_plan(h, c, &ep, &own))) return rcode;
    const wire::BlobPlan &pl = ep ? ep->plan : own;
    if (cap < pl.size) return fail(h, SHELFI_ERR_ARG, "destination buffer too small: need " + std::to_string(pl.size) + " bytes");
    cudaStream_t st = h->stream;
    auto bail = [&](int code) {
        cudaStreamSynchronize(st);
        return code;
    };
    CUB(h->s_image.reserve(image_bytes(pl.size)));
    timed(h, kKPack, st, [&] {
        launch_pack(h->rc, to_runs(h, pl.run0, pl.run1, pl.stride, c->limbs), c->d, h->s_image.as<u64>(), c->n_ct, st);
    });
    const bool dev_hdr = ep && c->n_ct >= kDeviceHeaderMinCts;
    if (dev_hdr) launch_write_segments(ep->d_desc.as<SegDesc>(), ep->d_bytes.as<uint8_t>(), h->s_image.as<uint8_t>(), ep->n_seg, st);
    CUB(cudaGetLastError());
    if ((rcode = d2h_sync(h, dst, h->s_image.p, pl.size, st))) return bail(rcode);
    if (!dev_hdr) write_segments(pl, dst);
    *len = pl.size;
    return 0;
}

}  // namespace

// =============================================================================================
// C ABI
// =============================================================================================
extern "C" {

const char *shelfi_version(void) { return "0.1.0-b200"; }

int shelfi_ckks_create(const char *scheme, uint32_t batch_size, uint32_t scale_factor_bits, const char *cryptodir,
                       int device, shelfi_ckks **out) {
    if (!out) return SHELFI_ERR_ARG;
    shelfi_ckks *h = new shelfi_ckks();
    h->scheme = scheme ? scheme : "ckks";
    h->batch = batch_size;
    h->scale_bits = scale_factor_bits;
    h->cryptodir = cryptodir ? cryptodir : "";
    h->device = device;
    if (const char *e = std::getenv("SHELFI_EVALMULT_ROUND")) h->opt_evalmult_round = std::atoi(e) == 1;
    if (const char *e = std::getenv("SHELFI_DECODE_NOISE")) h->opt_decode_noise = std::atoi(e) == 1;
    *out = h;
    return 0;
}

void shelfi_ckks_destroy(shelfi_ckks *h) {
    if (!h) return;
    if (h->cuda_ready) {
        cudaSetDevice(h->device);
        cudaDeviceSynchronize();
        h->prof.release();
        free_params(h);
        for (shelfi_ct *c : h->live) {   // vectors die with their handle
            if (c->d) cudaFree(c->d);
            if (c->buf_ev) cudaEventDestroy(c->buf_ev);
            if (c->ready) cudaEventDestroy(c->ready);
            delete c;
        }
        h->live.clear();
        h->pool.trim();
        for (cudaEvent_t e : h->spare_events) cudaEventDestroy(e);
        h->spare_events.clear();
        for (int i = 0; i < kEncStreams; i++) {
            h->r_x[i].release();
            if (h->ev_x_in[i]) cudaEventDestroy(h->ev_x_in[i]);
            if (h->ev_x_free[i]) cudaEventDestroy(h->ev_x_free[i]);
            if (h->enc_streams[i]) cudaStreamDestroy(h->enc_streams[i]);
            if (h->ev_part_done[i]) cudaEventDestroy(h->ev_part_done[i]);
        }
        if (h->ev_fork) cudaEventDestroy(h->ev_fork);
        h->r_out.release();
        if (h->ev_r_out) cudaEventDestroy(h->ev_r_out);
        if (h->ev_scratch) cudaEventDestroy(h->ev_scratch);
        for (DeviceBuffer *b : {&h->s_x, &h->s_m, &h->s_ve, &h->s_coeff, &h->s_dense, &h->s_image, &h->s_image2, &h->s_image3,
                                &h->s_out, &h->s_vals, &h->s_zs, &h->s_rows, &h->s_pt, &h->s_resc, &h->s_dense2})
            b->release();
        for (int i = 0; i < 2; i++) { cudaEventDestroy(h->ev_copy[i]); cudaEventDestroy(h->ev_done[i]); }
        for (int i = 0; i < 2; i++) {
            if (h->stage[i]) cudaFreeHost(h->stage[i]);
            if (h->ev_stage[i]) cudaEventDestroy(h->ev_stage[i]);
        }
        for (int i = 0; i < kStagePieces; i++) {
            if (h->ev_piece[i]) cudaEventDestroy(h->ev_piece[i]);
        }
        if (h->pin_x) cudaFreeHost(h->pin_x);
        delete h->copier;
        cudaStreamDestroy(h->stream);
        cudaStreamDestroy(h->copy_stream);
        if (h->d2h_stream) { cudaStreamDestroy(h->d2h_stream); cudaEventDestroy(h->ev_stage_out); }
        if (h->enc_stream) {
            cudaStreamDestroy(h->enc_stream);
            cudaEventDestroy(h->ev_enc_fork);
            for (int i = 0; i < 2; i++) { cudaEventDestroy(h->ev_enc_ready[i]); cudaEventDestroy(h->ev_enc_free[i]); }
        }
    }
    delete h;
}

const char *shelfi_last_error(const shelfi_ckks *h) { return h ? h->err.c_str() : "null handle"; }

int shelfi_load_crypto_params_from_memory(shelfi_ckks *h, const uint8_t *cc, size_t cc_len, const uint8_t *pk,
                                          size_t pk_len, const uint8_t *sk, size_t sk_len) {
    if (!h) return SHELFI_ERR_ARG;
    std::lock_guard<std::mutex> lock(h->mu);
    int rcode = ensure_cuda(h);
    if (rcode) return rcode;
    h->have_cc = h->have_pk = h->have_sk = false;
    drop_enc_plans(h);
    int missing = 0;
    std::string detail;
    wire::ContextParams ctx;
    std::string e = cc ? wire::parse_context(cc, cc_len, &ctx) : std::string("file missing");
    if (e.empty()) {
        rcode = install_context(h, ctx);
        if (rcode) return rcode;
    } else {
        missing |= 1;
        detail += "cryptocontext: " + e + "; ";
    }
    wire::KeyFile key;
    e = !h->have_cc ? std::string("no context") : pk ? wire::parse_public_key(pk, pk_len, &key) : std::string("file missing");
    if (e.empty() && !same_ring(key.ctx, h->ctx)) e = "key belongs to different parameters";
    if (e.empty()) {
        size_t words = (size_t)h->ctx.limbs * h->ctx.ring_dim;
        if ((rcode = upload_key_poly(h, &h->d_pk_b, key.polys.data()))) return rcode;
        if ((rcode = upload_key_poly(h, &h->d_pk_a, key.polys.data() + words))) return rcode;
        h->key_tag = key.tag;
        h->have_pk = true;
    } else {
        missing |= 2;
        detail += "key-public: " + e + "; ";
    }
    wire::KeyFile skey;
    e = !h->have_cc ? std::string("no context") : sk ? wire::parse_private_key(sk, sk_len, &skey) : std::string("file missing");
    if (e.empty() && !same_ring(skey.ctx, h->ctx)) e = "key belongs to different parameters";
    if (e.empty()) {
        if ((rcode = upload_key_poly(h, &h->d_sk, skey.polys.data()))) return rcode;
        if (!h->have_pk) h->key_tag = skey.tag;
        h->have_sk = true;
    } else {
        missing |= 4;
        detail += "key-private: " + e + "; ";
    }
    h->err = detail;
    return missing;
}

int shelfi_load_crypto_params(shelfi_ckks *h) {
    if (!h) return SHELFI_ERR_ARG;
    std::vector<uint8_t> f[3];
    bool ok[3];
    static const char *names[3] = {"cryptocontext.txt", "key-public.txt", "key-private.txt"};
    for (int i = 0; i < 3; i++) ok[i] = read_file(h->cryptodir + names[i], &f[i]);  // ckks.cpp:12,16,20
    return shelfi_load_crypto_params_from_memory(h, ok[0] ? f[0].data() : nullptr, f[0].size(),
                                                 ok[1] ? f[1].data() : nullptr, f[1].size(),
                                                 ok[2] ? f[2].data() : nullptr, f[2].size());
}

int shelfi_gen_crypto_context_and_keygen(shelfi_ckks *h) {
    return shelfi_gen_crypto_context_and_keygen_ex(h, 1, 0);   // ckks.cpp:26: multDepth = 1, automatic ring dimension
}

int shelfi_gen_crypto_context_and_keygen_ex(shelfi_ckks *h, uint32_t mult_depth, uint32_t ring_dim) {
    // ckks.cpp:25-59: genCryptoContextCKKS(multDepth, scaleFactorBits, batchSize), Enable(ENCRYPTION | SHE),
    // KeyGen, three SerializeToFile calls; the object's own context/keys stay untouched.
    if (!h) return 0;
    std::lock_guard<std::mutex> lock(h->mu);
    if (ensure_cuda(h)) return 0;
    const uint32_t bits = h->scale_bits, batch = h->batch;
    if (bits < 14 || bits > 59) { fail(h, SHELFI_ERR_ARG, "scaleFactorBits must be in [14, 59]"); return 0; }
    if (mult_depth < 1 || mult_depth + 1 > (uint32_t)kMaxLimbs) { fail(h, SHELFI_ERR_ARG, "multDepth must be in [1, 7]"); return 0; }
    const uint32_t N = ring_dim ? ring_dim : nt::select_ring_dim(bits, mult_depth);
    if (N != 8192 && N != 16384 && N != 32768) {
        fail(h, SHELFI_ERR_UNSUPPORTED, "parameter set needs ring dimension " + std::to_string(N) + "; supported: 8192, 16384, 32768");
        return 0;
    }
    if (batch == 0 || (batch & (batch - 1)) || batch > N / 2) {
        fail(h, SHELFI_ERR_ARG, "batchSize must be a power of two <= N/2");
        return 0;
    }
    wire::ContextParams p;
    p.ring_dim = N; p.cyclotomic_order = 2 * N;
    const int L = nt::modulus_chain(bits, mult_depth, N, p.modulus);   // q_last = scaling prime, q_0 60 bits
    p.limbs = (uint32_t)L;
    for (int l = 0; l < L; l++) p.root[l] = nt::min_primitive_root(2 * N, p.modulus[l]);
    wire::set_big_modulus(&p);
    p.plaintext_modulus = bits; p.batch_size = batch;
    p.sigma = 3.19f; p.assurance = 9.0f; p.hermite = 1.006f;
    // relinWindow, depth, maxDepth, [extra word of newer PALISADE], mode OPTIMIZED, HEStd_128_classic,
    // EXACTRESCALE, HYBRID, dnum
    const uint32_t dnum = nt::num_large_digits(mult_depth);
    const char *variant = std::getenv("SHELFI_WIRE_VARIANT");
    if (h->opt_wire_variant == 8 || (!h->opt_wire_variant && variant && std::atoi(variant) == 8)) {
        const uint32_t v[8] = {0, 1, 2, 1, 0, 2, 1, dnum};
        std::memcpy(p.rlwe_ints, v, sizeof v); p.rlwe_int_count = 8;
    } else {
        const uint32_t v[9] = {0, 1, 2, 0, 1, 0, 2, 1, dnum};
        std::memcpy(p.rlwe_ints, v, sizeof v); p.rlwe_int_count = 9;
    }
    p.enabled_features = 5;  // ENCRYPTION | SHE (ckks.cpp:32-33)
    if (!write_file(h->cryptodir + "cryptocontext.txt", wire::emit_context(p))) {  // ckks.cpp:41
        fail(h, SHELFI_ERR_IO, "cannot write cryptocontext.txt");
        return 0;
    }
    // [PAL] KeyGen: s <- ternary, e <- Gaussian, a <- uniform (EVALUATION), b = e - a*s.  The transforms
    // run on the device through a scratch handle so this object's loaded state is left alone.
    shelfi_ckks tmp;
    tmp.device = h->device; tmp.batch = batch; tmp.scale_bits = bits;
    auto cleanup = [&]() { free_params(&tmp); tmp.s_dense.release(); tmp.s_rows.release();
                           if (tmp.stream) cudaStreamDestroy(tmp.stream);
                           if (tmp.copy_stream) cudaStreamDestroy(tmp.copy_stream);
                           for (int i = 0; i < 2; i++) { if (tmp.ev_copy[i]) cudaEventDestroy(tmp.ev_copy[i]);
                                                         if (tmp.ev_done[i]) cudaEventDestroy(tmp.ev_done[i]); } };
    if (ensure_cuda(&tmp) || install_context(&tmp, p)) { h->err = tmp.err; cleanup(); return 0; }
    // Secret material (s, e) and public material (a, the key tag) are drawn under independent keys, so nothing that
    // is published carries information about the stream that produced the secret key.
    SamplerConst secret{};
    SamplerKey pubkey;
    if (call_key(h, kKeyDomainKeySecret, &secret.key) || call_key(h, kKeyDomainKeyPublic, &pubkey)) { cleanup(); return 0; }
    std::memcpy(secret.gauss_cdf, tmp.gauss_cdf, sizeof secret.gauss_cdf);
    const size_t LN = (size_t)L * N;
    std::vector<uint64_t> s_hat(LN), a_rows(LN), b_rows(LN);
    DeviceBuffer d_a, d_b, d_tag;
    bool ok = tmp.s_rows.reserve(2 * LN * 8) == cudaSuccess && tmp.s_dense.reserve(2 * LN * 8) == cudaSuccess &&
              d_a.reserve(LN * 8) == cudaSuccess && d_b.reserve(LN * 8) == cudaSuccess && d_tag.reserve(64) == cudaSuccess;
    uint32_t tag_words[16] = {0};
    if (ok) {
        launch_keygen_sample(tmp.rc, secret, pubkey, tmp.s_rows.as<u64>(), d_a.as<u64>(), tmp.stream);
        launch_ntt_forward(tmp.rc, tmp.s_rows.as<u64>(), tmp.s_dense.as<u64>(), 2 * (u32)L, tmp.stream);
        launch_keygen_combine(tmp.rc, tmp.s_dense.as<u64>(), d_a.as<u64>(), d_b.as<u64>(), tmp.stream);
        launch_chacha20_block(pubkey, 0, 0, 0, kStreamKeyTag, d_tag.as<u32>(), tmp.stream);
        ok = cudaStreamSynchronize(tmp.stream) == cudaSuccess &&
             cudaMemcpy(s_hat.data(), tmp.s_dense.p, LN * 8, cudaMemcpyDeviceToHost) == cudaSuccess &&
             cudaMemcpy(a_rows.data(), d_a.p, LN * 8, cudaMemcpyDeviceToHost) == cudaSuccess &&
             cudaMemcpy(b_rows.data(), d_b.p, LN * 8, cudaMemcpyDeviceToHost) == cudaSuccess &&
             cudaMemcpy(tag_words, d_tag.p, 64, cudaMemcpyDeviceToHost) == cudaSuccess;
        // the secret key and its error never stay behind in device memory
        cudaMemsetAsync(tmp.s_rows.p, 0, 2 * LN * 8, tmp.stream);
        cudaMemsetAsync(tmp.s_dense.p, 0, 2 * LN * 8, tmp.stream);
        cudaStreamSynchronize(tmp.stream);
    }
    d_a.release(); d_b.release(); d_tag.release();
    if (!ok) { fail(h, SHELFI_ERR_CUDA, "key generation kernels failed"); cleanup(); return 0; }
    wire::KeyFile pub, priv;
    pub.ctx = p; priv.ctx = p;
    pub.polys.assign(b_rows.begin(), b_rows.end());
    pub.polys.insert(pub.polys.end(), a_rows.begin(), a_rows.end());
    priv.polys.assign(s_hat.begin(), s_hat.end());
    char tag[33];   // [PAL] GenerateUniqueKeyID: 32 hex characters
    std::snprintf(tag, sizeof tag, "%08x%08x%08x%08x", tag_words[0], tag_words[1], tag_words[2], tag_words[3]);
    pub.tag = tag; priv.tag = tag;
    cleanup();
    if (!write_file(h->cryptodir + "key-public.txt", wire::emit_public_key(pub))) {  // ckks.cpp:48
        fail(h, SHELFI_ERR_IO, "cannot write key-public.txt");
        return 0;
    }
    if (!write_file(h->cryptodir + "key-private.txt", wire::emit_private_key(priv))) {  // ckks.cpp:53
        fail(h, SHELFI_ERR_IO, "cannot write key-private.txt");
        return 0;
    }
    return 1;
}

int shelfi_set_sampler_seed(shelfi_ckks *h, uint64_t seed) {
    if (!h) return SHELFI_ERR_ARG;
    std::lock_guard<std::mutex> lock(h->mu);
    h->seeded = true; h->seed = seed;
    return 0;
}
int shelfi_clear_sampler_seed(shelfi_ckks *h) {
    if (!h) return SHELFI_ERR_ARG;
    std::lock_guard<std::mutex> lock(h->mu);
    h->seeded = false;
    return 0;
}
int shelfi_set_option(shelfi_ckks *h, int option, int64_t value) {
    if (!h) return SHELFI_ERR_ARG;
    std::lock_guard<std::mutex> lock(h->mu);
    switch (option) {
        case SHELFI_OPT_EVALMULT_ROUND:
            if (value != 0 && value != 1) return fail(h, SHELFI_ERR_ARG, "evalmult_round: 0 or 1");
            h->opt_evalmult_round = (int)value;
            return 0;
        case SHELFI_OPT_DECODE_NOISE:
            if (value != 0 && value != 1) return fail(h, SHELFI_ERR_ARG, "decode_noise: 0 or 1");
            h->opt_decode_noise = (int)value;
            return 0;
        case SHELFI_OPT_WIRE_VARIANT:
            if (value != 0 && value != 8 && value != 9) return fail(h, SHELFI_ERR_ARG, "wire_variant: 0, 8 or 9");
            h->opt_wire_variant = (int)value;
            return 0;
        default:
            return fail(h, SHELFI_ERR_ARG, "unknown option");
    }
}
int64_t shelfi_get_option(const shelfi_ckks *h, int option) {
    if (!h) return -1;
    switch (option) {
        case SHELFI_OPT_EVALMULT_ROUND: return h->opt_evalmult_round;
        case SHELFI_OPT_DECODE_NOISE: return h->opt_decode_noise;
        case SHELFI_OPT_WIRE_VARIANT: return h->opt_wire_variant ? h->opt_wire_variant : (h->have_cc ? h->ctx.rlwe_int_count : 9);
        default: return -1;
    }
}
// known-answer tap of the device sampler: one ChaCha20 block (key 8 words, counter, nonce 3 words -> 16 words)
int shelfi_chacha20_block_device(shelfi_ckks *h, const uint32_t key[8], uint32_t counter, const uint32_t nonce[3],
                                 uint32_t out[16]) {
    if (!h) return SHELFI_ERR_ARG;
    std::lock_guard<std::mutex> lock(h->mu);
    int rcode = ensure_cuda(h);
    if (rcode) return rcode;
    if (!key || !nonce || !out) return fail(h, SHELFI_ERR_ARG, "null argument");
    SamplerKey k;
    std::memcpy(k.k, key, sizeof k.k);
    DeviceBuffer d;
    CU(d.reserve(64));
    launch_chacha20_block(k, counter, nonce[0], nonce[1], nonce[2], d.as<u32>(), h->stream);
    cudaError_t e = cudaStreamSynchronize(h->stream);
    if (e == cudaSuccess) e = cudaMemcpy(out, d.p, 64, cudaMemcpyDeviceToHost);
    d.release();
    if (e != cudaSuccess) return cuda_fail(h, e, "chacha20 block");
    return 0;
}

int shelfi_profile_enable(shelfi_ckks *h, int on) {
    if (!h) return SHELFI_ERR_ARG;
    std::lock_guard<std::mutex> lock(h->mu);
    h->prof.on = on != 0;
    return 0;
}
int shelfi_profile_reset(shelfi_ckks *h) {
    if (!h) return SHELFI_ERR_ARG;
    std::lock_guard<std::mutex> lock(h->mu);
    h->prof.drain();
    for (int i = 0; i < kKernelCount; i++) { h->prof.total_ms[i] = 0; h->prof.launches[i] = 0; }
    return 0;
}
int shelfi_profile_get(shelfi_ckks *h, int kernel, double *ms_total, uint64_t *launches) {
    if (!h || kernel < 0 || kernel >= kKernelCount) return SHELFI_ERR_ARG;
    std::lock_guard<std::mutex> lock(h->mu);
    h->prof.drain();
    if (ms_total) *ms_total = h->prof.total_ms[kernel];
    if (launches) *launches = h->prof.launches[kernel];
    return 0;
}

uint32_t shelfi_ring_dim(const shelfi_ckks *h) { return h && h->have_cc ? h->ctx.ring_dim : 0; }
uint32_t shelfi_num_limbs(const shelfi_ckks *h) { return h && h->have_cc ? h->ctx.limbs : 0; }
uint32_t shelfi_slots(const shelfi_ckks *h) { return h && h->have_cc ? h->ctx.batch_size : 0; }
uint32_t shelfi_batch_size(const shelfi_ckks *h) { return h ? h->batch : 0; }
uint64_t shelfi_modulus(const shelfi_ckks *h, uint32_t l) { return h && h->have_cc && l < h->ctx.limbs ? h->ctx.modulus[l] : 0; }
uint64_t shelfi_root_of_unity(const shelfi_ckks *h, uint32_t l) { return h && h->have_cc && l < h->ctx.limbs ? h->ctx.root[l] : 0; }
double shelfi_scaling_factor(const shelfi_ckks *h) { return h && h->have_cc ? (double)h->ctx.modulus[h->ctx.limbs - 1] : 0.0; }
int shelfi_key_tag(const shelfi_ckks *h, char out[33]) {
    if (!h || !out) return SHELFI_ERR_ARG;
    std::memset(out, 0, 33);
    std::memcpy(out, h->key_tag.c_str(), std::min<size_t>(32, h->key_tag.size()));
    return 0;
}
uint64_t shelfi_num_ciphertexts(const shelfi_ckks *h, uint64_t n) { return h && h->batch ? (n + h->batch - 1) / h->batch : 0; }
uint64_t shelfi_ciphertext_words(const shelfi_ckks *h) { return h && h->have_cc ? 2ull * h->ctx.limbs * h->ctx.ring_dim : 0; }
uint64_t shelfi_blob_size_level(const shelfi_ckks *h, uint64_t n_ct, uint32_t limbs) {
    return h && h->have_cc && n_ct && limbs >= 1 && limbs <= h->ctx.limbs ? wire::ciphertext_blob_size(h->ctx, n_ct, limbs) : 0;
}
uint64_t shelfi_blob_size(const shelfi_ckks *h, uint64_t n_ct) {
    return h && h->have_cc && n_ct ? wire::ciphertext_blob_size(h->ctx, n_ct) : 0;
}
int64_t shelfi_weight_constant(const shelfi_ckks *h, float w) { return h && h->have_cc ? weight_constant(h, w) : 0; }

int shelfi_encrypt(shelfi_ckks *h, const double *x, uint64_t n, uint8_t **blob, uint64_t *blob_len) {
    if (!h || !blob) return SHELFI_ERR_ARG;
    std::lock_guard<std::mutex> lock(h->mu);
    return encrypt_host(h, &x, &n, 1, nullptr, 0, blob, nullptr, blob_len, nullptr);
}
int shelfi_encrypt_into(shelfi_ckks *h, const double *x, uint64_t n, uint8_t *dst, uint64_t cap, uint64_t *blob_len) {
    if (!h) return SHELFI_ERR_ARG;
    std::lock_guard<std::mutex> lock(h->mu);
    if (!blob_len) return fail(h, SHELFI_ERR_ARG, "null argument");
    return encrypt_host(h, &x, &n, 1, dst, cap, nullptr, nullptr, blob_len, nullptr);
}
int shelfi_compute_weighted_average(shelfi_ckks *h, const uint8_t *const *blobs, const uint64_t *blob_lens,
                                    const float *weights, uint32_t k, uint8_t **blob, uint64_t *blob_len) {
    if (!h || !blob) return SHELFI_ERR_ARG;
    std::lock_guard<std::mutex> lock(h->mu);
    if (!blob_len) return fail(h, SHELFI_ERR_ARG, "null argument");
    return wavg_host(h, blobs, blob_lens, weights, k, 1, nullptr, 0, blob, nullptr, blob_len, nullptr);
}
int shelfi_compute_weighted_average_into(shelfi_ckks *h, const uint8_t *const *blobs, const uint64_t *blob_lens,
                                         const float *weights, uint32_t k, uint8_t *dst, uint64_t cap,
                                         uint64_t *blob_len) {
    if (!h) return SHELFI_ERR_ARG;
    std::lock_guard<std::mutex> lock(h->mu);
    if (!blob_len) return fail(h, SHELFI_ERR_ARG, "null argument");
    return wavg_host(h, blobs, blob_lens, weights, k, 1, dst, cap, nullptr, nullptr, blob_len, nullptr);
}
int shelfi_compute_weighted_average_rescaled(shelfi_ckks *h, const uint8_t *const *blobs, const uint64_t *blob_lens,
                                             const float *weights, uint32_t k, uint8_t **blob, uint64_t *blob_len) {
    if (!h || !blob) return SHELFI_ERR_ARG;
    std::lock_guard<std::mutex> lock(h->mu);
    if (!blob_len) return fail(h, SHELFI_ERR_ARG, "null argument");
    return wavg_host(h, blobs, blob_lens, weights, k, 1, nullptr, 0, blob, nullptr, blob_len, nullptr, true);
}
int shelfi_compute_weighted_average_rescaled_into(shelfi_ckks *h, const uint8_t *const *blobs, const uint64_t *blob_lens,
                                                  const float *weights, uint32_t k, uint8_t *dst, uint64_t cap,
                                                  uint64_t *blob_len) {
    if (!h) return SHELFI_ERR_ARG;
    std::lock_guard<std::mutex> lock(h->mu);
    if (!blob_len) return fail(h, SHELFI_ERR_ARG, "null argument");
    return wavg_host(h, blobs, blob_lens, weights, k, 1, dst, cap, nullptr, nullptr, blob_len, nullptr, true);
}
int shelfi_rescale(shelfi_ckks *h, const uint8_t *blob, uint64_t blob_len, uint8_t **out, uint64_t *out_len) {
    if (!h || !out) return SHELFI_ERR_ARG;
    std::lock_guard<std::mutex> lock(h->mu);
    return rescale_host(h, blob, blob_len, nullptr, 0, out, out_len);
}
int shelfi_rescale_into(shelfi_ckks *h, const uint8_t *blob, uint64_t blob_len, uint8_t *dst, uint64_t cap, uint64_t *out_len) {
    if (!h) return SHELFI_ERR_ARG;
    std::lock_guard<std::mutex> lock(h->mu);
    return rescale_host(h, blob, blob_len, dst, cap, nullptr, out_len);
}
int shelfi_decrypt(shelfi_ckks *h, const uint8_t *blob, uint64_t blob_len, uint64_t n, double *out) {
    if (!h) return SHELFI_ERR_ARG;
    std::lock_guard<std::mutex> lock(h->mu);
    return decrypt_host(h, &blob, &blob_len, &n, 1, out);
}

// ---- many independent vectors per call (the layers of a state_dict: code/benchmark.py:490-493,
// code/benchmark_selection.py:152-193 call encrypt / computeWeightedAverage / decrypt once per layer) ----
int shelfi_encrypt_layers(shelfi_ckks *h, const double *const *x, const uint64_t *layer_len, uint32_t n_layers, uint8_t **blobs,
                          uint64_t *offsets, uint64_t *lens, uint64_t *total_len) {
    if (!h || !blobs) return SHELFI_ERR_ARG;
    std::lock_guard<std::mutex> lock(h->mu);
    return encrypt_host(h, x, layer_len, n_layers, nullptr, 0, blobs, offsets, lens, total_len);
}
int shelfi_compute_weighted_average_layers(shelfi_ckks *h, const uint8_t *const *blobs, const uint64_t *blob_lens,
                                           const float *weights, uint32_t k, uint32_t n_layers, uint8_t **out,
                                           uint64_t *offsets, uint64_t *lens, uint64_t *total_len) {
    if (!h || !out) return SHELFI_ERR_ARG;
    std::lock_guard<std::mutex> lock(h->mu);
    return wavg_host(h, blobs, blob_lens, weights, k, n_layers, nullptr, 0, out, offsets, lens, total_len);
}
int shelfi_encrypt_layers_into(shelfi_ckks *h, const double *const *x, const uint64_t *layer_len, uint32_t n_layers, uint8_t *dst,
                               uint64_t cap, uint64_t *offsets, uint64_t *lens, uint64_t *total_len) {
    if (!h) return SHELFI_ERR_ARG;
    std::lock_guard<std::mutex> lock(h->mu);
    return encrypt_host(h, x, layer_len, n_layers, dst, cap, nullptr, offsets, lens, total_len);
}
int shelfi_compute_weighted_average_layers_into(shelfi_ckks *h, const uint8_t *const *blobs, const uint64_t *blob_lens,
                                                const float *weights, uint32_t k, uint32_t n_layers, uint8_t *dst,
                                                uint64_t cap, uint64_t *offsets, uint64_t *lens, uint64_t *total_len) {
    if (!h) return SHELFI_ERR_ARG;
    std::lock_guard<std::mutex> lock(h->mu);
    return wavg_host(h, blobs, blob_lens, weights, k, n_layers, dst, cap, nullptr, offsets, lens, total_len);
}
// the layer-wise average followed by ModReduce on every ciphertext of every layer (rescale-before-transport, opt-in)
int shelfi_compute_weighted_average_layers_rescaled(shelfi_ckks *h, const uint8_t *const *blobs, const uint64_t *blob_lens,
                                                    const float *weights, uint32_t k, uint32_t n_layers, uint8_t **out,
                                                    uint64_t *offsets, uint64_t *lens, uint64_t *total_len) {
    if (!h || !out) return SHELFI_ERR_ARG;
    std::lock_guard<std::mutex> lock(h->mu);
    return wavg_host(h, blobs, blob_lens, weights, k, n_layers, nullptr, 0, out, offsets, lens, total_len, true);
}
int shelfi_compute_weighted_average_layers_rescaled_into(shelfi_ckks *h, const uint8_t *const *blobs,
                                                         const uint64_t *blob_lens, const float *weights, uint32_t k,
                                                         uint32_t n_layers, uint8_t *dst, uint64_t cap, uint64_t *offsets,
                                                         uint64_t *lens, uint64_t *total_len) {
    if (!h) return SHELFI_ERR_ARG;
    std::lock_guard<std::mutex> lock(h->mu);
    return wavg_host(h, blobs, blob_lens, weights, k, n_layers, dst, cap, nullptr, offsets, lens, total_len, true);
}
int shelfi_decrypt_layers(shelfi_ckks *h, const uint8_t *const *blobs, const uint64_t *blob_lens,
                          const uint64_t *data_dimensions, uint32_t n_layers, double *out) {
    if (!h) return SHELFI_ERR_ARG;
    std::lock_guard<std::mutex> lock(h->mu);
    return decrypt_host(h, blobs, blob_lens, data_dimensions, n_layers, out);
}

// ---- resident ciphertext vectors ----
int shelfi_ct_encrypt(shelfi_ckks *h, const void *x, uint64_t n, int dtype, shelfi_ct **out) {
    if (!h) return SHELFI_ERR_ARG;
    std::lock_guard<std::mutex> lock(h->mu);
    return ct_encrypt(h, x, n, dtype, out);
}
int shelfi_ct_from_blob(shelfi_ckks *h, const uint8_t *blob, uint64_t blob_len, shelfi_ct **out) {
    if (!h) return SHELFI_ERR_ARG;
    std::lock_guard<std::mutex> lock(h->mu);
    return ct_from_blob(h, blob, blob_len, out);
}
int shelfi_ct_weighted_average(shelfi_ckks *h, shelfi_ct *const *cts, const float *weights, uint32_t k, int rescale,
                               shelfi_ct **out) {
    if (!h) return SHELFI_ERR_ARG;
    std::lock_guard<std::mutex> lock(h->mu);
    return ct_weighted_average(h, cts, weights, k, rescale, out);
}
int shelfi_ct_decrypt(shelfi_ckks *h, const shelfi_ct *ct, uint64_t n, double *out) {
    if (!h) return SHELFI_ERR_ARG;
    std::lock_guard<std::mutex> lock(h->mu);
    return ct_decrypt(h, ct, n, out);
}
uint64_t shelfi_ct_blob_size(const shelfi_ct *ct) { return ct ? ct_blob_size(ct) : 0; }
int shelfi_ct_to_blob(shelfi_ckks *h, const shelfi_ct *ct, uint8_t *dst, uint64_t cap, uint64_t *blob_len) {
    if (!h) return SHELFI_ERR_ARG;
    std::lock_guard<std::mutex> lock(h->mu);
    return ct_to_blob(h, ct, dst, cap, blob_len);
}
int shelfi_ct_info(const shelfi_ct *ct, uint64_t *n_ct, uint32_t *limbs, uint64_t *depth, uint64_t *level,
                   double *scaling_factor) {
    if (!ct || ct->meta.empty()) return SHELFI_ERR_ARG;
    if (n_ct) *n_ct = ct->n_ct;
    if (limbs) *limbs = ct->limbs;
    if (depth) *depth = ct->meta[0].depth;
    if (level) *level = ct->meta[0].level;
    if (scaling_factor) *scaling_factor = ct->meta[0].scaling_factor;
    return 0;
}
const uint64_t *shelfi_ct_device_ptr(const shelfi_ct *ct) {
    if (!ct) return nullptr;
    if (ct->ready_pending) cudaEventSynchronize(ct->ready);   // the caller will read it on a stream of its own
    return ct->d;
}
void shelfi_ct_free(shelfi_ckks *h, shelfi_ct *ct) {
    if (!h || !ct) return;
    std::lock_guard<std::mutex> lock(h->mu);
    drop_ct(h, ct);
}
int shelfi_resident_stats(shelfi_ckks *h, uint64_t *in_use, uint64_t *pooled, uint64_t *limit) {
    if (!h) return SHELFI_ERR_ARG;
    std::lock_guard<std::mutex> lock(h->mu);
    if (in_use) *in_use = h->pool.in_use;
    if (pooled) *pooled = h->pool.pooled;
    if (limit) *limit = h->pool.limit;
    return 0;
}
int shelfi_resident_set_limit(shelfi_ckks *h, uint64_t bytes) {
    if (!h) return SHELFI_ERR_ARG;
    std::lock_guard<std::mutex> lock(h->mu);
    int rcode = ensure_cuda(h);
    if (rcode) return rcode;
    CU(cudaDeviceSynchronize());   // parked buffers may still be read by queued kernels
    h->pool.trim();
    h->pool.limit = bytes;
    return 0;
}
int shelfi_synchronize(shelfi_ckks *h) {
    if (!h) return SHELFI_ERR_ARG;
    std::lock_guard<std::mutex> lock(h->mu);
    if (!h->cuda_ready) return 0;
    CU(cudaSetDevice(h->device));
    CU(cudaStreamSynchronize(h->stream));
    CU(cudaStreamSynchronize(h->copy_stream));
    if (h->d2h_stream) CU(cudaStreamSynchronize(h->d2h_stream));
    for (cudaStream_t es : h->enc_streams)
        if (es) CU(cudaStreamSynchronize(es));
    return 0;
}

void shelfi_free(void *p) { std::free(p); }
void *shelfi_host_alloc(size_t bytes) {
    void *p = nullptr;
    return cudaHostAlloc(&p, bytes ? bytes : 1, cudaHostAllocDefault) == cudaSuccess ? p : nullptr;
}
void shelfi_host_free(void *p) { if (p) cudaFreeHost(p); }

int shelfi_encrypt_device(shelfi_ckks *h, const double *d_x, uint64_t n, uint64_t ct_index0, uint64_t *d_ct,
                          void *stream) {
    if (!h) return SHELFI_ERR_ARG;
    std::lock_guard<std::mutex> lock(h->mu);
    int rcode = check_stage(h, true, false);
    if (rcode) return rcode;
    if (!d_x || !d_ct || n == 0) return fail(h, SHELFI_ERR_ARG, "null or empty argument");
    return encrypt_dense(h, d_x, n, ct_index0, d_ct, static_cast<cudaStream_t>(stream));
}

int shelfi_encrypt_many_device(shelfi_ckks *h, const double *const *d_x, uint32_t k, uint64_t n, uint64_t ct_index0,
                               uint64_t *const *d_ct, void *stream) {
    if (!h) return SHELFI_ERR_ARG;
    std::lock_guard<std::mutex> lock(h->mu);
    int rcode = check_stage(h, true, false);
    if (rcode) return rcode;
    if (!d_x || !d_ct || n == 0 || k == 0) return fail(h, SHELFI_ERR_ARG, "null or empty argument");
    for (uint32_t i = 0; i < k; i++)
        if (!d_x[i] || !d_ct[i]) return fail(h, SHELFI_ERR_ARG, "null learner buffer");
    return encrypt_dense_many(h, d_x, k, n, ct_index0, d_ct, static_cast<cudaStream_t>(stream));
}

static int weighted_sum_device(shelfi_ckks *h, const uint64_t *const *d_cts, const float *weights, uint32_t k,
                               uint64_t n_ct, uint64_t *d_out, bool accumulate, cudaStream_t s) {
    int rcode = check_stage(h, false, false);
    if (rcode) return rcode;
    if (!d_cts || !weights || !d_out || k == 0 || n_ct == 0) return fail(h, SHELFI_ERR_ARG, "null or empty argument");
    return weighted_sum_dense(h, d_cts, weights, k, n_ct, d_out, accumulate, s);
}

int shelfi_weighted_sum_device(shelfi_ckks *h, const uint64_t *const *d_cts, const float *weights, uint32_t k,
                               uint64_t n_ct, uint64_t *d_out, void *stream) {
    if (!h) return SHELFI_ERR_ARG;
    std::lock_guard<std::mutex> lock(h->mu);
    return weighted_sum_device(h, d_cts, weights, k, n_ct, d_out, false, static_cast<cudaStream_t>(stream));
}

int shelfi_weighted_sum_accumulate_device(shelfi_ckks *h, const uint64_t *const *d_cts, const float *weights, uint32_t k,
                                          uint64_t n_ct, uint64_t *d_acc, void *stream) {
    if (!h) return SHELFI_ERR_ARG;
    std::lock_guard<std::mutex> lock(h->mu);
    return weighted_sum_device(h, d_cts, weights, k, n_ct, d_acc, true, static_cast<cudaStream_t>(stream));
}

int shelfi_decrypt_device(shelfi_ckks *h, const uint64_t *d_ct, uint64_t n_ct, double scaling_factor, double *d_out,
                          uint64_t n, void *stream) {
    if (!h) return SHELFI_ERR_ARG;
    std::lock_guard<std::mutex> lock(h->mu);
    int rcode = check_stage(h, false, true);
    if (rcode) return rcode;
    if (!d_ct || !d_out || n_ct == 0) return fail(h, SHELFI_ERR_ARG, "null or empty argument");
    if (n == 0 || n > n_ct * h->batch || n + h->batch <= n_ct * h->batch)
        return fail(h, SHELFI_ERR_ARG, "n inconsistent with the number of ciphertexts");
    return decrypt_dense(h, reinterpret_cast<const u64 *>(d_ct), n_ct, scaling_factor, d_out, n,
                         static_cast<cudaStream_t>(stream));
}

// Decrypt with the all-gather folded into the decode kernel's stores: value i goes to d_outs[p][offset + i] for every p.
int shelfi_decrypt_gather_device(shelfi_ckks *h, const uint64_t *d_ct, uint64_t n_ct, double scaling_factor,
                                 double *const *d_outs, uint32_t n_outs, uint64_t offset, uint64_t n, void *stream) {
    if (!h) return SHELFI_ERR_ARG;
    std::lock_guard<std::mutex> lock(h->mu);
    int rcode = check_stage(h, false, true);
    if (rcode) return rcode;
    if (!d_ct || !d_outs || n_ct == 0) return fail(h, SHELFI_ERR_ARG, "null or empty argument");
    if (n_outs == 0 || n_outs > (uint32_t)kMaxPeerOuts + 1)
        return fail(h, SHELFI_ERR_ARG, "between 1 and " + std::to_string(kMaxPeerOuts + 1) + " destinations");
    for (uint32_t i = 0; i < n_outs; i++)
        if (!d_outs[i]) return fail(h, SHELFI_ERR_ARG, "null destination");
    if (n == 0 || n > n_ct * h->batch || n + h->batch <= n_ct * h->batch)
        return fail(h, SHELFI_ERR_ARG, "n inconsistent with the number of ciphertexts");
    double *peers[kMaxPeerOuts] = {nullptr};
    for (uint32_t i = 1; i < n_outs; i++) peers[i - 1] = d_outs[i] + offset;
    return decrypt_dense(h, reinterpret_cast<const u64 *>(d_ct), n_ct, scaling_factor, d_outs[0] + offset, n,
                         static_cast<cudaStream_t>(stream), nullptr, 0, peers, n_outs - 1);
}

// Decrypt on the device with the scratch in a file-scope device array (kernels.cu: g_static_scratch) instead of this
// handle's grow-only buffers.  No counterpart in the reference: added for callers
// that may not allocate (stream capture, serving loops, harnesses that forbid cudaMalloc).  Nothing here allocates or
// sets anything up lazily: the handle must already hold its context and secret key on the device, and a request beyond
// the array is refused.  Same results as shelfi_decrypt_device.  Every call on a device shares the one array: calls
// that overlap on the GPU (two streams, two handles, two threads) race on it.
int shelfi_decrypt_static_device(shelfi_ckks *h, const uint64_t *d_ct, uint64_t n_ct, double scaling_factor, double *d_out,
                                 uint64_t n, void *stream) {
    if (!h) return SHELFI_ERR_ARG;
    std::lock_guard<std::mutex> lock(h->mu);
    if (!h->cuda_ready || !h->have_cc) return fail(h, SHELFI_ERR_NOT_LOADED, "crypto context not loaded (call loadCryptoParams)");
    if (!h->have_sk) return fail(h, SHELFI_ERR_NOT_LOADED, "private key not loaded");
    CU(cudaSetDevice(h->device));
    if (h->opt_decode_noise)
        return fail(h, SHELFI_ERR_ARG, "SHELFI_OPT_DECODE_NOISE is not served by the static-scratch decrypt; use shelfi_decrypt_device");
    if (h->batch == 0 || h->batch > h->ctx.batch_size)
        return fail(h, SHELFI_ERR_PARAMS, "batchSize exceeds the slots of the loaded context");
    if (!d_ct || !d_out || n_ct == 0) return fail(h, SHELFI_ERR_ARG, "null or empty argument");
    if (n == 0 || n > n_ct * h->batch || n + h->batch <= n_ct * h->batch)
        return fail(h, SHELFI_ERR_ARG, "n inconsistent with the number of ciphertexts");
    const bool general = general_decode(h);
    const uint64_t per_ct = static_scratch_words(h->rc, h->ec, 1, general);
    if (n_ct > kStaticScratchWords / per_ct)
        return fail(h, SHELFI_ERR_ARG,
                    std::to_string(n_ct) + " ciphertexts need " + std::to_string(n_ct * per_ct * sizeof(uint64_t)) +
                        " bytes of scratch; the static array holds " + std::to_string(kStaticScratchWords * sizeof(uint64_t)) +
                        " (at most " + std::to_string(kStaticScratchWords / per_ct) + " ciphertexts of this context per call)");
    DecodeConst dc{};
    dc.crt = h->crt;
    const double p = (double)h->ctx.plaintext_modulus;
    dc.pre = std::pow(scaling_factor, -1) * std::pow(2, p);   // as decrypt_dense
    dc.pow_p = std::pow(2, -p);
    CU(launch_decrypt_static(h->rc, h->ec, dc, h->crtn, general, reinterpret_cast<const u64 *>(d_ct), h->d_sk.as<tw_t>(),
                             (u32)n_ct, d_out, n, static_cast<cudaStream_t>(stream)));
    return 0;
}

// Host-only self-test of the copy threads (no device needed): `rounds` copies of `bytes` patterned bytes with `threads`
// workers, each verified; returns 0, or the 1-based round that came back wrong.
int shelfi_selftest_parallel_copy(uint64_t bytes, uint32_t threads, uint32_t rounds) {
    if (bytes == 0 || threads == 0 || threads > 64) return SHELFI_ERR_ARG;
    std::vector<uint8_t> src(bytes), dst(bytes);
    ParallelCopier pc(threads);
    for (uint32_t r = 0; r < rounds; r++) {
        for (uint64_t i = 0; i < bytes; i += 4096) src[i] = (uint8_t)(i / 4096 + r);
        src[bytes - 1] = (uint8_t)(r * 7 + 1);
        std::fill(dst.begin(), dst.end(), (uint8_t)0xEE);
        pc.copy(dst.data(), src.data(), bytes);
        if (std::memcmp(dst.data(), src.data(), bytes) != 0) return (int)r + 1;
    }
    return 0;
}

// Stream-ordered barrier between the ranks of one box; d_flags[r] = rank r's flag array (9 zero-initialised u64 in a peer
// buffer).  Every rank calls it with the same, growing epoch.
static int peer_barrier(shelfi_ckks *h, uint64_t *const *d_flags, uint32_t n_ranks, uint32_t my_rank, uint64_t epoch,
                        double *poison, uint64_t poison_stride, void *stream) {
    int rcode = ensure_cuda(h);
    if (rcode) return rcode;
    if (!d_flags || n_ranks == 0 || n_ranks > 8 || my_rank >= n_ranks || epoch == 0)
        return fail(h, SHELFI_ERR_ARG, "1..8 ranks, my_rank below that, epoch from 1");
    PeerFlags pf{};
    for (uint32_t i = 0; i < n_ranks; i++) {
        if (!d_flags[i]) return fail(h, SHELFI_ERR_ARG, "null flag array");
        pf.flags[i] = reinterpret_cast<unsigned long long *>(d_flags[i]);
    }
    uint64_t timeout_ns = 20ull * 1000000000ull;
    if (const char *e = std::getenv("SHELFI_PEER_TIMEOUT_MS")) timeout_ns = (uint64_t)std::max(1, std::atoi(e)) * 1000000ull;
    launch_peer_barrier(pf, n_ranks, my_rank, epoch, timeout_ns, poison, poison_stride, static_cast<cudaStream_t>(stream));
    CU(cudaGetLastError());
    return 0;
}
int shelfi_peer_barrier_device(shelfi_ckks *h, uint64_t *const *d_flags, uint32_t n_ranks, uint32_t my_rank, uint64_t epoch,
                               void *stream) {
    if (!h) return SHELFI_ERR_ARG;
    std::lock_guard<std::mutex> lock(h->mu);
    return peer_barrier(h, d_flags, n_ranks, my_rank, epoch, nullptr, 0, stream);
}
// the same, guarding a gathered result: if the wait for rank r times out, d_values[r * block] (the first value of the
// slice rank r owed this GPU) is set to NaN
int shelfi_peer_barrier_guard_device(shelfi_ckks *h, uint64_t *const *d_flags, uint32_t n_ranks, uint32_t my_rank,
                                     uint64_t epoch, double *d_values, uint64_t block, void *stream) {
    if (!h) return SHELFI_ERR_ARG;
    std::lock_guard<std::mutex> lock(h->mu);
    if (!d_values || block == 0) return fail(h, SHELFI_ERR_ARG, "null result buffer or zero block");
    return peer_barrier(h, d_flags, n_ranks, my_rank, epoch, d_values, block, stream);
}

// Buffers other processes of the box can map (CUDA IPC): the destinations of shelfi_decrypt_gather_device.
int shelfi_peer_buffer_create(shelfi_ckks *h, uint64_t bytes, void **d_ptr, uint8_t handle[SHELFI_PEER_HANDLE_BYTES]) {
    if (!h) return SHELFI_ERR_ARG;
    std::lock_guard<std::mutex> lock(h->mu);
    int rcode = ensure_cuda(h);
    if (rcode) return rcode;
    if (!d_ptr || !handle || bytes == 0) return fail(h, SHELFI_ERR_ARG, "null or empty argument");
    static_assert(sizeof(cudaIpcMemHandle_t) == SHELFI_PEER_HANDLE_BYTES, "handle size");
    void *p = nullptr;
    CU(cudaMalloc(&p, bytes));   // a whole allocation of its own: the handle maps from its first byte
    cudaError_t e = cudaMemset(p, 0, bytes);   // barrier flags start at epoch 0
    if (e == cudaSuccess) e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { cudaFree(p); return cuda_fail(h, e, "cudaMemset"); }
    cudaIpcMemHandle_t ipc;
    e = cudaIpcGetMemHandle(&ipc, p);
    if (e != cudaSuccess) { cudaFree(p); return cuda_fail(h, e, "cudaIpcGetMemHandle"); }
    std::memcpy(handle, &ipc, sizeof ipc);
    *d_ptr = p;
    return 0;
}
int shelfi_peer_buffer_open(shelfi_ckks *h, const uint8_t handle[SHELFI_PEER_HANDLE_BYTES], void **d_ptr) {
    if (!h) return SHELFI_ERR_ARG;
    std::lock_guard<std::mutex> lock(h->mu);
    int rcode = ensure_cuda(h);
    if (rcode) return rcode;
    if (!d_ptr || !handle) return fail(h, SHELFI_ERR_ARG, "null argument");
    cudaIpcMemHandle_t ipc;
    std::memcpy(&ipc, handle, sizeof ipc);
    CU(cudaIpcOpenMemHandle(d_ptr, ipc, cudaIpcMemLazyEnablePeerAccess));
    return 0;
}
int shelfi_peer_buffer_close(shelfi_ckks *h, void *d_ptr) {
    if (!h) return SHELFI_ERR_ARG;
    std::lock_guard<std::mutex> lock(h->mu);
    int rcode = ensure_cuda(h);
    if (rcode) return rcode;
    if (d_ptr) CU(cudaIpcCloseMemHandle(d_ptr));
    return 0;
}
int shelfi_peer_buffer_destroy(shelfi_ckks *h, void *d_ptr) {
    if (!h) return SHELFI_ERR_ARG;
    std::lock_guard<std::mutex> lock(h->mu);
    int rcode = ensure_cuda(h);
    if (rcode) return rcode;
    if (d_ptr) CU(cudaFree(d_ptr));
    return 0;
}

int shelfi_rescale_device(shelfi_ckks *h, const uint64_t *d_ct, uint64_t n_ct, uint32_t limbs_in, uint64_t *d_out, void *stream) {
    if (!h) return SHELFI_ERR_ARG;
    std::lock_guard<std::mutex> lock(h->mu);
    int rcode = check_stage(h, false, false);
    if (rcode) return rcode;
    if (!d_ct || !d_out || n_ct == 0) return fail(h, SHELFI_ERR_ARG, "null or empty argument");
    return rescale_dense(h, reinterpret_cast<const u64 *>(d_ct), n_ct, limbs_in ? limbs_in : h->ctx.limbs,
                         reinterpret_cast<u64 *>(d_out), static_cast<cudaStream_t>(stream));
}
int shelfi_decrypt_level_device(shelfi_ckks *h, const uint64_t *d_ct, uint64_t n_ct, uint32_t limbs, double scaling_factor,
                                double *d_out, uint64_t n, void *stream) {
    if (!h) return SHELFI_ERR_ARG;
    std::lock_guard<std::mutex> lock(h->mu);
    int rcode = check_stage(h, false, true);
    if (rcode) return rcode;
    if (!d_ct || !d_out || n_ct == 0) return fail(h, SHELFI_ERR_ARG, "null or empty argument");
    if (limbs < 1 || limbs > h->ctx.limbs) return fail(h, SHELFI_ERR_ARG, "tower count outside the chain");
    if (n == 0 || n > n_ct * h->batch || n + h->batch <= n_ct * h->batch)
        return fail(h, SHELFI_ERR_ARG, "n inconsistent with the number of ciphertexts");
    return decrypt_dense(h, reinterpret_cast<const u64 *>(d_ct), n_ct, scaling_factor, d_out, n,
                         static_cast<cudaStream_t>(stream), nullptr, limbs);
}
int shelfi_decrypt_words_level_device(shelfi_ckks *h, const uint64_t *d_ct, uint64_t n_ct, uint32_t limbs, uint64_t *d_words,
                                      void *stream) {
    if (!h) return SHELFI_ERR_ARG;
    std::lock_guard<std::mutex> lock(h->mu);
    int rcode = check_stage(h, false, true);
    if (rcode) return rcode;
    if (!d_ct || !d_words || n_ct == 0) return fail(h, SHELFI_ERR_ARG, "null or empty argument");
    if (limbs < 1 || limbs > h->ctx.limbs) return fail(h, SHELFI_ERR_ARG, "tower count outside the chain");
    const LevelView lv = level_view(h, limbs);
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    if ((rcode = scratch_begin(h, s))) return rcode;
    CU(h->s_coeff.reserve(n_ct * limbs * h->ctx.ring_dim * sizeof(uint64_t)));
    launch_decrypt_intt(lv.rc, reinterpret_cast<const u64 *>(d_ct), h->d_sk.as<tw_t>(), h->s_coeff.as<u64>(), (u32)n_ct, s);
    DecodeConst dc{};
    dc.crt = lv.crt;
    launch_crt(lv.rc, dc, lv.crtn, h->s_coeff.as<u64>(), (u32)n_ct, nullptr, reinterpret_cast<u64 *>(d_words), nullptr,
               nullptr, s);
    CU(cudaGetLastError());
    return scratch_end(h, s);
}

// ---- per-slot plaintext weights: out = sum_l ct_l (*) Encode(w_l)  ([PAL] EvalMult(ct, plaintext) + EvalAdd; the
// TenSEAL/FLARE-style aggregation of code/benchmark_nvidia.py:439,461-465) ----
int shelfi_weighted_sum_vector_device(shelfi_ckks *h, const uint64_t *const *d_cts, const double *const *d_w, uint32_t k,
                                      uint64_t n, uint64_t *d_out, void *stream) {
    if (!h) return SHELFI_ERR_ARG;
    std::lock_guard<std::mutex> lock(h->mu);
    int rcode = check_stage(h, false, false);
    if (rcode) return rcode;
    if (!d_cts || !d_w || !d_out || k == 0 || n == 0) return fail(h, SHELFI_ERR_ARG, "null or empty argument");
    for (uint32_t i = 0; i < k; i++)
        if (!d_cts[i] || !d_w[i]) return fail(h, SHELFI_ERR_ARG, "null learner buffer");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const uint64_t n_ct = (n + h->batch - 1) / h->batch, N = h->ctx.ring_dim;
    if (n_ct > (0x7fffffffull / (2 * kMaxLimbs)) >> h->rc.log_blocks)
        return fail(h, SHELFI_ERR_ARG, "too many ciphertexts for one call");
    if ((rcode = scratch_begin(h, s))) return rcode;
    CU(h->s_m.reserve(n_ct * N * sizeof(int64_t)));
    CU(h->s_pt.reserve(n_ct * h->ctx.limbs * N * sizeof(uint64_t)));
    // Consecutive learners that share one weight vector (the TenSEAL script's constant vector) are summed first:
    // sum_l (ct_l (*) pt) == (sum_l ct_l) (*) pt as canonical residues, one encoding and one multiply per run.
    const uint32_t L = h->ctx.limbs;
    const uint64_t ct_words = 2ull * L * N;
    for (uint32_t i0 = 0; i0 < k;) {
        uint32_t i1 = i0 + 1;
        while (i1 < k && d_w[i1] == d_w[i0]) i1++;
        timed(h, kKEncodeSample, s, [&] {
            launch_encode_only(h->rc, h->ec, d_w[i0], n, (u32)n_ct, h->s_m.as<i64>(), s);
            launch_plain_ntt(h->rc, h->s_m.as<i64>(), h->s_pt.as<u64>(), (u32)n_ct, s);
        });
        const u64 *src = reinterpret_cast<const u64 *>(d_cts[i0]);
        if (i1 - i0 > 1) {
            CU(h->s_dense.reserve(n_ct * ct_words * 8));
            for (uint32_t base = i0; base < i1; base += kMaxLearnersPerLaunch) {
                WavgArgs wa{};
                wa.k = std::min<uint32_t>(kMaxLearnersPerLaunch, i1 - base);
                wa.accumulate = base > i0 ? 1u : 0u;
                for (uint32_t j = 0; j < wa.k; j++) {
                    wa.in[j] = reinterpret_cast<const u64 *>(d_cts[base + j]);
                    for (uint32_t l = 0; l < L; l++) wa.c[j][l] = 1;
                }
                timed(h, kKWeightedSum, s, [&] { launch_weighted_sum(h->rc, wa, h->s_dense.as<u64>(), n_ct, s); });
            }
            src = h->s_dense.as<u64>();
        }
        timed(h, kKWeightedSum, s, [&] {
            launch_ct_pt_mac(h->rc, src, h->s_pt.as<u64>(), reinterpret_cast<u64 *>(d_out), n_ct, i0 > 0, s);
        });
        i0 = i1;
    }
    CU(cudaGetLastError());
    return scratch_end(h, s);
}

// ---- selective encryption (benchmark_selection.py:152-193, benchmark_selection_rate.py:133-141): only the chosen
// parameters go through CKKS, the rest is averaged in plaintext on the same device ----
int shelfi_encrypt_selected_many_device(shelfi_ckks *h, const double *const *d_x, uint32_t k, const uint32_t *d_sel,
                                        uint64_t n_sel, uint64_t ct_index0, uint64_t *const *d_ct, void *stream) {
    if (!h) return SHELFI_ERR_ARG;
    std::lock_guard<std::mutex> lock(h->mu);
    int rcode = check_stage(h, true, false);
    if (rcode) return rcode;
    if (!d_x || !d_ct || !d_sel || n_sel == 0 || k == 0) return fail(h, SHELFI_ERR_ARG, "null or empty argument");
    for (uint32_t i = 0; i < k; i++)
        if (!d_x[i] || !d_ct[i]) return fail(h, SHELFI_ERR_ARG, "null learner buffer");
    return encrypt_dense_many(h, d_x, k, n_sel, ct_index0, d_ct, static_cast<cudaStream_t>(stream), d_sel);
}

int shelfi_decrypt_scatter_device(shelfi_ckks *h, const uint64_t *d_ct, uint64_t n_ct, double scaling_factor,
                                  double *d_out, const uint32_t *d_sel, uint64_t n_sel, void *stream) {
    if (!h) return SHELFI_ERR_ARG;
    std::lock_guard<std::mutex> lock(h->mu);
    int rcode = check_stage(h, false, true);
    if (rcode) return rcode;
    if (!d_ct || !d_out || !d_sel || n_ct == 0) return fail(h, SHELFI_ERR_ARG, "null or empty argument");
    if (n_sel == 0 || n_sel > n_ct * h->batch || n_sel + h->batch <= n_ct * h->batch)
        return fail(h, SHELFI_ERR_ARG, "n_sel inconsistent with the number of ciphertexts");
    return decrypt_dense(h, reinterpret_cast<const u64 *>(d_ct), n_ct, scaling_factor, d_out, n_sel,
                         static_cast<cudaStream_t>(stream), d_sel);
}

int shelfi_plain_weighted_sum_device(shelfi_ckks *h, const double *const *d_x, const float *weights, uint32_t k,
                                     uint64_t n, double *d_out, void *stream) {
    if (!h) return SHELFI_ERR_ARG;
    std::lock_guard<std::mutex> lock(h->mu);
    int rcode = ensure_cuda(h);
    if (rcode) return rcode;
    if (!d_x || !weights || !d_out || k == 0 || n == 0) return fail(h, SHELFI_ERR_ARG, "null or empty argument");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    for (uint32_t base = 0; base < k; base += kMaxLearnersPerLaunch) {
        PlainWavgArgs pa{};
        pa.k = std::min<uint32_t>(kMaxLearnersPerLaunch, k - base);
        pa.accumulate = base ? 1u : 0u;
        for (uint32_t i = 0; i < pa.k; i++) {
            if (!d_x[base + i]) return fail(h, SHELFI_ERR_ARG, "null learner buffer");
            pa.in[i] = d_x[base + i];
            pa.w[i] = (double)weights[base + i];   // the same float32 -> double widening as the encrypted part
        }
        launch_plain_weighted_sum(pa, n, d_out, s);
    }
    CU(cudaGetLastError());
    return 0;
}

int shelfi_pack_device(shelfi_ckks *h, const uint64_t *d_dense, uint64_t n_ct, uint8_t *d_image, void *stream) {
    if (!h) return SHELFI_ERR_ARG;
    std::lock_guard<std::mutex> lock(h->mu);
    int rcode = check_stage(h, false, false);
    if (rcode) return rcode;
    std::vector<wire::CiphertextMeta> meta(1);
    wire::BlobPlan plan = wire::plan_ciphertext_blob(h->ctx, n_ct, meta);
    launch_pack(h->rc, to_runs(h, plan.run0, plan.run1, plan.stride), reinterpret_cast<const u64 *>(d_dense),
                reinterpret_cast<u64 *>(d_image), n_ct, static_cast<cudaStream_t>(stream));
    CU(cudaGetLastError());
    return 0;
}
int shelfi_unpack_device(shelfi_ckks *h, const uint8_t *d_image, uint64_t n_ct, uint64_t *d_dense, void *stream) {
    if (!h) return SHELFI_ERR_ARG;
    std::lock_guard<std::mutex> lock(h->mu);
    int rcode = check_stage(h, false, false);
    if (rcode) return rcode;
    std::vector<wire::CiphertextMeta> meta(1);
    wire::BlobPlan plan = wire::plan_ciphertext_blob(h->ctx, n_ct, meta);
    launch_unpack(h->rc, to_runs(h, plan.run0, plan.run1, plan.stride), reinterpret_cast<const u64 *>(d_image),
                  reinterpret_cast<u64 *>(d_dense), n_ct, static_cast<cudaStream_t>(stream));
    CU(cudaGetLastError());
    return 0;
}

int shelfi_ntt_forward_device(shelfi_ckks *h, uint64_t *d_rows, uint32_t rows, void *stream) {
    if (!h) return SHELFI_ERR_ARG;
    std::lock_guard<std::mutex> lock(h->mu);
    int rcode = check_stage(h, false, false);
    if (rcode) return rcode;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const u64 *src = reinterpret_cast<const u64 *>(d_rows);
    if (h->rc.log_blocks > 0) {   // a tile's loader reads its whole row: transform from a copy
        if ((rcode = scratch_begin(h, s))) return rcode;
        size_t bytes = (size_t)rows * h->ctx.ring_dim * sizeof(uint64_t);
        CU(h->s_rows.reserve(bytes));
        CU(cudaMemcpyAsync(h->s_rows.p, d_rows, bytes, cudaMemcpyDeviceToDevice, s));
        src = h->s_rows.as<u64>();
    }
    timed(h, kKNttForward, s, [&] { launch_ntt_forward(h->rc, src, reinterpret_cast<u64 *>(d_rows), rows, s); });
    CU(cudaGetLastError());
    if (h->rc.log_blocks > 0) return scratch_end(h, s);
    return 0;
}
// measurement tap: the forward transform with its last stage done by a lane exchange (ring dimension 8192 only)
int shelfi_ntt_forward_shfl_device(shelfi_ckks *h, uint64_t *d_rows, uint32_t rows, void *stream) {
    if (!h) return SHELFI_ERR_ARG;
    std::lock_guard<std::mutex> lock(h->mu);
    int rcode = check_stage(h, false, false);
    if (rcode) return rcode;
    if (h->rc.log_blocks) return fail(h, SHELFI_ERR_UNSUPPORTED, "the lane-exchange variant covers ring dimension 8192");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    timed(h, kKNttForward, s, [&] {
        launch_ntt_forward_shfl(h->rc, reinterpret_cast<const u64 *>(d_rows), reinterpret_cast<u64 *>(d_rows), rows, s);
    });
    CU(cudaGetLastError());
    return 0;
}
int shelfi_ntt_inverse_device(shelfi_ckks *h, uint64_t *d_rows, uint32_t rows, void *stream) {
    if (!h) return SHELFI_ERR_ARG;
    std::lock_guard<std::mutex> lock(h->mu);
    int rcode = check_stage(h, false, false);
    if (rcode) return rcode;
    timed(h, kKNttInverse, static_cast<cudaStream_t>(stream), [&] {
        launch_ntt_inverse(h->rc, reinterpret_cast<u64 *>(d_rows), rows, static_cast<cudaStream_t>(stream));
    });
    CU(cudaGetLastError());
    return 0;
}
int shelfi_encode_device(shelfi_ckks *h, const double *d_x, uint64_t n, int64_t *d_coeff, void *stream) {
    if (!h) return SHELFI_ERR_ARG;
    std::lock_guard<std::mutex> lock(h->mu);
    int rcode = check_stage(h, false, false);
    if (rcode) return rcode;
    if (!d_x || !d_coeff || n == 0) return fail(h, SHELFI_ERR_ARG, "null or empty argument");
    uint64_t n_ct = (n + h->batch - 1) / h->batch;
    launch_encode_only(h->rc, h->ec, d_x, n, (u32)n_ct, reinterpret_cast<i64 *>(d_coeff), static_cast<cudaStream_t>(stream));
    CU(cudaGetLastError());
    return 0;
}
int shelfi_decrypt_integer_device(shelfi_ckks *h, const uint64_t *d_ct, uint64_t n_ct, int64_t *d_hi, uint64_t *d_lo,
                                  void *stream) {
    if (!h) return SHELFI_ERR_ARG;
    std::lock_guard<std::mutex> lock(h->mu);
    int rcode = check_stage(h, false, true);
    if (rcode) return rcode;
    if (h->ctx.limbs > 2) return fail(h, SHELFI_ERR_UNSUPPORTED, "128-bit tap covers 1 and 2 limbs; use shelfi_decrypt_words_device");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    if ((rcode = scratch_begin(h, s))) return rcode;
    CU(h->s_coeff.reserve(n_ct * h->ctx.limbs * h->ctx.ring_dim * sizeof(uint64_t)));
    launch_decrypt_intt(h->rc, reinterpret_cast<const u64 *>(d_ct), h->d_sk.as<tw_t>(), h->s_coeff.as<u64>(), (u32)n_ct, s);
    DecodeConst dc{};
    dc.crt = h->crt;
    if (!general_decode(h))
        launch_crt_only(h->rc, dc, h->s_coeff.as<u64>(), (u32)n_ct, reinterpret_cast<i64 *>(d_hi), reinterpret_cast<u64 *>(d_lo), s);
    else
        launch_crt(h->rc, dc, h->crtn, h->s_coeff.as<u64>(), (u32)n_ct, nullptr, nullptr, reinterpret_cast<i64 *>(d_hi),
                   reinterpret_cast<u64 *>(d_lo), s);
    CU(cudaGetLastError());
    return scratch_end(h, s);
}

int shelfi_decrypt_words_device(shelfi_ckks *h, const uint64_t *d_ct, uint64_t n_ct, uint64_t *d_words, void *stream) {
    if (!h) return SHELFI_ERR_ARG;
    std::lock_guard<std::mutex> lock(h->mu);
    int rcode = check_stage(h, false, true);
    if (rcode) return rcode;
    if (!d_ct || !d_words || n_ct == 0) return fail(h, SHELFI_ERR_ARG, "null or empty argument");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    if ((rcode = scratch_begin(h, s))) return rcode;
    CU(h->s_coeff.reserve(n_ct * h->ctx.limbs * h->ctx.ring_dim * sizeof(uint64_t)));
    launch_decrypt_intt(h->rc, reinterpret_cast<const u64 *>(d_ct), h->d_sk.as<tw_t>(), h->s_coeff.as<u64>(), (u32)n_ct, s);
    DecodeConst dc{};
    dc.crt = h->crt;
    launch_crt(h->rc, dc, h->crtn, h->s_coeff.as<u64>(), (u32)n_ct, nullptr, reinterpret_cast<u64 *>(d_words), nullptr,
               nullptr, s);
    CU(cudaGetLastError());
    return scratch_end(h, s);
}

}  // extern "C"
