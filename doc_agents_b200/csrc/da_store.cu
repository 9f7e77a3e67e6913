// Host side of libdocagents_b200.so: the device-resident embedding store and the C ABI
// declared in include/docagents_b200.h.  Mirrors, for one path only, what
// internal/store/postgres.go does for the reference: SaveChunks (:142-164, chunk -> document),
// SaveEmbeddings (:176-201, upsert by chunk id) and TopK (:218-285).
//
// There is no CPU fallback anywhere in this file: every search is the CUDA scan, and a
// missing device, a failed allocation or a failed launch is an error to the caller.
#include "../../include/docagents_b200.h"
#include "da_internal.h"

#include <nccl.h>

#include <algorithm>
#include <chrono>
#include <cmath>
#include <condition_variable>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cerrno>
#include <cstring>
#include <new>
#include <limits>
#include <atomic>
#include <memory>
#include <mutex>
#include <shared_mutex>
#include <string>
#include <thread>
#include <unordered_map>
#include <vector>

#include <fcntl.h>
#include <immintrin.h>
#include <pthread.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

namespace {

thread_local std::string g_err;

// DA_TRACE_SEARCH=1: host-side phase times of the general search path on stderr (measurements only)
bool trace_search()
{
    static const bool on = [] { const char *v = getenv("DA_TRACE_SEARCH"); return v && v[0] == '1'; }();
    return on;
}
double now_ms() { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

int fail(int code, const char *fmt, ...) __attribute__((format(printf, 2, 3)));
int fail(int code, const char *fmt, ...)
{
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    g_err = buf;
    return code;
}

#define CU(expr)                                                                                   \
    do {                                                                                           \
        cudaError_t e_ = (expr);                                                                   \
        if (e_ != cudaSuccess)                                                                     \
            return fail(DA_ERR_CUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(e_), __FILE__, __LINE__); \
    } while (0)
#define NC(expr)                                                                                   \
    do {                                                                                           \
        ncclResult_t r_ = (expr);                                                                  \
        if (r_ != ncclSuccess)                                                                     \
            return fail(DA_ERR_NCCL, "%s failed: %s (%s:%d)", #expr, ncclGetErrorString(r_), __FILE__, __LINE__); \
    } while (0)

struct Id16 {
    uint64_t a, b;
    bool operator==(const Id16 &o) const { return a == o.a && b == o.b; }
};
struct Id16Hash {
    size_t operator()(const Id16 &k) const
    {
        uint64_t z = k.a ^ (k.b * 0x9E3779B97F4A7C15ull);
        z ^= z >> 32; z *= 0xD6E8FEB86659FD93ull; z ^= z >> 32;
        return (size_t)z;
    }
};
inline Id16 load_id(const uint8_t *p) { Id16 k; memcpy(&k, p, 16); return k; }

constexpr uint32_t kNoRow = 0xFFFFFFFFu;
constexpr uint32_t kXchgSlotBytes = da::kMaxNQ * DA_MAX_K * 24;      // one rank's records of one pass (<= 8 queries, k <= 128)
constexpr uint64_t kBatchWorthBytes = 256ull << 20;  // searches reading less than this skip the request-batching window (~35 us of HBM time)
constexpr size_t kCheckWhileCopying = 4u << 20;    // query batches above this many bytes are finite-checked while they are staged
constexpr size_t kRecTrailer = 16;      // bytes after a rank's record block in a multi-process store: its overflow flag

struct ChunkInfo { uint32_t slot; int32_t shard; uint32_t row; };
// rows = embeddings resident for this document.  `ranges` = where they are (first local row, count) on the document's
// shard: one SaveEmbeddings call carries all chunks of one document (cmd/analysis/main.go:98-106), so a document is
// one run of rows, and a filtered search can name its rows by a handful of ranges instead of testing every row of
// the store.  `fragmented`: the ranges are not kept (too many pieces, a chunk moved between documents, rows spread
// over several shards) and the document is found through the per-row document slots.
constexpr size_t kMaxDocRanges = 8;
struct DocInfo {
    Id16 id; int32_t shard; uint32_t rows;
    bool fragmented = false;
    std::vector<std::pair<uint32_t, uint32_t>> ranges;
    void add_row(uint32_t row)
    {
        if (fragmented) return;
        if (!ranges.empty() && ranges.back().first + ranges.back().second == row) { ranges.back().second++; return; }
        if (ranges.size() >= kMaxDocRanges) { fragmented = true; ranges.clear(); ranges.shrink_to_fit(); return; }
        ranges.emplace_back(row, 1u);
    }
    void add_run(uint32_t first, uint32_t count)
    {
        if (fragmented || count == 0) return;
        if (!ranges.empty() && ranges.back().first + ranges.back().second == first) { ranges.back().second += count; return; }
        if (ranges.size() >= kMaxDocRanges) { fragment(); return; }
        ranges.emplace_back(first, count);
    }
    void fragment() { fragmented = true; ranges.clear(); ranges.shrink_to_fit(); }
};

template <typename T>
struct DevBuf {   // grow-only device buffer
    T *p = nullptr; size_t cap = 0;
    cudaError_t ensure(size_t n)
    {
        if (n <= cap) return cudaSuccess;
        if (p) cudaFree(p);
        p = nullptr; cap = 0;
        cudaError_t e = cudaMalloc((void **)&p, n * sizeof(T));
        if (e == cudaSuccess) cap = n;
        return e;
    }
    void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
};
template <typename T>
struct PinBuf {   // grow-only pinned host buffer
    T *p = nullptr; size_t cap = 0;
    cudaError_t ensure(size_t n)
    {
        if (n <= cap) return cudaSuccess;
        if (p) cudaFreeHost(p);
        p = nullptr; cap = 0;
        cudaError_t e = cudaMallocHost((void **)&p, n * sizeof(T));
        if (e == cudaSuccess) cap = n;
        return e;
    }
    void release() { if (p) cudaFreeHost(p); p = nullptr; cap = 0; }
};

struct Shard {
    int device = 0, sm_count = 148;
    bool ipc_mapped = false;                    // mat / inv_norm / doc_slot / chunk_ids are another process's memory (da_store_attach)
    uint64_t cap_rows = 0;
    uint32_t n_rows = 0, row_base = 0;
    float *mat = nullptr, *inv_norm = nullptr;
    uint16_t *hi = nullptr, *lo = nullptr;      // split-plane layout: upper / lower 16 bits of every element
    int8_t *q8 = nullptr;                       // int8 sidecar (DA_STORE_Q8_SIDECAR): rint(x / scale) of every element
    float *q8_factor = nullptr, *q8_err = nullptr;   // ... per row: scale / |row|, and its quantisation error bound
    float *q8_errmax = nullptr;                 // ... device: the largest bound of any row written so far (bits ordered as u32)
    float q8_err_max = 0.f;                     // ... its host copy, refreshed before a search when rows were written
    bool q8_dirty = false;
    uint32_t *doc_slot = nullptr;
    uint8_t *chunk_ids = nullptr;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev_scan0 = nullptr, ev_scan1 = nullptr, ev_all0 = nullptr, ev_all1 = nullptr;
    std::vector<cudaEvent_t> ev_pool;   // extra pairs for per-launch timing inside a bench loop
    cudaEvent_t ev_done = nullptr;      // peer-gather mode: this shard's records have been written
    ncclComm_t comm = nullptr;
    // fused cross-GPU exchange (one process per GPU): own buffer, every rank's buffer as mapped here (CUDA IPC)
    unsigned char *xg_buf = nullptr;
    unsigned char *xg_peer[da::kXchgMaxRanks] = {};
    uint32_t *h_xg_timeout = nullptr, *d_xg_timeout = nullptr;   // in-process exchange: mapped word a kernel sets if a shard never delivered
    // search state
    DevBuf<float> d_queries, d_qinv;
    DevBuf<uint32_t> d_bitmap;
    DevBuf<uint8_t> d_out, d_gather, d_final;
    DevBuf<unsigned long long> d_cand;          // batched tensor-core path: candidate lists
    DevBuf<uint32_t> d_cand_count, d_overflow, d_hist, d_gate, d_surv_count, d_group_sync;
    DevBuf<unsigned long long> d_surv;
    DevBuf<float> d_approx;                     // split-plane layout: approximate scores of every row
    DevBuf<uint16_t> d_qhi;                     // ... and the queries' upper halves for the bf16 contraction
    DevBuf<unsigned long long> d_akeys;         // ... and the scan's approximate top-k
    DevBuf<uint32_t> d_flt_rows, d_flt_n;       // narrow document filter: the requested rows, compacted, and their count
    DevBuf<unsigned long long> d_probe_keys;    // IVFFlat: the lists each query of the batch probes (key row = list id)
    DevBuf<unsigned long long> d_ivf_cta;       // ... per-query CTA candidates of the probe-mode scan
    DevBuf<unsigned int> d_ivf_counter;         // ... per-query tickets (left at zero by the kernel)
    DevBuf<float> d_ivf_scores;                 // ... partial query x centroid scores of the tensor-core ranking
    DevBuf<int8_t> d_q8q;                       // int8 sidecar: quantised queries, their factors and error bounds
    DevBuf<float> d_q8qf, d_q8qe;
    PinBuf<uint32_t> h_overflow;
    da::key_t64 *d_cta = nullptr;
    unsigned int *d_counter = nullptr;
    unsigned long long *d_rows_read = nullptr;
    PinBuf<unsigned long long> h_rows_read;
    PinBuf<float> h_queries;
    PinBuf<uint32_t> h_bitmap;
    PinBuf<uint8_t> h_result;
    // upsert staging
    DevBuf<uint32_t> d_up_rows, d_up_slots;
    DevBuf<uint8_t> d_up_ids;
    PinBuf<uint32_t> h_up_rows, h_up_slots;
    PinBuf<uint8_t> h_up_ids;
    DevBuf<float> d_up_stage;                 // vectors of one upsert chunk, validated before they are committed
    DevBuf<unsigned long long> d_up_bad;
    PinBuf<unsigned long long> h_up_bad;
    uint64_t hbm_bytes = 0;
    // IVFFlat index over this shard's rows (centroids are identical on every shard)
    struct Ivf {
        float *cent = nullptr, *cent_inv = nullptr, *sums = nullptr, *counts_f = nullptr;
        uint32_t *assign = nullptr, *offsets = nullptr, *cursor = nullptr, *counts = nullptr;
        uint32_t *list_rows = nullptr;
        uint32_t *sample = nullptr, *sample_assign = nullptr;
        void *split_scratch = nullptr;            // sort keys and (donor, receiver) pairs of the balancing rounds
        float *sample_mat = nullptr;              // the k-means sample, gathered once (tensor-core path)
        unsigned long long *packed = nullptr;     // per-row packed (score, list) for the tensor-core argmax
        uint32_t n_sample = 0;
        uint64_t bytes = 0;
    } ivf;
};

void free_ivf(Shard &s)
{
    Shard::Ivf &v = s.ivf;
    cudaFree(v.cent); cudaFree(v.cent_inv); cudaFree(v.sums); cudaFree(v.counts_f);
    cudaFree(v.assign); cudaFree(v.offsets); cudaFree(v.cursor); cudaFree(v.counts);
    cudaFree(v.list_rows); cudaFree(v.split_scratch);
    cudaFree(v.sample); cudaFree(v.sample_assign);
    cudaFree(v.sample_mat); cudaFree(v.packed);
    s.hbm_bytes -= v.bytes;
    v = Shard::Ivf();
}

// ---------------------------------------------------------------------------
// Sharing one device store between processes.  The writer of the reference (analysis agent,
// cmd/analysis/main.go:106) and its reader (query agent, cmd/query/main.go:101) are different processes
// (docker-compose.yml:87-128).  The writer exports its store: CUDA IPC handles of the four resident arrays plus,
// in POSIX shared memory, a log of what it has stored (documents, runs of rows) and a process-shared rwlock.  A
// reader attaches: it maps the SAME HBM, replays the log into its own id maps and row ranges, and before every
// search replays what was appended since -- so a document saved by the writer is found by the reader's next TopK.
// ---------------------------------------------------------------------------
constexpr uint64_t kShmMagic = 0x4441495043763031ull;          // "DAIPCv01"
enum : uint32_t { kRecDoc = 1, kRecRows = 2, kRecFragment = 3 };
struct ShmRecord {                 // 32 bytes
    uint32_t type, slot, first, count;
    uint8_t id[16];                // kRecDoc: the document id of `slot`
};
struct ShmHeader {
    uint64_t magic;
    uint32_t version, dim;
    uint64_t capacity_rows;
    float min_sim;
    uint32_t row_base;
    char device_bus_id[32];        // the writer's GPU: a reader must open the handles on the same one
    cudaIpcMemHandle_t h_mat, h_inv, h_slot, h_ids;
    pthread_rwlock_t lock;         // writer: while rows or log change; readers: while a search runs
    uint64_t max_records;
    std::atomic<uint64_t> n_records;   // published with release order after the records and the rows are in place
    std::atomic<uint64_t> n_rows;
    std::atomic<uint32_t> overflow;    // the log ran out of room: readers can no longer follow (export again with a larger store)
};
struct Shared {
    ShmHeader *hdr = nullptr;
    ShmRecord *rec = nullptr;
    size_t bytes = 0;
    std::string name;
    bool owner = false;            // the writer created it (and unlinks it)
    uint64_t pending = 0;          // writer: records written, published at the end of the operation
    std::atomic<uint64_t> replayed{0};   // reader: records already applied to its maps
};

// One small search in flight: its own stream, request block (query + row ranges, one host-to-device copy), scan
// scratch, and a result block in MAPPED host memory that the scan's last CTA writes directly, followed by a
// completion word the caller polls -- no device-to-host copy, no stream synchronisation.  Concurrent TopK calls
// (one goroutine per /api/query: cmd/query/main.go:34) each take a slot, so they overlap instead of queueing.
struct Slot {
    cudaStream_t stream = nullptr;
    uint8_t *h_req = nullptr, *d_req = nullptr;          // [dim floats][kMaxRanges x (first row, count) | document bitmap]
    size_t req_cap = 0;
    da::key_t64 *d_probe_keys = nullptr;                 // IVFFlat: the lists the query probes
    float *d_scores = nullptr; size_t scores_cap = 0;    // ... and its score against every centroid
    uint8_t *h_out = nullptr, *d_out = nullptr;          // mapped: [k ids of 16 B][k keys][rows read][completion word]
    da::key_t64 *d_cta = nullptr;
    unsigned int *d_counter = nullptr;
    unsigned long long *d_rows_read = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;            // resident timing only
    unsigned long long ticket = 0;
};
constexpr int kMaxRanges = DA_IVF_MAX_PROBES;            // the scan's range table
constexpr int kMaxSlots = 8;
constexpr size_t kSlotOutBytes = (size_t)DA_MAX_K * 24 + 32;    // records, then: exchange-timeout word, rows read, completion word

}  // namespace

struct da_store {
    std::shared_mutex mu;          // searches of the small-search path share it; everything else takes it alone
    int dim = 0;
    float min_sim = DA_DEFAULT_MIN_SIMILARITY;
    int world = 1, rank = 0, total_ranks = 1;
    std::vector<Shard> shards;
    std::unordered_map<Id16, uint32_t, Id16Hash> doc_slot_of;
    std::vector<DocInfo> docs;                 // slot -> document
    std::unordered_map<Id16, ChunkInfo, Id16Hash> chunks;
    // statistics
    float last_scan_ms = 0, last_search_ms = 0;
    uint64_t last_rows_read = 0, last_bytes = 0, launches = 0, scan_launches = 0;
    double scan_ms_sum = 0;
    int scan_grid = 0;
    int ivf_lists = 0, ivf_default_probes = 1;
    uint64_t tc_overflows = 0;
    // request batching for da_store_topk
    struct Pending {
        const uint8_t *doc_ids; size_t n_docs; const float *query; int k;
        uint8_t *out_ids; float *out_scores; int32_t *out_n;
        int rc = 0; std::string err; bool done = false;
    };
    uint32_t batch_window_us = 0, batch_max = 64;
    std::mutex bmu;
    std::condition_variable bcv;
    std::vector<Pending *> bqueue;
    bool bleader = false;
    uint64_t topk_calls = 0, topk_passes = 0;
    bool split16 = false;          // split-plane resident format (DA_STORE_SPLIT16)
    bool q8 = false;               // int8 sidecar next to the fp32 rows (DA_STORE_Q8_SIDECAR)
    bool p2p_gather = false;       // one process, several GPUs with peer access: records go straight to shard 0
    bool xg_inproc = false;        // one process, several GPUs, all-pairs peer access: the same exchange between this store's own shards
    bool xg_ok = false;            // one process per GPU, every peer's exchange buffer mapped: the scan exchanges and merges itself
    uint64_t xg_seq = 0;           // fused searches so far (the same on every rank: searches are collective)
    std::mutex xg_mu;              // small-search path on a multi-process store: one collective search at a time
    // small-search path (single shard, single process): slots and the statistics they update
    std::mutex slot_mu;
    std::condition_variable slot_cv;
    std::vector<std::unique_ptr<Slot>> slots;
    std::vector<Slot *> free_slots;
    std::mutex stats_mu;
    uint64_t fast_searches = 0;
    Shared shm;                    // exported (writer) or attached (reader) state; hdr == nullptr: a private store
    bool attached = false;         // reader of another process's store: read-only, device arrays are IPC mappings
    const unsigned long long *last_probe_keys = nullptr;   // device: where the last IVFFlat search left its probe lists
    size_t last_probe_count = 0;
    uint64_t bench_launches = 0;   // kernels launched inside the timed region of the last da_store_bench_resident
    float ivf_build_ms = 0, ivf_kmeans_ms = 0, ivf_assign_ms = 0, ivf_assign_tflops = 0;
};

namespace {

void free_shard(Shard &s)
{
    cudaSetDevice(s.device);
    free_ivf(s);
    if (!s.h_xg_timeout)            // (in-process: the peers' buffers are the other shards' own allocations, nothing to close)
        for (int r = 0; r < da::kXchgMaxRanks; r++)
            if (s.xg_peer[r] && s.xg_peer[r] != s.xg_buf) cudaIpcCloseMemHandle(s.xg_peer[r]);
    cudaFree(s.xg_buf);
    if (s.h_xg_timeout) cudaFreeHost(s.h_xg_timeout);
    if (s.comm) ncclCommDestroy(s.comm);
    if (s.stream) cudaStreamSynchronize(s.stream);
    if (s.ipc_mapped) {
        cudaIpcCloseMemHandle(s.mat); cudaIpcCloseMemHandle(s.inv_norm); cudaIpcCloseMemHandle(s.doc_slot); cudaIpcCloseMemHandle(s.chunk_ids);
        s.mat = nullptr; s.inv_norm = nullptr; s.doc_slot = nullptr; s.chunk_ids = nullptr;
    }
    cudaFree(s.mat); cudaFree(s.hi); cudaFree(s.lo); cudaFree(s.q8); cudaFree(s.q8_factor); cudaFree(s.q8_err); cudaFree(s.q8_errmax); cudaFree(s.inv_norm); cudaFree(s.doc_slot); cudaFree(s.chunk_ids);
    cudaFree(s.d_cta); cudaFree(s.d_counter); cudaFree(s.d_rows_read);
    s.d_queries.release(); s.d_qinv.release(); s.d_bitmap.release();
    s.d_out.release(); s.d_gather.release(); s.d_final.release();
    s.d_cand.release(); s.d_cand_count.release(); s.d_overflow.release(); s.h_overflow.release();
    s.d_approx.release(); s.d_akeys.release(); s.d_qhi.release(); s.d_probe_keys.release(); s.d_ivf_cta.release(); s.d_ivf_counter.release(); s.d_ivf_scores.release(); s.d_q8q.release(); s.d_q8qf.release(); s.d_q8qe.release(); s.d_flt_rows.release(); s.d_flt_n.release();
    s.d_hist.release(); s.d_gate.release(); s.d_surv_count.release(); s.d_surv.release(); s.d_group_sync.release();
    s.h_queries.release(); s.h_bitmap.release(); s.h_result.release(); s.h_rows_read.release();
    s.d_up_rows.release(); s.d_up_slots.release(); s.d_up_ids.release();
    s.h_up_rows.release(); s.h_up_slots.release(); s.h_up_ids.release();
    s.d_up_stage.release(); s.d_up_bad.release(); s.h_up_bad.release();
    if (s.ev_scan0) cudaEventDestroy(s.ev_scan0);
    if (s.ev_scan1) cudaEventDestroy(s.ev_scan1);
    if (s.ev_all0) cudaEventDestroy(s.ev_all0);
    if (s.ev_all1) cudaEventDestroy(s.ev_all1);
    for (cudaEvent_t e : s.ev_pool) cudaEventDestroy(e);
    if (s.ev_done) cudaEventDestroy(s.ev_done);
    if (s.stream) cudaStreamDestroy(s.stream);
    s = Shard();
}

// ---- shared-store plumbing (see "Sharing one device store between processes") ----
struct ShmWriteGuard {      // the exporting process takes the lock alone while rows or the log change
    ShmHeader *h;
    explicit ShmWriteGuard(da_store *st) : h(st->shm.hdr && st->shm.owner ? st->shm.hdr : nullptr) { if (h) pthread_rwlock_wrlock(&h->lock); }
    ~ShmWriteGuard() { if (h) pthread_rwlock_unlock(&h->lock); }
};
struct ShmReadGuard {       // an attached process holds it shared while it replays the log and searches
    ShmHeader *h;
    explicit ShmReadGuard(da_store *st) : h(st->attached && st->shm.hdr ? st->shm.hdr : nullptr) { if (h) pthread_rwlock_rdlock(&h->lock); }
    ~ShmReadGuard() { if (h) pthread_rwlock_unlock(&h->lock); }
};

// Writer: one more record (kept private until shm_publish).  A full log is remembered; readers then fail loudly.
void shm_append(da_store *st, uint32_t type, uint32_t slot, uint32_t first, uint32_t count, const Id16 *id = nullptr)
{
    Shared &sh = st->shm;
    if (!sh.hdr || !sh.owner) return;
    if (sh.pending >= sh.hdr->max_records) { sh.hdr->overflow.store(1u, std::memory_order_release); return; }
    ShmRecord &r = sh.rec[sh.pending++];
    r.type = type; r.slot = slot; r.first = first; r.count = count;
    if (id) memcpy(r.id, id, 16); else memset(r.id, 0, 16);
}

// Writer: make everything appended so far (and the rows it describes, already complete on the device) visible.
void shm_publish(da_store *st)
{
    Shared &sh = st->shm;
    if (!sh.hdr || !sh.owner) return;
    sh.hdr->n_rows.store(st->shards[0].n_rows, std::memory_order_release);
    sh.hdr->n_records.store(sh.pending, std::memory_order_release);
}

// Reader: apply the records appended since the last look.  Called with st->mu held alone.
int shm_replay(da_store *st)
{
    Shared &sh = st->shm;
    if (sh.hdr->overflow.load(std::memory_order_acquire))
        return fail(DA_ERR_STATE, "the exporting process's log is full; this reader can no longer follow it");
    const uint64_t n = sh.hdr->n_records.load(std::memory_order_acquire);
    for (uint64_t i = sh.replayed.load(std::memory_order_relaxed); i < n; i++) {
        const ShmRecord &r = sh.rec[i];
        if (r.type == kRecDoc) {
            if (r.slot != st->docs.size()) return fail(DA_ERR_STATE, "shared log out of order (document slot %u)", r.slot);
            const Id16 id = load_id(r.id);
            st->docs.push_back(DocInfo{id, 0, 0, false, {}});
            st->doc_slot_of.emplace(id, r.slot);
        } else if (r.slot >= st->docs.size()) {
            return fail(DA_ERR_STATE, "shared log names an unknown document slot %u", r.slot);
        } else if (r.type == kRecRows) {
            st->docs[r.slot].rows += r.count;
            st->docs[r.slot].add_run(r.first, r.count);
        } else if (r.type == kRecFragment) {
            st->docs[r.slot].fragment();
            st->docs[r.slot].rows = r.count;
        }
    }
    sh.replayed.store(n, std::memory_order_release);
    st->shards[0].n_rows = (uint32_t)sh.hdr->n_rows.load(std::memory_order_acquire);
    return DA_OK;
}

// Reader: before a search.  Cheap when nothing was appended.
int shm_refresh(da_store *st)
{
    if (!st->attached) return DA_OK;
    if (st->shm.hdr->n_records.load(std::memory_order_acquire) == st->shm.replayed.load(std::memory_order_acquire) &&
        !st->shm.hdr->overflow.load(std::memory_order_acquire))
        return DA_OK;
    std::unique_lock<std::shared_mutex> lock(st->mu);
    return shm_replay(st);
}

void shm_close(da_store *st)
{
    Shared &sh = st->shm;
    if (!sh.hdr) return;
    munmap(sh.hdr, sh.bytes);
    if (sh.owner) shm_unlink(sh.name.c_str());
    sh.hdr = nullptr; sh.rec = nullptr;
}

void free_slot(Slot &sl)
{
    if (sl.stream) cudaStreamSynchronize(sl.stream);
    cudaFree(sl.d_req); cudaFree(sl.d_cta); cudaFree(sl.d_counter); cudaFree(sl.d_rows_read); cudaFree(sl.d_probe_keys); cudaFree(sl.d_scores);
    if (sl.h_req) cudaFreeHost(sl.h_req);
    if (sl.h_out) cudaFreeHost(sl.h_out);
    if (sl.ev0) cudaEventDestroy(sl.ev0);
    if (sl.ev1) cudaEventDestroy(sl.ev1);
    if (sl.stream) cudaStreamDestroy(sl.stream);
    sl = Slot();
}

// Called with the shard's device current.
int make_slot(da_store *st, Slot &sl)
{
    const Shard &s = st->shards[0];
    const size_t req = (size_t)st->dim * 4 + (size_t)kMaxRanges * 8;
    CU(cudaStreamCreateWithFlags(&sl.stream, cudaStreamNonBlocking));
    CU(cudaMallocHost((void **)&sl.h_req, req));
    CU(cudaMalloc((void **)&sl.d_req, req));
    sl.req_cap = req;
    CU(cudaMalloc((void **)&sl.d_probe_keys, (size_t)DA_IVF_MAX_PROBES * sizeof(da::key_t64)));
    CU(cudaHostAlloc((void **)&sl.h_out, kSlotOutBytes, cudaHostAllocMapped));
    CU(cudaHostGetDevicePointer((void **)&sl.d_out, sl.h_out, 0));
    memset(sl.h_out, 0, kSlotOutBytes);
    CU(cudaMalloc((void **)&sl.d_cta, (size_t)da::scan_grid_for(s.sm_count) * DA_MAX_K * sizeof(da::key_t64)));
    CU(cudaMalloc((void **)&sl.d_counter, sizeof(unsigned int)));
    CU(cudaMalloc((void **)&sl.d_rows_read, sizeof(unsigned long long)));
    CU(cudaMemsetAsync(sl.d_counter, 0, sizeof(unsigned int), sl.stream));
    CU(cudaMemsetAsync(sl.d_rows_read, 0, sizeof(unsigned long long), sl.stream));
    CU(cudaEventCreate(&sl.ev0)); CU(cudaEventCreate(&sl.ev1));
    CU(cudaStreamSynchronize(sl.stream));
    return DA_OK;
}

// The request block grows with the store's document count (an IVFFlat search with a document filter carries a bitmap).
int slot_reserve_request(Slot &sl, size_t bytes)
{
    if (bytes <= sl.req_cap) return DA_OK;
    CU(cudaStreamSynchronize(sl.stream));
    cudaFree(sl.d_req); sl.d_req = nullptr;
    if (sl.h_req) cudaFreeHost(sl.h_req);
    sl.h_req = nullptr; sl.req_cap = 0;
    bytes += bytes / 2;
    CU(cudaMallocHost((void **)&sl.h_req, bytes));
    CU(cudaMalloc((void **)&sl.d_req, bytes));
    sl.req_cap = bytes;
    return DA_OK;
}

// A free slot, a new one while fewer than kMaxSlots exist, else the next one to be released.
Slot *acquire_slot(da_store *st, int *rc)
{
    std::unique_lock<std::mutex> lk(st->slot_mu);
    for (;;) {
        if (!st->free_slots.empty()) {
            Slot *sl = st->free_slots.back();
            st->free_slots.pop_back();
            return sl;
        }
        if ((int)st->slots.size() < kMaxSlots) {
            std::unique_ptr<Slot> sl(new Slot());
            *rc = make_slot(st, *sl);
            if (*rc != DA_OK) { free_slot(*sl); return nullptr; }
            st->slots.push_back(std::move(sl));
            return st->slots.back().get();
        }
        st->slot_cv.wait(lk);
    }
}

void release_slot(da_store *st, Slot *sl)
{
    { std::lock_guard<std::mutex> lk(st->slot_mu); st->free_slots.push_back(sl); }
    st->slot_cv.notify_one();
}

int init_shard(Shard &s, int device, int dim, uint64_t cap_rows, bool split16, bool q8, bool attach_only = false)
{
    s.device = device;
    s.cap_rows = cap_rows;
    s.ipc_mapped = attach_only;
    CU(cudaSetDevice(device));
    cudaDeviceProp prop;
    CU(cudaGetDeviceProperties(&prop, device));
    if (prop.major < 10)
        return fail(DA_ERR_CUDA, "device %d is sm_%d%d; this library is built for sm_100a only", device, prop.major, prop.minor);
    s.sm_count = prop.multiProcessorCount;
    size_t free_b = 0, total_b = 0;
    CU(cudaMemGetInfo(&free_b, &total_b));
    const uint64_t need = attach_only ? 0 : cap_rows * ((uint64_t)dim * 4 + 4 + 4 + 16 + (q8 ? (uint64_t)dim + 8 : 0));
    if (!attach_only && need + (512ull << 20) > free_b)
        return fail(DA_ERR_CAPACITY, "capacity_rows=%llu x dim=%d needs %.2f GB on device %d, %.2f GB free (no CPU fallback)",
                    (unsigned long long)cap_rows, dim, need / 1e9, device, free_b / 1e9);
    CU(cudaStreamCreateWithFlags(&s.stream, cudaStreamNonBlocking));
    CU(cudaEventCreate(&s.ev_scan0)); CU(cudaEventCreate(&s.ev_scan1));
    CU(cudaEventCreate(&s.ev_all0)); CU(cudaEventCreate(&s.ev_all1));
    CU(cudaEventCreateWithFlags(&s.ev_done, cudaEventDisableTiming));
    if (attach_only) {
        // the resident arrays come from the exporting process (cudaIpcOpenMemHandle, by the caller)
    } else if (split16) {
        CU(cudaMalloc((void **)&s.hi, cap_rows * (uint64_t)dim * 2));
        CU(cudaMalloc((void **)&s.lo, cap_rows * (uint64_t)dim * 2));
    } else {
        CU(cudaMalloc((void **)&s.mat, cap_rows * (uint64_t)dim * 4));
    }
    if (q8) {
        CU(cudaMalloc((void **)&s.q8, cap_rows * (uint64_t)dim));
        CU(cudaMalloc((void **)&s.q8_factor, cap_rows * 4));
        CU(cudaMalloc((void **)&s.q8_err, cap_rows * 4));
        CU(cudaMalloc((void **)&s.q8_errmax, 4));
        CU(cudaMemset(s.q8_errmax, 0, 4));
    }
    if (!attach_only) {
        CU(cudaMalloc((void **)&s.inv_norm, cap_rows * 4));
        CU(cudaMalloc((void **)&s.doc_slot, cap_rows * 4));
        CU(cudaMalloc((void **)&s.chunk_ids, cap_rows * 16));
    }
    const size_t cta_keys = (size_t)da::scan_grid_for(s.sm_count) * da::kMaxNQ * DA_MAX_K;
    CU(cudaMalloc((void **)&s.d_cta, cta_keys * sizeof(da::key_t64)));
    CU(cudaMalloc((void **)&s.d_counter, sizeof(unsigned int)));
    CU(cudaMalloc((void **)&s.d_rows_read, sizeof(unsigned long long)));
    CU(s.h_rows_read.ensure(1));
    CU(cudaMemsetAsync(s.d_counter, 0, sizeof(unsigned int), s.stream));
    CU(cudaMemsetAsync(s.d_rows_read, 0, sizeof(unsigned long long), s.stream));
    CU(cudaStreamSynchronize(s.stream));
    s.hbm_bytes = need + cta_keys * sizeof(da::key_t64);
    return DA_OK;
}

struct SearchPlan {
    int batch, k;
    float min_sim;
    bool all_docs, read_all;
    size_t bitmap_words;
    bool ivf;
    int probes;
    bool fused_xchg;   // several processes: the scan's last CTA exchanges the records with the other GPUs and merges (no collective)
    bool need_qinv;    // some kernel of the plan reads 1/|q| from memory (candidate paths); the plain scans compute it themselves
    bool ivf_select;   // IVFFlat: fp32 centroid scores + the selection kernel (small batches; lists fit the selection kernel)
    bool ivf_tc;       // IVFFlat: rank the centroids for the whole batch with one tensor-core GEMM (batches of 8+)
    int ivf_ks;        // ... slices of the dimension of that GEMM
    bool use_tc;       // batched queries through the tcgen05 candidate kernel + exact re-score
    bool use_q8;       // int8 sidecar scan for the candidates + exact re-score from the fp32 rows
    bool tc_i8;        // tensor path over the int8 sidecar (kind::i8) instead of the fp32 rows (TF32)
    bool compact;      // narrow document filter: compact the requested rows first, then scan that list
    bool none_requested;   // the filter names no document this store knows: zero rows, as `= ANY('{}')` (postgres.go:237-239)
    uint32_t bitmap_stride;   // words between per-query bitmaps; 0 = one document set for the batch
};

// Enqueue one complete search on every local shard: scan passes, collective, merge.
// Queries, query norms and bitmap are already on the devices.
int enqueue_search(da_store *st, const SearchPlan &pl, cudaEvent_t e0 = nullptr, cudaEvent_t e1 = nullptr)
{
    const int dim = st->dim;
    const size_t rec_bytes = (size_t)pl.batch * pl.k * 24;
    const size_t trailer = st->world > 1 ? kRecTrailer : 0;     // per-rank flags travel with the records (below)
    for (Shard &s : st->shards) {
        CU(cudaSetDevice(s.device));
        const bool first = (&s == &st->shards[0]);
        CU(cudaEventRecord(first && e0 ? e0 : s.ev_scan0, s.stream));
        // record block: [batch*k ids of 16 B][batch*k keys of 8 B]  (ids first keeps them 16-byte aligned).
        // Peer-gather mode: the scan's last CTA stores the block straight into shard 0's gather buffer over
        // NVLink -- compute and "all-to-one gather" are the same kernel, no collective is launched.
        const size_t shard_index = (size_t)(&s - &st->shards[0]);
        uint8_t *rec_base = st->p2p_gather ? st->shards[0].d_gather.p + shard_index * rec_bytes : s.d_out.p;
        if (st->p2p_gather && !first && !pl.fused_xchg)      // block `shard_index` is free again once shard 0 merged the previous search
            CU(cudaStreamWaitEvent(s.stream, st->shards[0].ev_done, 0));
        uint8_t *out_ids = rec_base;
        da::key_t64 *out_keys = reinterpret_cast<da::key_t64 *>(rec_base + (size_t)pl.batch * pl.k * 16);
        if ((s.n_rows == 0 || pl.none_requested) && !pl.fused_xchg) {   // an empty shard, or no known document requested: an empty record block
            CU(cudaMemsetAsync(rec_base, 0, rec_bytes, s.stream));
            CU(cudaEventRecord(first && e1 ? e1 : s.ev_scan1, s.stream));
            if (st->p2p_gather && !first) CU(cudaEventRecord(s.ev_done, s.stream));
            continue;
        }
        if (pl.ivf) {
            // (1) rank the centroids for every query of the batch: one tcgen05 GEMM (queries x centroids^T) + a
            //     per-query selection of the `probes` best lists, or -- small batches -- the fused SIMT scan over the
            //     lists x dim centroid matrix (up to 8 queries per pass, top `probes`);
            // (2) ONE probe-mode scan for the whole batch: blockIdx.y = query, each query walks its own lists.
            const float ninf = -std::numeric_limits<float>::infinity();
            const int L = st->ivf_lists;
            da::key_t64 *probe_keys = reinterpret_cast<da::key_t64 *>(s.d_probe_keys.p);
            if (first) { st->last_probe_keys = s.d_probe_keys.p; st->last_probe_count = (size_t)pl.batch * pl.probes; }
            if (pl.ivf_tc) {
                CU(da::tc_launch_scores(s.d_queries.p, pl.batch, s.ivf.cent, L, dim, pl.ivf_ks, s.d_ivf_scores.p, s.stream));
                CU(da::ivf_launch_select_probes(s.d_ivf_scores.p, pl.ivf_ks, pl.batch, L, pl.probes, probe_keys, s.stream));
                st->launches += 2;
            } else if (pl.ivf_select) {
                // fp32 scores of up to 8 queries per pass over the centroid matrix, then the same selection kernel
                for (int q0 = 0; q0 < pl.batch; q0 += 8) {
                    CU(da::ivf_launch_centroid_scores(s.ivf.cent, L, dim, s.d_queries.p + (size_t)q0 * dim, std::min(8, pl.batch - q0),
                                                      s.d_ivf_scores.p + (size_t)q0 * L, s.sm_count, s.stream));
                    st->launches++;
                }
                CU(da::ivf_launch_select_probes(s.d_ivf_scores.p, 1, pl.batch, L, pl.probes, probe_keys, s.stream));
                st->launches++;
            } else {            // more lists than the selection kernel holds in shared memory: the fused top-k scan ranks them
                const int step = da::scan_pick_nq(dim, pl.probes, pl.batch);
                for (int q0 = 0; q0 < pl.batch; q0 += step) {
                    da::ScanArgs c{};
                    c.mat = s.ivf.cent; c.inv_norm = s.ivf.cent_inv;
                    c.queries = s.d_queries.p + (size_t)q0 * dim; c.q_inv_norm = nullptr;
                    c.n_rows = (uint32_t)L; c.row_base = 0;
                    c.dim = dim; c.nq = std::min(step, pl.batch - q0); c.k = pl.probes; c.min_sim = ninf;
                    c.cta_out = s.d_cta; c.counter = s.d_counter; c.out_keys = probe_keys + (size_t)q0 * pl.probes;
                    CU(da::launch_scan(c, s.sm_count, s.stream, nullptr));
                    st->launches++;
                }
            }
            da::ScanArgs a{};
            a.mat = s.mat; a.inv_norm = s.inv_norm; a.doc_slot = s.doc_slot; a.chunk_ids = s.chunk_ids;
            a.doc_bitmap = pl.all_docs ? nullptr : s.d_bitmap.p; a.bitmap_stride = pl.bitmap_stride;
            a.queries = s.d_queries.p; a.q_inv_norm = nullptr;
            a.probe_keys = probe_keys; a.probes = pl.probes; a.probe_batch = pl.batch;
            a.list_offsets = s.ivf.offsets; a.list_rows = s.ivf.list_rows;
            a.n_rows = 0; a.row_base = s.row_base;
            a.dim = dim; a.nq = 1; a.k = pl.k; a.min_sim = pl.min_sim; a.read_all = pl.read_all ? 1 : 0;
            a.cta_out = reinterpret_cast<da::key_t64 *>(s.d_ivf_cta.p); a.counter = s.d_ivf_counter.p;
            a.out_keys = out_keys; a.out_ids = out_ids;
            a.rows_read = s.d_rows_read;
            int grid = 0;
            CU(da::launch_scan(a, s.sm_count, s.stream, &grid));
            st->launches++;
            if (first) st->scan_grid = grid;
            CU(cudaEventRecord(first && e1 ? e1 : s.ev_scan1, s.stream));
            if (st->p2p_gather && !first) CU(cudaEventRecord(s.ev_done, s.stream));
            continue;
        }
        if (pl.use_tc && pl.tc_i8) {
            const da::TcInt8Operands ops{s.d_q8q.p, s.q8, s.d_q8qe.p};
            CU(da::tc_launch_batch_candidates(s.d_queries.p, pl.batch, s.d_q8qf.p, s.mat, s.n_rows, dim, s.q8_factor,
                                              s.doc_slot, pl.all_docs ? nullptr : s.d_bitmap.p, pl.bitmap_stride, pl.k, pl.min_sim,
                                              s.q8_err_max, s.sm_count, s.d_cand.p, s.d_cand_count.p, s.d_hist.p, s.d_gate.p,
                                              s.d_overflow.p, s.d_group_sync.p, s.stream, nullptr, nullptr, &ops));
            CU(da::launch_rescore_select(s.d_cand.p, s.d_cand_count.p, da::tc_batch_cap(pl.k), s.d_gate.p, s.d_surv.p,
                                         s.d_surv_count.p, s.d_overflow.p, pl.batch, s.mat, dim, s.inv_norm, s.d_queries.p,
                                         s.d_qinv.p, s.chunk_ids, s.row_base, pl.min_sim, pl.k, out_keys, out_ids, s.stream));
            st->launches += 4;
            if (first) st->scan_grid = s.sm_count;
            CU(cudaEventRecord(first && e1 ? e1 : s.ev_scan1, s.stream));
            if (st->p2p_gather && !first) CU(cudaEventRecord(s.ev_done, s.stream));
            continue;
        }
        if (pl.use_tc) {
            CU(da::tc_launch_batch_candidates(s.d_queries.p, pl.batch, s.d_qinv.p, s.mat, s.n_rows, dim, s.inv_norm,
                                              s.doc_slot, pl.all_docs ? nullptr : s.d_bitmap.p, pl.bitmap_stride, pl.k, pl.min_sim,
                                              st->split16 ? da::kSplitBf16Eps : da::kTf32Eps, s.sm_count, s.d_cand.p,
                                              s.d_cand_count.p, s.d_hist.p, s.d_gate.p, s.d_overflow.p, s.d_group_sync.p, s.stream,
                                              st->split16 ? s.d_qhi.p : nullptr, st->split16 ? s.hi : nullptr));
            CU(da::launch_rescore_select(s.d_cand.p, s.d_cand_count.p, da::tc_batch_cap(pl.k), s.d_gate.p, s.d_surv.p,
                                         s.d_surv_count.p, s.d_overflow.p, pl.batch, s.mat, dim, s.inv_norm, s.d_queries.p,
                                         s.d_qinv.p, s.chunk_ids, s.row_base, pl.min_sim, pl.k, out_keys, out_ids, s.stream,
                                         s.hi, s.lo));
            st->launches += 4;
            if (first) st->scan_grid = s.sm_count;
            CU(cudaEventRecord(first && e1 ? e1 : s.ev_scan1, s.stream));
            if (st->p2p_gather && !first) CU(cudaEventRecord(s.ev_done, s.stream));
            continue;
        }
        if (pl.compact) {
            // one list per document set: shared by the batch (up to kMaxNQ queries per pass over it), or one per query
            const int step = pl.bitmap_stride ? 1 : da::scan_pick_nq(dim, pl.k, pl.batch);
            for (int q0 = 0; q0 < pl.batch; q0 += step) {
                if (q0 == 0 || pl.bitmap_stride) {
                    CU(da::launch_filter_compact(s.doc_slot, s.d_bitmap.p + (size_t)q0 * pl.bitmap_stride, s.n_rows,
                                                 s.d_flt_rows.p, s.d_flt_n.p, s.stream));
                    st->launches++;
                }
                da::ScanArgs a{};
                a.mat = s.mat; a.inv_norm = s.inv_norm; a.doc_slot = s.doc_slot; a.chunk_ids = s.chunk_ids;
                a.doc_bitmap = nullptr;                         // every listed row is requested
                a.queries = s.d_queries.p + (size_t)q0 * dim; a.q_inv_norm = nullptr;
                a.row_list = s.d_flt_rows.p; a.n_rows_dev = s.d_flt_n.p; a.n_rows = 0; a.row_base = s.row_base;
                a.dim = dim; a.nq = std::min(step, pl.batch - q0); a.k = pl.k; a.min_sim = pl.min_sim; a.read_all = 0;
                a.cta_out = s.d_cta; a.counter = s.d_counter;
                a.out_keys = out_keys + (size_t)q0 * pl.k; a.out_ids = out_ids + (size_t)q0 * pl.k * 16;
                a.rows_read = s.d_rows_read;
                int grid = 0;
                CU(da::launch_scan(a, s.sm_count, s.stream, &grid));
                st->launches++;
                if (first) st->scan_grid = grid;
            }
            CU(cudaEventRecord(first && e1 ? e1 : s.ev_scan1, s.stream));
            if (st->p2p_gather && !first) CU(cudaEventRecord(s.ev_done, s.stream));
            continue;
        }
        const int nq_max = da::scan_pick_nq(dim, pl.k, pl.batch);
        for (int q0 = 0; q0 < pl.batch; q0 += nq_max) {
            da::ScanArgs a{};
            a.mat = s.mat; a.inv_norm = s.inv_norm; a.doc_slot = s.doc_slot; a.chunk_ids = s.chunk_ids;
            a.doc_bitmap = pl.all_docs ? nullptr : s.d_bitmap.p + (size_t)q0 * pl.bitmap_stride;
            a.bitmap_stride = pl.bitmap_stride;
            a.queries = s.d_queries.p + (size_t)q0 * dim;
            a.q_inv_norm = pl.need_qinv ? s.d_qinv.p + q0 : nullptr;
            a.row_list = nullptr;
            a.n_rows = s.n_rows; a.row_base = s.row_base;
            a.dim = dim; a.nq = std::min(nq_max, pl.batch - q0); a.k = pl.k;
            a.min_sim = pl.min_sim; a.read_all = pl.read_all ? 1 : 0;
            a.cta_out = s.d_cta; a.counter = s.d_counter;
            a.out_keys = out_keys + (size_t)q0 * pl.k;
            a.out_ids = out_ids + (size_t)q0 * pl.k * 16;
            a.rows_read = s.d_rows_read;
            if (pl.fused_xchg) {
                // the records go to this rank's (or shard's) slot of its own exchange buffer; the last CTA copies them into
                // every other buffer, waits for the others' records, and writes the merged answer to d_final
                da::XchgArgs &x = a.xg;
                const int N = st->xg_inproc ? (int)st->shards.size() : st->total_ranks;
                const int me = st->xg_inproc ? (int)shard_index : st->rank;
                for (int r = 0; r < N; r++) x.peer[r] = s.xg_peer[r];
                x.n_ranks = N; x.rank = me; x.slot_bytes = kXchgSlotBytes;
                x.parity = (uint32_t)(st->xg_seq & 1);
                x.target = (unsigned long long)N * (st->xg_seq / 2 + 1);
                unsigned char *slot = s.xg_buf + 256 + ((size_t)x.parity * N + me) * kXchgSlotBytes;
                a.out_ids = slot;
                a.out_keys = reinterpret_cast<da::key_t64 *>(slot + (size_t)a.nq * pl.k * 16);
                x.final_ids = s.d_final.p;
                x.final_keys = reinterpret_cast<da::key_t64 *>(s.d_final.p + (size_t)pl.batch * pl.k * 16);
                x.timeout_flag = st->xg_inproc ? s.d_xg_timeout : reinterpret_cast<uint32_t *>(s.d_final.p + rec_bytes + 4);
            }
            if (st->split16) {             // scan the upper plane for candidates, then the exact finish from both planes
                a.mat = nullptr; a.hi = s.hi; a.approx = s.d_approx.p;
                a.min_sim = pl.min_sim - da::kSplitEps;
                a.out_keys = s.d_akeys.p; a.out_ids = nullptr;
            }
            int grid = 0;
            if (pl.use_q8) {   // int8 sidecar: intervals for every row, candidates, exact re-score from the fp32 rows
                const uint32_t cap = (uint32_t)std::min<uint64_t>(std::max<uint32_t>(s.n_rows, 1), 1u << 20);
                a.approx = s.d_approx.p; a.out_keys = s.d_akeys.p; a.out_ids = nullptr;
                int q8_launches = 0;
                CU(da::launch_q8_candidates(a, s.q8, s.q8_factor, s.q8_err, s.d_q8q.p, s.d_q8qf.p, s.d_q8qe.p, s.d_cand.p, cap,
                                            s.d_cand_count.p, s.d_gate.p, s.d_overflow.p, s.sm_count, s.stream, &grid, &q8_launches));
                CU(da::launch_rescore_select(s.d_cand.p, s.d_cand_count.p, (int)cap, s.d_gate.p, s.d_surv.p, s.d_surv_count.p,
                                             s.d_overflow.p, a.nq, s.mat, dim, s.inv_norm, a.queries, a.q_inv_norm, s.chunk_ids,
                                             s.row_base, pl.min_sim, pl.k, out_keys + (size_t)q0 * pl.k,
                                             out_ids + (size_t)q0 * pl.k * 16, s.stream, nullptr, nullptr, true));
                st->launches += (uint64_t)q8_launches + 2;
                if (first) st->scan_grid = grid;
                continue;
            }
            CU(da::launch_scan(a, s.sm_count, s.stream, &grid));
            st->launches++;
            if (first) st->scan_grid = grid;
            if (st->split16) {
                const uint32_t cap = (uint32_t)std::min<uint64_t>(std::max<uint32_t>(s.n_rows, 1), 1u << 20);
                CU(da::launch_split_finish(s.d_approx.p, s.n_rows, a.nq, pl.k, s.d_akeys.p, pl.min_sim, da::kSplitEps,
                                           s.d_cand.p, cap, s.d_cand_count.p, s.d_overflow.p, s.hi, s.lo, dim, s.inv_norm,
                                           a.queries, a.q_inv_norm, s.chunk_ids, s.row_base,
                                           out_keys + (size_t)q0 * pl.k, out_ids + (size_t)q0 * pl.k * 16, s.stream));
                st->launches += 3;
            }
        }
        CU(cudaEventRecord(first && e1 ? e1 : s.ev_scan1, s.stream));
        if (st->p2p_gather && !first) CU(cudaEventRecord(s.ev_done, s.stream));
    }
    st->scan_launches++;        // one timed unit per search: scan_ms_sum and last_bytes both cover ALL passes of a search
    if (pl.fused_xchg) st->xg_seq++;                  // one exchange per search, the same number on every rank / shard
    if (st->total_ranks > 1 && !pl.fused_xchg) {
        Shard &s0 = st->shards[0];
        if (st->p2p_gather) {
            CU(cudaSetDevice(s0.device));
            for (size_t g = 1; g < st->shards.size(); g++) CU(cudaStreamWaitEvent(s0.stream, st->shards[g].ev_done, 0));
        } else {
            // Several processes: a rank sees only its own candidate lists, so each block carries a trailer with the
            // rank's overflow flag; after the gather every rank holds every flag and takes the same decision.
            const bool lists = pl.use_tc || pl.use_q8 || st->split16;
            if (trailer && lists)
                for (Shard &s : st->shards) {
                    CU(cudaSetDevice(s.device));
                    CU(cudaMemcpyAsync(s.d_out.p + rec_bytes, s.d_overflow.p, 4, cudaMemcpyDeviceToDevice, s.stream));
                }
            NC(ncclGroupStart());
            for (Shard &s : st->shards)
                NC(ncclAllGather(s.d_out.p, s.d_gather.p, rec_bytes + trailer, ncclChar, s.comm, s.stream));
            NC(ncclGroupEnd());
        }
        CU(cudaSetDevice(s0.device));
        CU(da::launch_merge(s0.d_gather.p, st->total_ranks, pl.batch, pl.k,
                            reinterpret_cast<da::key_t64 *>(s0.d_final.p + (size_t)pl.batch * pl.k * 16),
                            s0.d_final.p, s0.stream, rec_bytes + trailer,
                            trailer ? reinterpret_cast<uint32_t *>(s0.d_final.p + rec_bytes) : nullptr));
        st->launches++;
        if (st->p2p_gather) CU(cudaEventRecord(s0.ev_done, s0.stream));   // gather blocks may be overwritten after this
    }
    return DA_OK;
}

// Smallest batch that goes to the tensor path.  DA_TC_MIN_BATCH overrides (measurements).
int tc_min_batch(const da_store *st)
{
    static const int env = [] { const char *v = getenv("DA_TC_MIN_BATCH"); return v ? atoi(v) : 0; }();
    if (env > 0) return env;
    // Measured at 10M x 3072 (ms per batch, fp32 SIMT scan vs tensor path, rows as the M operand): b=2 16.7 / 17.2,
    // b=3 21.0 / 17.3, b=4 20.9 / 17.3, b=8 34.3 / 17.2 -- the SIMT scan pays 2*dim FMAs and dim shared-memory words per
    // row and query, the tensor path reads the rows once whatever the batch.  The int8 sidecar scan serves 8 queries in 12.6 ms.
    if (st->split16) return 16;
    return 3;                      // (int8 SIMT scan: 2 queries 4.9 ms, 4 queries 6.4; int8 tensor path: 5.2-5.8 ms for 4..16)
}

int ivf_tc_min_batch()
{
    static const int env = [] { const char *v = getenv("DA_IVF_TC_MIN_BATCH"); return v ? atoi(v) : 0; }();
    return env > 0 ? env : 8;
}

// One process driving several GPUs (the Go query agent on an 8 x B200 box): with peer access between every pair of
// its devices the shards exchange exactly as ranks of a multi-process store do -- every shard's last CTA stores its
// records into every shard's buffer, the merged answer is taken from the first shard.
int setup_inprocess_exchange(da_store *st)
{
    const int N = (int)st->shards.size();
    if (N < 2 || N > da::kXchgMaxRanks || st->world != 1) return DA_OK;
    if (const char *v = getenv("DA_DISABLE_FUSED_EXCHANGE")) if (v[0] == '1') return DA_OK;
    for (int i = 0; i < N; i++)
        for (int j = 0; j < N; j++) {
            if (i == j) continue;
            int can = 0;
            if (cudaDeviceCanAccessPeer(&can, st->shards[i].device, st->shards[j].device) != cudaSuccess || !can) return DA_OK;
        }
    const size_t bytes = 256 + (size_t)2 * N * kXchgSlotBytes;
    for (int i = 0; i < N; i++) {
        Shard &s = st->shards[i];
        CU(cudaSetDevice(s.device));
        for (int j = 0; j < N; j++) {
            if (i == j) continue;
            cudaError_t e = cudaDeviceEnablePeerAccess(st->shards[j].device, 0);
            if (e == cudaErrorPeerAccessAlreadyEnabled) { cudaGetLastError(); e = cudaSuccess; }
            if (e != cudaSuccess) { cudaGetLastError(); return DA_OK; }
        }
        CU(cudaMalloc((void **)&s.xg_buf, bytes));
        CU(cudaMemset(s.xg_buf, 0, bytes));
        CU(cudaHostAlloc((void **)&s.h_xg_timeout, 64, cudaHostAllocMapped | cudaHostAllocPortable));
        memset(s.h_xg_timeout, 0, 64);
        CU(cudaHostGetDevicePointer((void **)&s.d_xg_timeout, s.h_xg_timeout, 0));
    }
    for (int i = 0; i < N; i++)
        for (int j = 0; j < N; j++) st->shards[i].xg_peer[j] = st->shards[j].xg_buf;
    st->xg_inproc = true;
    return DA_OK;
}

// One process per GPU: give every rank a mapping of every rank's exchange buffer (CUDA IPC handles travel through
// the NCCL communicator that exists anyway).  All ranks agree on the outcome; without it searches keep the
// all-gather + merge kernels.
int setup_fused_exchange(da_store *st)
{
    Shard &s = st->shards[0];
    const int N = st->total_ranks, me = st->rank;
    if (N < 2 || N > da::kXchgMaxRanks || st->shards.size() != 1) return DA_OK;
    if (const char *v = getenv("DA_DISABLE_FUSED_EXCHANGE")) if (v[0] == '1') return DA_OK;
    CU(cudaSetDevice(s.device));
    const size_t bytes = 256 + (size_t)2 * N * kXchgSlotBytes;
    CU(cudaMalloc((void **)&s.xg_buf, bytes));
    CU(cudaMemsetAsync(s.xg_buf, 0, bytes, s.stream));
    cudaIpcMemHandle_t mine;
    CU(cudaIpcGetMemHandle(&mine, s.xg_buf));
    unsigned char *d_handles = nullptr;
    int *d_ok = nullptr;
    CU(cudaMalloc((void **)&d_handles, (size_t)N * sizeof mine));
    CU(cudaMalloc((void **)&d_ok, sizeof(int)));
    CU(cudaMemcpyAsync(d_handles + (size_t)me * sizeof mine, &mine, sizeof mine, cudaMemcpyHostToDevice, s.stream));
    NC(ncclAllGather(d_handles + (size_t)me * sizeof mine, d_handles, sizeof mine, ncclChar, s.comm, s.stream));
    std::vector<cudaIpcMemHandle_t> all(N);
    CU(cudaMemcpyAsync(all.data(), d_handles, (size_t)N * sizeof mine, cudaMemcpyDeviceToHost, s.stream));
    CU(cudaStreamSynchronize(s.stream));
    int ok = 1;
    for (int r = 0; r < N; r++) {
        if (r == me) { s.xg_peer[r] = s.xg_buf; continue; }
        void *p = nullptr;
        if (cudaIpcOpenMemHandle(&p, all[r], cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { cudaGetLastError(); ok = 0; break; }
        s.xg_peer[r] = static_cast<unsigned char *>(p);
    }
    CU(cudaMemcpyAsync(d_ok, &ok, sizeof ok, cudaMemcpyHostToDevice, s.stream));
    NC(ncclAllReduce(d_ok, d_ok, 1, ncclInt, ncclMin, s.comm, s.stream));
    CU(cudaMemcpyAsync(&ok, d_ok, sizeof ok, cudaMemcpyDeviceToHost, s.stream));
    CU(cudaStreamSynchronize(s.stream));
    cudaFree(d_handles); cudaFree(d_ok);
    st->xg_ok = ok == 1;
    return DA_OK;
}

// Queries: caller's buffer -> pinned staging (once, on the first shard) -> every local device.  A large batch (4 096
// queries x 1 536 floats = 25 MB) is cut into slices: several threads copy slices into the pinned buffer while the
// slices already there are on their way to the device, so the staging costs about one pass of the slower of the two
// instead of their sum.
int upload_queries(da_store *st, Shard &s, const float *queries, size_t n_elems)
{
    Shard &s0 = st->shards[0];
    const size_t bytes = n_elems * sizeof(float);
    if (&s != &s0) {        // the pinned copy already exists (shards are staged in order)
        CU(cudaMemcpyAsync(s.d_queries.p, s0.h_queries.p, bytes, cudaMemcpyHostToDevice, s.stream));
        return DA_OK;
    }
    constexpr size_t kSlice = 2u << 20;
    static_assert(2 * kSlice == kCheckWhileCopying, "batches that skip validate_search's check must take the sliced path");
    if (bytes <= 2 * kSlice) {
        memcpy(s0.h_queries.p, queries, bytes);
        CU(cudaMemcpyAsync(s0.d_queries.p, s0.h_queries.p, bytes, cudaMemcpyHostToDevice, s0.stream));
        return DA_OK;
    }
    const size_t n_slices = (bytes + kSlice - 1) / kSlice;
    const int n_threads = (int)std::min<size_t>(4, n_slices);
    std::atomic<size_t> next{0};
    std::atomic<int> any_bad{0};
    std::vector<std::atomic<int>> ready(n_slices);
    for (auto &r : ready) r.store(0);
    const char *src = reinterpret_cast<const char *>(queries);
    char *pin = reinterpret_cast<char *>(s0.h_queries.p);
    std::vector<std::thread> workers;
    for (int t = 0; t < n_threads; t++)
        workers.emplace_back([&] {
            for (size_t i = next.fetch_add(1); i < n_slices; i = next.fetch_add(1)) {
                const size_t off = i * kSlice, len = std::min(kSlice, bytes - off);
                memcpy(pin + off, src + off, len);
                const uint32_t *w = reinterpret_cast<const uint32_t *>(pin + off);      // NaN/Inf: exponent all ones
                uint32_t bad = 0;
                for (size_t j = 0; j < len / 4; j++) bad |= (uint32_t)((w[j] & 0x7F800000u) == 0x7F800000u);
                if (bad) any_bad.store(1);
                ready[i].store(1, std::memory_order_release);
            }
        });
    cudaError_t ce = cudaSuccess;
    for (size_t i = 0; i < n_slices; i++) {
        while (!ready[i].load(std::memory_order_acquire)) _mm_pause();
        const size_t off = i * kSlice, len = std::min(kSlice, bytes - off);
        if (ce == cudaSuccess)
            ce = cudaMemcpyAsync(reinterpret_cast<char *>(s0.d_queries.p) + off, pin + off, len, cudaMemcpyHostToDevice, s0.stream);
    }
    for (auto &w : workers) w.join();
    if (ce != cudaSuccess) return fail(DA_ERR_CUDA, "query upload failed: %s", cudaGetErrorString(ce));
    if (any_bad.load()) {               // nothing has been launched yet; the copies in flight are harmless
        const uint32_t *w = reinterpret_cast<const uint32_t *>(queries);
        for (size_t i = 0; i < n_elems; i++)
            if ((w[i] & 0x7F800000u) == 0x7F800000u) return fail(DA_ERR_NONFINITE, "query element %zu is not finite", i);
    }
    return DA_OK;
}

// The smallest float f with (double)f >= d: for a float score s, (double)s >= d holds exactly when s >= f, so the
// kernels keep comparing in fp32 and still decide like Postgres' float8 comparison (postgres.go:240).
float float_at_least(double d)
{
    if (std::isnan(d)) return std::numeric_limits<float>::quiet_NaN();
    float f = (float)d;
    if ((double)f < d) f = std::nextafterf(f, std::numeric_limits<float>::infinity());
    return f;
}
float search_floor(const da_search_params *p)
{
    return p->min_similarity_f64 != 0.0 ? float_at_least(p->min_similarity_f64) : p->min_similarity;
}

// What every search checks before anything is staged (the errors of the reference's statement: dimension mismatch,
// negative LIMIT, NaN/Inf in the vector literal).
int validate_search(const da_store *st, const da_search_params *p)
{
    if (!p) return fail(DA_ERR_INVALID, "search params are NULL");
    if (p->struct_size != sizeof(da_search_params))
        return fail(DA_ERR_INVALID, "da_search_params.struct_size=%u, expected %zu", p->struct_size, sizeof(da_search_params));
    if (p->dim != st->dim) return fail(DA_ERR_DIM, "expected %d dimensions, not %d", st->dim, p->dim);
    if (p->k < 0) return fail(DA_ERR_INVALID, "LIMIT must not be negative (k=%d)", p->k);
    if (p->k > DA_MAX_K) return fail(DA_ERR_INVALID, "k=%d exceeds DA_MAX_K=%d", p->k, DA_MAX_K);
    if (p->batch < 1 || p->batch > DA_MAX_BATCH) return fail(DA_ERR_INVALID, "batch=%d out of range [1,%d]", p->batch, DA_MAX_BATCH);
    if (!p->queries) return fail(DA_ERR_INVALID, "query vector is NULL");
    const bool ivf = (p->flags & DA_SEARCH_IVF) != 0;
    if (ivf && st->split16) return fail(DA_ERR_STATE, "IVFFlat is not available with the split-plane layout");
    if (ivf && st->ivf_lists == 0) return fail(DA_ERR_STATE, "IVFFlat index has not been built on this store (da_store_build_ivf)");
    const bool all_docs = (p->flags & DA_SEARCH_ALL_DOCS) != 0;
    if (!all_docs && p->n_docs && !p->doc_ids) return fail(DA_ERR_INVALID, "doc_ids is NULL with n_docs=%zu", p->n_docs);
    // finite check on the raw bits: exponent all ones = Inf or NaN.  One branch-free pass (it vectorises), then the
    // exact position only when something was found.
    const size_t nq_elems = (size_t)p->batch * st->dim;
    if (nq_elems * 4 > kCheckWhileCopying) return DA_OK;       // a large batch is checked by the threads that copy it (upload_queries)
    const uint32_t *w = reinterpret_cast<const uint32_t *>(p->queries);
    uint32_t bad = 0;
    for (size_t i = 0; i < nq_elems; i++) bad |= (uint32_t)((w[i] & 0x7F800000u) == 0x7F800000u);
    if (bad)
        for (size_t i = 0; i < nq_elems; i++)
            if ((w[i] & 0x7F800000u) == 0x7F800000u) return fail(DA_ERR_NONFINITE, "query element %zu is not finite", i);
    return DA_OK;
}

// Validate + stage the inputs of a search on every local shard.
int stage_search(da_store *st, const da_search_params *p, SearchPlan &pl)
{
    {
        const int vrc = validate_search(st, p);
        if (vrc != DA_OK) return vrc;
    }
    const bool ivf = (p->flags & DA_SEARCH_IVF) != 0;
    const bool all_docs = (p->flags & DA_SEARCH_ALL_DOCS) != 0;
    const size_t nq_elems = (size_t)p->batch * st->dim;

    pl.batch = p->batch; pl.k = p->k; pl.min_sim = search_floor(p);
    pl.all_docs = all_docs; pl.read_all = (p->flags & DA_SEARCH_READ_ALL_ROWS) != 0;
    pl.bitmap_words = (st->docs.size() + 31) / 32 + 1;
    pl.ivf = ivf;
    pl.use_tc = false;
    pl.probes = 0;
    if (ivf) {
        int probes = p->ivf_probes > 0 ? p->ivf_probes : st->ivf_default_probes;
        pl.probes = std::max(1, std::min(probes, std::min(st->ivf_lists, DA_IVF_MAX_PROBES)));
        // batches of 8 or more rank the centroids as one dense contraction on the tensor cores (TF32 only steers
        // which lists are probed; every returned score comes from the fp32 scan of the probed rows)
        pl.ivf_tc = p->batch >= ivf_tc_min_batch() && !(p->flags & DA_SEARCH_NO_TENSOR) && st->ivf_lists <= da::kIvfSelectMaxLists;
        pl.ivf_ks = pl.ivf_tc ? da::tc_scores_splits(p->batch, st->ivf_lists, st->dim, st->shards[0].sm_count) : 0;
        pl.ivf_select = !pl.ivf_tc && st->ivf_lists <= da::kIvfSelectMaxLists && (size_t)8 * st->dim * 4 <= 200 * 1024;
    }

    // document filter -> bit per slot (document_id = ANY($2): postgres.go:239); unknown ids match nothing.
    // With doc_offsets every query has its own set: one bitmap per query, bitmap_words apart.
    const bool per_query = !all_docs && p->doc_offsets != nullptr;
    if (per_query) {
        if (p->doc_offsets[0] != 0 || p->doc_offsets[p->batch] != p->n_docs)
            return fail(DA_ERR_INVALID, "doc_offsets must start at 0 and end at n_docs");
        for (int q = 0; q < p->batch; q++)
            if (p->doc_offsets[q] > p->doc_offsets[q + 1]) return fail(DA_ERR_INVALID, "doc_offsets must not decrease");
    }
    pl.bitmap_stride = per_query ? (uint32_t)pl.bitmap_words : 0u;
    std::vector<uint32_t> bitmap, uni;
    if (!all_docs) {
        bitmap.assign(pl.bitmap_words * (per_query ? (size_t)p->batch : 1), 0u);
        uni.assign(pl.bitmap_words, 0u);
        int q = 0;
        for (size_t i = 0; i < p->n_docs; i++) {
            if (per_query) while (i >= p->doc_offsets[q + 1]) q++;
            auto it = st->doc_slot_of.find(load_id(p->doc_ids + 16 * i));
            if (it == st->doc_slot_of.end()) continue;
            bitmap[(per_query ? (size_t)q * pl.bitmap_words : 0) + (it->second >> 5)] |= 1u << (it->second & 31);
            uni[it->second >> 5] |= 1u << (it->second & 31);
        }
    }
    if (pl.k == 0) return DA_OK;
    // Batches of tc_min_batch() or more queries become a dense contraction: tensor cores generate the candidates, fp32 re-scores
    // them.  Not for narrow document filters, where the SIMT scan skips the rows it does not need.
    if (!ivf && !pl.read_all && !(p->flags & DA_SEARCH_NO_TENSOR) && p->batch >= tc_min_batch(st) && (!st->split16 || st->dim % 8 == 0)) {
        size_t selected = st->docs.size();
        if (!all_docs) {
            selected = 0;                              // documents requested by at least one query
            for (uint32_t w : uni) selected += (size_t)__builtin_popcount(w);
        }
        pl.use_tc = selected * 4 >= st->docs.size() && selected > 0;
        // The int8 contraction's slack (2 x ~0.02) is wide: it needs a similarity floor to bound the candidate set from
        // the first row tile on.  Without one the running k-th best sits at noise level for a long while and half the
        // rows would pass (measured: configs[3], no floor, k = 100, overflows every list); those searches keep TF32.
        pl.tc_i8 = pl.use_tc && st->q8 && !(p->flags & DA_SEARCH_NO_Q8) && st->dim % 16 == 0 && pl.min_sim >= 0.5f;
        if (pl.tc_i8) {
            // The contraction's slack is ONE bound for all rows, the largest of any resident row: a single row that
            // quantises badly (one huge element) would widen it for everybody, so such stores keep TF32 for batches.
            for (Shard &s : st->shards) {
                if (s.q8_dirty) {        // rows were written since the last look: fetch the rows' error bound (4 bytes)
                    CU(cudaSetDevice(s.device));
                    CU(cudaMemcpyAsync(&s.q8_err_max, s.q8_errmax, 4, cudaMemcpyDeviceToHost, s.stream));
                    CU(cudaStreamSynchronize(s.stream));
                    s.q8_dirty = false;
                }
                if (s.q8_err_max > 0.05f) pl.tc_i8 = false;
            }
        }
    }
    if (!all_docs) {
        size_t selected = 0;
        for (uint32_t w : uni) selected += (size_t)__builtin_popcount(w);
        pl.none_requested = selected == 0 && !pl.read_all;     // (the mask-as-predicate mode reads every row by definition)
    }
    // A filter that names few documents (the reference's case): the requested rows are compacted into a list first.
    if (!ivf && !all_docs && !pl.read_all && !pl.use_tc && !st->split16 && !(p->flags & DA_SEARCH_NO_COMPACT)) {
        size_t selected = 0;
        for (uint32_t w : uni) selected += (size_t)__builtin_popcount(w);
        pl.compact = selected * 16 <= st->docs.size();
    }
    // The int8 sidecar pays when most rows are read; a narrow document filter reads so few rows that the fp32 scan
    // (one kernel, no candidate passes) is the shorter way.
    if (st->q8 && !ivf && !pl.use_tc && !(p->flags & DA_SEARCH_NO_Q8)) {
        size_t selected = st->docs.size();
        if (!all_docs && !pl.read_all) {
            selected = 0;
            for (uint32_t w : uni) selected += (size_t)__builtin_popcount(w);
        }
        pl.use_q8 = selected * 4 >= st->docs.size() && selected > 0;
    }
    pl.need_qinv = pl.use_tc || (pl.use_q8 && !da::q8_one_kernel_form(pl.min_sim)) || st->split16;
    // The scan exchanges and merges across GPUs itself when every rank is certain to run the same plan: that must
    // follow from the call's arguments and the store's configuration alone (every document requested, one SIMT
    // pass, no candidate path), never from what a rank happens to hold.
    pl.fused_xchg = st->xg_ok && all_docs && !ivf && !st->split16 && p->batch <= da::scan_pick_nq(st->dim, pl.k, p->batch) &&
                    (p->batch < tc_min_batch(st) || (p->flags & DA_SEARCH_NO_TENSOR)) && (!st->q8 || (p->flags & DA_SEARCH_NO_Q8)) &&
                    !pl.use_tc && !pl.use_q8;
    // One process: its shards share this very plan, so any search answered by one plain scan pass per shard qualifies
    // (document filters included; the compacted-list and candidate paths keep the gather into the first shard).
    if (st->xg_inproc && !ivf && !st->split16 && !pl.use_tc && !pl.use_q8 && !pl.compact && !pl.none_requested &&
        p->batch <= da::scan_pick_nq(st->dim, pl.k, p->batch))
        pl.fused_xchg = true;
    const size_t rec_bytes = (size_t)pl.batch * pl.k * 24;
    for (Shard &s : st->shards) {
        CU(cudaSetDevice(s.device));
        if (pl.compact) { CU(s.d_flt_rows.ensure(std::max<size_t>(s.n_rows, 1))); CU(s.d_flt_n.ensure(1)); }
        if (pl.ivf && pl.k > 0) {
            CU(s.d_probe_keys.ensure((size_t)p->batch * pl.probes));
            CU(s.d_ivf_cta.ensure((size_t)p->batch * da::scan_probe_ctas_per_query(s.sm_count, p->batch) * pl.k));
            if (s.d_ivf_counter.cap < (size_t)p->batch) {        // tickets start at zero and every launch leaves them there
                CU(s.d_ivf_counter.ensure((size_t)p->batch));
                CU(cudaMemsetAsync(s.d_ivf_counter.p, 0, s.d_ivf_counter.cap * sizeof(unsigned int), s.stream));
            }
            if (pl.ivf_tc) CU(s.d_ivf_scores.ensure((size_t)pl.ivf_ks * p->batch * st->ivf_lists));
            if (pl.ivf_select) CU(s.d_ivf_scores.ensure((size_t)p->batch * st->ivf_lists));
        }
        if (pl.use_q8) {
            const size_t nq = (size_t)da::scan_pick_nq(st->dim, pl.k, p->batch);
            CU(s.d_approx.ensure(nq * std::max<size_t>(s.n_rows, 1)));
            CU(s.d_akeys.ensure(nq * (size_t)pl.k));
            CU(s.d_cand.ensure(nq * (size_t)std::min<uint64_t>(std::max<uint32_t>(s.n_rows, 1), 1u << 20)));
            CU(s.d_cand_count.ensure(nq)); CU(s.d_gate.ensure(nq)); CU(s.d_surv_count.ensure(nq));
            CU(s.d_surv.ensure(nq * (size_t)da::kSurvivorCap));
            CU(s.d_q8q.ensure(nq * (size_t)st->dim)); CU(s.d_q8qf.ensure(nq)); CU(s.d_q8qe.ensure(nq));
            CU(s.d_overflow.ensure(1)); CU(s.h_overflow.ensure(1));
            CU(cudaMemsetAsync(s.d_overflow.p, 0, 4, s.stream));
            *s.h_overflow.p = 0;
        }
        if (pl.tc_i8) {
            CU(s.d_q8q.ensure(nq_elems)); CU(s.d_q8qf.ensure(p->batch)); CU(s.d_q8qe.ensure(p->batch));
        }
        if (pl.use_tc) {
            CU(s.d_cand.ensure((size_t)p->batch * da::tc_batch_cap(pl.k)));
            CU(s.d_cand_count.ensure(p->batch)); CU(s.d_gate.ensure(p->batch)); CU(s.d_surv_count.ensure(p->batch));
            CU(s.d_hist.ensure(da::tc_batch_hist_words(p->batch)));
            CU(s.d_surv.ensure((size_t)p->batch * da::kSurvivorCap));
            CU(s.d_group_sync.ensure(1024));
            CU(s.d_overflow.ensure(1)); CU(s.h_overflow.ensure(1));
            CU(cudaMemsetAsync(s.d_overflow.p, 0, 4, s.stream));
            *s.h_overflow.p = 0;
        }
        if (st->split16) {
            const size_t nq = (size_t)da::scan_pick_nq(st->dim, pl.k, p->batch);
            CU(s.d_approx.ensure(nq * std::max<size_t>(s.n_rows, 1)));
            CU(s.d_akeys.ensure(nq * (size_t)pl.k));
            CU(s.d_cand.ensure(nq * (size_t)std::min<uint64_t>(std::max<uint32_t>(s.n_rows, 1), 1u << 20)));
            CU(s.d_cand_count.ensure(nq));
            CU(s.d_overflow.ensure(1)); CU(s.h_overflow.ensure(1));
            CU(cudaMemsetAsync(s.d_overflow.p, 0, 4, s.stream));
            *s.h_overflow.p = 0;
        }
        CU(s.d_queries.ensure(nq_elems)); CU(s.d_qinv.ensure(p->batch));
        if (&s == &st->shards[0]) CU(s.h_queries.ensure(nq_elems));      // one pinned copy serves every local device
        const size_t trailer = st->world > 1 ? kRecTrailer : 0;
        CU(s.d_out.ensure(rec_bytes + trailer)); CU(s.h_result.ensure(rec_bytes + trailer));
        if (st->total_ranks > 1) { CU(s.d_gather.ensure((rec_bytes + trailer) * st->total_ranks)); CU(s.d_final.ensure(rec_bytes + trailer)); }
        if (trailer) {
            CU(cudaMemsetAsync(s.d_final.p + rec_bytes, 0, trailer, s.stream));     // flags: overflow (merge kernel), exchange timeout
            CU(cudaMemsetAsync(s.d_out.p + rec_bytes, 0, trailer, s.stream));
            if (!s.d_overflow.p) { CU(s.d_overflow.ensure(1)); CU(s.h_overflow.ensure(1)); CU(cudaMemsetAsync(s.d_overflow.p, 0, 4, s.stream)); *s.h_overflow.p = 0; }
        }
        {
            int rc_up = upload_queries(st, s, p->queries, nq_elems);
            if (rc_up != DA_OK) return rc_up;
        }
        if (!all_docs) {
            CU(s.d_bitmap.ensure(bitmap.size())); CU(s.h_bitmap.ensure(bitmap.size()));
            memcpy(s.h_bitmap.p, bitmap.data(), bitmap.size() * 4);
            CU(cudaMemcpyAsync(s.d_bitmap.p, s.h_bitmap.p, bitmap.size() * 4, cudaMemcpyHostToDevice, s.stream));
        }
        if (pl.need_qinv) {
            CU(da::launch_query_inv_norm(s.d_queries.p, st->dim, p->batch, s.d_qinv.p, s.stream));
            st->launches++;
        }
        if (pl.tc_i8) {                       // the int8 contraction needs the queries as int8 with their scales and bounds
            CU(da::launch_q8_quantize_queries(s.d_queries.p, p->batch, st->dim, s.d_qinv.p, s.d_q8q.p, s.d_q8qf.p, s.d_q8qe.p, s.stream));
            st->launches++;
        }
        if (pl.use_tc && st->split16) {       // the bf16 contraction needs the queries' upper halves
            CU(s.d_qhi.ensure(nq_elems));
            CU(da::tc_launch_truncate_bf16(s.d_queries.p, nq_elems, s.d_qhi.p, s.stream));
            st->launches++;
        }
    }
    return DA_OK;
}

// Queue the read-back (and reset) of the rows-read counters behind the search on every stream.
int enqueue_stats_readback(da_store *st)
{
    for (Shard &s : st->shards) {
        CU(cudaSetDevice(s.device));
        CU(cudaMemcpyAsync(s.h_rows_read.p, s.d_rows_read, sizeof(unsigned long long), cudaMemcpyDeviceToHost, s.stream));
        CU(cudaMemsetAsync(s.d_rows_read, 0, sizeof(unsigned long long), s.stream));
        if (s.d_overflow.p && s.h_overflow.p)
            CU(cudaMemcpyAsync(s.h_overflow.p, s.d_overflow.p, 4, cudaMemcpyDeviceToHost, s.stream));
    }
    return DA_OK;
}

// After the streams were synchronised: fold the counters into the statistics.
int collect_stats(da_store *st, const SearchPlan &pl, int searches = 1)
{
    Shard &s0 = st->shards[0];
    CU(cudaSetDevice(s0.device));
    uint64_t rows_read = 0, rows_scanned = 0;
    for (Shard &s : st->shards) {
        rows_read += *s.h_rows_read.p;
        rows_scanned += s.n_rows;
    }
    const int nq = da::scan_pick_nq(st->dim, pl.k, pl.batch);
    const int passes = (pl.batch + nq - 1) / nq;
    rows_read /= (uint64_t)searches;
    if (pl.use_tc) rows_read = rows_scanned;        // the contraction reads every resident row once per 128 queries
    st->last_rows_read = rows_read;
    // bytes the algorithm must move: vectors of the rows read + inv_norm (4 B) of every row visited
    // + doc_slot (4 B) of every row visited when a document filter is present.
    if (pl.ivf) {  // centroid matrix once per pass of the ranking step + the probed rows (vector, inv_norm, doc slot if filtered, row number)
        const int cstep = pl.ivf_select ? 8 : da::scan_pick_nq(st->dim, pl.probes, pl.batch);
        const uint64_t cpasses = pl.ivf_tc ? 1 : (uint64_t)((pl.batch + cstep - 1) / cstep);
        st->last_bytes = cpasses * st->shards.size() * st->ivf_lists * ((uint64_t)st->dim * 4 + 4) +
                         rows_read * ((uint64_t)st->dim * 4 + (pl.all_docs ? 8 : 12));
    }
    else if (st->split16 && pl.use_tc)   // the upper plane once, whatever the batch
        st->last_bytes = rows_scanned * ((uint64_t)st->dim * 2 + (pl.all_docs ? 4 : 8));
    else if (st->split16)   // upper plane of the rows read + per row: 1/|row|, approximate score out and in (+ doc slot)
        st->last_bytes = rows_read * (uint64_t)st->dim * 2 + rows_scanned * (uint64_t)passes * (pl.all_docs ? 12 : 16);
    else if (pl.use_tc && pl.tc_i8)     // the int8 plane once, whatever the batch (+ factor, doc slot)
        st->last_bytes = rows_scanned * ((uint64_t)st->dim + (pl.all_docs ? 4 : 8));
    else if (pl.use_q8)     // int8 rows read + per row: factor, error bound, upper end out and in (+ doc slot)
        st->last_bytes = rows_read * (uint64_t)st->dim + rows_scanned * (uint64_t)passes * (pl.all_docs ? 16 : 20);
    else if (pl.compact)    // filter pass: every row's slot once per document set; then the listed rows (vector, row number, 1/|row|)
        st->last_bytes = rows_scanned * 4ull * (uint64_t)(pl.bitmap_stride ? pl.batch : 1) + rows_read * ((uint64_t)st->dim * 4 + 8);
    else if (pl.use_tc)
        st->last_bytes = rows_scanned * ((uint64_t)st->dim * 4 + (pl.all_docs ? 4 : 8));   // every row once, whatever the batch
    else
        st->last_bytes = rows_read * (uint64_t)st->dim * 4 +
                         rows_scanned * (uint64_t)passes * (pl.all_docs ? 4 : 8);
    return DA_OK;
}

// ---------------------------------------------------------------------------
// Small-search path: one query on a single-shard, single-process store -- the reference's own regime (one question,
// 1..16 documents: cmd/query/main.go:20-24) and the headline single-query scan.  One host-to-device copy (query +
// row ranges), ONE kernel (the scan computes 1/|q| itself and walks the requested documents' row ranges: no bitmap,
// no document slots, no compaction pass), results written by the last CTA into mapped host memory and picked up by
// polling a completion word.  Calls hold the store's lock SHARED and each uses its own slot, so they overlap.
// ---------------------------------------------------------------------------
struct FastPlan {
    int k = 0;
    float min_sim = 0.f;
    bool ranged = false;        // false: every row of the shard (all documents requested)
    int n_ranges = 0;
    uint32_t total_rows = 0;    // rows the scan will read
    bool nothing = false;       // no requested document holds a row: zero rows, as `= ANY('{}')`
    bool ivf = false;           // probe the IVFFlat lists: centroid scan (top `probes`), then the probe-mode scan
    int probes = 0;
    bool bitmap = false;        // IVFFlat with a document filter: bit per document slot, tested on the probed rows
    bool xchg = false;          // one process per GPU: the scan exchanges its records with the other ranks and merges (collective)
};

bool fast_path_enabled()
{
    static const bool off = [] { const char *v = getenv("DA_NO_FAST_PATH"); return v && v[0] == '1'; }();
    return !off;
}

// 1: eligible (fp and ranges filled), 0: not eligible (the general path answers), < 0: error.  Lock held (shared).
int plan_fast(da_store *st, const da_search_params *p, FastPlan &fp, uint2 *ranges)
{
    if (!fast_path_enabled() || st->shards.size() != 1 || st->split16) return 0;
    if (!p || p->struct_size != sizeof(da_search_params) || p->batch != 1 || p->doc_offsets) return 0;
    if (p->flags & (DA_SEARCH_READ_ALL_ROWS | DA_SEARCH_NO_COMPACT)) return 0;
    const bool all_docs = (p->flags & DA_SEARCH_ALL_DOCS) != 0;
    const bool ivf = (p->flags & DA_SEARCH_IVF) != 0;
    const bool q8_ok = st->q8 && !(p->flags & DA_SEARCH_NO_Q8);
    if (st->total_ranks != 1) {
        // several processes: only the search every rank is certain to plan the same way (see SearchPlan::fused_xchg)
        if (!(st->xg_ok && all_docs && !ivf && !q8_ok)) return 0;
        fp.xchg = true;
    }
    if (all_docs && q8_ok && !ivf) return 0;               // the int8 sidecar scan reads a quarter of the bytes
    const int vrc = validate_search(st, p);
    if (vrc != DA_OK) return vrc;
    const Shard &s = st->shards[0];
    fp.k = p->k; fp.min_sim = search_floor(p);
    if (ivf) {
        fp.ivf = true;
        const int probes = p->ivf_probes > 0 ? p->ivf_probes : st->ivf_default_probes;
        fp.probes = std::max(1, std::min(probes, std::min(st->ivf_lists, DA_IVF_MAX_PROBES)));
        fp.bitmap = !all_docs;
        fp.total_rows = 0;
        fp.nothing = s.n_rows == 0;
        if (fp.bitmap) {            // filters apply to the probed rows (pgvector filters after the index scan)
            bool any = false;
            for (size_t i = 0; i < p->n_docs && !any; i++) {
                auto it = st->doc_slot_of.find(load_id(p->doc_ids + 16 * i));
                any = it != st->doc_slot_of.end() && st->docs[it->second].rows > 0;
            }
            if (!any) fp.nothing = true;
        }
        return 1;
    }
    if (all_docs) {
        fp.ranged = false; fp.total_rows = s.n_rows; fp.nothing = s.n_rows == 0 && !fp.xchg;    // a rank of a collective search always takes part
        return 1;
    }
    fp.ranged = true;
    uint32_t seen_small[64];
    size_t n_seen = 0, selected = 0;
    std::vector<uint32_t> seen_big;
    for (size_t i = 0; i < p->n_docs; i++) {
        auto it = st->doc_slot_of.find(load_id(p->doc_ids + 16 * i));
        if (it == st->doc_slot_of.end()) continue;          // unknown ids match nothing
        const uint32_t slot = it->second;
        bool dup = false;                                   // a repeated id is harmless (postgres.go:239)
        for (size_t j = 0; j < n_seen && !dup; j++) dup = seen_small[j] == slot;
        for (size_t j = 0; j < seen_big.size() && !dup; j++) dup = seen_big[j] == slot;
        if (dup) continue;
        if (n_seen < 64) seen_small[n_seen++] = slot; else seen_big.push_back(slot);
        const DocInfo &d = st->docs[slot];
        if (d.rows == 0) continue;
        if (d.fragmented) return 0;
        selected++;
        for (const auto &r : d.ranges) {
            if (fp.n_ranges == kMaxRanges) return 0;        // too many pieces for the range table: the bitmap paths
            ranges[fp.n_ranges++] = make_uint2(r.first, r.second);
            fp.total_rows += r.second;
        }
    }
    if (q8_ok && selected * 4 >= st->docs.size() && selected > 0) return 0;   // most of the corpus: the sidecar scan pays
    fp.nothing = fp.total_rows == 0;
    return 1;
}

// Bytes of the request block: the query, then the range table or (IVFFlat with a filter) the document bitmap.
size_t fast_request_bytes(const da_store *st, const FastPlan &fp)
{
    const size_t qbytes = (size_t)st->dim * 4;
    if (fp.ivf) return qbytes + (fp.bitmap ? ((st->docs.size() + 31) / 32 + 1) * 4 : 0);
    return qbytes + (fp.ranged ? (size_t)fp.n_ranges * 8 : 0);
}

// Fill the slot's pinned request block and queue its one host-to-device copy.
int upload_fast(da_store *st, Slot *sl, const FastPlan &fp, const da_search_params *p, const uint2 *ranges)
{
    const size_t qbytes = (size_t)st->dim * 4, bytes = fast_request_bytes(st, fp);
    const int rc = slot_reserve_request(*sl, bytes);
    if (rc != DA_OK) return rc;
    memcpy(sl->h_req, p->queries, qbytes);
    if (fp.ivf && fp.bitmap) {
        uint32_t *bm = reinterpret_cast<uint32_t *>(sl->h_req + qbytes);
        memset(bm, 0, bytes - qbytes);
        for (size_t i = 0; i < p->n_docs; i++) {
            auto it = st->doc_slot_of.find(load_id(p->doc_ids + 16 * i));
            if (it != st->doc_slot_of.end()) bm[it->second >> 5] |= 1u << (it->second & 31);
        }
    } else if (fp.ranged) {
        memcpy(sl->h_req + qbytes, ranges, (size_t)fp.n_ranges * 8);
    }
    CU(cudaMemcpyAsync(sl->d_req, sl->h_req, bytes, cudaMemcpyHostToDevice, sl->stream));
    return DA_OK;
}

// Queue the kernels of one planned search on a slot whose request block is (being) uploaded: ONE scan, or for
// IVFFlat the centroid scan and the probe-mode scan.  `flag` = the last kernel writes the completion word.
int enqueue_fast(da_store *st, Slot *sl, const FastPlan &fp, bool flag, int *grid_out, int *launches)
{
    Shard &s = st->shards[0];
    const size_t qbytes = (size_t)st->dim * 4;
    da::ScanArgs a{};
    a.mat = s.mat; a.inv_norm = s.inv_norm; a.doc_slot = s.doc_slot; a.chunk_ids = s.chunk_ids;
    a.doc_bitmap = nullptr;
    a.queries = reinterpret_cast<const float *>(sl->d_req); a.q_inv_norm = nullptr;
    a.n_rows = fp.total_rows; a.row_base = s.row_base;
    a.dim = st->dim; a.nq = 1; a.k = fp.k; a.min_sim = fp.min_sim; a.read_all = 0;
    a.cta_out = sl->d_cta; a.counter = sl->d_counter;
    if (fp.ivf) {
        const int L = st->ivf_lists;
        if (L <= da::kIvfSelectMaxLists && (size_t)8 * st->dim * 4 <= 200 * 1024) {
            if (sl->scores_cap < (size_t)L) {
                CU(cudaStreamSynchronize(sl->stream));
                cudaFree(sl->d_scores); sl->d_scores = nullptr; sl->scores_cap = 0;
                CU(cudaMalloc((void **)&sl->d_scores, (size_t)L * sizeof(float)));
                sl->scores_cap = (size_t)L;
            }
            CU(da::ivf_launch_rank_one_query(s.ivf.cent, L, st->dim, a.queries, sl->d_scores, fp.probes, sl->d_probe_keys, sl->d_counter,
                                             s.sm_count, sl->stream));
            (*launches)++;
        } else {
            da::ScanArgs c{};
            c.mat = s.ivf.cent; c.inv_norm = s.ivf.cent_inv;
            c.queries = a.queries; c.q_inv_norm = nullptr;
            c.n_rows = (uint32_t)L; c.row_base = 0;
            c.dim = st->dim; c.nq = 1; c.k = fp.probes; c.min_sim = -std::numeric_limits<float>::infinity();
            c.cta_out = sl->d_cta; c.counter = sl->d_counter; c.out_keys = sl->d_probe_keys;
            CU(da::launch_scan(c, s.sm_count, sl->stream, nullptr));
            (*launches)++;
        }
        a.probe_keys = sl->d_probe_keys; a.probes = fp.probes; a.probe_batch = 1;
        a.list_offsets = s.ivf.offsets; a.list_rows = s.ivf.list_rows;
        a.n_rows = 0;
        if (fp.bitmap) a.doc_bitmap = reinterpret_cast<const uint32_t *>(sl->d_req + qbytes);
    } else if (fp.ranged) {
        a.ranges = reinterpret_cast<const uint2 *>(sl->d_req + qbytes); a.probes = fp.n_ranges; a.probe_batch = 1;
    }
    a.out_ids = sl->d_out;
    a.out_keys = reinterpret_cast<da::key_t64 *>(sl->d_out + (size_t)fp.k * 16);
    if (fp.xchg) {      // the records go to this rank's slot of its exchange buffer; the merged answer to the mapped block
        da::XchgArgs &x = a.xg;
        const int N = st->total_ranks;
        for (int r = 0; r < N; r++) x.peer[r] = s.xg_peer[r];
        x.n_ranks = N; x.rank = st->rank; x.slot_bytes = kXchgSlotBytes;
        x.parity = (uint32_t)(st->xg_seq & 1);
        x.target = (unsigned long long)N * (st->xg_seq / 2 + 1);
        unsigned char *slot = s.xg_buf + 256 + ((size_t)x.parity * N + st->rank) * kXchgSlotBytes;
        x.final_ids = a.out_ids; x.final_keys = a.out_keys;
        a.out_ids = slot;
        a.out_keys = reinterpret_cast<da::key_t64 *>(slot + (size_t)fp.k * 16);
        x.timeout_flag = reinterpret_cast<uint32_t *>(sl->d_out + kSlotOutBytes - 24);
        st->xg_seq++;
    }
    a.rows_read = sl->d_rows_read;
    a.rows_read_out = reinterpret_cast<unsigned long long *>(sl->d_out + kSlotOutBytes - 16);
    if (flag) {
        a.done_flag = reinterpret_cast<volatile unsigned long long *>(sl->d_out + kSlotOutBytes - 8);
        a.done_value = sl->ticket;
    }
    CU(da::launch_scan(a, s.sm_count, sl->stream, grid_out));
    (*launches)++;
    return DA_OK;
}

// Algorithmic bytes of one small search that read `rows_read` rows.
uint64_t fast_bytes(const da_store *st, const FastPlan &fp, uint64_t rows_read)
{
    const uint64_t dim4 = (uint64_t)st->dim * 4;
    if (fp.ivf)     // the centroid matrix, then the probed rows: vector, 1/|row|, row number (+ document slot with a filter)
        return (uint64_t)st->ivf_lists * (dim4 + 4) + rows_read * (dim4 + (fp.bitmap ? 12 : 8));
    return rows_read * (dim4 + 4) + (uint64_t)fp.n_ranges * 8;      // no document slots, no bitmap on the range path
}

// Wait for the completion word of the slot's current ticket.
int wait_fast(Slot *sl)
{
    volatile unsigned long long *flag = reinterpret_cast<volatile unsigned long long *>(sl->h_out + kSlotOutBytes - 8);
    for (uint64_t spins = 1; *flag != sl->ticket; spins++) {
        _mm_pause();
        if ((spins & 0x3FFF) == 0) {               // now and then: did the stream fail, or finish without the word?
            const cudaError_t q = cudaStreamQuery(sl->stream);
            if (q == cudaSuccess) {
                if (*flag == sl->ticket) break;
                return fail(DA_ERR_CUDA, "the scan finished without delivering its result");
            }
            if (q != cudaErrorNotReady) return fail(DA_ERR_CUDA, "search failed on the device: %s", cudaGetErrorString(q));
            if (spins > (1u << 20)) std::this_thread::yield();
        }
    }
    std::atomic_thread_fence(std::memory_order_acquire);
    return DA_OK;
}

// The whole small search.  Returns 1 when it answered, 0 when the general path must, < 0 on error.
int search_fast(da_store *st, const da_search_params *p, uint8_t *out_chunk_ids, float *out_scores, uint32_t *out_rows,
                int32_t *out_n)
{
    std::shared_lock<std::shared_mutex> lock(st->mu);
    FastPlan fp;
    uint2 ranges[kMaxRanges];
    const int pr = plan_fast(st, p, fp, ranges);
    if (pr <= 0) return pr;
    out_n[0] = 0;
    if (fp.k == 0) return 1;
    if (!out_chunk_ids || !out_scores) return fail(DA_ERR_INVALID, "output buffers are NULL");
    const auto t0 = std::chrono::steady_clock::now();
    uint64_t rows_read = 0;
    int grid = 0, launches = 0;
    const unsigned long long *probe_src = nullptr;
    if (!fp.nothing) {
        Shard &s = st->shards[0];
        CU(cudaSetDevice(s.device));
        int rc = DA_OK;
        std::unique_lock<std::mutex> collective(st->xg_mu, std::defer_lock);
        if (fp.xchg) collective.lock();                  // searches of a multi-process store are collective: one at a time, in call order
        Slot *sl = acquire_slot(st, &rc);
        if (!sl) return rc;
        sl->ticket++;
        *reinterpret_cast<volatile uint32_t *>(sl->h_out + kSlotOutBytes - 24) = 0u;
        rc = upload_fast(st, sl, fp, p, ranges);
        if (rc == DA_OK) rc = enqueue_fast(st, sl, fp, true, &grid, &launches);
        if (rc == DA_OK) rc = wait_fast(sl);
        if (rc == DA_OK && fp.xchg && *reinterpret_cast<volatile uint32_t *>(sl->h_out + kSlotOutBytes - 24) != 0u)
            rc = fail(DA_ERR_STATE, "a rank did not deliver its candidates within 4 s (searches on a multi-process store are collective)");
        if (rc != DA_OK) { cudaStreamSynchronize(sl->stream); release_slot(st, sl); return rc; }
        const uint8_t *ids = sl->h_out;
        const da::key_t64 *keys = reinterpret_cast<const da::key_t64 *>(sl->h_out + (size_t)fp.k * 16);
        int c = 0;
        for (int i = 0; i < fp.k; i++) {
            const da::key_t64 key = keys[i];
            if (!key) break;
            out_scores[c] = da::key_score(key);
            if (out_rows) out_rows[c] = da::key_row(key);
            memcpy(out_chunk_ids + (size_t)c * 16, ids + (size_t)i * 16, 16);
            c++;
        }
        out_n[0] = c;
        rows_read = *reinterpret_cast<const unsigned long long *>(sl->h_out + kSlotOutBytes - 16);
        probe_src = sl->d_probe_keys;
        release_slot(st, sl);
    }
    const float ms = std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - t0).count();
    {
        std::lock_guard<std::mutex> sl(st->stats_mu);
        st->fast_searches++;
        st->last_search_ms = ms; st->last_scan_ms = ms;         // host clock around copy + scan + pick-up on this path
        st->last_rows_read = rows_read;
        st->last_bytes = fast_bytes(st, fp, rows_read);
        if (!fp.nothing) { st->launches += (uint64_t)launches; st->scan_grid = grid; }
        if (fp.ivf && !fp.nothing) { st->last_probe_keys = probe_src; st->last_probe_count = (size_t)fp.probes; }
    }
    return 1;
}

}  // namespace

namespace da {
// Lets other translation units of the library report through da_last_error().
void set_last_error(const std::string &msg) { g_err = msg; }
}  // namespace da

// ---------------------------------------------------------------------------
// C ABI
// ---------------------------------------------------------------------------
extern "C" {

const char *da_last_error(void) { return g_err.c_str(); }
int da_abi_version(void) { return DA_ABI_VERSION; }

int da_nccl_unique_id(uint8_t out[DA_NCCL_ID_BYTES])
{
    static_assert(sizeof(ncclUniqueId) == DA_NCCL_ID_BYTES, "ncclUniqueId size");
    if (!out) return fail(DA_ERR_INVALID, "out is NULL");
    ncclUniqueId id;
    NC(ncclGetUniqueId(&id));
    memcpy(out, &id, sizeof id);
    return DA_OK;
}

int da_store_create(const da_config *cfg, da_store **out)
{
    if (!cfg || !out) return fail(DA_ERR_INVALID, "cfg/out is NULL");
    *out = nullptr;
    if (cfg->struct_size != sizeof(da_config)) return fail(DA_ERR_INVALID, "da_config.struct_size=%u, expected %zu", cfg->struct_size, sizeof(da_config));
    if (cfg->dim <= 0 || (cfg->dim & 3)) return fail(DA_ERR_DIM, "dim=%d must be a positive multiple of 4", cfg->dim);
    if (cfg->dim > 16384) return fail(DA_ERR_DIM, "dim=%d exceeds 16384", cfg->dim);
    if ((cfg->flags & DA_STORE_Q8_SIDECAR) && (cfg->dim % 32)) return fail(DA_ERR_DIM, "the int8 sidecar needs dim %% 32 == 0 (dim=%d)", cfg->dim);
    if ((cfg->flags & DA_STORE_Q8_SIDECAR) && (cfg->flags & DA_STORE_SPLIT16)) return fail(DA_ERR_INVALID, "DA_STORE_Q8_SIDECAR and DA_STORE_SPLIT16 exclude each other");
    if ((cfg->flags & DA_STORE_SPLIT16) && (cfg->dim % 16)) return fail(DA_ERR_DIM, "the split-plane layout needs dim %% 16 == 0 (dim=%d)", cfg->dim);
    if (cfg->n_devices < 1 || cfg->n_devices > DA_MAX_DEVICES) return fail(DA_ERR_INVALID, "n_devices=%d out of range", cfg->n_devices);
    if (cfg->capacity_rows == 0) return fail(DA_ERR_INVALID, "capacity_rows is 0");
    const int world = cfg->world_size > 1 ? cfg->world_size : 1;
    const int rank = world > 1 ? cfg->rank : 0;
    if (rank < 0 || rank >= world) return fail(DA_ERR_INVALID, "rank=%d outside world_size=%d", rank, world);
    const uint64_t total = (uint64_t)world * cfg->n_devices;
    if (total * cfg->capacity_rows > 0xFFFFFFF0ull) return fail(DA_ERR_CAPACITY, "total capacity exceeds 2^32 rows");
    int n_dev = 0;
    cudaError_t ce = cudaGetDeviceCount(&n_dev);
    if (ce != cudaSuccess || n_dev == 0)
        return fail(DA_ERR_CUDA, "no CUDA device: %s (this library has no CPU fallback)", cudaGetErrorString(ce));
    for (int i = 0; i < cfg->n_devices; i++)
        if (cfg->device_ids[i] < 0 || cfg->device_ids[i] >= n_dev) return fail(DA_ERR_INVALID, "device id %d not present (%d devices)", cfg->device_ids[i], n_dev);

    da_store *st = new da_store();
    st->dim = cfg->dim;
    st->min_sim = cfg->min_similarity_f64 != 0.0 ? float_at_least(cfg->min_similarity_f64) : cfg->min_similarity;
    st->world = world; st->rank = rank; st->total_ranks = (int)total;
    st->batch_window_us = cfg->batch_window_us;
    st->split16 = (cfg->flags & DA_STORE_SPLIT16) != 0;
    st->q8 = (cfg->flags & DA_STORE_Q8_SIDECAR) != 0;
    st->batch_max = cfg->batch_max ? std::min<uint32_t>(cfg->batch_max, DA_MAX_BATCH) : 64;
    st->shards.resize(cfg->n_devices);
    for (int i = 0; i < cfg->n_devices; i++) {
        int rc = init_shard(st->shards[i], cfg->device_ids[i], cfg->dim, cfg->capacity_rows, st->split16, st->q8);
        if (rc != DA_OK) { da_store_destroy(st); return rc; }
        st->shards[i].row_base = (uint32_t)(((uint64_t)rank * cfg->n_devices + i) * cfg->capacity_rows);
    }
    if (world == 1 && cfg->n_devices > 1 && !getenv("DA_DISABLE_P2P")) {
        // One process driving several GPUs (what the Go query agent does): if every shard can store into
        // shard 0's memory, the scan epilogue delivers its records there directly and no collective is needed.
        bool ok = true;
        const int d0 = cfg->device_ids[0];
        for (int i = 1; i < cfg->n_devices && ok; i++) {
            int can = 0;
            ok = cudaDeviceCanAccessPeer(&can, cfg->device_ids[i], d0) == cudaSuccess && can;
            if (!ok) break;
            cudaSetDevice(cfg->device_ids[i]);
            cudaError_t e = cudaDeviceEnablePeerAccess(d0, 0);
            if (e == cudaErrorPeerAccessAlreadyEnabled) { cudaGetLastError(); e = cudaSuccess; }
            ok = (e == cudaSuccess);
        }
        st->p2p_gather = ok;
        if (ok) {
            int rc = setup_inprocess_exchange(st);
            if (rc != DA_OK) { da_store_destroy(st); return rc; }
        }
    }
    if (total > 1) {
        int rc = DA_OK;
        if (world == 1) {
            std::vector<ncclComm_t> comms(cfg->n_devices);
            std::vector<int> devs(cfg->device_ids, cfg->device_ids + cfg->n_devices);
            ncclResult_t r = ncclCommInitAll(comms.data(), cfg->n_devices, devs.data());
            if (r != ncclSuccess) rc = fail(DA_ERR_NCCL, "ncclCommInitAll failed: %s", ncclGetErrorString(r));
            else for (int i = 0; i < cfg->n_devices; i++) st->shards[i].comm = comms[i];
        } else {
            ncclUniqueId id;
            memcpy(&id, cfg->nccl_id, sizeof id);
            ncclResult_t r = ncclGroupStart();
            for (int i = 0; i < cfg->n_devices && r == ncclSuccess; i++) {
                cudaSetDevice(cfg->device_ids[i]);
                r = ncclCommInitRank(&st->shards[i].comm, (int)total, id, rank * cfg->n_devices + i);
            }
            ncclResult_t r2 = ncclGroupEnd();
            if (r == ncclSuccess) r = r2;
            if (r != ncclSuccess) rc = fail(DA_ERR_NCCL, "ncclCommInitRank failed: %s", ncclGetErrorString(r));
        }
        if (rc != DA_OK) { da_store_destroy(st); return rc; }
        if (world > 1 && cfg->n_devices == 1) {
            rc = setup_fused_exchange(st);
            if (rc != DA_OK) { da_store_destroy(st); return rc; }
        }
    }
    *out = st;
    return DA_OK;
}

void da_store_destroy(da_store *st)
{
    if (!st) return;
    if (!st->shards.empty()) cudaSetDevice(st->shards[0].device);
    for (auto &sl : st->slots) free_slot(*sl);
    for (Shard &s : st->shards) free_shard(s);
    shm_close(st);
    delete st;
}

int da_store_save_chunks(da_store *st, const uint8_t doc_id[16], const uint8_t *chunk_ids, size_t n)
{
    if (!st || !doc_id) return fail(DA_ERR_INVALID, "store/doc_id is NULL");
    if (n && !chunk_ids) return fail(DA_ERR_INVALID, "chunk_ids is NULL");
    if (st->attached) return fail(DA_ERR_STATE, "an attached store is read-only: chunks are announced in the exporting process");
    std::unique_lock<std::shared_mutex> lock(st->mu);
    ShmWriteGuard shared_writer(st);
    const Id16 did = load_id(doc_id);
    uint32_t slot;
    auto it = st->doc_slot_of.find(did);
    if (it == st->doc_slot_of.end()) {
        slot = (uint32_t)st->docs.size();
        st->docs.push_back(DocInfo{did, -1, 0, false, {}});
        st->doc_slot_of.emplace(did, slot);
        shm_append(st, kRecDoc, slot, 0, 0, &did);
    } else slot = it->second;
    for (size_t i = 0; i < n; i++) {
        const Id16 cid = load_id(chunk_ids + 16 * i);
        auto c = st->chunks.find(cid);
        if (c == st->chunks.end()) { st->chunks.emplace(cid, ChunkInfo{slot, -1, kNoRow}); continue; }
        if (c->second.slot == slot) continue;
        if (c->second.row != kNoRow) {
            st->docs[c->second.slot].rows--; st->docs[slot].rows++;
            st->docs[c->second.slot].fragment(); st->docs[slot].fragment();      // their rows are no longer whole runs
            shm_append(st, kRecFragment, c->second.slot, 0, st->docs[c->second.slot].rows);
            shm_append(st, kRecFragment, slot, 0, st->docs[slot].rows);
        }
        c->second.slot = slot;
        if (c->second.row != kNoRow) {   // the chunk already has a vector: move its row to the new document
            Shard &s = st->shards[c->second.shard];
            CU(cudaSetDevice(s.device));
            CU(cudaMemcpyAsync(s.doc_slot + c->second.row, &slot, 4, cudaMemcpyHostToDevice, s.stream));
            CU(cudaStreamSynchronize(s.stream));
        }
    }
    shm_publish(st);
    return DA_OK;
}

int da_store_save_embeddings(da_store *st, const uint8_t *chunk_ids, const float *vecs, size_t n, int32_t dim)
{
    if (!st) return fail(DA_ERR_INVALID, "store is NULL");
    if (n == 0) return DA_OK;                                   // postgres.go:177-179
    if (!chunk_ids || !vecs) return fail(DA_ERR_INVALID, "chunk_ids/vecs is NULL");
    if (dim != st->dim) return fail(DA_ERR_DIM, "expected %d dimensions, not %d", st->dim, dim);
    if (st->attached) return fail(DA_ERR_STATE, "an attached store is read-only: embeddings are saved in the exporting process");
    std::unique_lock<std::shared_mutex> lock(st->mu);
    ShmWriteGuard shared_writer(st);

    // Validate everything before the store changes: the reference's statement is atomic.  Ids on the host,
    // vector elements on the device (staged, checked, and only then committed).
    std::vector<ChunkInfo *> info(n);
    {
        std::unordered_map<Id16, size_t, Id16Hash> seen;
        seen.reserve(n * 2);
        for (size_t i = 0; i < n; i++) {
            const Id16 cid = load_id(chunk_ids + 16 * i);
            if (!seen.emplace(cid, i).second)
                return fail(DA_ERR_INVALID, "ON CONFLICT DO UPDATE command cannot affect row a second time (embedding %zu repeats a chunk id)", i);
            auto c = st->chunks.find(cid);
            if (c == st->chunks.end())
                return fail(DA_ERR_UNKNOWN_CHUNK, "embedding %zu violates foreign key: chunk not announced by save_chunks", i);
            info[i] = &c->second;
        }
    }
    // Placement: a document lives on one shard (keeps its rows together); new documents go to
    // the emptiest shard.
    const int S = (int)st->shards.size();
    std::vector<uint32_t> pending(S, 0);
    std::vector<int32_t> tgt_shard(n);
    std::vector<uint32_t> tgt_row(n);
    std::vector<std::pair<uint32_t, int32_t>> doc_assign;       // undo log if capacity check fails
    for (size_t i = 0; i < n; i++) {
        ChunkInfo &ci = *info[i];
        if (ci.row != kNoRow) { tgt_shard[i] = ci.shard; tgt_row[i] = ci.row; continue; }
        DocInfo &d = st->docs[ci.slot];
        if (d.shard < 0) {
            int best = 0;
            for (int s = 1; s < S; s++)
                if (st->shards[s].n_rows + pending[s] < st->shards[best].n_rows + pending[best]) best = s;
            d.shard = best;
            doc_assign.emplace_back(ci.slot, -1);
        }
        tgt_shard[i] = d.shard;
        tgt_row[i] = st->shards[d.shard].n_rows + pending[d.shard]++;
    }
    for (int s = 0; s < S; s++)
        if ((uint64_t)st->shards[s].n_rows + pending[s] > st->shards[s].cap_rows) {
            for (auto &u : doc_assign) st->docs[u.first].shard = u.second;
            return fail(DA_ERR_CAPACITY, "shard %d is full (%llu rows); raise capacity_rows (no CPU fallback)", s,
                        (unsigned long long)st->shards[s].cap_rows);
        }

    // Vectors travel host -> staging buffer -> (finite check) -> final rows.  A call that fits one staging chunk is
    // uploaded once; a larger one is streamed twice (check pass, commit pass) so that a bad element still leaves
    // the store untouched.
    size_t kChunkRows = std::max<size_t>(1, ((size_t)256 << 20) / ((size_t)dim * 4));
    if (const char *e = getenv("DA_UPSERT_CHUNK_ROWS")) kChunkRows = std::max<size_t>(1, (size_t)atoll(e));   // test knob
    std::vector<std::vector<size_t>> idx(S);
    for (size_t i = 0; i < n; i++) idx[tgt_shard[i]].push_back(i);
    auto upload = [&](Shard &sh, const std::vector<size_t> &ix, size_t b, size_t e) -> int {
        for (size_t j = b; j < e;) {                    // runs of consecutive inputs -> one copy each
            size_t r = j + 1;
            while (r < e && ix[r] == ix[r - 1] + 1) r++;
            CU(cudaMemcpyAsync(sh.d_up_stage.p + (j - b) * (size_t)dim, vecs + ix[j] * (size_t)dim, (r - j) * (size_t)dim * 4,
                               cudaMemcpyHostToDevice, sh.stream));
            j = r;
        }
        return DA_OK;
    };
    for (int s = 0; s < S; s++) {                       // ---- check pass
        Shard &sh = st->shards[s];
        const size_t cnt = idx[s].size();
        if (!cnt) continue;
        CU(cudaSetDevice(sh.device));
        CU(sh.d_up_stage.ensure(std::min(cnt, kChunkRows) * (size_t)dim));
        CU(sh.d_up_bad.ensure(1)); CU(sh.h_up_bad.ensure(1));
        CU(cudaMemsetAsync(sh.d_up_bad.p, 0xFF, sizeof(unsigned long long), sh.stream));
        for (size_t b = 0; b < cnt; b += kChunkRows) {
            const size_t e = std::min(cnt, b + kChunkRows);
            int rc = upload(sh, idx[s], b, e);
            if (rc != DA_OK) return rc;
            CU(da::launch_validate_stage(sh.d_up_stage.p, (e - b) * (uint64_t)dim, b * (uint64_t)dim, sh.d_up_bad.p, sh.stream));
            st->launches++;
        }
        CU(cudaMemcpyAsync(sh.h_up_bad.p, sh.d_up_bad.p, sizeof(unsigned long long), cudaMemcpyDeviceToHost, sh.stream));
    }
    for (int s = 0; s < S; s++) {
        Shard &sh = st->shards[s];
        if (idx[s].empty()) continue;
        CU(cudaSetDevice(sh.device));
        CU(cudaStreamSynchronize(sh.stream));
        if (*sh.h_up_bad.p != ~0ull) {
            for (auto &u : doc_assign) st->docs[u.first].shard = u.second;
            const size_t j = (size_t)(*sh.h_up_bad.p / (uint64_t)dim);
            return fail(DA_ERR_NONFINITE, "element %llu of embedding %zu is not finite",
                        (unsigned long long)(*sh.h_up_bad.p % (uint64_t)dim), idx[s][j]);
        }
    }
    for (int s = 0; s < S; s++) {                       // ---- commit pass
        Shard &sh = st->shards[s];
        const size_t cnt = idx[s].size();
        if (!cnt) continue;
        CU(cudaSetDevice(sh.device));
        CU(sh.d_up_rows.ensure(cnt)); CU(sh.d_up_slots.ensure(cnt)); CU(sh.d_up_ids.ensure(cnt * 16));
        CU(sh.h_up_rows.ensure(cnt)); CU(sh.h_up_slots.ensure(cnt)); CU(sh.h_up_ids.ensure(cnt * 16));
        for (size_t j = 0; j < cnt; j++) {
            const size_t t = idx[s][j];
            sh.h_up_rows.p[j] = tgt_row[t];
            sh.h_up_slots.p[j] = info[t]->slot;
            memcpy(sh.h_up_ids.p + 16 * j, chunk_ids + 16 * t, 16);
        }
        CU(cudaMemcpyAsync(sh.d_up_rows.p, sh.h_up_rows.p, cnt * 4, cudaMemcpyHostToDevice, sh.stream));
        CU(cudaMemcpyAsync(sh.d_up_slots.p, sh.h_up_slots.p, cnt * 4, cudaMemcpyHostToDevice, sh.stream));
        CU(cudaMemcpyAsync(sh.d_up_ids.p, sh.h_up_ids.p, cnt * 16, cudaMemcpyHostToDevice, sh.stream));
        for (size_t b = 0; b < cnt; b += kChunkRows) {
            const size_t e = std::min(cnt, b + kChunkRows);
            if (cnt > kChunkRows) {                     // the staging buffer was reused by the check pass: stream again
                int rc = upload(sh, idx[s], b, e);
                if (rc != DA_OK) return rc;
            }
            if (st->split16)
                CU(da::launch_commit_rows_split(sh.d_up_stage.p, dim, sh.d_up_rows.p + b, sh.d_up_slots.p + b, sh.d_up_ids.p + 16 * b,
                                                (uint32_t)(e - b), sh.hi, sh.lo, sh.inv_norm, sh.doc_slot, sh.chunk_ids, sh.stream));
            else
                CU(da::launch_commit_rows(sh.d_up_stage.p, dim, sh.d_up_rows.p + b, sh.d_up_slots.p + b, sh.d_up_ids.p + 16 * b,
                                          (uint32_t)(e - b), sh.mat, sh.inv_norm, sh.doc_slot, sh.chunk_ids, sh.stream));
            st->launches++;
            if (st->q8) {
                CU(da::launch_q8_quantize_rows(sh.mat, dim, sh.d_up_rows.p + b, 0, (uint32_t)(e - b), sh.inv_norm, sh.q8,
                                               sh.q8_factor, sh.q8_err, sh.q8_errmax, sh.stream));
                sh.q8_dirty = true;
                st->launches++;
            }
        }
        if (st->ivf_lists) {   // keep the index current, as pgvector does on insert: nearest list for the touched rows
            CU(da::ivf_launch_assign(sh.mat, dim, sh.d_up_rows.p, (uint32_t)cnt, sh.ivf.cent, st->ivf_lists, sh.ivf.assign, 1, sh.stream));
            CU(da::ivf_launch_build_lists(sh.ivf.assign, sh.n_rows + pending[s], st->ivf_lists, sh.ivf.counts, sh.ivf.offsets,
                                          sh.ivf.cursor, sh.ivf.list_rows, sh.stream));
            st->launches += 4;
        }
    }
    for (int s = 0; s < S; s++) {
        Shard &sh = st->shards[s];
        CU(cudaSetDevice(sh.device));
        CU(cudaStreamSynchronize(sh.stream));
        sh.n_rows += pending[s];
    }
    uint32_t run_slot = 0, run_first = 0, run_count = 0;
    for (size_t i = 0; i < n; i++) {
        if (info[i]->row == kNoRow) {                                        // a new embedding (an overwrite keeps its row)
            st->docs[info[i]->slot].rows++;
            st->docs[info[i]->slot].add_row(tgt_row[i]);
            // shared log: runs of new rows of one document (one record per document in the usual call)
            if (run_count && run_slot == info[i]->slot && run_first + run_count == tgt_row[i]) run_count++;
            else {
                if (run_count) shm_append(st, kRecRows, run_slot, run_first, run_count);
                run_slot = info[i]->slot; run_first = tgt_row[i]; run_count = 1;
            }
        }
        info[i]->shard = tgt_shard[i]; info[i]->row = tgt_row[i];
    }
    if (run_count) shm_append(st, kRecRows, run_slot, run_first, run_count);
    shm_publish(st);
    return DA_OK;
}

int da_store_bulk_load(da_store *st, const uint8_t *chunk_ids, const uint8_t *doc_ids, const float *vecs, size_t n, int32_t dim)
{
    if (!st) return fail(DA_ERR_INVALID, "store is NULL");
    if (n == 0) return DA_OK;
    if (!chunk_ids || !doc_ids || !vecs) return fail(DA_ERR_INVALID, "chunk_ids/doc_ids/vecs is NULL");
    if (dim != st->dim) return fail(DA_ERR_DIM, "expected %d dimensions, not %d", st->dim, dim);
    // announce the chunks document by document (runs of equal document ids), then one upsert for all rows
    for (size_t i = 0; i < n;) {
        size_t e = i + 1;
        while (e < n && memcmp(doc_ids + 16 * e, doc_ids + 16 * i, 16) == 0) e++;
        int rc = da_store_save_chunks(st, doc_ids + 16 * i, chunk_ids + 16 * i, e - i);
        if (rc != DA_OK) return rc;
        i = e;
    }
    return da_store_save_embeddings(st, chunk_ids, vecs, n, dim);
}

int da_store_search(da_store *st, const da_search_params *p, uint8_t *out_chunk_ids, float *out_scores,
                    uint32_t *out_rows, int32_t *out_n)
{
    if (!st) return fail(DA_ERR_INVALID, "store is NULL");
    if (!out_n) return fail(DA_ERR_INVALID, "out_n is NULL");
    ShmReadGuard shared_reader(st);                // an attached store: the exporting process is not mid-write while we look
    {
        const int rr = shm_refresh(st);            // ... and what it has saved since our last search becomes visible
        if (rr != DA_OK) return rr;
    }
    {
        const int fr = search_fast(st, p, out_chunk_ids, out_scores, out_rows, out_n);     // one query, one shard: the short way
        if (fr != 0) return fr < 0 ? fr : DA_OK;
    }
    std::unique_lock<std::shared_mutex> lock(st->mu);
    Shard &s0 = st->shards[0];
    SearchPlan pl{};
    CU(cudaSetDevice(s0.device));
    CU(cudaEventRecord(s0.ev_all0, s0.stream));
    const double t_begin = now_ms();
    int rc = stage_search(st, p, pl);
    const double t_staged = now_ms();
    double t_waited = t_staged;
    if (rc != DA_OK) return rc;
    for (int q = 0; q < pl.batch; q++) out_n[q] = 0;
    if (pl.k == 0) return DA_OK;
    if (!out_chunk_ids || !out_scores) return fail(DA_ERR_INVALID, "output buffers are NULL");
    const size_t nk = (size_t)pl.batch * pl.k, rec_bytes = nk * 24;
    for (int attempt = 0; attempt < 3; attempt++) {
        rc = enqueue_search(st, pl);
        if (rc != DA_OK) return rc;
        CU(cudaSetDevice(s0.device));
        const uint8_t *src = st->total_ranks > 1 ? s0.d_final.p : s0.d_out.p;
        const size_t trailer = st->world > 1 ? kRecTrailer : 0;
        CU(cudaMemcpyAsync(s0.h_result.p, src, rec_bytes + trailer, cudaMemcpyDeviceToHost, s0.stream));
        CU(cudaEventRecord(s0.ev_all1, s0.stream));
        rc = enqueue_stats_readback(st);
        if (rc != DA_OK) return rc;
        for (Shard &s : st->shards) { CU(cudaSetDevice(s.device)); CU(cudaStreamSynchronize(s.stream)); }
        t_waited = now_ms();
        bool overflow = false;
        if (trailer && *reinterpret_cast<const uint32_t *>(s0.h_result.p + rec_bytes + 4) != 0)
            return fail(DA_ERR_STATE, "a rank did not deliver its candidates within 4 s (searches on a multi-process store are collective)");
        if (st->xg_inproc)
            for (Shard &s : st->shards)
                if (*reinterpret_cast<volatile uint32_t *>(s.h_xg_timeout) != 0u) {
                    *reinterpret_cast<volatile uint32_t *>(s.h_xg_timeout) = 0u;
                    return fail(DA_ERR_STATE, "a shard did not deliver its candidates within 4 s");
                }
        if (trailer)         // several processes: the OR of every rank's flag, the same word on every rank
            overflow = *reinterpret_cast<const uint32_t *>(s0.h_result.p + rec_bytes) != 0;
        else if (pl.use_tc || pl.use_q8 || st->split16)
            for (Shard &s : st->shards) overflow = overflow || (*s.h_overflow.p != 0);
        if (!overflow) break;
        if (st->split16 && !pl.use_tc)
            return fail(DA_ERR_STATE, "split-plane search: more than 2^20 rows lie within the truncation slack of the k-th best");
        // A candidate list stayed full after pruning (a mass of scores within the TF32 slack of the k-th best) on
        // this rank or, in a multi-process store, on any rank: every rank answers the batch again another way
        // (the collective call sequence stays the same on all of them: one gather per attempt).
        st->tc_overflows++;
        if (pl.use_tc && pl.tc_i8) {
            pl.tc_i8 = false;              // int8 slack too wide for this batch: the TF32 contraction of the fp32 rows next
        } else {
            pl.use_tc = false;             // answer this batch with the SIMT scan instead
            pl.tc_i8 = false;
        }
        pl.use_q8 = false;                 // ... of the fp32 rows
        for (Shard &s : st->shards) {      // the flag is sticky on the device: clear it for the second attempt
            CU(cudaSetDevice(s.device));
            CU(cudaMemsetAsync(s.d_overflow.p, 0, 4, s.stream));
            if (trailer) CU(cudaMemsetAsync(s.d_out.p + rec_bytes, 0, trailer, s.stream));
            *s.h_overflow.p = 0;
        }
    }
    const uint8_t *ids = s0.h_result.p;
    const da::key_t64 *keys = reinterpret_cast<const da::key_t64 *>(s0.h_result.p + nk * 16);
    // records are best first and zero-padded: a query's hits are a prefix of its k slots
    auto unpack = [&](int q_lo, int q_hi) {
        for (int q = q_lo; q < q_hi; q++) {
            const da::key_t64 *kq = keys + (size_t)q * pl.k;
            int c = 0;
            while (c < pl.k && kq[c]) c++;
            float *sc = out_scores + (size_t)q * pl.k;
            for (int i = 0; i < c; i++) sc[i] = da::key_score(kq[i]);
            if (out_rows) {
                uint32_t *rw = out_rows + (size_t)q * pl.k;
                for (int i = 0; i < c; i++) rw[i] = da::key_row(kq[i]);
            }
            memcpy(out_chunk_ids + (size_t)q * pl.k * 16, ids + (size_t)q * pl.k * 16, (size_t)c * 16);
            out_n[q] = c;
        }
    };
    if (nk >= (1u << 16)) {             // a large batch: four threads share the queries
        std::vector<std::thread> th;
        const int per = (pl.batch + 3) / 4;
        for (int t = 0; t < 4; t++) th.emplace_back(unpack, std::min(pl.batch, t * per), std::min(pl.batch, (t + 1) * per));
        for (auto &t : th) t.join();
    } else {
        unpack(0, pl.batch);
    }
    if (trace_search())
        fprintf(stderr, "da_store_search trace: batch %d k %d: stage %.3f ms, enqueue+device+readback %.3f ms, unpack %.3f ms\n", pl.batch, pl.k,
                t_staged - t_begin, t_waited - t_staged, now_ms() - t_waited);
    CU(cudaSetDevice(s0.device));
    float ms = 0;
    CU(cudaEventElapsedTime(&ms, s0.ev_all0, s0.ev_all1));
    st->last_search_ms = ms;
    CU(cudaEventElapsedTime(&ms, s0.ev_scan0, s0.ev_scan1));
    st->last_scan_ms = ms;
    rc = collect_stats(st, pl);
    if (rc != DA_OK) return rc;
    st->scan_ms_sum += st->last_scan_ms;
    return DA_OK;
}

static int topk_direct(da_store *st, const uint8_t *doc_ids, size_t n_docs, const float *queries, int32_t batch,
                       int32_t dim, int32_t k, uint8_t *out_chunk_ids, float *out_scores, int32_t *out_n,
                       const uint64_t *doc_offsets = nullptr)
{
    da_search_params p{};
    p.struct_size = sizeof p;
    p.flags = 0;
    p.doc_ids = doc_ids; p.n_docs = n_docs; p.doc_offsets = doc_offsets;
    p.queries = queries; p.batch = batch; p.dim = dim;
    p.k = k;
    p.min_similarity = st->min_sim;
    return da_store_search(st, &p, out_chunk_ids, out_scores, nullptr, out_n);
}

int da_store_topk(da_store *st, const uint8_t *doc_ids, size_t n_docs, const float *query, int32_t dim,
                  int32_t k, uint8_t *out_chunk_ids, float *out_scores, int32_t *out_n)
{
    if (!st) return fail(DA_ERR_INVALID, "store is NULL");
    if (st->batch_window_us == 0 || k <= 0 || k > DA_MAX_K || dim != st->dim || !query || !out_n ||
        !out_chunk_ids || !out_scores || (n_docs && !doc_ids)) {
        {
            std::lock_guard<std::mutex> lk(st->bmu);
            st->topk_calls++; st->topk_passes++;
        }
        return topk_direct(st, doc_ids, n_docs, query, 1, dim, k, out_chunk_ids, out_scores, out_n);
    }
    // A call is checked on its own before it may join a batch, so that a bad request fails alone (pgvector rejects
    // NaN/Inf for the statement that carries them, not for its neighbours).
    for (int32_t i = 0; i < dim; i++)
        if (!std::isfinite(query[i])) {
            { std::lock_guard<std::mutex> lk(st->bmu); st->topk_calls++; }
            return fail(DA_ERR_NONFINITE, "query element %d is not finite", (int)i);
        }
    // A search that reads little (the reference's own regime: a handful of documents) gains nothing from sharing a pass
    // and would only pay the window: it takes the small-search path at once, on its own slot, concurrently with the
    // others.  The window is for searches that stream a large part of the corpus, where one pass can serve them all.
    {
        da_search_params sp{};
        sp.struct_size = sizeof sp;
        sp.doc_ids = doc_ids; sp.n_docs = n_docs; sp.queries = query; sp.batch = 1; sp.dim = dim; sp.k = k;
        sp.min_similarity = st->min_sim;
        bool light = false;
        {
            ShmReadGuard shared_reader(st);
            if (shm_refresh(st) == DA_OK) {
                std::shared_lock<std::shared_mutex> shared(st->mu);
                FastPlan fp;
                uint2 ranges[kMaxRanges];
                light = plan_fast(st, &sp, fp, ranges) == 1 && (uint64_t)fp.total_rows * (uint64_t)dim * 4 <= kBatchWorthBytes;
            }
        }
        if (light) {
            { std::lock_guard<std::mutex> lk(st->bmu); st->topk_calls++; st->topk_passes++; }
            return topk_direct(st, doc_ids, n_docs, query, 1, dim, k, out_chunk_ids, out_scores, out_n);
        }
    }
    // Request batching: the first caller to find no leader becomes one, waits for the window (or a full
    // batch), then answers every queued call that asks for the same document set with one batched pass.
    da_store::Pending me{doc_ids, n_docs, query, k, out_chunk_ids, out_scores, out_n};
    std::unique_lock<std::mutex> lk(st->bmu);
    st->bqueue.push_back(&me);
    st->topk_calls++;
    st->bcv.notify_all();
    while (!me.done) {
        if (st->bleader) { st->bcv.wait(lk); continue; }
        st->bleader = true;
        const auto deadline = std::chrono::steady_clock::now() + std::chrono::microseconds(st->batch_window_us);
        while (st->bqueue.size() < st->batch_max && st->bcv.wait_until(lk, deadline) != std::cv_status::timeout) {}
        // group = the oldest queued calls, each keeping its own document set (per-query masks on the device)
        std::vector<da_store::Pending *> group, rest;
        for (da_store::Pending *r : st->bqueue) {
            if (group.size() < st->batch_max) group.push_back(r); else rest.push_back(r);
        }
        st->bqueue.swap(rest);
        st->topk_passes++;
        lk.unlock();
        int kmax = 0;
        size_t total_docs = 0;
        for (da_store::Pending *r : group) { kmax = std::max(kmax, r->k); total_docs += r->n_docs; }
        const size_t g = group.size();
        std::vector<float> qs(g * (size_t)dim), sc(g * (size_t)kmax);
        std::vector<uint8_t> ids(g * (size_t)kmax * 16), docs(total_docs * 16 + 16);
        std::vector<uint64_t> offs(g + 1, 0);
        std::vector<int32_t> cnt(g);
        for (size_t i = 0; i < g; i++) {
            memcpy(&qs[i * dim], group[i]->query, (size_t)dim * 4);
            if (group[i]->n_docs) memcpy(&docs[16 * offs[i]], group[i]->doc_ids, 16 * group[i]->n_docs);
            offs[i + 1] = offs[i] + group[i]->n_docs;
        }
        const int rc = topk_direct(st, docs.data(), total_docs, qs.data(), (int32_t)g, dim, kmax, ids.data(), sc.data(),
                                   cnt.data(), offs.data());
        const std::string err = rc == DA_OK ? std::string() : g_err;
        uint64_t extra_passes = 0;
        for (size_t i = 0; i < g; i++) {            // a smaller k is a prefix of the batch's top-kmax
            da_store::Pending *r = group[i];
            r->rc = rc; r->err = err;
            if (rc != DA_OK && g > 1) {
                // the merged pass failed: every member is answered on its own, so each gets its own verdict and a
                // fault that belongs to one request (or to the batched path) does not fail the others
                r->rc = topk_direct(st, r->doc_ids, r->n_docs, r->query, 1, dim, r->k, r->out_ids, r->out_scores, r->out_n);
                r->err = r->rc == DA_OK ? std::string() : g_err;
                extra_passes++;
                continue;
            }
            if (rc == DA_OK) {
                const int n = std::min(cnt[i], r->k);
                memcpy(r->out_ids, &ids[i * (size_t)kmax * 16], (size_t)n * 16);
                memcpy(r->out_scores, &sc[i * (size_t)kmax], (size_t)n * 4);
                *r->out_n = n;
            }
        }
        lk.lock();
        st->topk_passes += extra_passes;
        for (da_store::Pending *r : group) r->done = true;
        st->bleader = false;
        st->bcv.notify_all();
    }
    if (me.rc != DA_OK) g_err = me.err;
    return me.rc;
}

int da_store_bench_resident(da_store *st, const da_search_params *p, int32_t iters, float *out_total_ms)
{
    if (!st || !out_total_ms) return fail(DA_ERR_INVALID, "store/out is NULL");
    if (iters < 1) return fail(DA_ERR_INVALID, "iters must be >= 1");
    ShmReadGuard shared_reader(st);
    {
        const int rr = shm_refresh(st);
        if (rr != DA_OK) return rr;
    }
    {   // one query on one shard: the small-search path, its inputs uploaded once, `iters` scans back to back
        std::shared_lock<std::shared_mutex> shared(st->mu);
        FastPlan fp;
        uint2 ranges[kMaxRanges];
        const int pr = plan_fast(st, p, fp, ranges);
        if (pr < 0) return pr;
        if (pr == 1) {
            if (fp.k == 0) return fail(DA_ERR_INVALID, "k must be > 0");
            *out_total_ms = 0.f;
            if (fp.nothing) return DA_OK;
            Shard &s = st->shards[0];
            CU(cudaSetDevice(s.device));
            int rc = DA_OK, grid = 0;
            std::unique_lock<std::mutex> collective(st->xg_mu, std::defer_lock);
            if (fp.xchg) collective.lock();
            Slot *sl = acquire_slot(st, &rc);
            if (!sl) return rc;
            sl->ticket++;
            *reinterpret_cast<volatile uint32_t *>(sl->h_out + kSlotOutBytes - 24) = 0u;
            int launches = 0;
            rc = upload_fast(st, sl, fp, p, ranges);
            if (rc == DA_OK && cudaEventRecord(sl->ev0, sl->stream) != cudaSuccess) rc = fail(DA_ERR_CUDA, "event record failed");
            for (int i = 0; i < iters && rc == DA_OK; i++) rc = enqueue_fast(st, sl, fp, i == iters - 1, &grid, &launches);
            if (rc == DA_OK && cudaEventRecord(sl->ev1, sl->stream) != cudaSuccess) rc = fail(DA_ERR_CUDA, "event record failed");
            if (rc == DA_OK) rc = wait_fast(sl);
            if (rc == DA_OK && fp.xchg && *reinterpret_cast<volatile uint32_t *>(sl->h_out + kSlotOutBytes - 24) != 0u)
                rc = fail(DA_ERR_STATE, "a rank did not deliver its candidates within 4 s (searches on a multi-process store are collective)");
            if (rc == DA_OK && cudaEventSynchronize(sl->ev1) != cudaSuccess) rc = fail(DA_ERR_CUDA, "event sync failed");
            if (rc == DA_OK && cudaEventElapsedTime(out_total_ms, sl->ev0, sl->ev1) != cudaSuccess) rc = fail(DA_ERR_CUDA, "event time failed");
            if (rc != DA_OK) cudaStreamSynchronize(sl->stream);
            const uint64_t rows_read = rc == DA_OK ? *reinterpret_cast<const unsigned long long *>(sl->h_out + kSlotOutBytes - 16) : 0;
            release_slot(st, sl);
            if (rc != DA_OK) return rc;
            std::lock_guard<std::mutex> g(st->stats_mu);
            st->fast_searches += (uint64_t)iters;
            st->launches += (uint64_t)launches;
            st->scan_launches += (uint64_t)iters;
            st->scan_ms_sum += *out_total_ms;
            st->last_scan_ms = *out_total_ms / (float)iters;
            st->last_rows_read = rows_read;                          // every scan hands over (and clears) its own count
            st->last_bytes = fast_bytes(st, fp, st->last_rows_read);
            st->scan_grid = grid;
            st->bench_launches = (uint64_t)launches;
            return DA_OK;
        }
    }
    std::unique_lock<std::shared_mutex> lock(st->mu);
    SearchPlan pl{};
    int rc = stage_search(st, p, pl);
    if (rc != DA_OK) return rc;
    if (pl.k == 0) return fail(DA_ERR_INVALID, "k must be > 0");
    for (Shard &s : st->shards) { CU(cudaSetDevice(s.device)); CU(cudaStreamSynchronize(s.stream)); }
    Shard &s0 = st->shards[0];
    CU(cudaSetDevice(s0.device));
    while (s0.ev_pool.size() < 2 * (size_t)iters) {
        cudaEvent_t e;
        CU(cudaEventCreate(&e));
        s0.ev_pool.push_back(e);
    }
    CU(cudaEventRecord(s0.ev_all0, s0.stream));
    const uint64_t launches_before = st->launches;
    for (int i = 0; i < iters; i++) {
        rc = enqueue_search(st, pl, s0.ev_pool[2 * i], s0.ev_pool[2 * i + 1]);
        if (rc != DA_OK) return rc;
    }
    st->bench_launches = st->launches - launches_before;
    CU(cudaSetDevice(s0.device));
    CU(cudaEventRecord(s0.ev_all1, s0.stream));
    rc = enqueue_stats_readback(st);
    if (rc != DA_OK) return rc;
    for (Shard &s : st->shards) { CU(cudaSetDevice(s.device)); CU(cudaStreamSynchronize(s.stream)); }
    CU(cudaSetDevice(s0.device));
    CU(cudaEventElapsedTime(out_total_ms, s0.ev_all0, s0.ev_all1));
    double scan_ms = 0;
    for (int i = 0; i < iters; i++) {
        float ms = 0;
        CU(cudaEventElapsedTime(&ms, s0.ev_pool[2 * i], s0.ev_pool[2 * i + 1]));
        scan_ms += ms;
        st->last_scan_ms = ms;
    }
    st->scan_ms_sum += scan_ms;
    return collect_stats(st, pl, iters);
}

int da_store_build_ivf(da_store *st, const da_ivf_params *p)
{
    if (!st || !p) return fail(DA_ERR_INVALID, "store/params is NULL");
    if (p->struct_size != sizeof(da_ivf_params)) return fail(DA_ERR_INVALID, "da_ivf_params.struct_size mismatch");
    if (p->lists < 1 || p->lists > 65536) return fail(DA_ERR_INVALID, "lists=%d out of range [1, 65536]", p->lists);
    if (st->split16) return fail(DA_ERR_STATE, "IVFFlat is not available with the split-plane layout");
    if (st->attached) return fail(DA_ERR_STATE, "an attached store is read-only: build the index in the exporting process");
    std::unique_lock<std::shared_mutex> lock(st->mu);
    const int L = p->lists, D = st->dim;
    const int iters = p->iters <= 0 ? 10 : std::min(p->iters, 20);
    const bool use_tc = !(p->flags & DA_IVF_SIMT_ASSIGN);
    for (Shard &s : st->shards)
        if (s.n_rows == 0) return fail(DA_ERR_STATE, "build_ivf: a shard holds no rows");
    if ((uint64_t)L > st->shards[0].n_rows) return fail(DA_ERR_INVALID, "lists=%d exceeds the %u rows of the first shard", L, st->shards[0].n_rows);
    const uint64_t sample_total = p->sample_rows ? p->sample_rows : std::max<uint64_t>((uint64_t)L * 50, 10000);   // pgvector's rule
    const uint64_t per_shard = (sample_total + st->total_ranks - 1) / st->total_ranks;
    st->ivf_lists = 0;
    cudaEvent_t ev[4];
    {
        Shard &s0 = st->shards[0];
        CU(cudaSetDevice(s0.device));
        while (s0.ev_pool.size() < 4) { cudaEvent_t e; CU(cudaEventCreate(&e)); s0.ev_pool.push_back(e); }
        for (int i = 0; i < 4; i++) ev[i] = s0.ev_pool[i];
        CU(cudaEventRecord(ev[0], s0.stream));
    }
    for (Shard &s : st->shards) {
        CU(cudaSetDevice(s.device));
        free_ivf(s);
        Shard::Ivf &v = s.ivf;
        v.n_sample = (uint32_t)std::min<uint64_t>(s.n_rows, std::max<uint64_t>(per_shard, (uint64_t)L));
        const size_t cd = (size_t)L * D * 4;
        CU(cudaMalloc((void **)&v.cent, cd)); CU(cudaMalloc((void **)&v.sums, cd));
        CU(cudaMalloc((void **)&v.cent_inv, (size_t)L * 4)); CU(cudaMalloc((void **)&v.counts_f, (size_t)L * 4));
        CU(cudaMalloc((void **)&v.assign, s.cap_rows * 4)); CU(cudaMalloc((void **)&v.list_rows, s.cap_rows * 4));
        CU(cudaMalloc(&v.split_scratch, da::ivf_split_scratch_bytes(L)));
        CU(cudaMalloc((void **)&v.offsets, (size_t)(L + 1) * 4)); CU(cudaMalloc((void **)&v.cursor, (size_t)L * 4));
        CU(cudaMalloc((void **)&v.counts, (size_t)L * 4));
        CU(cudaMalloc((void **)&v.sample, (size_t)v.n_sample * 4)); CU(cudaMalloc((void **)&v.sample_assign, (size_t)v.n_sample * 4));
        v.bytes = 2 * cd + (size_t)L * 20 + s.cap_rows * 8 + (size_t)v.n_sample * 8 + da::ivf_split_scratch_bytes(L);
        if (use_tc) {
            CU(cudaMalloc((void **)&v.sample_mat, (size_t)v.n_sample * D * 4));
            CU(cudaMalloc((void **)&v.packed, da::tc_assign_scratch_bytes((uint32_t)s.cap_rows)));
            v.bytes += (size_t)v.n_sample * D * 4 + s.cap_rows * 8;
        }
        s.hbm_bytes += v.bytes;
        CU(da::ivf_launch_sample(s.n_rows, v.n_sample, p->seed + 0x51ED * (uint64_t)(&s - &st->shards[0] + 1 + st->rank * st->shards.size()),
                                 v.sample, s.stream));
        st->launches++;
        if (use_tc) {
            CU(da::tc_launch_gather_rows(s.mat, D, v.sample, v.n_sample, v.sample_mat, s.stream));
            st->launches++;
        }
    }
    // initial centroids: sample rows of global rank 0, then the same on every shard
    {
        Shard &s0 = st->shards[0];
        CU(cudaSetDevice(s0.device));
        if (st->rank == 0) {
            CU(da::ivf_launch_init_centroids(s0.mat, D, s0.inv_norm, s0.ivf.sample, s0.ivf.n_sample, L, p->seed, s0.ivf.cent,
                                             s0.ivf.cent_inv, s0.stream));
            st->launches += 2;
        }
        if (st->total_ranks > 1) {
            NC(ncclGroupStart());
            for (Shard &s : st->shards) NC(ncclBroadcast(s.ivf.cent, s.ivf.cent, (size_t)L * D, ncclFloat, 0, s.comm, s.stream));
            NC(ncclGroupEnd());
            NC(ncclGroupStart());
            for (Shard &s : st->shards) NC(ncclBroadcast(s.ivf.cent_inv, s.ivf.cent_inv, (size_t)L, ncclFloat, 0, s.comm, s.stream));
            NC(ncclGroupEnd());
        }
    }
    { Shard &s0 = st->shards[0]; CU(cudaSetDevice(s0.device)); CU(cudaEventRecord(ev[1], s0.stream)); }
    for (int it = 0; it < iters; it++) {
        for (Shard &s : st->shards) {
            Shard::Ivf &v = s.ivf;
            CU(cudaSetDevice(s.device));
            CU(cudaMemsetAsync(v.sums, 0, (size_t)L * D * 4, s.stream));
            CU(cudaMemsetAsync(v.counts_f, 0, (size_t)L * 4, s.stream));
            if (use_tc) CU(da::tc_launch_assign(v.sample_mat, v.n_sample, D, v.cent, L, v.packed, v.sample_assign, s.stream));
            else CU(da::ivf_launch_assign(s.mat, D, v.sample, v.n_sample, v.cent, L, v.sample_assign, 0, s.stream));
            CU(da::ivf_launch_accumulate(s.mat, D, s.inv_norm, v.sample, v.n_sample, v.sample_assign, v.sums, v.counts_f, s.stream));
            st->launches += 2;
        }
        if (st->total_ranks > 1) {   // data-parallel k-means: partial sums and counts meet here
            NC(ncclGroupStart());
            for (Shard &s : st->shards) {
                NC(ncclAllReduce(s.ivf.sums, s.ivf.sums, (size_t)L * D, ncclFloat, ncclSum, s.comm, s.stream));
                NC(ncclAllReduce(s.ivf.counts_f, s.ivf.counts_f, (size_t)L, ncclFloat, ncclSum, s.comm, s.stream));
            }
            NC(ncclGroupEnd());
        }
        for (Shard &s : st->shards) {
            CU(cudaSetDevice(s.device));
            CU(da::ivf_launch_finalize(s.ivf.sums, s.ivf.counts_f, L, D, s.ivf.cent, s.stream));
            st->launches++;
        }
    }
    // Balancing rounds (da_ivf.cu, "Splitting heavy lists"): count the sample per list under the current centroids
    // (globally: all-reduce), let the heaviest lists split into the empty ones on global rank 0, hand the centroids
    // to everybody.  A list of c x the mean needs log2(c) rounds; eight cover 256 x (each costs one assignment of the sample).
    static const int rounds_env = [] { const char *v = getenv("DA_IVF_SPLIT_ROUNDS"); return v ? atoi(v) : 8; }();
    static const float split_heavy_late = [] { const char *v = getenv("DA_IVF_SPLIT_HEAVY_LATE"); return v ? (float)atof(v) : 1.5f; }();
    static const float split_light_late = [] { const char *v = getenv("DA_IVF_SPLIT_LIGHT_LATE"); return v ? (float)atof(v) : 0.25f; }();
    const int split_rounds = (p->flags & DA_IVF_NO_SPLIT) ? 0 : rounds_env;
    // The rounds judge the lists on rows the centroids have NOT been averaged from.  On the training sample a starved
    // stub looks alive: its centroid is the mean of its dozen sample members, each of which therefore scores highest with it,
    // while every other row of the corpus prefers the big sibling -- measured: lists of 13 to 40 rows (their sample members,
    // exactly) next to lists of 9 600, invisible to a count over the training sample.  So a second sample is drawn.
    if (split_rounds > 0)
        for (Shard &s : st->shards) {
            Shard::Ivf &v = s.ivf;
            CU(cudaSetDevice(s.device));
            CU(da::ivf_launch_sample(s.n_rows, v.n_sample,
                                     p->seed + 0x7A11DA7Eull + 0x51ED * (uint64_t)(&s - &st->shards[0] + 1 + st->rank * st->shards.size()),
                                     v.sample, s.stream));
            st->launches++;
            if (use_tc) {
                CU(da::tc_launch_gather_rows(s.mat, D, v.sample, v.n_sample, v.sample_mat, s.stream));
                st->launches++;
            }
        }
    for (int round = 0; round < split_rounds; round++) {
        for (Shard &s : st->shards) {
            Shard::Ivf &v = s.ivf;
            CU(cudaSetDevice(s.device));
            if (use_tc) CU(da::tc_launch_assign(v.sample_mat, v.n_sample, D, v.cent, L, v.packed, v.sample_assign, s.stream));
            else CU(da::ivf_launch_assign(s.mat, D, v.sample, v.n_sample, v.cent, L, v.sample_assign, 0, s.stream));
            CU(cudaMemsetAsync(v.counts_f, 0, (size_t)L * 4, s.stream));
            CU(da::ivf_launch_count(v.sample_assign, v.n_sample, v.counts_f, s.stream));
            st->launches += 2;
        }
        if (st->total_ranks > 1) {
            NC(ncclGroupStart());
            for (Shard &s : st->shards) NC(ncclAllReduce(s.ivf.counts_f, s.ivf.counts_f, (size_t)L, ncclFloat, ncclSum, s.comm, s.stream));
            NC(ncclGroupEnd());
        }
        if (st->rank == 0) {
            Shard &s0 = st->shards[0];
            CU(cudaSetDevice(s0.device));
            CU(da::ivf_launch_split_round(s0.ivf.counts_f, L, s0.ivf.split_scratch, s0.mat, D, s0.inv_norm, s0.ivf.sample,
                                          s0.ivf.n_sample, s0.ivf.sample_assign, p->seed + 0xB5 * (uint64_t)(round + 1),
                                          s0.ivf.cent, s0.stream, round < split_rounds - 2 ? 1.5f : split_heavy_late,
                                          round < split_rounds - 2 ? 0.25f : split_light_late));
            st->launches += 2;
            if (getenv("DA_IVF_TRACE")) {           // measurements: what each balancing round saw and did
                CU(cudaStreamSynchronize(s0.stream));
                std::vector<float> cnt(L);
                uint32_t n_pairs = 0;
                CU(cudaMemcpy(cnt.data(), s0.ivf.counts_f, (size_t)L * 4, cudaMemcpyDeviceToHost));
                int Pp = 1; while (Pp < L) Pp <<= 1;
                CU(cudaMemcpy(&n_pairs, reinterpret_cast<const uint8_t *>(s0.ivf.split_scratch) + (size_t)Pp * 8 + (size_t)(L / 2) * 8, 4, cudaMemcpyDeviceToHost));
                double tot = 0; for (float c : cnt) tot += c;
                const double mean = tot / L;
                int heavy = 0, light = 0, empty = 0; float mx = 0;
                for (float c : cnt) { heavy += c > 1.5 * mean; light += c < 0.25 * mean; empty += c == 0; mx = std::max(mx, c); }
                fprintf(stderr, "ivf balance round %d: sample mean %.1f max %.0f, lists > 1.5x mean %d, < 0.25x mean %d (empty %d), pairs split %u\n",
                        round, mean, mx, heavy, light, empty, n_pairs);
            }
        }
        if (st->total_ranks > 1) {
            NC(ncclGroupStart());
            for (Shard &s : st->shards) NC(ncclBroadcast(s.ivf.cent, s.ivf.cent, (size_t)L * D, ncclFloat, 0, s.comm, s.stream));
            NC(ncclGroupEnd());
        }
    }
    // every resident row joins the list of its nearest centroid
    for (Shard &s : st->shards) {
        Shard::Ivf &v = s.ivf;
        CU(cudaSetDevice(s.device));
        if (&s == &st->shards[0]) CU(cudaEventRecord(ev[2], s.stream));
        if (use_tc) CU(da::tc_launch_assign(s.mat, s.n_rows, D, v.cent, L, v.packed, v.assign, s.stream));
        else CU(da::ivf_launch_assign(s.mat, D, nullptr, s.n_rows, v.cent, L, v.assign, 0, s.stream));
        if (&s == &st->shards[0]) CU(cudaEventRecord(ev[3], s.stream));
        CU(da::ivf_launch_build_lists(v.assign, s.n_rows, L, v.counts, v.offsets, v.cursor, v.list_rows, s.stream));
        st->launches += 4;
    }
    for (Shard &s : st->shards) { CU(cudaSetDevice(s.device)); CU(cudaStreamSynchronize(s.stream)); }
    {
        Shard &s0 = st->shards[0];
        CU(cudaSetDevice(s0.device));
        cudaEvent_t done = s0.ev_all1;
        CU(cudaEventRecord(done, s0.stream));
        CU(cudaEventSynchronize(done));
        CU(cudaEventElapsedTime(&st->ivf_build_ms, ev[0], done));
        CU(cudaEventElapsedTime(&st->ivf_kmeans_ms, ev[1], ev[2]));
        CU(cudaEventElapsedTime(&st->ivf_assign_ms, ev[2], ev[3]));
        st->ivf_assign_tflops = st->ivf_assign_ms > 0
            ? (float)(2.0 * s0.n_rows * (double)L * D / (st->ivf_assign_ms * 1e-3) / 1e12) : 0.f;
    }
    st->ivf_lists = L;
    st->ivf_default_probes = p->default_probes > 0 ? p->default_probes : 1;
    return DA_OK;
}

int da_store_generate_synthetic(da_store *st, const da_synth_params *p)
{
    if (!st || !p) return fail(DA_ERR_INVALID, "store/params is NULL");
    if (p->struct_size != sizeof(da_synth_params)) return fail(DA_ERR_INVALID, "da_synth_params.struct_size mismatch");
    if (!p->chunks_per_doc || !p->n_clusters) return fail(DA_ERR_INVALID, "chunks_per_doc/n_clusters must be > 0");
    if (p->row_lo + p->n_rows > 0xFFFFFFF0ull) return fail(DA_ERR_CAPACITY, "rows exceed 2^32");
    if (st->attached) return fail(DA_ERR_STATE, "an attached store is read-only");
    std::unique_lock<std::shared_mutex> lock(st->mu);
    ShmWriteGuard shared_writer(st);
    const int S = (int)st->shards.size();
    for (Shard &s : st->shards)
        if (s.n_rows) return fail(DA_ERR_STATE, "generate_synthetic needs an empty store");
    if (!st->docs.empty()) return fail(DA_ERR_STATE, "generate_synthetic needs an empty store");
    // documents [doc_lo, doc_hi] -> slots 0..
    const uint64_t doc_lo = p->row_lo / p->chunks_per_doc;
    const uint64_t doc_hi = p->n_rows ? (p->row_lo + p->n_rows - 1) / p->chunks_per_doc : doc_lo;
    // split rows over shards on document boundaries where possible
    uint64_t per = (p->n_rows + S - 1) / S;
    if (per % p->chunks_per_doc) per += p->chunks_per_doc - per % p->chunks_per_doc;
    uint64_t lo = p->row_lo, end = p->row_lo + p->n_rows;
    for (int i = 0; i < S; i++) {
        Shard &s = st->shards[i];
        uint64_t hi = std::min(end, lo + per);
        if (i == S - 1) hi = end;
        if (hi - lo > s.cap_rows) return fail(DA_ERR_CAPACITY, "shard %d: %llu rows exceed capacity_rows=%llu", i,
                                              (unsigned long long)(hi - lo), (unsigned long long)s.cap_rows);
        da::SynthArgs a{};
        a.seed_data = p->seed_data; a.seed_centroid = p->seed_centroid; a.seed_ids = p->seed_ids;
        a.chunks_per_doc = p->chunks_per_doc; a.n_clusters = p->n_clusters;
        a.cmul = p->cmul; a.nmul = p->nmul; a.smul = p->smul; a.super_group = p->super_group;
        a.row_lo = lo; a.n_rows = (uint32_t)(hi - lo); a.doc_slot_base = (uint32_t)doc_lo; a.dim = st->dim;
        CU(cudaSetDevice(s.device));
        if (st->split16) {                 // generate fp32 rows chunk by chunk into the staging buffer, then split them
            const uint32_t chunk = 32768;
            CU(s.d_up_stage.ensure((size_t)std::min<uint32_t>(a.n_rows ? a.n_rows : 1, chunk) * st->dim));
            const uint32_t total_rows = a.n_rows;
            const uint64_t lo0 = a.row_lo;
            for (uint32_t off = 0; off < total_rows; off += chunk) {
                a.row_lo = lo0 + off; a.n_rows = std::min(chunk, total_rows - off);
                CU(da::launch_synth(a, s.d_up_stage.p, s.inv_norm + off, s.doc_slot + off, s.chunk_ids + (size_t)off * 16, s.stream));
                CU(da::launch_split_rows(s.d_up_stage.p, a.n_rows, st->dim, s.hi + (size_t)off * st->dim, s.lo + (size_t)off * st->dim, s.stream));
                st->launches += 2;
            }
            a.row_lo = lo0; a.n_rows = total_rows;
        } else {
            CU(da::launch_synth(a, s.mat, s.inv_norm, s.doc_slot, s.chunk_ids, s.stream));
            st->launches++;
            if (st->q8) {
                CU(da::launch_q8_quantize_rows(s.mat, st->dim, nullptr, 0, a.n_rows, s.inv_norm, s.q8, s.q8_factor, s.q8_err,
                                               s.q8_errmax, s.stream));
                s.q8_dirty = true;
                st->launches++;
            }
        }
        s.row_base = (uint32_t)lo;
        s.n_rows = (uint32_t)(hi - lo);
        lo = hi;
    }
    if (p->n_rows) {
        st->docs.reserve(doc_hi - doc_lo + 1);
        st->doc_slot_of.reserve((doc_hi - doc_lo + 1) * 2);
        for (uint64_t d = doc_lo; d <= doc_hi; d++) {
            uint8_t id[16];
            da::synth_doc_id(p->seed_ids, d, id);
            const Id16 k = load_id(id);
            st->doc_slot_of.emplace(k, (uint32_t)(d - doc_lo));
            {
                const uint64_t first = std::max<uint64_t>(p->row_lo, d * p->chunks_per_doc);
                const uint64_t last = std::min<uint64_t>(p->row_lo + p->n_rows, (d + 1) * p->chunks_per_doc);
                DocInfo di{k, 0, (uint32_t)(last - first), false, {}};
                if (S == 1) di.ranges.emplace_back((uint32_t)(first - p->row_lo), (uint32_t)(last - first));
                else di.fragmented = true;             // a synthetic fill over several shards does not track ranges
                shm_append(st, kRecDoc, (uint32_t)st->docs.size(), 0, 0, &k);
                if (S == 1) shm_append(st, kRecRows, (uint32_t)st->docs.size(), di.ranges[0].first, di.ranges[0].second);
                st->docs.push_back(std::move(di));
            }
        }
    }
    for (Shard &s : st->shards) { CU(cudaSetDevice(s.device)); CU(cudaStreamSynchronize(s.stream)); }
    shm_publish(st);
    return DA_OK;
}

int da_synth_query(const da_synth_params *p, uint64_t seed_query, uint64_t qi, uint32_t cluster, int32_t dim, float *out)
{
    if (!p || !out) return fail(DA_ERR_INVALID, "params/out is NULL");
    if (p->struct_size != sizeof(da_synth_params)) return fail(DA_ERR_INVALID, "da_synth_params.struct_size mismatch");
    if (dim <= 0) return fail(DA_ERR_DIM, "dim=%d", dim);
    da::synth_vector_host(seed_query, qi, p->seed_centroid, cluster, p->cmul, p->nmul, dim, out, p->smul, p->super_group);
    return DA_OK;
}

int da_synth_ids(const da_synth_params *p, int32_t kind, uint64_t first, uint64_t n, uint8_t *out)
{
    if (!p || (n && !out)) return fail(DA_ERR_INVALID, "params/out is NULL");
    if (p->struct_size != sizeof(da_synth_params)) return fail(DA_ERR_INVALID, "da_synth_params.struct_size mismatch");
    if (kind != 0 && kind != 1) return fail(DA_ERR_INVALID, "kind must be 0 (chunk) or 1 (document)");
    for (uint64_t i = 0; i < n; i++) da::synth_id_host(p->seed_ids, kind, first + i, out + 16 * i);
    return DA_OK;
}

int da_store_stats(da_store *st, da_stats *out)
{
    if (!st || !out) return fail(DA_ERR_INVALID, "store/out is NULL");
    {
        ShmReadGuard shared_reader(st);
        const int rr = shm_refresh(st);
        if (rr != DA_OK) return rr;
    }
    std::unique_lock<std::shared_mutex> lock(st->mu);
    std::lock_guard<std::mutex> sg(st->stats_mu);
    memset(out, 0, sizeof *out);
    out->struct_size = sizeof *out;
    out->dim = st->dim; out->n_devices = (int)st->shards.size();
    out->world_size = st->world; out->rank = st->rank;
    for (size_t i = 0; i < st->shards.size(); i++) {
        out->rows += st->shards[i].n_rows;
        out->rows_per_device[i] = st->shards[i].n_rows;
        out->hbm_bytes += st->shards[i].hbm_bytes;
    }
    out->n_docs = st->docs.size();
    out->last_scan_ms = st->last_scan_ms; out->last_search_ms = st->last_search_ms;
    out->last_rows_read = st->last_rows_read; out->last_bytes_algorithmic = st->last_bytes;
    out->kernel_launches = st->launches;
    out->ivf_lists = st->ivf_lists;
    out->scan_grid = st->scan_grid;
    out->scan_ms_sum = st->scan_ms_sum; out->scan_launches = st->scan_launches;
    out->ivf_build_ms = st->ivf_build_ms; out->ivf_kmeans_ms = st->ivf_kmeans_ms;
    out->ivf_assign_ms = st->ivf_assign_ms; out->ivf_assign_tflops = st->ivf_assign_tflops;
    out->peer_gather = st->p2p_gather ? 1 : st->xg_ok ? 2 : 0; out->tensor_overflows = st->tc_overflows;
    { std::lock_guard<std::mutex> bl(st->bmu); out->topk_calls = st->topk_calls; out->topk_passes = st->topk_passes; }
    out->small_searches = st->fast_searches;
    out->last_bench_launches = st->bench_launches;
    return DA_OK;
}

int da_store_debug_read_rows(da_store *st, int32_t device_index, uint64_t row_lo, uint64_t n, float *out_vecs,
                             float *out_inv_norm, uint32_t *out_doc_slot, uint8_t *out_chunk_ids)
{
    if (!st) return fail(DA_ERR_INVALID, "store is NULL");
    std::unique_lock<std::shared_mutex> lock(st->mu);
    if (device_index < 0 || device_index >= (int)st->shards.size()) return fail(DA_ERR_INVALID, "device_index out of range");
    Shard &s = st->shards[device_index];
    if (row_lo + n > s.n_rows) return fail(DA_ERR_INVALID, "rows [%llu,%llu) outside the %u resident rows",
                                           (unsigned long long)row_lo, (unsigned long long)(row_lo + n), s.n_rows);
    CU(cudaSetDevice(s.device));
    if (out_vecs && st->split16 && n) {
        CU(s.d_up_stage.ensure(n * st->dim));
        CU(da::launch_join_rows(s.hi + row_lo * st->dim, s.lo + row_lo * st->dim, (uint32_t)n, st->dim, s.d_up_stage.p, s.stream));
        CU(cudaStreamSynchronize(s.stream));
        CU(cudaMemcpy(out_vecs, s.d_up_stage.p, n * st->dim * 4, cudaMemcpyDeviceToHost));
    } else if (out_vecs) {
        CU(cudaMemcpy(out_vecs, s.mat + row_lo * st->dim, n * st->dim * 4, cudaMemcpyDeviceToHost));
    }
    if (out_inv_norm) CU(cudaMemcpy(out_inv_norm, s.inv_norm + row_lo, n * 4, cudaMemcpyDeviceToHost));
    if (out_doc_slot) CU(cudaMemcpy(out_doc_slot, s.doc_slot + row_lo, n * 4, cudaMemcpyDeviceToHost));
    if (out_chunk_ids) CU(cudaMemcpy(out_chunk_ids, s.chunk_ids + row_lo * 16, n * 16, cudaMemcpyDeviceToHost));
    return DA_OK;
}

int da_store_debug_read_ivf(da_store *st, int32_t device_index, uint32_t *out_offsets, uint32_t *out_list_rows,
                            float *out_centroids)
{
    if (!st) return fail(DA_ERR_INVALID, "store is NULL");
    std::unique_lock<std::shared_mutex> lock(st->mu);
    if (device_index < 0 || device_index >= (int)st->shards.size()) return fail(DA_ERR_INVALID, "device_index out of range");
    if (!st->ivf_lists) return fail(DA_ERR_STATE, "IVFFlat index has not been built on this store");
    Shard &s = st->shards[device_index];
    const int L = st->ivf_lists;
    CU(cudaSetDevice(s.device));
    if (out_offsets) CU(cudaMemcpy(out_offsets, s.ivf.offsets, (size_t)(L + 1) * 4, cudaMemcpyDeviceToHost));
    if (out_list_rows) CU(cudaMemcpy(out_list_rows, s.ivf.list_rows, (size_t)s.n_rows * 4, cudaMemcpyDeviceToHost));
    if (out_centroids) CU(cudaMemcpy(out_centroids, s.ivf.cent, (size_t)L * st->dim * 4, cudaMemcpyDeviceToHost));
    return DA_OK;
}

int da_store_doc_rows(da_store *st, const uint8_t *doc_ids, size_t n_docs, uint32_t *out_rows)
{
    if (!st) return fail(DA_ERR_INVALID, "store is NULL");
    if (n_docs && (!doc_ids || !out_rows)) return fail(DA_ERR_INVALID, "doc_ids/out_rows is NULL");
    ShmReadGuard shared_reader(st);
    {
        const int rr = shm_refresh(st);
        if (rr != DA_OK) return rr;
    }
    std::unique_lock<std::shared_mutex> lock(st->mu);
    for (size_t i = 0; i < n_docs; i++) {
        auto it = st->doc_slot_of.find(load_id(doc_ids + 16 * i));
        out_rows[i] = it == st->doc_slot_of.end() ? 0u : st->docs[it->second].rows;
    }
    return DA_OK;
}

int da_store_export(da_store *st, const char *shm_name)
{
    if (!st || !shm_name || shm_name[0] != '/') return fail(DA_ERR_INVALID, "store is NULL or the name does not start with '/'");
    std::unique_lock<std::shared_mutex> lock(st->mu);
    if (st->attached || st->shm.hdr) return fail(DA_ERR_STATE, "this store is already shared");
    if (st->shards.size() != 1 || st->total_ranks != 1 || st->split16)
        return fail(DA_ERR_STATE, "only a single-GPU store in the default fp32 layout can be exported");
    Shard &s = st->shards[0];
    CU(cudaSetDevice(s.device));
    // one record per document plus one per run of rows in the usual flow; room for a sixteenth of the rows, at least 64 Ki
    const uint64_t max_records = std::max<uint64_t>(65536, s.cap_rows / 16 + 2 * (uint64_t)st->docs.size() + 8 * kMaxDocRanges);
    const size_t bytes = sizeof(ShmHeader) + (size_t)max_records * sizeof(ShmRecord);
    shm_unlink(shm_name);
    const int fd = shm_open(shm_name, O_CREAT | O_EXCL | O_RDWR, 0600);
    if (fd < 0) return fail(DA_ERR_STATE, "shm_open(%s) failed: %s", shm_name, strerror(errno));
    if (ftruncate(fd, (off_t)bytes) != 0) { close(fd); shm_unlink(shm_name); return fail(DA_ERR_STATE, "ftruncate failed: %s", strerror(errno)); }
    void *mem = mmap(nullptr, bytes, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
    close(fd);
    if (mem == MAP_FAILED) { shm_unlink(shm_name); return fail(DA_ERR_STATE, "mmap failed: %s", strerror(errno)); }
    ShmHeader *h = new (mem) ShmHeader();
    h->version = 1; h->dim = (uint32_t)st->dim; h->capacity_rows = s.cap_rows; h->min_sim = st->min_sim; h->row_base = s.row_base;
    h->max_records = max_records;
    h->n_records.store(0); h->n_rows.store(0); h->overflow.store(0);
    cudaError_t ce = cudaDeviceGetPCIBusId(h->device_bus_id, (int)sizeof h->device_bus_id, s.device);
    if (ce == cudaSuccess) ce = cudaIpcGetMemHandle(&h->h_mat, s.mat);
    if (ce == cudaSuccess) ce = cudaIpcGetMemHandle(&h->h_inv, s.inv_norm);
    if (ce == cudaSuccess) ce = cudaIpcGetMemHandle(&h->h_slot, s.doc_slot);
    if (ce == cudaSuccess) ce = cudaIpcGetMemHandle(&h->h_ids, s.chunk_ids);
    pthread_rwlockattr_t at;
    pthread_rwlockattr_init(&at);
    pthread_rwlockattr_setpshared(&at, PTHREAD_PROCESS_SHARED);
    pthread_rwlockattr_setkind_np(&at, PTHREAD_RWLOCK_PREFER_WRITER_NONRECURSIVE_NP);     // searches must not starve the writer
    const int lrc = pthread_rwlock_init(&h->lock, &at);
    pthread_rwlockattr_destroy(&at);
    if (ce != cudaSuccess || lrc != 0) {
        munmap(mem, bytes); shm_unlink(shm_name);
        return fail(DA_ERR_CUDA, "export failed: %s", ce != cudaSuccess ? cudaGetErrorString(ce) : strerror(lrc));
    }
    st->shm.hdr = h; st->shm.rec = reinterpret_cast<ShmRecord *>(h + 1); st->shm.bytes = bytes;
    st->shm.name = shm_name; st->shm.owner = true; st->shm.pending = 0;
    // what the store already holds, as the same records a reader replays
    for (uint32_t slot = 0; slot < st->docs.size(); slot++) {
        const DocInfo &d = st->docs[slot];
        shm_append(st, kRecDoc, slot, 0, 0, &d.id);
        if (d.fragmented) shm_append(st, kRecFragment, slot, 0, d.rows);
        else for (const auto &r : d.ranges) shm_append(st, kRecRows, slot, r.first, r.second);
    }
    CU(cudaStreamSynchronize(s.stream));
    h->magic = kShmMagic;                      // last: a reader that sees the magic sees a complete header
    shm_publish(st);
    return DA_OK;
}

int da_store_attach(const char *shm_name, da_store **out)
{
    if (!shm_name || !out) return fail(DA_ERR_INVALID, "name/out is NULL");
    *out = nullptr;
    const int fd = shm_open(shm_name, O_RDWR, 0600);
    if (fd < 0) return fail(DA_ERR_STATE, "shm_open(%s) failed: %s (has the other process exported its store?)", shm_name, strerror(errno));
    struct stat sb;
    if (fstat(fd, &sb) != 0 || (size_t)sb.st_size < sizeof(ShmHeader)) { close(fd); return fail(DA_ERR_STATE, "%s is not an exported store", shm_name); }
    void *mem = mmap(nullptr, (size_t)sb.st_size, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
    close(fd);
    if (mem == MAP_FAILED) return fail(DA_ERR_STATE, "mmap failed: %s", strerror(errno));
    ShmHeader *h = static_cast<ShmHeader *>(mem);
    if (h->magic != kShmMagic || h->version != 1 ||
        (size_t)sb.st_size < sizeof(ShmHeader) + (size_t)h->max_records * sizeof(ShmRecord)) {
        munmap(mem, (size_t)sb.st_size);
        return fail(DA_ERR_STATE, "%s is not an exported store of this library version", shm_name);
    }
    int device = -1;
    if (cudaDeviceGetByPCIBusId(&device, h->device_bus_id) != cudaSuccess || device < 0) {
        munmap(mem, (size_t)sb.st_size);
        return fail(DA_ERR_CUDA, "the exporting process's GPU (%s) is not visible to this process", h->device_bus_id);
    }
    da_store *st = new da_store();
    st->dim = (int)h->dim; st->min_sim = h->min_sim;
    st->attached = true;
    st->shm.hdr = h; st->shm.rec = reinterpret_cast<ShmRecord *>(h + 1); st->shm.bytes = (size_t)sb.st_size;
    st->shm.name = shm_name; st->shm.owner = false;
    st->shards.resize(1);
    Shard &s = st->shards[0];
    int rc = init_shard(s, device, st->dim, h->capacity_rows, false, false, true);
    if (rc == DA_OK) {
        s.row_base = h->row_base;
        cudaError_t ce = cudaIpcOpenMemHandle((void **)&s.mat, h->h_mat, cudaIpcMemLazyEnablePeerAccess);
        if (ce == cudaSuccess) ce = cudaIpcOpenMemHandle((void **)&s.inv_norm, h->h_inv, cudaIpcMemLazyEnablePeerAccess);
        if (ce == cudaSuccess) ce = cudaIpcOpenMemHandle((void **)&s.doc_slot, h->h_slot, cudaIpcMemLazyEnablePeerAccess);
        if (ce == cudaSuccess) ce = cudaIpcOpenMemHandle((void **)&s.chunk_ids, h->h_ids, cudaIpcMemLazyEnablePeerAccess);
        if (ce != cudaSuccess) rc = fail(DA_ERR_CUDA, "cudaIpcOpenMemHandle failed: %s (attach from a process other than the exporter)", cudaGetErrorString(ce));
    }
    if (rc == DA_OK) {
        pthread_rwlock_rdlock(&h->lock);
        rc = shm_replay(st);
        pthread_rwlock_unlock(&h->lock);
    }
    if (rc != DA_OK) { da_store_destroy(st); return rc; }
    *out = st;
    return DA_OK;
}

int da_store_debug_read_probes(da_store *st, int32_t device_index, int32_t batch, int32_t probes, uint32_t *out_lists)
{
    if (!st || !out_lists) return fail(DA_ERR_INVALID, "store/out is NULL");
    std::unique_lock<std::shared_mutex> lock(st->mu);
    if (device_index < 0 || device_index >= (int)st->shards.size()) return fail(DA_ERR_INVALID, "device_index out of range");
    Shard &s = st->shards[device_index];
    const unsigned long long *src = device_index == 0 ? st->last_probe_keys : s.d_probe_keys.p;
    const size_t have = device_index == 0 ? st->last_probe_count : s.d_probe_keys.cap;
    if (batch < 1 || probes < 1 || !src || (size_t)batch * probes > have)
        return fail(DA_ERR_INVALID, "no IVFFlat search of that shape has run on this shard");
    std::vector<unsigned long long> keys((size_t)batch * probes);
    CU(cudaSetDevice(s.device));
    CU(cudaDeviceSynchronize());
    CU(cudaMemcpy(keys.data(), src, keys.size() * 8, cudaMemcpyDeviceToHost));
    for (size_t i = 0; i < keys.size(); i++) out_lists[i] = keys[i] ? da::key_row(keys[i]) : 0xFFFFFFFFu;
    return DA_OK;
}

}  // extern "C"
