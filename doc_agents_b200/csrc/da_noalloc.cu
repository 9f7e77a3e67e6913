// da_scan_topk_noalloc: the fused exact scan for callers that are not allowed to allocate.
//
// Who that is.  The store behind da_store_search reserves its HBM up front and grows small
// device buffers on demand with cudaMalloc.  Some hosts cannot tolerate that: a serving loop
// that must never hit the implicit device-wide synchronisation of cudaMalloc/cudaFree; a
// process whose whole GPU was already handed to another arena (a 10M x 3072 corpus leaves
// little of the 180 GB), where a late cudaMalloc simply fails; code that runs at a point
// where its host framework forbids calls into the CUDA allocator.  Such a caller owns its
// device buffers already -- the embedding matrix, the per-row arrays, the query, the outputs --
// and only needs the kernel.
//
// What this entry point does differently.  The scan needs scratch in global memory: one
// ticket counter, k result keys, and k candidate keys per CTA for the last-CTA merge.
// Here that scratch is a fixed file-scope __device__ array (g_scratch below), part of the
// library image: every GPU that loads the library carries it (about 258 KB) until the process
// exits, and nothing is ever allocated or freed -- not by this file, not by the launch helpers
// it calls (they only set a function attribute and launch).
//
// Things a user must know
//   - limits: one query per call, k <= DA_NOALLOC_MAX_K (128), dim a multiple of 4 and at most
//     16384, the device must have at most DA_NOALLOC_MAX_CTAS (256) SMs.  Anything beyond is
//     rejected with DA_ERR_CAPACITY / DA_ERR_INVALID and a message; nothing is launched.
//   - the first call on each device looks up the device address of the scratch array
//     (cudaGetSymbolAddress) and the SM count, and loads every kernel variant this entry point can
//     ever launch (the scan picks a variant by dim % 8, pointer alignment and tile shape; CUDA
//     loads a function lazily on first use, which allocates and synchronises -- so they are all
//     loaded here, not by some later call with another dim).  Make that first call during
//     start-up, before any stream capture begins on that device.
//   - the calling thread's current CUDA device is left as it was found.
//   - all calls on one device share the one scratch array, and it keeps whatever the last call
//     left in it.  Each call therefore clears what has to start from zero (the ticket counter
//     and the result slots) on the caller's stream before the scan.  Two calls whose kernels
//     may overlap on the same GPU (different streams, different threads) race on the scratch
//     and give wrong answers: serialise them, e.g. by using one stream.
//   - the call only enqueues work on `stream`; outputs are valid once the stream has reached
//     that point.  It never synchronises.
#include "../../include/docagents_b200.h"
#include "da_internal.h"

#include <mutex>
#include <string>

namespace da {
void set_last_error(const std::string &msg);   // da_store.cu: feeds da_last_error()
}

namespace {

constexpr int kMaxCtas = DA_NOALLOC_MAX_CTAS;
constexpr int kMaxK = DA_NOALLOC_MAX_K;
constexpr size_t kOffCounter = 0;        // unsigned int
constexpr size_t kOffFinalKeys = 256;    // kMaxK keys (1 KB)
constexpr size_t kOffCtaKeys = 2048;     // kMaxCtas * kMaxK keys
constexpr size_t kScratchBytes = kOffCtaKeys + (size_t)kMaxCtas * kMaxK * sizeof(da::key_t64);

// The scratch itself.  In host code this name is only a handle for cudaGetSymbolAddress --
// it is NOT a device address and must never be passed to a kernel as a pointer.
static __device__ __align__(256) unsigned char g_scratch[kScratchBytes];

// Device code may name the array directly.  Clears what the scan expects to find at zero.
__global__ void noalloc_reset_kernel(int k)
{
    if (threadIdx.x == 0) *reinterpret_cast<unsigned int *>(g_scratch + kOffCounter) = 0u;
    da::key_t64 *fin = reinterpret_cast<da::key_t64 *>(g_scratch + kOffFinalKeys);
    for (int i = threadIdx.x; i < k; i += blockDim.x) fin[i] = 0ull;
}

// Turns the k final keys (left in the scratch by the scan) into plain outputs.
__global__ void noalloc_decode_kernel(int k, float *out_scores, uint32_t *out_rows, int32_t *out_n)
{
    const da::key_t64 *fin = reinterpret_cast<const da::key_t64 *>(g_scratch + kOffFinalKeys);
    int n = 0;
    for (int i = 0; i < k; i++) n += (fin[i] != 0ull);          // keys are best-first, zero-padded
    for (int i = threadIdx.x; i < k; i += blockDim.x) {
        const da::key_t64 key = fin[i];
        out_scores[i] = key ? da::key_score(key) : 0.f;
        if (out_rows) out_rows[i] = key ? da::key_row(key) : 0u;
    }
    if (threadIdx.x == 0) *out_n = n;
}

struct DeviceSetup {
    bool ready = false;
    unsigned char *scratch = nullptr;    // device address of g_scratch on this device
    int sm_count = 0;
};
constexpr int kMaxVisibleDevices = 64;
DeviceSetup g_setup[kMaxVisibleDevices];
std::mutex g_setup_mu;

int fail(int code, const std::string &msg)
{
    da::set_last_error(msg);
    return code;
}

}  // namespace

extern "C" int da_scan_topk_noalloc(int32_t device, const float *d_mat, const float *d_inv_norm,
                                    const uint32_t *d_doc_slot, const uint8_t *d_chunk_ids,
                                    const uint32_t *d_doc_bitmap, const float *d_query, uint64_t n_rows,
                                    uint32_t row_base, int32_t dim, int32_t k, float min_similarity,
                                    void *stream, float *d_out_scores, uint32_t *d_out_rows,
                                    uint8_t *d_out_chunk_ids, int32_t *d_out_n)
{
    if (device < 0 || device >= kMaxVisibleDevices) return fail(DA_ERR_INVALID, "device index " + std::to_string(device) + " out of range");
    if (!d_mat || !d_inv_norm || !d_chunk_ids || !d_query || !d_out_scores || !d_out_chunk_ids || !d_out_n)
        return fail(DA_ERR_INVALID, "a required device pointer is NULL");
    if (d_doc_bitmap && !d_doc_slot) return fail(DA_ERR_INVALID, "d_doc_bitmap given without d_doc_slot");
    if (dim <= 0 || (dim & 3) || dim > 16384) return fail(DA_ERR_DIM, "dim=" + std::to_string(dim) + " must be a multiple of 4 in (0, 16384]");
    if (k < 1) return fail(DA_ERR_INVALID, "k=" + std::to_string(k) + " must be at least 1");
    if (k > kMaxK)
        return fail(DA_ERR_CAPACITY, "k=" + std::to_string(k) + " exceeds the no-allocation limit DA_NOALLOC_MAX_K=" +
                                         std::to_string(kMaxK) + " (the scratch array is fixed at build time)");
    if (n_rows > 0xFFFFFFF0ull || (uint64_t)row_base + n_rows > 0xFFFFFFF0ull)
        return fail(DA_ERR_CAPACITY, "row_base + n_rows exceeds 2^32 - 16 (rows are numbered in 32 bits)");
    if ((reinterpret_cast<uintptr_t>(d_mat) & 15) || (reinterpret_cast<uintptr_t>(d_query) & 15) ||
        (reinterpret_cast<uintptr_t>(d_chunk_ids) & 15) || (reinterpret_cast<uintptr_t>(d_out_chunk_ids) & 15))
        return fail(DA_ERR_INVALID, "d_mat, d_query, d_chunk_ids and d_out_chunk_ids must be 16-byte aligned");

    // the caller's current device is restored on every way out
    struct DeviceGuard {
        int prev = -1;
        ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
    } guard;
    cudaError_t e = cudaGetDevice(&guard.prev);
    if (e != cudaSuccess) { guard.prev = -1; cudaGetLastError(); }
    if (guard.prev == device) guard.prev = -1;           // nothing to restore
    else {
        e = cudaSetDevice(device);
        if (e != cudaSuccess) return fail(DA_ERR_CUDA, std::string("cudaSetDevice failed: ") + cudaGetErrorString(e));
    }

    DeviceSetup su;
    {
        std::lock_guard<std::mutex> lock(g_setup_mu);
        DeviceSetup &slot = g_setup[device];
        if (!slot.ready) {
            // once per device, with that device current: the array's address on THIS device
            unsigned char *ptr = nullptr;
            e = cudaGetSymbolAddress((void **)&ptr, g_scratch);
            if (e != cudaSuccess || !ptr)
                return fail(DA_ERR_CUDA, std::string("cudaGetSymbolAddress(g_scratch) failed: ") + cudaGetErrorString(e));
            int sms = 0;
            e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
            if (e != cudaSuccess) return fail(DA_ERR_CUDA, std::string("cudaDeviceGetAttribute failed: ") + cudaGetErrorString(e));
            // load every function a later call may launch, now (lazy loading allocates and synchronises)
            cudaFuncAttributes fa;
            e = cudaFuncGetAttributes(&fa, noalloc_reset_kernel);
            if (e == cudaSuccess) e = cudaFuncGetAttributes(&fa, noalloc_decode_kernel);
            if (e == cudaSuccess) e = da::scan_preload_single_query_variants();
            if (e != cudaSuccess) return fail(DA_ERR_CUDA, std::string("loading the scan kernels failed: ") + cudaGetErrorString(e));
            slot.scratch = ptr;
            slot.sm_count = sms;
            slot.ready = true;
        }
        su = slot;
    }
    if (da::scan_grid_for(su.sm_count) > kMaxCtas)
        return fail(DA_ERR_CAPACITY, "device has " + std::to_string(su.sm_count) + " SMs; the scratch array holds candidates for at most " +
                                         std::to_string(kMaxCtas) + " CTAs");

    cudaStream_t s = static_cast<cudaStream_t>(stream);
    noalloc_reset_kernel<<<1, 128, 0, s>>>(k);
    e = cudaGetLastError();
    if (e != cudaSuccess) return fail(DA_ERR_CUDA, std::string("reset launch failed: ") + cudaGetErrorString(e));

    da::ScanArgs a{};
    a.mat = d_mat; a.inv_norm = d_inv_norm; a.doc_slot = d_doc_slot; a.chunk_ids = d_chunk_ids;
    a.doc_bitmap = d_doc_bitmap;
    a.queries = d_query; a.q_inv_norm = nullptr;         // the scan computes 1/|q| itself
    a.row_list = nullptr;
    a.n_rows = (uint32_t)n_rows; a.row_base = row_base;
    a.dim = dim; a.nq = 1; a.k = k;
    a.min_sim = min_similarity; a.read_all = 0;
    a.cta_out = reinterpret_cast<da::key_t64 *>(su.scratch + kOffCtaKeys);
    a.counter = reinterpret_cast<unsigned int *>(su.scratch + kOffCounter);
    a.out_keys = reinterpret_cast<da::key_t64 *>(su.scratch + kOffFinalKeys);
    a.out_ids = d_out_chunk_ids;
    a.rows_read = nullptr;
    e = da::launch_scan(a, su.sm_count, s, nullptr);
    if (e != cudaSuccess) return fail(DA_ERR_CUDA, std::string("scan launch failed: ") + cudaGetErrorString(e));

    noalloc_decode_kernel<<<1, 128, 0, s>>>(k, d_out_scores, d_out_rows, d_out_n);
    e = cudaGetLastError();
    if (e != cudaSuccess) return fail(DA_ERR_CUDA, std::string("decode launch failed: ") + cudaGetErrorString(e));
    return DA_OK;
}
