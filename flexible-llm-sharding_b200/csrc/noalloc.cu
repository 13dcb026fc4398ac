// fls_op_head_noalloc: the softmax head for callers that must not allocate device memory.
//
// Who this is for.  `fls_op_head` / `fls_lm_head` need a context, and a context reserves its scratch
// with cudaMalloc in fls_configure.  Some hosts of this library cannot do that: a scoring plugin
// loaded into a serving process whose own pool has already claimed all of HBM (a cudaMalloc from the
// plugin fails or fragments that pool), or a caller on a latency-critical thread where cudaMalloc's
// implicit device-wide synchronisation would stall the weight-copy stream and every other stream of
// the process.  Such callers own their input and output buffers already; the only thing missing is
// the kernel's scratch - the per-tile (max, sum-exp) partials and the per-row log-sum-exp that
// connect the two lm_head GEMM passes.  Here that scratch is a fixed array compiled into the
// library image, so the call itself allocates nothing (no cudaMalloc / cudaFree, no context, no
// allocating library routine).
//
// Limits and rules (also in include/fls_b200.h):
//  * at most FLS_NOALLOC_MAX_ROWS rows and FLS_NOALLOC_MAX_VOCAB vocabulary entries per call; larger
//    requests are rejected with FLS_ERR_CAPACITY.  The array is ~1 MiB and every GPU that loads the
//    library holds it until the process exits.
//  * the first call on a given device looks up the array's device address (and opts the GEMM kernels
//    into their shared-memory size).  Make that first call before starting any stream capture.
//  * all calls on a device share the one array and it keeps whatever the previous call left.  Nothing
//    in it has to start at zero: each call writes every partial and every log-sum-exp entry it later
//    reads, so stale contents are never observed and no clearing pass is needed.  Because the array is
//    shared, two calls whose kernels overlap on the GPU (different streams, or different host threads)
//    race on it - issue them on one stream or order them yourself.
#include <mutex>

#include "../../include/fls_b200.h"
#include "common.cuh"

namespace {

constexpr int kMaxRows = FLS_NOALLOC_MAX_ROWS;
constexpr int kMaxBlocks = (FLS_NOALLOC_MAX_VOCAB + 255) / 256;  // lm_head GEMM tiles are 256 columns wide
constexpr size_t kPartialBytes = sizeof(float2) * kMaxRows * kMaxBlocks;
constexpr size_t kLseBytes = sizeof(float) * kMaxRows;

// The scratch itself.  In host code this name is only a handle for the runtime's symbol table, not a
// device address: it must go through cudaGetSymbolAddress (below) before a kernel can use it.
static __device__ __align__(256) unsigned char g_head_scratch[kPartialBytes + kLseBytes];

constexpr int kMaxDevices = 64;
std::mutex g_setup_mu;
void* g_scratch_addr[kMaxDevices] = {};  // device address of g_head_scratch, per device
int g_num_sms[kMaxDevices] = {};

// Once per device, with that device current.
int device_setup(int dev, unsigned char** scratch, int* num_sms) {
    std::lock_guard<std::mutex> lk(g_setup_mu);
    if (!g_scratch_addr[dev]) {
        void* p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, g_head_scratch);
        if (e != cudaSuccess || !p)
            return fls::fail(fls::kErrCuda, "fls_op_head_noalloc: cudaGetSymbolAddress failed on device %d: %s", dev,
                             cudaGetErrorString(e));
        int sms = 0;
        e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        if (e != cudaSuccess)
            return fls::fail(fls::kErrCuda, "fls_op_head_noalloc: cudaDeviceGetAttribute: %s", cudaGetErrorString(e));
        int major = 0;
        cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev);
        if (major != 10)
            return fls::fail(fls::kErrCuda, "fls_op_head_noalloc: device %d is not sm_100 (this library is sm_100a only)",
                             dev);
        FLS_TRY(fls::gemm_init());
        g_num_sms[dev] = sms;
        g_scratch_addr[dev] = p;
    }
    *scratch = static_cast<unsigned char*>(g_scratch_addr[dev]);
    *num_sms = g_num_sms[dev];
    return fls::kOk;
}

}  // namespace

extern "C" int fls_op_head_noalloc(const uint16_t* x_dev, const uint16_t* w_dev, int n_rows, int n, int k,
                                   uint16_t* probs_dev, const int32_t* answer_ids_dev, int n_ids, float* logprob_dev,
                                   float* lse_dev, void* stream) {
    using namespace fls;
    if (!x_dev || !w_dev || n_rows <= 0 || n <= 0 || k <= 0)
        return fail(kErrInvalid, "fls_op_head_noalloc: null input or empty problem (%d x %d x %d)", n_rows, n, k);
    if (n_rows > kMaxRows || n > FLS_NOALLOC_MAX_VOCAB)
        return fail(kErrCapacity,
                    "fls_op_head_noalloc: %d rows x %d vocabulary entries exceeds the built-in scratch (at most %d rows "
                    "x %d entries per call); split the rows into several calls or use fls_op_head with a context",
                    n_rows, n, kMaxRows, FLS_NOALLOC_MAX_VOCAB);
    if (n_ids > 0 && (!answer_ids_dev || !logprob_dev))
        return fail(kErrInvalid, "fls_op_head_noalloc: n_ids > 0 needs answer_ids_dev and logprob_dev");
    if (n % 8 || k % 8 || ((reinterpret_cast<uintptr_t>(x_dev) | reinterpret_cast<uintptr_t>(w_dev) |
                            reinterpret_cast<uintptr_t>(probs_dev)) & 15))
        return fail(kErrInvalid, "fls_op_head_noalloc: n and k must be multiples of 8 and x, w, probs 16-byte aligned "
                    "(n=%d k=%d)", n, k);
    int dev = 0;
    FLS_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kMaxDevices) return fail(kErrInvalid, "fls_op_head_noalloc: device ordinal %d not supported", dev);
    unsigned char* scratch = nullptr;
    int num_sms = 0;
    FLS_TRY(device_setup(dev, &scratch, &num_sms));
    float2* partial = reinterpret_cast<float2*>(scratch);
    float* lse = reinterpret_cast<float*>(scratch + kPartialBytes);
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);

    GemmArgs a{};
    a.A = reinterpret_cast<const __half*>(x_dev), a.lda = k;
    a.B = reinterpret_cast<const __half*>(w_dev), a.ldb = k, a.b_rows = n;
    a.M = n_rows, a.N = n, a.K = k;
    a.partial = partial;
    FLS_TRY(gemm_launch(kEpiLsePartial, a, num_sms, st));                              // writes every partial
    FLS_TRY(lse_reduce_launch(partial, n_rows, gemm_lse_blocks(n), lse, st));           // writes lse[0, n_rows)
    if (probs_dev) {
        a.C = reinterpret_cast<__half*>(probs_dev), a.ldc = n, a.lse = lse;
        FLS_TRY(gemm_launch(kEpiProbs, a, num_sms, st));
    }
    if (n_ids > 0)
        FLS_TRY(gather_logprob_launch(a.A, k, a.B, k, k, answer_ids_dev, n_ids, lse, n_rows, n, logprob_dev, st));
    if (lse_dev) FLS_CUDA(cudaMemcpyAsync(lse_dev, lse, sizeof(float) * n_rows, cudaMemcpyDeviceToDevice, st));
    return kOk;
}
