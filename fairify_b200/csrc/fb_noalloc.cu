// fb_noalloc.cu -- fb_prune_noalloc: the pre-solver stage for callers that must not allocate.
//
// fb_prune() asks the CUDA allocator for its [n][N] uint32 activation-count scratch on every call.  Some
// callers cannot do that: code running inside a latency-critical service loop that has promised its host
// never to touch the device allocator after start-up (a cudaMalloc can take the device-wide allocator lock
// and stall every other stream), hosts that pre-partition the whole HBM between tenants and hand this
// library no pool to draw from, and embedders whose memory is accounted by an outer framework that treats
// any allocation it did not make itself as a leak.  For them the scratch lives in the library image: a
// fixed file-scope __device__ array that each GPU gets a copy of when the library is loaded on it and keeps
// until the process exits.
//
// What that implies for a user:
//   * size limit: n * N <= FB_NOALLOC_MAX_WORDS (4 Mi counters = 16 MiB per GPU).  Larger batches are
//     rejected with FB_ERR_UNSUPPORTED; split them or use fb_prune.
//   * the first call on each device looks the array's device address up (cudaGetSymbolAddress, once per
//     device, with that device current).  Make that first call during start-up, before any stream capture.
//   * every call on a device uses the SAME array.  Two calls that overlap on the GPU (different streams,
//     different host threads) race on it and corrupt each other's counts; serialise them.  The array keeps
//     the previous call's contents: the profile kernel overwrites every counter it is responsible for, and
//     when it accumulates atomically (few partitions, samples split over warps) the range is zeroed first.
#include <cstdio>
#include <mutex>

#include "../../include/fairify_b200.h"

#define FB_NOALLOC_MAX_WORDS (4u * 1024u * 1024u)
#define FB_NOALLOC_MAX_DEVICES 64

// In device code this name is the array.  In HOST code it is only a handle for cudaGetSymbolAddress and
// friends -- never a pointer a kernel may be given.
static __device__ __align__(256) uint32_t fb_noalloc_scratch[FB_NOALLOC_MAX_WORDS];

static std::mutex fb_noalloc_mu;
static uint32_t *fb_noalloc_ptr[FB_NOALLOC_MAX_DEVICES];      // per-device address of the array above

void fb_set_error(const char *msg);      // fb_api.cu: message returned by fb_last_error()
extern "C" int fb_model_device(const fb_model *m, int *device);

#define FB_NOALLOC_FAIL(code, ...)                          \
    do {                                                    \
        char msg_[384];                                     \
        snprintf(msg_, sizeof(msg_), __VA_ARGS__);          \
        fb_set_error(msg_);                                 \
        return (code);                                      \
    } while (0)

// device address of the scratch array on the current device (looked up once per device)
static int fb_noalloc_scratch_ptr(uint32_t **out)
{
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) FB_NOALLOC_FAIL(FB_ERR_CUDA, "fb_prune_noalloc: cudaGetDevice failed: %s", cudaGetErrorString(e));
    if (dev < 0 || dev >= FB_NOALLOC_MAX_DEVICES)
        FB_NOALLOC_FAIL(FB_ERR_UNSUPPORTED, "fb_prune_noalloc: device index %d not supported", dev);
    std::lock_guard<std::mutex> lock(fb_noalloc_mu);
    if (!fb_noalloc_ptr[dev]) {
        uint32_t *p = nullptr;
        e = cudaGetSymbolAddress((void **)&p, fb_noalloc_scratch);
        if (e != cudaSuccess || !p)
            FB_NOALLOC_FAIL(FB_ERR_CUDA, "fb_prune_noalloc: cudaGetSymbolAddress failed on device %d: %s", dev,
                            cudaGetErrorString(e));
        fb_noalloc_ptr[dev] = p;
    }
    *out = fb_noalloc_ptr[dev];
    return FB_OK;
}

extern "C" int fb_prune_noalloc(const fb_model *m, int64_t n, const int32_t *boxes, const int32_t *samples, int S,
                                uint64_t seed, int64_t first_index, const double *tol, const fb_prune_out *out,
                                void *stream)
{
    int N = 0;
    if (!m || fb_model_info(m, nullptr, &N, nullptr, nullptr) != FB_OK)
        FB_NOALLOC_FAIL(FB_ERR_INVALID, "fb_prune_noalloc: null model");
    if (n < 0) FB_NOALLOC_FAIL(FB_ERR_INVALID, "fb_prune_noalloc: n = %lld", (long long)n);
    if ((unsigned long long)n > FB_NOALLOC_MAX_WORDS / (unsigned)N)          // (a division: n * N could wrap for huge n)
        FB_NOALLOC_FAIL(FB_ERR_UNSUPPORTED,
                        "fb_prune_noalloc: %lld partitions x %d neurons exceed the built-in scratch of %u counters "
                        "(at most %u partitions per call for this model); split the batch or use fb_prune",
                        (long long)n, N, FB_NOALLOC_MAX_WORDS, FB_NOALLOC_MAX_WORDS / (unsigned)N);
    if (!out) FB_NOALLOC_FAIL(FB_ERR_INVALID, "fb_prune_noalloc: null output block");
    if (n == 0) return FB_OK;
    // the scratch address below is the current device's copy of the array: it must be the device the model's
    // weights live on, or kernels on one GPU would be handed a pointer into another
    int mdev = -1, cur = -1;
    if (fb_model_device(m, &mdev) != FB_OK || cudaGetDevice(&cur) != cudaSuccess || mdev != cur)
        FB_NOALLOC_FAIL(FB_ERR_INVALID, "fb_prune_noalloc: the model lives on device %d but device %d is current", mdev, cur);
    uint32_t *counts = nullptr;
    if (int rc = fb_noalloc_scratch_ptr(&counts)) return rc;
    // the two stages of fb_prune, on the built-in scratch; neither allocates (fb_profile zeroes the
    // counters itself whenever it is going to accumulate into them)
    int rc = fb_profile(m, n, boxes, samples, S, seed, first_index, counts, stream);
    if (rc == FB_OK) rc = fb_ibp(m, n, boxes, counts, tol, out, stream);
    return rc;
}
