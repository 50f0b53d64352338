// fb_api.cu -- C ABI of libfairify_b200.so (declared in include/fairify_b200.h) and the runtime-shape
// kernel launchers.  No torch types; device pointers, sizes and a cudaStream_t (as void*) only.
#include <algorithm>
#include <climits>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <new>
#include <mutex>
#include <vector>

#include "../../include/fairify_b200.h"
#include "fb_generic.cuh"
#include "fb_launch.cuh"
#include "fb_mma.cuh"
#include "fb_registry.h"

cudaError_t fb_profile_mma(const FbMmaNet &net, int64_t n, const int32_t *boxes, const int32_t *samples, uint64_t seed,
                           int64_t first_index, int nsamp, int chunks, int chunk_len, uint32_t *counts, int num_sms,
                           cudaStream_t st);      // fb_mma.cu

static_assert(sizeof(fb_prune_out) == sizeof(FbPruneOut), "fb_prune_out and FbPruneOut must match");

// ---------------------------------------------------------------------------------------------------
// errors
// ---------------------------------------------------------------------------------------------------
static thread_local char fb_err[512] = "";

static int fb_fail(int code, const char *fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(fb_err, sizeof(fb_err), fmt, ap);
    va_end(ap);
    return code;
}

#define FB_CUDA(expr)                                                                                      \
    do {                                                                                                   \
        cudaError_t e_ = (expr);                                                                           \
        if (e_ != cudaSuccess)                                                                             \
            return fb_fail(FB_ERR_CUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(e_), __FILE__, __LINE__); \
    } while (0)

extern "C" const char *fb_last_error(void) { return fb_err; }
// for the library's other translation units (not part of the C ABI)
void fb_set_error(const char *msg) { snprintf(fb_err, sizeof(fb_err), "%s", msg); }
extern "C" int fb_version(void) { return 210; }

extern "C" int fb_num_sms(int *out)
{
    int dev = 0, sms = 0;
    FB_CUDA(cudaGetDevice(&dev));
    FB_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    *out = sms;
    return FB_OK;
}

// ---------------------------------------------------------------------------------------------------
// model
// ---------------------------------------------------------------------------------------------------
struct fb_model {
    int device = 0, num_sms = 0;
    GenNet gen{};
    std::vector<double> hW64, hB64;
    std::vector<float> hW32, hB32;
    double *dW64 = nullptr, *dB64 = nullptr;
    float *dW32 = nullptr, *dB32 = nullptr;
    const FbSpecFwdEntry *spec = nullptr;               // compile-time-specialised kernels for this shape, if any
    const FbSpecProfileEntry *spec_profile = nullptr;
    const FbSpecIbpEntry *spec_ibp = nullptr;           // (its launcher is null for shapes left to the runtime-shape kernel)
    const FbSpecMmaEntry *spec_mma = nullptr;           // shape-specialised DMMA profile kernel (the wide networks)
    void *p64 = nullptr, *p32 = nullptr;
    FbTileNet tile{};               // padded float64 weights for the tiled profile kernel
    double *dWpad = nullptr;
    float *dWpad32 = nullptr;
    float *dWiv32 = nullptr;
    FbMmaNet mma{};                 // fragment-major weights for the DMMA profile kernel
    double *dWf = nullptr, *dBp = nullptr;
    uint32_t *dUnb = nullptr;       // per neuron: a non-zero weight on an input the singular-verification query leaves unbounded
};

// The specialised profile kernel tests "pre-activation > 0" on the high word of the double, which misreads positive
// subnormals below 2^-1042.  They cannot occur when every hidden pre-activation is a multiple of some 2^q with
// q >= -1022: sample features are integers, so layer 0's pre-activations are multiples of 2^q0, q0 = the smallest
// exponent of a least-significant set bit among W0 and b0; a layer's products are multiples of 2^(q_prev + its
// weights' smallest such exponent), and rounding a multiple of 2^q to nearest gives a multiple of 2^q again.  A non-zero
// multiple of 2^q with q >= -1022 is a normal number.  Keras float32 weights give q >= -149 per layer, so only
// pathological models (subnormal weights through many layers) fail this and take the runtime-shape kernels.
static int fb_lsb_exponent(double v)
{
    if (v == 0.0) return INT32_MAX;
    int e;
    const double m = std::frexp(std::fabs(v), &e);                 // v = m * 2^e, 0.5 <= m < 1
    const uint64_t M = (uint64_t)std::ldexp(m, 53);                // 53-bit integer mantissa
    return e - 53 + __builtin_ctzll(M);
}
static bool fb_model_subnormal_free(const GenNet &g, const std::vector<double> &W, const std::vector<double> &B)
{
    int64_t q = 0;                                                 // inputs are integers
    for (int l = 0; l + 1 < g.L; ++l) {                            // hidden layers only: the logit uses the full test
        int wmin = INT32_MAX, bmin = INT32_MAX;
        for (int k = g.woff[l]; k < g.woff[l + 1]; ++k) wmin = std::min(wmin, fb_lsb_exponent(W[k]));
        for (int k = g.boff[l]; k < g.boff[l + 1]; ++k) bmin = std::min(bmin, fb_lsb_exponent(B[k]));
        const int64_t prod = (wmin == INT32_MAX) ? INT32_MAX : q + wmin;
        q = std::min<int64_t>(prod, bmin);
        if (q == INT32_MAX) q = 0;                                 // an all-zero layer: everything after it restarts from the biases
        if (q < -1022) return false;
    }
    return true;
}

template <class E, size_t NT>
static const E *fb_find_spec(const FbSpecTableT<E> *const (&tables)[NT], int L, const int32_t *dims)
{
    for (const FbSpecTableT<E> *t : tables)
        for (int k = 0; k < t->count; ++k) {
            const FbSpecShape &e = t->entries[k].shape;
            if (e.nd != L + 1) continue;
            bool same = true;
            for (int i = 0; i <= L; ++i) same = same && (e.dims[i] == dims[i]);
            if (same) return &t->entries[k];
        }
    return nullptr;
}
#define FB_TAB_F(U) &fb_spec_fwd_table_##U,
#define FB_TAB_P(U) &fb_spec_profile_table_##U,
#define FB_TAB_I(U) &fb_spec_ibp_table_##U,
static const FbSpecFwdTable *const fb_fwd_tables[] = {FB_SPEC_FOR_EACH_UNIT(FB_TAB_F)};
static const FbSpecProfileTable *const fb_profile_tables[] = {FB_SPEC_FOR_EACH_UNIT(FB_TAB_P)};
static const FbSpecIbpTable *const fb_ibp_tables[] = {FB_SPEC_FOR_EACH_UNIT(FB_TAB_I)};
static const FbSpecMmaTable *const fb_mma_tables[] = {&fb_spec_mma_table_a, &fb_spec_mma_table_b, &fb_spec_mma_table_c, &fb_spec_mma_table_d};

extern "C" void fb_model_destroy(fb_model *m)
{
    if (!m) return;
    if (m->spec) m->spec->free_params(m->p64, m->p32);
    cudaFree(m->dW64); cudaFree(m->dB64); cudaFree(m->dW32); cudaFree(m->dB32); cudaFree(m->dWpad); cudaFree(m->dWpad32); cudaFree(m->dWiv32); cudaFree(m->dUnb); cudaFree(m->dWf); cudaFree(m->dBp);
    delete m;
}

extern "C" int fb_model_create(int L, const int32_t *dims, const float *w, const float *b, fb_model **out)
{
    if (!dims || !w || !b || !out) return fb_fail(FB_ERR_INVALID, "fb_model_create: null argument");
    if (L < 1 || L > FB_MAX_LAYERS) return fb_fail(FB_ERR_UNSUPPORTED, "fb_model_create: %d layers (supported: 1..%d)", L, FB_MAX_LAYERS);
    if (dims[L] != 1) return fb_fail(FB_ERR_UNSUPPORTED, "fb_model_create: output width %d (the path is defined for a single logit)", dims[L]);
    fb_model *m = new (std::nothrow) fb_model;
    if (!m) return fb_fail(FB_ERR_NOMEM, "fb_model_create: out of host memory");
    GenNet &g = m->gen;
    g.L = L; g.maxw = dims[0];
    int wo = 0, bo = 0;
    for (int l = 0; l <= L; ++l) {
        if (dims[l] < 1 || dims[l] > FB_MAXW) { delete m; return fb_fail(FB_ERR_UNSUPPORTED, "fb_model_create: layer width %d (supported: 1..%d)", dims[l], FB_MAXW); }
        g.dims[l] = dims[l];
        if (dims[l] > g.maxw) g.maxw = dims[l];
    }
    for (int l = 0; l < L; ++l) { g.woff[l] = wo; g.boff[l] = bo; wo += dims[l] * dims[l + 1]; bo += dims[l + 1]; }
    g.woff[L] = wo; g.boff[L] = bo; g.P = wo; g.N = bo;
    if ((g.N + 31) / 32 > FB_MAX_NW) { delete m; return fb_fail(FB_ERR_UNSUPPORTED, "fb_model_create: %d neurons (supported: <= %d)", g.N, FB_MAX_NW * 32); }
    m->hW32.assign(w, w + g.P); m->hB32.assign(b, b + g.N);
    m->hW64.resize(g.P); m->hB64.resize(g.N);
    for (int k = 0; k < g.P; ++k) {
        if (!std::isfinite(w[k])) { delete m; return fb_fail(FB_ERR_INVALID, "fb_model_create: weight %d is not finite", k); }
        // -0.0 -> +0.0: the reference's `weight < 0` test treats both as non-negative and the products are
        // zero either way, so results are unchanged while the device can test the sign bit alone
        m->hW32[k] = (w[k] == 0.0f) ? 0.0f : w[k];
        m->hW64[k] = (double)m->hW32[k];
    }
    for (int k = 0; k < g.N; ++k) {
        if (!std::isfinite(b[k])) { delete m; return fb_fail(FB_ERR_INVALID, "fb_model_create: bias %d is not finite", k); }
        m->hB64[k] = (double)b[k];
    }
    cudaError_t e = cudaGetDevice(&m->device);
    if (e == cudaSuccess) e = cudaDeviceGetAttribute(&m->num_sms, cudaDevAttrMultiProcessorCount, m->device);
    if (e == cudaSuccess) e = cudaMalloc(&m->dW64, sizeof(double) * g.P);
    if (e == cudaSuccess) e = cudaMalloc(&m->dB64, sizeof(double) * g.N);
    if (e == cudaSuccess) e = cudaMalloc(&m->dW32, sizeof(float) * g.P);
    if (e == cudaSuccess) e = cudaMalloc(&m->dB32, sizeof(float) * g.N);
    if (e == cudaSuccess) e = cudaMemcpy(m->dW64, m->hW64.data(), sizeof(double) * g.P, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(m->dB64, m->hB64.data(), sizeof(double) * g.N, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(m->dW32, m->hW32.data(), sizeof(float) * g.P, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(m->dB32, m->hB32.data(), sizeof(float) * g.N, cudaMemcpyHostToDevice);
    if (e != cudaSuccess) {
        fb_model_destroy(m);
        return fb_fail(FB_ERR_CUDA, "fb_model_create: %s (is a CUDA device available? there is no CPU path)", cudaGetErrorString(e));
    }
    g.W64 = m->dW64; g.B64 = m->dB64; g.W32 = m->dW32; g.B32 = m->dB32;
    {
        // singular_verification* bounds only the first out_l inputs of a hidden layer l >= 1 (utils/prune.py:266-273 with
        // :302-345); a neuron with a non-zero weight on any later input can always be made positive (see fb_spec.cuh)
        std::vector<uint32_t> unb((size_t)(g.N + 31) / 32, 0u);
        for (int l = 1; l + 1 < L; ++l)
            for (int j = 0; j < dims[l + 1]; ++j)
                for (int k = dims[l + 1]; k < dims[l]; ++k)
                    if (m->hW64[(size_t)g.woff[l] + (size_t)k * dims[l + 1] + j] != 0.0) {
                        unb[(g.boff[l] + j) >> 5] |= 1u << ((g.boff[l] + j) & 31);
                        break;
                    }
        e = cudaMalloc(&m->dUnb, sizeof(uint32_t) * unb.size());
        if (e == cudaSuccess) e = cudaMemcpy(m->dUnb, unb.data(), sizeof(uint32_t) * unb.size(), cudaMemcpyHostToDevice);
        if (e != cudaSuccess) { fb_model_destroy(m); return fb_fail(FB_ERR_CUDA, "fb_model_create: %s", cudaGetErrorString(e)); }
        g.unb = m->dUnb;
    }
    {
        // weights with every row padded to a multiple of 32 columns (zeros): operand of fb_profile_tile_kernel
        FbTileNet &t = m->tile;
        t.L = L; t.N = g.N; t.maxw = g.maxw;
        int off = 0;
        for (int l = 0; l <= L; ++l) { t.dims[l] = g.dims[l]; t.dimsp[l] = (g.dims[l] + 31) & ~31; t.boff[l] = g.boff[l]; }
        for (int l = 0; l < L; ++l) { t.woffp[l] = off; off += g.dims[l] * t.dimsp[l + 1]; }
        t.woffp[L] = off;
        std::vector<double> wp((size_t)off, 0.0);
        for (int l = 0; l < L; ++l)
            for (int i = 0; i < g.dims[l]; ++i)
                for (int j = 0; j < g.dims[l + 1]; ++j)
                    wp[(size_t)t.woffp[l] + (size_t)i * t.dimsp[l + 1] + j] = m->hW64[(size_t)g.woff[l] + (size_t)i * g.dims[l + 1] + j];
        e = cudaMalloc(&m->dWpad, sizeof(double) * wp.size());
        if (e == cudaSuccess) e = cudaMemcpy(m->dWpad, wp.data(), sizeof(double) * wp.size(), cudaMemcpyHostToDevice);
        if (e != cudaSuccess) { fb_model_destroy(m); return fb_fail(FB_ERR_CUDA, "fb_model_create: %s", cudaGetErrorString(e)); }
        std::vector<float> wp32(wp.size());
        for (size_t k = 0; k < wp.size(); ++k) wp32[k] = (float)wp[k];                 // exact: the weights are float32 values
        e = cudaMalloc(&m->dWpad32, sizeof(float) * wp32.size());
        if (e == cudaSuccess) e = cudaMemcpy(m->dWpad32, wp32.data(), sizeof(float) * wp32.size(), cudaMemcpyHostToDevice);
        if (e != cudaSuccess) { fb_model_destroy(m); return fb_fail(FB_ERR_CUDA, "fb_model_create: %s", cudaGetErrorString(e)); }
        // the interval form of the float32 weights (fb_tile_forward_f32<.., true>): rows max(w_k, 0), min(w_k, 0) alternate
        std::vector<float> wiv(2 * wp32.size(), 0.0f);
        for (int l = 0; l < L; ++l)
            for (int i = 0; i < g.dims[l]; ++i)
                for (int j = 0; j < t.dimsp[l + 1]; ++j) {
                    const float w = wp32[(size_t)t.woffp[l] + (size_t)i * t.dimsp[l + 1] + j];
                    wiv[2 * (size_t)t.woffp[l] + (size_t)(2 * i) * t.dimsp[l + 1] + j] = w > 0.0f ? w : 0.0f;
                    wiv[2 * (size_t)t.woffp[l] + (size_t)(2 * i + 1) * t.dimsp[l + 1] + j] = w < 0.0f ? w : 0.0f;
                }
        e = cudaMalloc(&m->dWiv32, sizeof(float) * wiv.size());
        if (e == cudaSuccess) e = cudaMemcpy(m->dWiv32, wiv.data(), sizeof(float) * wiv.size(), cudaMemcpyHostToDevice);
        if (e != cudaSuccess) { fb_model_destroy(m); return fb_fail(FB_ERR_CUDA, "fb_model_create: %s", cudaGetErrorString(e)); }
        t.Wp = m->dWpad; t.B = m->dB64; t.Wp32 = m->dWpad32; t.Wiv32 = m->dWiv32; t.B32 = m->dB32;
    }
    {
        // weights in the operand order of the FP64 matrix instruction, columns of every tile permuted (fb_mma.cuh)
        FbMmaNet &q = m->mma;
        q.L = L; q.N = g.N;
        int wo = 0, bo2 = 0;
        for (int l = 0; l <= L; ++l) { q.dims[l] = g.dims[l]; q.boff[l] = g.boff[l]; }
        q.maxnb = ((g.dims[0] + 3) / 4 + 1) / 2;
        for (int l = 0; l < L; ++l) {
            q.kb[l] = (g.dims[l] + 3) / 4; q.nb[l] = (g.dims[l + 1] + 7) / 8;
            q.maxnb = std::max(q.maxnb, q.nb[l]);
            q.woff[l] = wo; q.bpoff[l] = bo2;
            wo += q.kb[l] * q.nb[l] * 32;
            bo2 += q.nb[l] * 8;
        }
        q.woff[L] = wo; q.bpoff[L] = bo2; q.wtotal = wo; q.btotal = bo2;
        std::vector<double> wf((size_t)wo, 0.0), bp((size_t)bo2, 0.0);
        for (int l = 0; l < L; ++l) {
            const int IN = g.dims[l], OUT = g.dims[l + 1];
            const double *W = m->hW64.data() + g.woff[l];
            const int KB = q.kb[l], NB = q.nb[l];
            for (int tile = 0; tile < NB; ++tile) {
                const int cg = tile / FB_MMA_NT, nt = tile % FB_MMA_NT, ntg = std::min(FB_MMA_NT, NB - cg * FB_MMA_NT);
                for (int c = 0; c < 8; ++c) {
                    const int j = 8 * tile + fb_mma_col_neuron(c);
                    if (j < OUT) bp[(size_t)q.bpoff[l] + 8 * tile + c] = m->hB64[(size_t)g.boff[l] + j];
                }
                for (int kb = 0; kb < KB; ++kb)
                    for (int lane = 0; lane < 32; ++lane) {
                        const int k = 4 * kb + (lane & 3), j = 8 * tile + fb_mma_col_neuron(lane >> 2);
                        if (k < IN && j < OUT)
                            wf[(size_t)q.woff[l] + ((size_t)cg * KB * FB_MMA_NT + (size_t)kb * ntg + nt) * 32 + lane] = W[(size_t)k * OUT + j];
                    }
            }
        }
        e = cudaMalloc(&m->dWf, sizeof(double) * wf.size());
        if (e == cudaSuccess) e = cudaMalloc(&m->dBp, sizeof(double) * bp.size());
        if (e == cudaSuccess) e = cudaMemcpy(m->dWf, wf.data(), sizeof(double) * wf.size(), cudaMemcpyHostToDevice);
        if (e == cudaSuccess) e = cudaMemcpy(m->dBp, bp.data(), sizeof(double) * bp.size(), cudaMemcpyHostToDevice);
        if (e != cudaSuccess) { fb_model_destroy(m); return fb_fail(FB_ERR_CUDA, "fb_model_create: %s", cudaGetErrorString(e)); }
        q.Wf = m->dWf; q.Bp = m->dBp;
    }
    m->spec = fb_find_spec(fb_fwd_tables, L, dims);
    m->spec_profile = fb_find_spec(fb_profile_tables, L, dims);
    m->spec_ibp = fb_find_spec(fb_ibp_tables, L, dims);
    m->spec_mma = fb_find_spec(fb_mma_tables, L, dims);
    if (!m->spec_profile || !m->spec_ibp) m->spec = nullptr;
    if (m->spec && !fb_model_subnormal_free(g, m->hW64, m->hB64)) m->spec = nullptr;   // runtime-shape kernels: exact sign test
    if (m->spec) {
        m->p64 = m->spec->make_params64(m->hW64.data(), m->hB64.data());
        m->p32 = m->spec->make_params32(m->hW32.data(), m->hB32.data());
        if (!m->p64 || !m->p32) { fb_model_destroy(m); return fb_fail(FB_ERR_NOMEM, "fb_model_create: out of host memory"); }
    }
    *out = m;
    return FB_OK;
}

extern "C" int fb_model_info(const fb_model *m, int *L, int *N, int *P, int *specialised)
{
    if (!m) return fb_fail(FB_ERR_INVALID, "fb_model_info: null model");
    if (L) *L = m->gen.L;
    if (N) *N = m->gen.N;
    if (P) *P = m->gen.P;
    if (specialised) *specialised = m->spec ? 1 : 0;
    return FB_OK;
}

// Scratch and staging buffers come from a stream-ordered memory pool that belongs to this library (one per device,
// created on first use), not from the device's default pool: freed blocks stay cached in it between calls -- a caller
// looping over fb_prune_host would otherwise re-acquire gigabytes from the driver per call -- without changing the
// release behaviour of a pool the embedding application may share.  fb_release_cache() hands the cached memory back.
static std::mutex fb_pool_mu;
static cudaMemPool_t fb_pools[64] = {};

static cudaError_t fb_pool_for(int device, cudaMemPool_t *out)
{
    if (device < 0 || device >= 64) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lock(fb_pool_mu);
    if (!fb_pools[device]) {
        cudaMemPoolProps props = {};
        props.allocType = cudaMemAllocationTypePinned;
        props.handleTypes = cudaMemHandleTypeNone;
        props.location.type = cudaMemLocationTypeDevice;
        props.location.id = device;
        cudaMemPool_t pool;
        cudaError_t e = cudaMemPoolCreate(&pool, &props);
        if (e != cudaSuccess) return e;
        unsigned long long keep = ~0ull;
        cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep);
        fb_pools[device] = pool;
    }
    *out = fb_pools[device];
    return cudaSuccess;
}

static cudaError_t fb_malloc_async(void **p, size_t bytes, int device, cudaStream_t st)
{
    cudaMemPool_t pool;
    cudaError_t e = fb_pool_for(device, &pool);
    if (e != cudaSuccess) return e;
    return cudaMallocFromPoolAsync(p, bytes ? bytes : 1, pool, st);
}

extern "C" int fb_release_cache(void)
{
    std::lock_guard<std::mutex> lock(fb_pool_mu);
    for (int d = 0; d < 64; ++d)
        if (fb_pools[d]) FB_CUDA(cudaMemPoolTrimTo(fb_pools[d], 0));
    return FB_OK;
}

// set FAIRIFY_B200_FORCE_GENERIC=1 to route every model through the runtime-shape kernels (tests)
static bool fb_use_spec(const fb_model *m)
{
    static const bool force_generic = [] { const char *v = getenv("FAIRIFY_B200_FORCE_GENERIC"); return v && v[0] == '1'; }();
    return m->spec != nullptr && !force_generic;
}

// ---------------------------------------------------------------------------------------------------
// partition grid
// ---------------------------------------------------------------------------------------------------
#define FB_MAX_D 64
struct FbGridDesc {
    int d;
    int col[FB_MAX_D];          // column of the k-th key (digit order: last key fastest)
    int64_t lo[FB_MAX_D], hi[FB_MAX_D], nch[FB_MAX_D];   // by column
    int64_t T;
};
struct fb_grid { FbGridDesc g; int64_t n; };

__global__ void __launch_bounds__(256)
fb_grid_boxes_kernel(const FbGridDesc g, int64_t first, int64_t count, int32_t *__restrict__ boxes)
{
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; p < count; p += stride) {
        int64_t idx = first + p;
        int2 *box = reinterpret_cast<int2 *>(boxes + p * (2 * g.d));
        for (int k = g.d - 1; k >= 0; --k) {
            const int c = g.col[k];
            const int64_t nc = g.nch[c];
            int64_t a = g.lo[c], b = g.hi[c];
            if (nc > 1) {
                const int64_t digit = idx % nc;
                idx /= nc;
                a = g.lo[c] + digit * g.T;
                b = min(a + g.T - 1, g.hi[c]);
            }
            box[c] = make_int2((int)a, (int)b);
        }
    }
}

extern "C" int fb_grid_create(int d, const int64_t *lo, const int64_t *hi, int64_t T, const int32_t *key_order, fb_grid **out)
{
    if (!lo || !hi || !key_order || !out) return fb_fail(FB_ERR_INVALID, "fb_grid_create: null argument");
    if (d < 1 || d > FB_MAX_D) return fb_fail(FB_ERR_UNSUPPORTED, "fb_grid_create: %d features (supported: 1..%d)", d, FB_MAX_D);
    if (T < 1) return fb_fail(FB_ERR_INVALID, "fb_grid_create: partition size %lld", (long long)T);
    fb_grid *g = new (std::nothrow) fb_grid;
    if (!g) return fb_fail(FB_ERR_NOMEM, "fb_grid_create: out of host memory");
    g->g.d = d; g->g.T = T; g->n = 1;
    std::vector<int> seen(d, 0);
    for (int k = 0; k < d; ++k) {
        const int c = key_order[k];
        if (c < 0 || c >= d || seen[c]) { delete g; return fb_fail(FB_ERR_INVALID, "fb_grid_create: key_order is not a permutation"); }
        seen[c] = 1; g->g.col[k] = c;
    }
    for (int c = 0; c < d; ++c) {
        if (hi[c] < lo[c] || lo[c] < INT32_MIN || hi[c] > INT32_MAX) { delete g; return fb_fail(FB_ERR_INVALID, "fb_grid_create: bad range for column %d", c); }
        const int64_t size = hi[c] - lo[c] + 1;
        g->g.lo[c] = lo[c]; g->g.hi[c] = hi[c];
        g->g.nch[c] = (size <= T) ? 1 : (size + T - 1) / T;     // utils/input_partition.py:24-27
        if (g->n > INT64_MAX / g->g.nch[c]) { delete g; return fb_fail(FB_ERR_UNSUPPORTED, "fb_grid_create: grid size overflows int64"); }
        g->n *= g->g.nch[c];
    }
    *out = g;
    return FB_OK;
}
extern "C" void fb_grid_destroy(fb_grid *g) { delete g; }
extern "C" int64_t fb_grid_size(const fb_grid *g) { return g ? g->n : -1; }

extern "C" int fb_grid_boxes(const fb_grid *g, int64_t first, int64_t count, int32_t *boxes, void *stream)
{
    if (!g) return fb_fail(FB_ERR_INVALID, "fb_grid_boxes: null grid");
    if (first < 0 || count < 0 || first + count > g->n) return fb_fail(FB_ERR_INVALID, "fb_grid_boxes: range [%lld, %lld) outside the grid of %lld", (long long)first, (long long)(first + count), (long long)g->n);
    if (count == 0) return FB_OK;
    if (!boxes) return fb_fail(FB_ERR_INVALID, "fb_grid_boxes: null output");
    int sms = 0;
    if (fb_num_sms(&sms)) return FB_ERR_CUDA;
    const int grid = (int)std::min<int64_t>((count + 255) / 256, (int64_t)sms * 8);
    fb_grid_boxes_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(g->g, first, count, boxes);
    FB_CUDA(cudaGetLastError());
    return FB_OK;
}

// point -> partition index (and per-partition row counts): the membership test of
// find_partition_result_for_point (src/AC/Verify-AC-experiment-new2.py:587-611) and p_list_density
// (utils/input_partition.py:184-218) as O(d) index arithmetic per row instead of a scan over all partitions
__global__ void __launch_bounds__(256)
fb_grid_locate_kernel(const FbGridDesc g, int64_t m, const double *__restrict__ X, int64_t *__restrict__ index,
                      unsigned int *__restrict__ density)
{
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; r < m; r += stride) {
        int64_t idx = 0;
        bool ok = true;
        for (int k = 0; k < g.d && ok; ++k) {
            const int c = g.col[k];
            const double x = X[r * g.d + c];
            if (!(x >= (double)g.lo[c] && x <= (double)g.hi[c])) { ok = false; break; }
            int64_t digit = 0;
            if (g.nch[c] > 1) {
                digit = (int64_t)floor(__ddiv_rn(__dadd_rn(x, -(double)g.lo[c]), (double)g.T));
                const int64_t top = min(g.lo[c] + digit * g.T + g.T - 1, g.hi[c]);
                if (x > (double)top) { ok = false; break; }
            }
            idx = idx * g.nch[c] + digit;
        }
        if (index) index[r] = ok ? idx : -1;
        if (density && ok) atomicAdd(&density[idx], 1u);
    }
}

extern "C" int fb_grid_locate(const fb_grid *g, int64_t m, const double *X, int64_t *index, uint32_t *density, void *stream)
{
    if (!g || m < 0) return fb_fail(FB_ERR_INVALID, "fb_grid_locate: bad argument");
    if (m == 0) return FB_OK;
    if (!X || (!index && !density)) return fb_fail(FB_ERR_INVALID, "fb_grid_locate: null argument");
    int sms = 0;
    if (fb_num_sms(&sms)) return FB_ERR_CUDA;
    const int grid = (int)std::min<int64_t>((m + 255) / 256, (int64_t)sms * 8);
    fb_grid_locate_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(g->g, m, X, index, density);
    FB_CUDA(cudaGetLastError());
    return FB_OK;
}

// ---------------------------------------------------------------------------------------------------
// profile
// ---------------------------------------------------------------------------------------------------
// Samples are split into `chunks` slices per partition only when there are too few partitions to fill
// the machine with one warp each.
static void fb_pick_chunks(const fb_model *m, int64_t n, int S, int *chunks, int *chunk_len)
{
    const int max_len = FB_PROFILE_MAX_STEPS * 32 * FB_PROFILE_SPT;
    const int64_t want_warps = (int64_t)m->num_sms * 64;
    int c = 1;
    const int gran = 32 * FB_PROFILE_SPT;       // samples one warp step covers in the specialised kernel
    if (n < want_warps) c = (int)std::min<int64_t>((S + gran - 1) / gran, (want_warps + n - 1) / n);
    c = std::max(c, (S + max_len - 1) / max_len);
    int len = (S + c - 1) / c;
    len = ((len + gran - 1) / gran) * gran;
    c = (S + len - 1) / len;
    *chunks = std::max(c, 1); *chunk_len = len;
}

// every launch uses the model's device buffers: the model's device must be the current one
static int fb_check_device(const char *fn, const fb_model *m)
{
    int dev = -1;
    FB_CUDA(cudaGetDevice(&dev));
    if (dev != m->device)
        return fb_fail(FB_ERR_INVALID, "%s: the model lives on device %d but device %d is current", fn, m->device, dev);
    return FB_OK;
}

extern "C" int fb_model_device(const fb_model *m, int *device)
{
    if (!m || !device) return fb_fail(FB_ERR_INVALID, "fb_model_device: null argument");
    *device = m->device;
    return FB_OK;
}

static int fb_check_batch(const char *fn, const fb_model *m, int64_t n, const void *boxes, int S, bool need_s)
{
    if (!m) return fb_fail(FB_ERR_INVALID, "%s: null model", fn);
    if (int rc = fb_check_device(fn, m)) return rc;
    if (n < 0) return fb_fail(FB_ERR_INVALID, "%s: n = %lld", fn, (long long)n);
    if (n > 0 && !boxes) return fb_fail(FB_ERR_INVALID, "%s: null boxes", fn);
    if (need_s && (S < 1 || S > 65535)) return fb_fail(FB_ERR_INVALID, "%s: S = %d (supported: 1..65535)", fn, S);
    return FB_OK;
}

// Runtime-shape profile and partition screen: the tiled kernels (fb_generic.cuh) for the big networks, the sample-per-
// lane kernels for the rest.  Measured on B200 (% of FP64 peak, tiled vs sample-per-lane profile): BM-4 30 vs 11, AC-4
// 28 vs 13, AC-5 / BM-7 37 vs 20, a 4004-weight funnel of mostly narrow layers (GC-5's shape) 23 vs 21; small networks forced
// onto this path lose (AC-1 6 %), hence the weight-count threshold.
// FAIRIFY_B200_TILED_PROFILE=1 / 0 forces one or the other (tests, comparisons).
static bool fb_use_tiled_profile(const fb_model *m)
{
    static const int forced = [] { const char *v = getenv("FAIRIFY_B200_TILED_PROFILE"); return !v ? -1 : (v[0] == '0' ? 0 : 1); }();
    if (forced >= 0) return forced == 1;
    return m->gen.P >= 4000;
}

// FAIRIFY_B200_MMA_PROFILE=1 sends every runtime-shape profile through the DMMA kernel, =0 none (tests, comparisons);
// by default the networks that used to take the tiled kernel do.
static bool fb_use_mma_profile(const fb_model *m)
{
    static const int forced = [] { const char *v = getenv("FAIRIFY_B200_MMA_PROFILE"); return !v ? -1 : (v[0] == '0' ? 0 : 1); }();
    if (forced >= 0) return forced == 1;
    return fb_use_tiled_profile(m);
}

template <int TM, bool RESIDENT>
static int fb_profile_tiled_launch(const fb_model *m, int64_t n, const int32_t *boxes, const int32_t *samples, int S,
                                   uint64_t seed, int64_t first_index, uint32_t *counts, cudaStream_t st)
{
    const size_t smem = RESIDENT ? fb_tile_smem_bytes_resident(m->tile.maxw, m->tile.woffp[m->tile.L], m->gen.N, TM)
                                 : fb_tile_smem_bytes(m->tile.maxw, (m->tile.maxw + 31) & ~31, m->gen.N, TM);
    const int64_t units = n * ((S + TM - 1) / TM);
    if (samples == nullptr) {
        auto k = fb_profile_tile_kernel<TM, true, RESIDENT>;
        FB_CUDA(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        k<<<fb_grid_for(k, FB_TILE_THREADS, smem, units, m->num_sms), FB_TILE_THREADS, smem, st>>>(
            m->tile, n, boxes, samples, seed, first_index, S, counts);
    } else {
        auto k = fb_profile_tile_kernel<TM, false, RESIDENT>;
        FB_CUDA(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        k<<<fb_grid_for(k, FB_TILE_THREADS, smem, units, m->num_sms), FB_TILE_THREADS, smem, st>>>(
            m->tile, n, boxes, samples, seed, first_index, S, counts);
    }
    FB_CUDA(cudaGetLastError());
    return FB_OK;
}

static int fb_profile_tiled(const fb_model *m, int64_t n, const int32_t *boxes, const int32_t *samples, int S, uint64_t seed,
                            int64_t first_index, uint32_t *counts, bool already_zeroed, cudaStream_t st)
{
    if (!already_zeroed) FB_CUDA(cudaMemsetAsync(counts, 0, sizeof(uint32_t) * n * m->gen.N, st));   // the kernel adds
    const size_t limit = 220 * 1024;
    static const int forced_tm = [] { const char *v = getenv("FAIRIFY_B200_TILE_TM"); return v ? atoi(v) : 0; }();   // tuning aid
    if (forced_tm == 32) return fb_profile_tiled_launch<32, false>(m, n, boxes, samples, S, seed, first_index, counts, st);
    if (forced_tm == 64) return fb_profile_tiled_launch<64, false>(m, n, boxes, samples, S, seed, first_index, counts, st);
    // 128 samples per CTA with all weights resident in shared memory when that fits (networks up to ~64 wide), else
    // 128 samples with staged weights, else 64 samples with staged weights (widths above 100)
    if (fb_tile_smem_bytes_resident(m->tile.maxw, m->tile.woffp[m->tile.L], m->gen.N, 128) <= limit && forced_tm != 128)
        return fb_profile_tiled_launch<128, true>(m, n, boxes, samples, S, seed, first_index, counts, st);
    if (fb_tile_smem_bytes(m->tile.maxw, (m->tile.maxw + 31) & ~31, m->gen.N, 128) <= limit)
        return fb_profile_tiled_launch<128, false>(m, n, boxes, samples, S, seed, first_index, counts, st);
    return fb_profile_tiled_launch<64, false>(m, n, boxes, samples, S, seed, first_index, counts, st);
}

template <int TM>
static int fb_screen_tiled_launch(const fb_model *m, int64_t n, const int32_t *boxes, const int32_t *samples, int S,
                                  uint64_t seed, int64_t first_index, int pa_col, int ra_col, int ra_thr,
                                  unsigned long long *key, cudaStream_t st)
{
    const size_t smem = fb_screen_tile_smem_bytes(m->tile.maxw, (m->tile.maxw + 31) & ~31, TM);
    if (samples == nullptr) {
        auto k = fb_screen_tile_kernel<TM, true>;
        FB_CUDA(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        k<<<fb_grid_for(k, FB_TILE_THREADS, smem, n, m->num_sms), FB_TILE_THREADS, smem, st>>>(
            m->tile, n, boxes, samples, seed, first_index, S, pa_col, ra_col, ra_thr, key);
    } else {
        auto k = fb_screen_tile_kernel<TM, false>;
        FB_CUDA(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        k<<<fb_grid_for(k, FB_TILE_THREADS, smem, n, m->num_sms), FB_TILE_THREADS, smem, st>>>(
            m->tile, n, boxes, samples, seed, first_index, S, pa_col, ra_col, ra_thr, key);
    }
    FB_CUDA(cudaGetLastError());
    return FB_OK;
}

static int fb_screen_tiled(const fb_model *m, int64_t n, const int32_t *boxes, const int32_t *samples, int S, uint64_t seed,
                           int64_t first_index, int pa_col, int ra_col, int ra_thr, unsigned long long *key, cudaStream_t st)
{
    // 128-row tiles where two CTAs of that size still share an SM (networks up to 64 wide: BM-7 38 vs 52 ms per 10^5
    // partitions), 64-row tiles otherwise (BM-4 42 vs 48 ms per 2*10^4, AC-4 90 vs 109 ms per 10^5 with 128)
    static const int forced_tm = [] { const char *v = getenv("FAIRIFY_B200_SCREEN_TM"); return v ? atoi(v) : 0; }();   // tuning aid
    const size_t smem128 = fb_screen_tile_smem_bytes(m->tile.maxw, (m->tile.maxw + 31) & ~31, 128);
    if (forced_tm != 64 && (smem128 <= 100 * 1024 || (forced_tm == 128 && smem128 <= 220 * 1024)))
        return fb_screen_tiled_launch<128>(m, n, boxes, samples, S, seed, first_index, pa_col, ra_col, ra_thr, key, st);
    return fb_screen_tiled_launch<64>(m, n, boxes, samples, S, seed, first_index, pa_col, ra_col, ra_thr, key, st);
}

static size_t fb_profile_gen_smem(const fb_model *m)
{
    return sizeof(double) * ((size_t)m->gen.P + m->gen.N) + sizeof(uint32_t) * (FB_PROFILE_THREADS / 32) * m->gen.N;
}

extern "C" int fb_profile(const fb_model *m, int64_t n, const int32_t *boxes, const int32_t *samples, int S, uint64_t seed,
                          int64_t first_index, uint32_t *counts, void *stream)
{
    if (int rc = fb_check_batch("fb_profile", m, n, boxes, S, true)) return rc;
    if (n == 0) return FB_OK;
    if (!counts) return fb_fail(FB_ERR_INVALID, "fb_profile: null counts");
    cudaStream_t st = (cudaStream_t)stream;
    int chunks, chunk_len;
    fb_pick_chunks(m, n, S, &chunks, &chunk_len);
    if (chunks > 1) FB_CUDA(cudaMemsetAsync(counts, 0, sizeof(uint32_t) * n * m->gen.N, st));
    // 1. shape-specialised DMMA kernel (the wide and mid-size reference networks; FAIRIFY_B200_MMA_PROFILE=0 turns the
    //    DMMA kernels off, FAIRIFY_B200_FORCE_GENERIC=1 the shape-specialised ones: tests, comparisons)
    // 2. shape-specialised FMA kernel (the small networks)   3. runtime-shape DMMA kernel   4. the older runtime-shape kernels
    static const int mma_env = [] { const char *v = getenv("FAIRIFY_B200_MMA_PROFILE"); return !v ? -1 : (v[0] == '0' ? 0 : 1); }();
    static const bool force_generic = [] { const char *v = getenv("FAIRIFY_B200_FORCE_GENERIC"); return v && v[0] == '1'; }();
    if (m->spec_mma && mma_env != 0 && !force_generic) {
        FB_CUDA(m->spec_mma->profile(m->mma, n, boxes, samples, seed, first_index, S, chunks, chunk_len, counts, m->num_sms, st));
        return FB_OK;
    }
    if (fb_use_spec(m)) {
        FB_CUDA(m->spec_profile->profile(m->p64, samples == nullptr, n, boxes, samples, seed, first_index, S, chunks, chunk_len,
                                         counts, m->num_sms, st));
        return FB_OK;
    }
    if (fb_use_mma_profile(m)) {
        FB_CUDA(fb_profile_mma(m->mma, n, boxes, samples, seed, first_index, S, chunks, chunk_len, counts, m->num_sms, st));
        return FB_OK;
    }
    if (fb_use_tiled_profile(m)) return fb_profile_tiled(m, n, boxes, samples, S, seed, first_index, counts, chunks > 1, st);
    const size_t smem = fb_profile_gen_smem(m);
    if (smem > 227 * 1024) return fb_fail(FB_ERR_UNSUPPORTED, "fb_profile: model needs %zu bytes of shared memory", smem);
    const int64_t ctas = (n * chunks + (FB_PROFILE_THREADS / 32) - 1) / (FB_PROFILE_THREADS / 32);
    if (samples == nullptr) {
        auto k = fb_profile_gen_kernel<true>;
        FB_CUDA(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        k<<<fb_grid_for(k, FB_PROFILE_THREADS, smem, ctas, m->num_sms), FB_PROFILE_THREADS, smem, st>>>(
            m->gen, n, boxes, samples, seed, first_index, S, chunks, chunk_len, counts);
    } else {
        auto k = fb_profile_gen_kernel<false>;
        FB_CUDA(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        k<<<fb_grid_for(k, FB_PROFILE_THREADS, smem, ctas, m->num_sms), FB_PROFILE_THREADS, smem, st>>>(
            m->gen, n, boxes, samples, seed, first_index, S, chunks, chunk_len, counts);
    }
    FB_CUDA(cudaGetLastError());
    return FB_OK;
}

// ---------------------------------------------------------------------------------------------------
// interval bounds + masks
// ---------------------------------------------------------------------------------------------------
extern "C" int fb_ibp(const fb_model *m, int64_t n, const int32_t *boxes, const uint32_t *counts, const double *tol,
                      const fb_prune_out *out, void *stream)
{
    if (int rc = fb_check_batch("fb_ibp", m, n, boxes, 1, false)) return rc;
    if (!out) return fb_fail(FB_ERR_INVALID, "fb_ibp: null output block");
    if (n == 0) return FB_OK;
    cudaStream_t st = (cudaStream_t)stream;
    const FbPruneOut *o = reinterpret_cast<const FbPruneOut *>(out);
    // (shapes above ~1200 weights have no specialised bounds kernel: fb_spec_ops_ibp.cuh)
    if (fb_use_spec(m) && m->spec_ibp->ibp != nullptr) {
        FB_CUDA(m->spec_ibp->ibp(m->p64, n, boxes, counts, tol, m->dUnb, o, m->num_sms, st));
        return FB_OK;
    }
    auto k = fb_ibp_gen_kernel;
    const int64_t ctas = (n + FB_IBP_GEN_WARPS - 1) / FB_IBP_GEN_WARPS;
    k<<<fb_grid_for(k, FB_IBP_GEN_WARPS * 32, 0, ctas, m->num_sms), FB_IBP_GEN_WARPS * 32, 0, st>>>(m->gen, n, boxes, counts,
                                                                                                   tol, *o);
    FB_CUDA(cudaGetLastError());
    return FB_OK;
}

extern "C" int fb_prune(const fb_model *m, int64_t n, const int32_t *boxes, const int32_t *samples, int S, uint64_t seed,
                        int64_t first_index, const double *tol, const fb_prune_out *out, void *stream)
{
    if (int rc = fb_check_batch("fb_prune", m, n, boxes, S, true)) return rc;
    if (!out) return fb_fail(FB_ERR_INVALID, "fb_prune: null output block");
    if (n == 0) return FB_OK;
    cudaStream_t st = (cudaStream_t)stream;
    uint32_t *counts = nullptr;
    FB_CUDA(fb_malloc_async((void **)&counts, sizeof(uint32_t) * n * m->gen.N, m->device, st));
    int rc = fb_profile(m, n, boxes, samples, S, seed, first_index, counts, stream);
    if (rc == FB_OK) rc = fb_ibp(m, n, boxes, counts, tol, out, stream);
    cudaFreeAsync(counts, st);
    return rc;
}

// host-buffer variant ------------------------------------------------------------------------------
namespace {
struct DevBuf {
    void *p = nullptr;
    int device;
    cudaStream_t st;
    DevBuf(int dev, cudaStream_t s) : device(dev), st(s) {}
    ~DevBuf() { if (p) cudaFreeAsync(p, st); }
    cudaError_t alloc(size_t bytes) { return fb_malloc_async(&p, bytes, device, st); }
};
}  // namespace

// Copy streams of fb_prune_host, one pair per device, created on first use and kept.
static int fb_copy_streams(int device, cudaStream_t *up, cudaStream_t *down)
{
    static cudaStream_t s_up[64] = {}, s_down[64] = {};
    static std::mutex mu;
    if (device < 0 || device >= 64) return fb_fail(FB_ERR_UNSUPPORTED, "fb_prune_host: device ordinal %d", device);
    std::lock_guard<std::mutex> lock(mu);
    if (!s_up[device]) {
        FB_CUDA(cudaStreamCreateWithFlags(&s_up[device], cudaStreamNonBlocking));
        FB_CUDA(cudaStreamCreateWithFlags(&s_down[device], cudaStreamNonBlocking));
    }
    *up = s_up[device]; *down = s_down[device];
    return FB_OK;
}

// Partitions per slice of the host pipeline: big enough that a slice fills the GPU many times over, small enough
// that the first upload and the last download (the only copies not hidden behind compute) are short.
#define FB_HOST_SLICE 1000000

extern "C" int fb_prune_host(const fb_model *m, int64_t n, const int32_t *boxes, const int32_t *samples, int S,
                             uint64_t seed, int64_t first_index, const double *tol, const fb_prune_out *out, void *stream)
{
    if (int rc = fb_check_batch("fb_prune_host", m, n, boxes, S, true)) return rc;
    if (!out) return fb_fail(FB_ERR_INVALID, "fb_prune_host: null output block");
    if (n == 0) return FB_OK;
    cudaStream_t st = (cudaStream_t)stream;
    const int N = m->gen.N, NW = (N + 31) / 32, d = m->gen.dims[0];
    cudaStream_t up, down;
    if (int rc = fb_copy_streams(m->device, &up, &down)) return rc;

    // device copies of the inputs and one device buffer per requested output (whole batch; freed on `st` at return)
    DevBuf dboxes(m->device, st), dsamp(m->device, st), dtol(m->device, st);
    FB_CUDA(dboxes.alloc(sizeof(int32_t) * n * d * 2));
    if (samples) FB_CUDA(dsamp.alloc(sizeof(int32_t) * n * S * d));
    if (tol) {
        FB_CUDA(dtol.alloc(sizeof(double) * N));
        FB_CUDA(cudaMemcpyAsync(dtol.p, tol, sizeof(double) * N, cudaMemcpyHostToDevice, st));
    }
    struct Slot { void *host; size_t row_bytes; void **dev_field; void *dev; };
    fb_prune_out dev{};
    const size_t mb = sizeof(uint32_t) * NW, fb = sizeof(double) * N;
    Slot slots[] = {
        {out->cand, mb, (void **)&dev.cand, nullptr},     {out->b_dead, mb, (void **)&dev.b_dead, nullptr},
        {out->dead, mb, (void **)&dev.dead, nullptr},     {out->sliver, mb, (void **)&dev.sliver, nullptr},
        {out->verdict, 1, (void **)&dev.verdict, nullptr}, {out->counts, sizeof(uint16_t) * N, (void **)&dev.counts, nullptr},
        {out->ws_lb, fb, (void **)&dev.ws_lb, nullptr},   {out->ws_ub, fb, (void **)&dev.ws_ub, nullptr},
        {out->pl_lb, fb, (void **)&dev.pl_lb, nullptr},   {out->pl_ub, fb, (void **)&dev.pl_ub, nullptr},
        {out->lb_rd, fb, (void **)&dev.lb_rd, nullptr},   {out->ub_ru, fb, (void **)&dev.ub_ru, nullptr},
        {out->s_dead, mb, (void **)&dev.s_dead, nullptr}, {out->s_cand, mb, (void **)&dev.s_cand, nullptr},
        {out->record, sizeof(uint32_t) * (2 * NW + 2), (void **)&dev.record, nullptr},
    };
    std::vector<DevBuf> bufs;
    bufs.reserve(sizeof(slots) / sizeof(slots[0]));
    for (Slot &s : slots) {
        if (!s.host) continue;
        bufs.emplace_back(m->device, st);
        FB_CUDA(bufs.back().alloc(s.row_bytes * n));
        s.dev = bufs.back().p;
    }

    // Three-stage pipeline over slices of the batch: upload on `up`, kernels on the caller's stream, download on
    // `down`; slice k's kernels overlap slice k+1's upload and slice k-1's download.  Results do not depend on the
    // slicing (every partition's samples are a function of its global index only).
    struct Events {
        std::vector<cudaEvent_t> v;
        ~Events() { for (cudaEvent_t e : v) cudaEventDestroy(e); }
        cudaError_t make(cudaEvent_t *e)
        {
            cudaError_t rc = cudaEventCreateWithFlags(e, cudaEventDisableTiming);
            if (rc == cudaSuccess) v.push_back(*e);
            return rc;
        }
    } events;
    // whatever way this function is left, nothing issued on the three streams may outlive the buffers
    struct Drain {
        cudaStream_t a, b, c;
        ~Drain() { cudaStreamSynchronize(a); cudaStreamSynchronize(b); cudaStreamSynchronize(c); }
    } drain{up, st, down};
    cudaEvent_t ready;                                   // allocations (and tol) are in place on `st`
    FB_CUDA(events.make(&ready));
    FB_CUDA(cudaEventRecord(ready, st));
    FB_CUDA(cudaStreamWaitEvent(up, ready, 0));
    FB_CUDA(cudaStreamWaitEvent(down, ready, 0));
    int rc = FB_OK;
    static const int64_t slice = [] {                    // FAIRIFY_B200_HOST_SLICE overrides the slice size (tuning aid)
        const char *v = getenv("FAIRIFY_B200_HOST_SLICE");
        const long long x = v ? atoll(v) : 0;
        return x > 0 ? (int64_t)x : (int64_t)FB_HOST_SLICE;
    }();
    for (int64_t off = 0; off < n && rc == FB_OK; off += slice) {
        const int64_t cnt = std::min<int64_t>(slice, n - off);
        cudaEvent_t uploaded, computed;
        FB_CUDA(events.make(&uploaded));
        FB_CUDA(events.make(&computed));
        int32_t *bx = (int32_t *)dboxes.p + off * d * 2;
        FB_CUDA(cudaMemcpyAsync(bx, boxes + off * d * 2, sizeof(int32_t) * cnt * d * 2, cudaMemcpyHostToDevice, up));
        int32_t *sm = nullptr;
        if (samples) {
            sm = (int32_t *)dsamp.p + off * S * d;
            FB_CUDA(cudaMemcpyAsync(sm, samples + off * S * d, sizeof(int32_t) * cnt * S * d, cudaMemcpyHostToDevice, up));
        }
        FB_CUDA(cudaEventRecord(uploaded, up));
        FB_CUDA(cudaStreamWaitEvent(st, uploaded, 0));
        for (Slot &s : slots) *s.dev_field = s.dev ? (char *)s.dev + s.row_bytes * off : nullptr;
        rc = fb_prune(m, cnt, bx, sm, S, seed, first_index + off, (const double *)dtol.p, &dev, stream);
        if (rc != FB_OK) break;
        FB_CUDA(cudaEventRecord(computed, st));
        FB_CUDA(cudaStreamWaitEvent(down, computed, 0));
        for (Slot &s : slots)
            if (s.host)
                FB_CUDA(cudaMemcpyAsync((char *)s.host + s.row_bytes * off, (char *)s.dev + s.row_bytes * off,
                                        s.row_bytes * cnt, cudaMemcpyDeviceToHost, down));
    }
    // everything issued must finish before the buffers are released (also on the error path)
    cudaError_t e1 = cudaStreamSynchronize(up), e2 = cudaStreamSynchronize(st), e3 = cudaStreamSynchronize(down);
    if (rc != FB_OK) return rc;
    FB_CUDA(e1); FB_CUDA(e2); FB_CUDA(e3);
    return FB_OK;
}

// ---------------------------------------------------------------------------------------------------
// logits, pair screen, partition screen
// ---------------------------------------------------------------------------------------------------
extern "C" int fb_forward_f64(const fb_model *m, int64_t rows, const double *X, const uint32_t *dead_bits, double *logits,
                              void *stream)
{
    if (!m || rows < 0 || (rows > 0 && (!X || !logits))) return fb_fail(FB_ERR_INVALID, "fb_forward_f64: bad argument");
    if (int rc = fb_check_device("fb_forward_f64", m)) return rc;
    if (rows == 0) return FB_OK;
    cudaStream_t st = (cudaStream_t)stream;
    if (fb_use_spec(m)) { FB_CUDA(m->spec->forward64(m->p64, rows, X, dead_bits, logits, m->num_sms, st)); return FB_OK; }
    auto k = fb_forward_gen_kernel<double>;
    const int64_t ctas = (rows + FB_FWD_THREADS - 1) / FB_FWD_THREADS;
    k<<<fb_grid_for(k, FB_FWD_THREADS, 0, ctas, m->num_sms), FB_FWD_THREADS, 0, st>>>(m->gen, rows, X, dead_bits, logits);
    FB_CUDA(cudaGetLastError());
    return FB_OK;
}

extern "C" int fb_forward_f32(const fb_model *m, int64_t rows, const float *X, const uint32_t *dead_bits, float *logits,
                              void *stream)
{
    if (!m || rows < 0 || (rows > 0 && (!X || !logits))) return fb_fail(FB_ERR_INVALID, "fb_forward_f32: bad argument");
    if (int rc = fb_check_device("fb_forward_f32", m)) return rc;
    if (rows == 0) return FB_OK;
    cudaStream_t st = (cudaStream_t)stream;
    if (fb_use_spec(m)) { FB_CUDA(m->spec->forward32(m->p32, rows, X, dead_bits, logits, m->num_sms, st)); return FB_OK; }
    auto k = fb_forward_gen_kernel<float>;
    const int64_t ctas = (rows + FB_FWD_THREADS - 1) / FB_FWD_THREADS;
    k<<<fb_grid_for(k, FB_FWD_THREADS, 0, ctas, m->num_sms), FB_FWD_THREADS, 0, st>>>(m->gen, rows, X, dead_bits, logits);
    FB_CUDA(cudaGetLastError());
    return FB_OK;
}

extern "C" int fb_pair_screen(const fb_model *m, int64_t rows, const float *X, int n_pa, const int32_t *pa_cols_host,
                              int k_alts, const float *alts, uint8_t *orig_pred, int32_t *first_flip, void *stream)
{
    if (!m || rows < 0 || !pa_cols_host || (rows > 0 && (!X || !orig_pred || !first_flip)))
        return fb_fail(FB_ERR_INVALID, "fb_pair_screen: bad argument");
    if (int rc = fb_check_device("fb_pair_screen", m)) return rc;
    if (n_pa < 1 || n_pa > FB_MAX_PA) return fb_fail(FB_ERR_UNSUPPORTED, "fb_pair_screen: %d protected columns (supported: 1..%d)", n_pa, FB_MAX_PA);
    if (k_alts < 0 || (k_alts > 0 && !alts)) return fb_fail(FB_ERR_INVALID, "fb_pair_screen: bad alternatives");
    FbPaCols pa{};
    pa.n = n_pa;
    for (int q = 0; q < n_pa; ++q) {
        if (pa_cols_host[q] < 0 || pa_cols_host[q] >= m->gen.dims[0]) return fb_fail(FB_ERR_INVALID, "fb_pair_screen: protected column %d out of range", pa_cols_host[q]);
        pa.col[q] = pa_cols_host[q];
    }
    if (rows == 0) return FB_OK;
    cudaStream_t st = (cudaStream_t)stream;
    if (fb_use_spec(m)) {
        FB_CUDA(m->spec->pair_screen(m->p32, rows, X, &pa, k_alts, alts, orig_pred, first_flip, m->num_sms, st));
        return FB_OK;
    }
    auto k = fb_pair_screen_gen_kernel;
    const int64_t ctas = (rows + FB_FWD_THREADS - 1) / FB_FWD_THREADS;
    k<<<fb_grid_for(k, FB_FWD_THREADS, 0, ctas, m->num_sms), FB_FWD_THREADS, 0, st>>>(m->gen, rows, X, pa, k_alts, alts,
                                                                                 orig_pred, first_flip);
    FB_CUDA(cudaGetLastError());
    return FB_OK;
}

__global__ void __launch_bounds__(256)
fb_ce_unpack_kernel(int64_t n, int d, int pa_col, bool relaxed, const int32_t *__restrict__ boxes,
                    const unsigned long long *__restrict__ key, int32_t *__restrict__ ce_sample,
                    int32_t *__restrict__ pa_pos, int32_t *__restrict__ pa_neg, int32_t *__restrict__ ra_dpos,
                    int32_t *__restrict__ ra_dneg, uint8_t *__restrict__ verdict, uint32_t *__restrict__ record, int rec_words)
{
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; p < n; p += stride) {
        const unsigned long long k = key[p];
        int32_t cs = -1, vp = -1, vn = -1, dp = 0, dn = 0;
        uint8_t flag = 0;
        if (k == FB_SCREEN_SKIPPED) flag = FB_VERDICT_SKIPPED;
        else if (k != ~0ull && k != FB_SCREEN_NOSEARCH) {
            const int lo = boxes[p * (2 * d) + 2 * pa_col];
            cs = (int32_t)(k >> 32);
            if (relaxed) {
                vp = lo + (int32_t)((k >> 24) & 0xffu); vn = lo + (int32_t)((k >> 16) & 0xffu);
                dp = (int32_t)((k >> 8) & 0xffu) - 128; dn = (int32_t)(k & 0xffu) - 128;
            } else {
                vp = lo + (int32_t)((k >> 16) & 0xffffu); vn = lo + (int32_t)(k & 0xffffu);
            }
            flag = FB_VERDICT_CE;
        }
        if (ce_sample) ce_sample[p] = cs;
        if (pa_pos) pa_pos[p] = vp;
        if (pa_neg) pa_neg[p] = vn;
        if (ra_dpos) ra_dpos[p] = dp;
        if (ra_dneg) ra_dneg[p] = dn;
        if (verdict && flag) verdict[p] |= flag;
        if (record) {
            uint32_t *r = record + p * rec_words;
            r[rec_words - 1] = (uint32_t)cs;
            if (flag) r[0] |= flag;
        }
    }
}

// key[p] = "leave this box alone" where the caller's mask says so
__global__ void __launch_bounds__(256)
fb_screen_mark_kernel(int64_t n, const uint8_t *__restrict__ skip, unsigned long long *__restrict__ key)
{
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; p < n; p += stride)
        key[p] = skip[p] ? FB_SCREEN_NOSEARCH : ~0ull;
}

// The float32 enclosure of the logit over a whole box, for the screen kernels that do not compute it themselves (the
// tiled and sample-per-lane ones, the specialised ones of the bigger architectures): one warp per box, lanes = neurons,
// the screens' own arithmetic -- multiply-adds in ascending input order from 0, round to nearest, bias last -- on both
// ends of the interval, which encloses every logit the screen would COMPUTE for a point of the box (rounding is
// monotone).  One-signed: no sample of the box can show two classes, key[p] = "leave alone".  For a relaxed query the
// relaxed column is widened by the threshold on both sides.  Generated samples only.
#define FB_BOX_WARPS 8
__global__ void __launch_bounds__(FB_BOX_WARPS * 32)
fb_screen_box_kernel(const GenNet net, int64_t n, const int32_t *__restrict__ boxes, int pa_col, int ra_col, int ra_thr,
                     unsigned long long *key)
{
    extern __shared__ __align__(16) unsigned char fb_smem[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int D0 = net.dims[0], maxw = net.maxw;
    float *s_lo = reinterpret_cast<float *>(fb_smem) + (size_t)warp * 4 * maxw;     // [2][maxw]
    float *s_hi = s_lo + 2 * maxw;
    const bool relaxed = ra_col >= 0;
    for (int64_t p = (int64_t)blockIdx.x * FB_BOX_WARPS + warp; p < n; p += (int64_t)gridDim.x * FB_BOX_WARPS) {
        if (key[p] == FB_SCREEN_NOSEARCH) continue;
        const int32_t *box = boxes + p * (2 * D0);
        const uint32_t W = (uint32_t)(__ldg(box + 2 * pa_col + 1) - __ldg(box + 2 * pa_col) + 1);
        if (W == 0 || W > (relaxed ? FB_SCREEN_MAX_W_RELAXED : FB_SCREEN_MAX_W_STRICT)) continue;   // the screen kernel flags it
        bool ok = true;
        __syncwarp();
        for (int i = lane; i < D0; i += 32) {
            int lo = __ldg(box + 2 * i), hi = __ldg(box + 2 * i + 1);
            if (lo < -(1 << 30) || hi > (1 << 30) || lo > hi) ok = false;
            else if (relaxed && i == ra_col) { lo -= ra_thr; hi += ra_thr; }
            s_lo[i] = (float)lo; s_hi[i] = (float)hi;
        }
        if (!__all_sync(0xffffffffu, ok)) continue;
        __syncwarp();
        for (int l = 0; l < net.L; ++l) {
            const int IN = net.dims[l], OUT = net.dims[l + 1];
            const bool last = (l == net.L - 1);
            const float *ilo = s_lo + (l & 1) * maxw, *ihi = s_hi + (l & 1) * maxw;
            float *olo = s_lo + ((l + 1) & 1) * maxw, *ohi = s_hi + ((l + 1) & 1) * maxw;
            const float *Wl = net.W32 + net.woff[l];
            for (int j = lane; j < OUT; j += 32) {
                float lo = 0.0f, hi = 0.0f;
                for (int k = 0; k < IN; ++k) {
                    const float w = __ldg(Wl + (size_t)k * OUT + j);
                    const bool neg = w < 0.0f;
                    const float xl = ilo[k], xh = ihi[k];
                    lo = __fmaf_rn(w, neg ? xh : xl, lo);
                    hi = __fmaf_rn(w, neg ? xl : xh, hi);
                }
                const float bj = __ldg(net.B32 + net.boff[l] + j);
                lo = __fadd_rn(lo, bj); hi = __fadd_rn(hi, bj);
                olo[j] = (last || fb_pos(lo)) ? lo : 0.0f;
                ohi[j] = (last || fb_pos(hi)) ? hi : 0.0f;
            }
            __syncwarp();
        }
        if (lane == 0) {
            const float ylo = s_lo[(net.L & 1) * maxw], yhi = s_hi[(net.L & 1) * maxw];
            if (ylo > 0.0f || yhi < 0.0f) key[p] = FB_SCREEN_NOSEARCH;
        }
    }
}

extern "C" int fb_partition_screen_ex(const fb_model *m, int64_t n, const int32_t *boxes, const int32_t *samples, int S,
                                      uint64_t seed, int64_t first_index, int pa_col, int ra_col, int ra_threshold,
                                      int32_t *ce_sample, int32_t *pa_pos, int32_t *pa_neg, int32_t *ra_dpos, int32_t *ra_dneg,
                                      uint8_t *verdict, uint32_t *record, const uint8_t *skip, void *stream)
{
    if (int rc = fb_check_batch("fb_partition_screen", m, n, boxes, S, true)) return rc;
    const int d = m->gen.dims[0];
    if (pa_col < 0 || pa_col >= d) return fb_fail(FB_ERR_INVALID, "fb_partition_screen: protected column %d out of range", pa_col);
    if (ra_col >= d || ra_col == pa_col) return fb_fail(FB_ERR_INVALID, "fb_partition_screen: relaxed column %d out of range (or equal to the protected column)", ra_col);
    if (ra_col < 0) { ra_col = -1; ra_threshold = 0; }
    if (ra_threshold < 0 || ra_threshold > FB_SCREEN_MAX_THR)
        return fb_fail(FB_ERR_UNSUPPORTED, "fb_partition_screen: relaxed threshold %d (supported: 0..%d)", ra_threshold, FB_SCREEN_MAX_THR);
    if (n == 0) return FB_OK;
    if (!ce_sample && !record && !verdict) return fb_fail(FB_ERR_INVALID, "fb_partition_screen: no output requested");
    cudaStream_t st = (cudaStream_t)stream;
    int chunks, chunk_len;
    fb_pick_chunks(m, n, S, &chunks, &chunk_len);
    DevBuf key(m->device, st);
    FB_CUDA(key.alloc(sizeof(unsigned long long) * n));
    if (skip) {
        const int mgrid = (int)std::min<int64_t>((n + 255) / 256, (int64_t)m->num_sms * 8);
        fb_screen_mark_kernel<<<mgrid, 256, 0, st>>>(n, skip, (unsigned long long *)key.p);
        FB_CUDA(cudaGetLastError());
    } else {
        FB_CUDA(cudaMemsetAsync(key.p, 0xff, sizeof(unsigned long long) * n, st));
    }
    const bool spec_screen = fb_use_spec(m) && (ra_col < 0 || m->gen.P <= FB_SCREEN_SPEC_RELAXED_MAXP);
    if (samples == nullptr && !(spec_screen && m->gen.P <= FB_SCREEN_SPEC_RELAXED_MAXP) && m->gen.maxw <= FB_MAXW) {
        // (the small specialised kernels run this pass themselves, per warp)
        const size_t bsmem = sizeof(float) * FB_BOX_WARPS * 4 * m->gen.maxw;
        const int bgrid = (int)std::min<int64_t>((n + FB_BOX_WARPS - 1) / FB_BOX_WARPS, (int64_t)m->num_sms * 8);
        fb_screen_box_kernel<<<bgrid, FB_BOX_WARPS * 32, bsmem, st>>>(m->gen, n, boxes, pa_col, ra_col, ra_threshold,
                                                                      (unsigned long long *)key.p);
        FB_CUDA(cudaGetLastError());
    }
    if (spec_screen) {
        FB_CUDA(m->spec->partition_screen(m->p32, samples == nullptr, n, boxes, samples, seed, first_index, S, chunks,
                                          chunk_len, pa_col, ra_col, ra_threshold, (unsigned long long *)key.p, m->num_sms, st));
    } else if (fb_use_tiled_profile(m) || fb_use_spec(m) || ra_col >= 0) {
        // (every relaxed query without a specialised kernel: the tiled screen has the interval pass, the sample-per-lane one not)
        if (int rc = fb_screen_tiled(m, n, boxes, samples, S, seed, first_index, pa_col, ra_col, ra_threshold,
                                     (unsigned long long *)key.p, st)) return rc;
    } else {
        const int64_t ctas = (n * chunks + (FB_PROFILE_THREADS / 32) - 1) / (FB_PROFILE_THREADS / 32);
        if (samples == nullptr) {
            auto k = fb_partition_screen_gen_kernel<true>;
            k<<<fb_grid_for(k, FB_PROFILE_THREADS, 0, ctas, m->num_sms), FB_PROFILE_THREADS, 0, st>>>(
                m->gen, n, boxes, samples, seed, first_index, S, chunks, chunk_len, pa_col, ra_col, ra_threshold,
                (unsigned long long *)key.p);
        } else {
            auto k = fb_partition_screen_gen_kernel<false>;
            k<<<fb_grid_for(k, FB_PROFILE_THREADS, 0, ctas, m->num_sms), FB_PROFILE_THREADS, 0, st>>>(
                m->gen, n, boxes, samples, seed, first_index, S, chunks, chunk_len, pa_col, ra_col, ra_threshold,
                (unsigned long long *)key.p);
        }
        FB_CUDA(cudaGetLastError());
    }
    const int grid = (int)std::min<int64_t>((n + 255) / 256, (int64_t)m->num_sms * 8);
    fb_ce_unpack_kernel<<<grid, 256, 0, st>>>(n, d, pa_col, ra_col >= 0, boxes, (const unsigned long long *)key.p, ce_sample,
                                              pa_pos, pa_neg, ra_dpos, ra_dneg, verdict, record, 2 * ((m->gen.N + 31) / 32) + 2);
    FB_CUDA(cudaGetLastError());
    return FB_OK;
}

extern "C" int fb_partition_screen(const fb_model *m, int64_t n, const int32_t *boxes, const int32_t *samples, int S,
                                   uint64_t seed, int64_t first_index, int pa_col, int32_t *ce_sample, int32_t *pa_pos,
                                   int32_t *pa_neg, uint8_t *verdict, void *stream)
{
    if (n > 0 && (!ce_sample || !pa_pos || !pa_neg)) return fb_fail(FB_ERR_INVALID, "fb_partition_screen: null output");
    return fb_partition_screen_ex(m, n, boxes, samples, S, seed, first_index, pa_col, -1, 0, ce_sample, pa_pos, pa_neg, nullptr,
                                  nullptr, verdict, nullptr, nullptr, stream);
}

// ---------------------------------------------------------------------------------------------------
// hybrid prediction: find_partition_result_for_point + hybrid_predict, src/AC/Verify-AC-experiment-new2.py:587-638.
// Per row: the partition that contains it (index arithmetic), that partition's recorded verdict, and the logit of the
// model the verdict selects -- the "fairer" model where the partition was found unfair ('sat'), the original one
// everywhere else ('unsat', 'unknown', not evaluated, outside every partition).
// ---------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
fb_hybrid_select_kernel(int64_t rows, const int64_t *__restrict__ index, const uint8_t *__restrict__ status,
                        const float *__restrict__ y_orig, const float *__restrict__ y_fair, float *__restrict__ logit,
                        uint8_t *__restrict__ route, unsigned int *__restrict__ counters)
{
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; r < rows; r += stride) {
        const int64_t idx = index[r];
        const uint8_t st = idx < 0 ? (uint8_t)FB_STATUS_NONE : status[idx];
        const uint8_t rt = st <= FB_STATUS_UNKNOWN ? st : (uint8_t)FB_STATUS_NONE;
        logit[r] = (rt == FB_STATUS_SAT) ? y_fair[r] : y_orig[r];
        if (route) route[r] = rt;
        if (counters) atomicAdd(&counters[rt], 1u);
    }
}

extern "C" int fb_hybrid_predict(const fb_grid *g, const fb_model *original, const fb_model *fairer, int64_t rows,
                                 const double *X64, const float *X32, const uint8_t *status, float *logit, uint8_t *route,
                                 uint32_t *counters, void *stream)
{
    if (!g || !original || !fairer || rows < 0) return fb_fail(FB_ERR_INVALID, "fb_hybrid_predict: bad argument");
    if (original->gen.dims[0] != g->g.d || fairer->gen.dims[0] != g->g.d)
        return fb_fail(FB_ERR_INVALID, "fb_hybrid_predict: the models take %d / %d features, the grid has %d", original->gen.dims[0],
                       fairer->gen.dims[0], g->g.d);
    if (rows == 0) return FB_OK;
    if (!X64 || !X32 || !status || !logit) return fb_fail(FB_ERR_INVALID, "fb_hybrid_predict: null argument");
    if (int rc = fb_check_device("fb_hybrid_predict", original)) return rc;
    if (int rc = fb_check_device("fb_hybrid_predict", fairer)) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    DevBuf idx(original->device, st), yo(original->device, st), yf(original->device, st);
    FB_CUDA(idx.alloc(sizeof(int64_t) * rows));
    FB_CUDA(yo.alloc(sizeof(float) * rows));
    FB_CUDA(yf.alloc(sizeof(float) * rows));
    if (int rc = fb_grid_locate(g, rows, X64, (int64_t *)idx.p, nullptr, stream)) return rc;
    if (int rc = fb_forward_f32(original, rows, X32, nullptr, (float *)yo.p, stream)) return rc;
    if (int rc = fb_forward_f32(fairer, rows, X32, nullptr, (float *)yf.p, stream)) return rc;
    const int grid = (int)std::min<int64_t>((rows + 255) / 256, (int64_t)original->num_sms * 8);
    fb_hybrid_select_kernel<<<grid, 256, 0, st>>>(rows, (const int64_t *)idx.p, status, (const float *)yo.p, (const float *)yf.p,
                                                  logit, route, counters);
    FB_CUDA(cudaGetLastError());
    return FB_OK;
}

// ---------------------------------------------------------------------------------------------------
// How close do the sampled pre-activations come to zero, relative to the size of their terms?  The activation counts
// depend only on signs; the reference obtains its pre-activations from numpy's BLAS dgemv, whose summation order is not
// specified, while every kernel here sums in ascending input order.  The two can disagree on a sign only if
// |pre-activation| is within a few rounding errors of sum_i |w_i x_i|.  This diagnostic pass (one sample per thread,
// weights from global memory: not a fast path) counts the (sample, neuron) pairs with
//     |pre| <= 2^-40 * (sum_i |w_i x_i| + |b|)
// and reports the smallest ratio seen, so that "no sign depends on the summation order" is a measured statement.
// ---------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128)
fb_nearzero_kernel(const GenNet net, int64_t n, const int32_t *__restrict__ boxes, const int32_t *__restrict__ samples,
                   uint64_t seed, int64_t first_index, int nsamp, unsigned long long *__restrict__ count,
                   unsigned long long *__restrict__ min_ratio_bits)
{
    const int D0 = net.dims[0];
    double a[FB_MAXW], y[FB_MAXW];
    unsigned long long mine = 0;
    double best = 1.0;
    const int64_t total = n * (int64_t)nsamp;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t u = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; u < total; u += stride) {
        const int64_t p = u / nsamp;
        const int s = (int)(u - p * nsamp);
        const int32_t *box = boxes + p * (2 * D0);
        if (samples == nullptr) fb_draw_row<double>(box, D0, fb_draw_group(fb_box_widest(box, D0)), (uint32_t)s, first_index + p, seed, a);
        else for (int i = 0; i < D0; ++i) a[i] = (double)__ldg(samples + ((int64_t)p * nsamp + s) * D0 + i);
        double *in = a, *out = y;
        for (int l = 0; l < net.L; ++l) {
            const int IN = net.dims[l], OUT = net.dims[l + 1];
            const double *w = net.W64 + net.woff[l], *b = net.B64 + net.boff[l];
            for (int j = 0; j < OUT; ++j) {
                double acc = 0.0, mag = 0.0;
                for (int i = 0; i < IN; ++i) {
                    const double wt = __ldg(w + (int64_t)i * OUT + j);
                    acc = __fma_rn(wt, in[i], acc);
                    mag = __fma_rn(fabs(wt), fabs(in[i]), mag);
                }
                const double pre = __dadd_rn(acc, __ldg(b + j));
                mag = __dadd_rn(mag, fabs(__ldg(b + j)));
                if (mag > 0.0) {
                    const double ratio = fabs(pre) / mag;
                    if (ratio <= 0x1p-40) ++mine;
                    best = fmin(best, ratio);
                }
                out[j] = (l == net.L - 1) ? pre : (pre > 0.0 ? pre : 0.0);
            }
            double *t = in; in = out; out = t;
        }
    }
    if (mine) atomicAdd(count, mine);
    atomicMin(min_ratio_bits, (unsigned long long)__double_as_longlong(best));       // non-negative doubles order like integers
}

extern "C" int fb_profile_nearzero(const fb_model *m, int64_t n, const int32_t *boxes, const int32_t *samples, int S,
                                   uint64_t seed, int64_t first_index, uint64_t *count, double *min_ratio, void *stream)
{
    if (int rc = fb_check_batch("fb_profile_nearzero", m, n, boxes, S, true)) return rc;
    if (!count || !min_ratio) return fb_fail(FB_ERR_INVALID, "fb_profile_nearzero: null output");
    cudaStream_t st = (cudaStream_t)stream;
    const double one = 1.0;
    FB_CUDA(cudaMemsetAsync(count, 0, sizeof(uint64_t), st));
    FB_CUDA(cudaMemcpyAsync(min_ratio, &one, sizeof(double), cudaMemcpyHostToDevice, st));
    if (n == 0) return FB_OK;
    const int grid = (int)std::min<int64_t>((n * S + 127) / 128, (int64_t)m->num_sms * 16);
    fb_nearzero_kernel<<<grid, 128, 0, st>>>(m->gen, n, boxes, samples, seed, first_index, S, (unsigned long long *)count,
                                             (unsigned long long *)min_ratio);
    FB_CUDA(cudaGetLastError());
    return FB_OK;
}

// ---------------------------------------------------------------------------------------------------
// heuristic_prune (reference utils/prune.py:862-939), one partition per thread.  Runs only for partitions on
// which Z3 answered `unknown`, so it is written for exactness, not speed: numpy's mean (pairwise sum with 8
// accumulators per block of <= 128), median and linear-interpolated percentile are restated operation by
// operation (same as oracle/fairify_oracle.c, which is checked bit for bit against numpy).
// ---------------------------------------------------------------------------------------------------
__device__ static double fb_np_pairwise_block(const double *a, int n)      // n <= 128
{
    if (n < 8) {
        double r = 0.0;
        for (int i = 0; i < n; ++i) r = __dadd_rn(r, a[i]);
        return r;
    }
    double r[8];
    int i;
    for (i = 0; i < 8; ++i) r[i] = a[i];
    for (i = 8; i < n - (n % 8); i += 8)
        for (int k = 0; k < 8; ++k) r[k] = __dadd_rn(r[k], a[i + k]);
    double res = __dadd_rn(__dadd_rn(__dadd_rn(r[0], r[1]), __dadd_rn(r[2], r[3])),
                           __dadd_rn(__dadd_rn(r[4], r[5]), __dadd_rn(r[6], r[7])));
    for (; i < n; ++i) res = __dadd_rn(res, a[i]);
    return res;
}
// numpy splits blocks larger than 128 in two (recursively); layers are at most FB_MAXW <= 256 wide, so one
// split suffices and no device recursion (and no dynamic stack) is needed
__device__ static double fb_np_pairwise(const double *a, int n)
{
    static_assert(FB_MAXW <= 256, "fb_np_pairwise handles one level of numpy's pairwise split");
    if (n <= 128) return fb_np_pairwise_block(a, n);
    int n2 = n / 2;
    n2 -= n2 % 8;
    return __dadd_rn(fb_np_pairwise_block(a, n2), fb_np_pairwise_block(a + n2, n - n2));
}
__device__ static double fb_np_mean(const double *a, int n) { return __ddiv_rn(fb_np_pairwise(a, n), (double)n); }
__device__ static void fb_insertion_sort(double *a, int n)
{
    for (int i = 1; i < n; ++i) {
        const double v = a[i];
        int k = i - 1;
        while (k >= 0 && a[k] > v) { a[k + 1] = a[k]; --k; }
        a[k + 1] = v;
    }
}
__device__ static double fb_np_median_sorted(const double *s, int n)
{
    if (n & 1) return s[n / 2];
    return __ddiv_rn(__dadd_rn(__dadd_rn(0.0, s[n / 2 - 1]), s[n / 2]), 2.0);
}
__device__ static double fb_np_percentile_sorted(const double *s, int n, double pct)
{
    const double q = __ddiv_rn(pct, 100.0);
    const double v = __dmul_rn((double)(n - 1), q);
    int prev, next;
    double gamma;
    if (v >= (double)(n - 1)) { prev = next = n - 1; gamma = __dadd_rn(v, -(double)(n - 1)); }
    else if (v < 0) { prev = next = 0; gamma = v; }
    else { prev = (int)floor(v); next = prev + 1; gamma = __dadd_rn(v, -(double)prev); }
    const double a = s[prev], b = s[next], diff = __dadd_rn(b, -a);
    double r = __dadd_rn(a, __dmul_rn(diff, gamma));
    if (gamma >= 0.5) r = __dadd_rn(b, -__dmul_rn(diff, __dadd_rn(1.0, -gamma)));
    return r;
}

__device__ static void fb_keep_one_bits(uint32_t *w, int bo, int out)
{
    for (int j = 0; j < out; ++j)
        if (!((w[(bo + j) >> 5] >> ((bo + j) & 31)) & 1u)) return;
    w[bo >> 5] &= ~(1u << (bo & 31));
}

__global__ void __launch_bounds__(64)
fb_heuristic_kernel(const GenNet net, int64_t n, const double *__restrict__ ws_lb, const double *__restrict__ ws_ub,
                    const uint32_t *__restrict__ cand, const uint32_t *__restrict__ s_cand,
                    const uint32_t *__restrict__ deads, double pct, uint32_t *__restrict__ new_deads,
                    uint32_t *__restrict__ merged)
{
    const int N = net.N, NW = (N + 31) / 32;
    double cv[FB_MAXW], nv[FB_MAXW];
    uint32_t nw[FB_MAX_NW], mw[FB_MAX_NW];
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; p < n; p += stride) {
        const double *lb = ws_lb + p * N, *ub = ws_ub + p * N;
        const uint32_t *cb = cand + p * NW, *sb = s_cand + p * NW, *db = deads + p * NW;
        for (int k = 0; k < NW; ++k) nw[k] = 0;
        for (int l = 0; l + 1 < net.L; ++l) {
            const int out = net.dims[l + 1], bo = net.boff[l];
            int nc = 0, nn = 0;
            for (int j = 0; j < out; ++j) {
                const int idx = bo + j;
                if ((cb[idx >> 5] >> (idx & 31)) & 1u) cv[nc++] = ub[idx];
                else nv[nn++] = ub[idx];
            }
            if (nn == 0) {
                for (int j = 0; j < out; ++j) nw[(bo + j) >> 5] |= 1u << ((bo + j) & 31);
                continue;
            }
            if (nc == 0) continue;
            const double cmean = fb_np_mean(cv, nc), nmean = fb_np_mean(nv, nn);
            fb_insertion_sort(cv, nc); fb_insertion_sort(nv, nn);
            const double cmed = fb_np_median_sorted(cv, nc), nmed = fb_np_median_sorted(nv, nn);
            const double n_lo = fb_np_percentile_sorted(nv, nn, pct);
            const double n_hi = fb_np_percentile_sorted(nv, nn, __dadd_rn(100.0, -pct));
            if (nmean > __dmul_rn(2.0, cmean) && nmed > __dmul_rn(2.0, cmed)) {
                const double hi10 = __dmul_rn(0.1, n_hi);
                for (int j = 0; j < out; ++j) {
                    const int idx = bo + j;
                    if (((sb[idx >> 5] >> (idx & 31)) & 1u) && ub[idx] < n_lo && ub[idx] < hi10 && ub[idx] < fabs(lb[idx]))
                        nw[idx >> 5] |= 1u << (idx & 31);
                }
            }
        }
        for (int l = 0; l < net.L; ++l) fb_keep_one_bits(nw, net.boff[l], net.dims[l + 1]);
        for (int k = 0; k < NW; ++k) mw[k] = db[k] | nw[k];
        for (int l = 0; l < net.L; ++l) fb_keep_one_bits(mw, net.boff[l], net.dims[l + 1]);
        for (int k = 0; k < NW; ++k) { new_deads[p * NW + k] = nw[k]; merged[p * NW + k] = mw[k]; }
    }
}

extern "C" int fb_heuristic(const fb_model *m, int64_t n, const double *ws_lb, const double *ws_ub, const uint32_t *cand,
                            const uint32_t *s_cand, const uint32_t *deads, double pct, uint32_t *new_deads,
                            uint32_t *merged, void *stream)
{
    if (!m || n < 0) return fb_fail(FB_ERR_INVALID, "fb_heuristic: bad argument");
    if (int rc = fb_check_device("fb_heuristic", m)) return rc;
    if (n == 0) return FB_OK;
    if (!ws_lb || !ws_ub || !cand || !s_cand || !deads || !new_deads || !merged)
        return fb_fail(FB_ERR_INVALID, "fb_heuristic: null argument");
    if (!(pct >= 0.0 && pct <= 100.0)) return fb_fail(FB_ERR_INVALID, "fb_heuristic: percentile %g outside [0, 100]", pct);
    const int grid = (int)std::min<int64_t>((n + 63) / 64, (int64_t)m->num_sms * 8);
    fb_heuristic_kernel<<<grid, 64, 0, (cudaStream_t)stream>>>(m->gen, n, ws_lb, ws_ub, cand, s_cand, deads, pct, new_deads,
                                                               merged);
    FB_CUDA(cudaGetLastError());
    return FB_OK;
}

// ---------------------------------------------------------------------------------------------------
// FP64 issue-rate microbenchmark: 8 independent DFMA chains per thread, enough warps to saturate the
// pipe.  Gives the measured denominator for "% of FP64 peak".
// ---------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) fb_dfma_kernel(int iters, double seed, double *sink)
{
    double a0 = seed + threadIdx.x, a1 = a0 + 1, a2 = a0 + 2, a3 = a0 + 3, a4 = a0 + 4, a5 = a0 + 5, a6 = a0 + 6, a7 = a0 + 7;
    const double m = 1.0000001, c = 1e-9;
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int u = 0; u < 16; ++u) {
            a0 = __fma_rn(a0, m, c); a1 = __fma_rn(a1, m, c); a2 = __fma_rn(a2, m, c); a3 = __fma_rn(a3, m, c);
            a4 = __fma_rn(a4, m, c); a5 = __fma_rn(a5, m, c); a6 = __fma_rn(a6, m, c); a7 = __fma_rn(a7, m, c);
        }
    }
    const double s = ((a0 + a1) + (a2 + a3)) + ((a4 + a5) + (a6 + a7));
    if (s == 12345.678) sink[0] = s;       // never true; keeps the chains alive
}

extern "C" int fb_fp64_peak(int iters, double *tflops, double *ms_out)
{
    if (iters < 1 || !tflops) return fb_fail(FB_ERR_INVALID, "fb_fp64_peak: bad argument");
    int sms = 0;
    if (fb_num_sms(&sms)) return FB_ERR_CUDA;
    double *sink = nullptr;
    FB_CUDA(cudaMalloc(&sink, sizeof(double)));
    cudaEvent_t e0, e1;
    FB_CUDA(cudaEventCreate(&e0)); FB_CUDA(cudaEventCreate(&e1));
    const int grid = sms * 8;
    fb_dfma_kernel<<<grid, 256>>>(iters / 8 + 1, 1.0, sink);           // warm-up
    FB_CUDA(cudaEventRecord(e0));
    fb_dfma_kernel<<<grid, 256>>>(iters, 1.0, sink);
    FB_CUDA(cudaEventRecord(e1));
    FB_CUDA(cudaEventSynchronize(e1));
    float ms = 0;
    FB_CUDA(cudaEventElapsedTime(&ms, e0, e1));
    const double flops = (double)grid * 256.0 * (double)iters * 16.0 * 8.0 * 2.0;
    *tflops = flops / (ms * 1e-3) / 1e12;
    if (ms_out) *ms_out = ms;
    cudaEventDestroy(e0); cudaEventDestroy(e1); cudaFree(sink);
    return FB_OK;
}
