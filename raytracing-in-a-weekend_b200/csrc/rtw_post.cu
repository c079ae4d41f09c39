// rtw_post.cu — the reference's bilateral filter (Rust2/src/postprocessing.rs:70-135) as four kernels.
//
//   1. gradient      one thread per interior pixel: |grad I| of the luminance (:75-85), to a scratch row-major array
//   2. params        ONE warp: the reference accumulates that array in a single f32, in row-major order (:75-87);
//                    f32 addition is not associative, so the sum has to be THAT sum: taken 1024 values at a time as
//                    integers of the running sum's binade, exact (see post_params_kernel). Then sigma_range/sigma_spatial.
//   3. weight table  exp(-inv_spatial * d2 - inv_range * (diff/255)^2) (:108-114) only depends on (|dx|, |dy|, |diff|):
//                    (size+1)^2 * 256 entries, computed once in double and rounded to f32
//   4. filter        one thread per (pixel, channel); walks the reference's window in the reference's order
//                    (x outer, y inner, both half-open: postprocessing.rs:30-61) with sequential f32 sums
//
// All f32 arithmetic that reaches the output is written with the round-to-nearest intrinsics so that no
// contraction to FMA can happen (rustc never contracts). Bytes in, bytes out: the result is compared byte for
// byte with oracle/postprocessing.hpp (exp_mode 1).
#include <mutex>
#include <string>
#include "rtw_post.h"
#include "../../include/rtw.h"

namespace rtw { void set_last_error(const std::string& msg); }

namespace rtw {

namespace {

constexpr uint32_t kTableMaxSize = 64;    // (64+1)^2 KiB = 4.1 MiB of weights; larger windows evaluate exp in place
constexpr uint32_t kChunk = 1024;         // gradient values per group of the ordered sum

__device__ __forceinline__ float post_intensity(const uint8_t* __restrict__ p) {       // postprocessing.rs:64-68
    const float r = __fdiv_rn(__fmul_rn(0.2989f, (float)p[0]), 255.0f);
    const float g = __fdiv_rn(__fmul_rn(0.5870f, (float)p[1]), 255.0f);
    const float b = __fdiv_rn(__fmul_rn(0.1140f, (float)p[2]), 255.0f);
    return __fadd_rn(__fadd_rn(r, g), b);
}

__global__ void post_gradient_kernel(const uint8_t* __restrict__ img, uint32_t w, uint32_t h, float* __restrict__ grad) {
    const uint32_t iw = w - 2u, n = iw * (h - 2u);
    for (uint32_t k = blockIdx.x * blockDim.x + threadIdx.x; k < n; k += gridDim.x * blockDim.x) {
        const uint32_t y = 1u + k / iw, x = 1u + k % iw;
        const float Iu = post_intensity(img + 3u * ((size_t)(y - 1u) * w + x));
        const float Id = post_intensity(img + 3u * ((size_t)(y + 1u) * w + x));
        const float Ir = post_intensity(img + 3u * ((size_t)y * w + x + 1u));
        const float Il = post_intensity(img + 3u * ((size_t)y * w + x - 1u));
        const float a = __fsub_rn(Iu, Id), b = __fsub_rn(Il, Ir);
        grad[k] = __fsqrt_rn(__fadd_rn(__fmul_rn(a, a), __fmul_rn(b, b)));
    }
}

// One warp takes the reference's f32 sum `for y { for x { gradient_sum += g } }` (:75-87) in the reference's order
// without adding one value at a time. While the running sum S stays inside one binade [2^e, 2^(e+1)] it is an integer
// I = S / u (u = 2^(e-23), 2^23 <= I <= 2^24) and fl(S + x) for x >= 0 is I + rint(x / u): unless x / u lies
// exactly half-way between two integers (a tie, rounded to even: depends on I) no addend's rounding depends on
// the sum so far, so a group of 1024 values adds as 1024 independent integers (one warp reduction), in any order. A group with a tie, with a value
// that is not a finite non-negative number below 2^21 u, or whose total would leave the binade is added the plain
// way, one __fadd_rn after the other in the reference's order (every lane computes the same sum): about one group per
// binade. Checked against the sequential sum bit for bit (tests/test_postprocessing.py through avg_gradient).
// `grad` is padded to whole groups; values past n count as +0, which changes neither form of the sum.
__device__ __forceinline__ float4 post_load4(const float* __restrict__ grad, uint32_t e, uint32_t n) {
    if (e + 3u < n) return *reinterpret_cast<const float4*>(grad + e);
    float4 v;
    v.x = e < n ? grad[e] : 0.0f; v.y = e + 1u < n ? grad[e + 1u] : 0.0f;
    v.z = e + 2u < n ? grad[e + 2u] : 0.0f; v.w = 0.0f;
    return v;
}

constexpr uint32_t kSumK = kChunk / 128u;                         // float4 per lane and group

// lane l holds float4 j of a group = its values (j * 32 + l) * 4 .. + 3
__device__ __forceinline__ void post_load_group(float4 (&x)[kSumK], const float* __restrict__ grad, uint32_t g, uint32_t n,
                                                uint32_t lane) {
#pragma unroll
    for (uint32_t j = 0; j < kSumK; ++j) x[j] = post_load4(grad, g * kChunk + (j * 32u + lane) * 4u, n);
}

__device__ __forceinline__ float post_add_group(float sum, const float4 (&x)[kSumK]) {
    const uint32_t FULL = 0xFFFFFFFFu;
    const uint32_t sb = __float_as_uint(sum), ex = (sb >> 23) & 0xFFu;
    const bool binade_ok = ex > 23u && ex < 0xFEu;                 // normal, and u = 2^(e-23) normal too
    const float scale = __uint_as_float((277u - ex) << 23);        // 1 / u
    const float unit = __uint_as_float((ex - 23u) << 23);          // u
    const uint32_t I = (sb & 0x7FFFFFu) | 0x800000u;
    uint32_t tot = 0u;
    bool good = true;
#pragma unroll
    for (uint32_t j = 0; j < kSumK; ++j) {
        const float v[4] = {x[j].x, x[j].y, x[j].z, x[j].w};
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            const float y = __fmul_rn(v[c], scale);                // exact: a power of two
            const float t = __fadd_rn(y, 8388608.0f);              // y < 2^21: the mantissa of t is rint(y)
            const float d = __fsub_rn(y, __fsub_rn(t, 8388608.0f));          // exact: y - rint(y)
            good = good && (y >= 0.0f) && (y < 2097152.0f) && (fabsf(d) != 0.5f);       // 2^21: a group's total < 2^31
            tot += __float_as_uint(t) & 0x7FFFFFu;
        }
    }
    const uint32_t total = __reduce_add_sync(FULL, tot);
    const bool all_good = __all_sync(FULL, good) != 0;
    if (binade_ok && all_good && I + total <= 0x1000000u) return __fmul_rn((float)(I + total), unit);   // exact
#pragma unroll 1
    for (uint32_t j = 0; j < kSumK; ++j) {
        float4 xj = x[0];
#pragma unroll
        for (uint32_t k = 1; k < kSumK; ++k) if (k == j) xj = x[k];
#pragma unroll 2
        for (uint32_t l = 0; l < 32u; ++l) {
            sum = __fadd_rn(sum, __shfl_sync(FULL, xj.x, l)); sum = __fadd_rn(sum, __shfl_sync(FULL, xj.y, l));
            sum = __fadd_rn(sum, __shfl_sync(FULL, xj.z, l)); sum = __fadd_rn(sum, __shfl_sync(FULL, xj.w, l));
        }
    }
    return sum;
}

__global__ void post_params_kernel(const float* __restrict__ grad, uint32_t n, uint32_t w, uint32_t h, float* __restrict__ params) {
    const uint32_t lane = threadIdx.x;
    const uint32_t n_groups = (n + kChunk - 1u) / kChunk;
    // three groups in registers: one being added, two in flight behind it (a single warp has nothing else to hide
    // the load latency with); values past n load as +0
    float4 a[kSumK], b[kSumK], c[kSumK];
    post_load_group(a, grad, 0u, n, lane); post_load_group(b, grad, 1u, n, lane); post_load_group(c, grad, 2u, n, lane);
    float sum = 0.0f;
    for (uint32_t g = 0; g < n_groups; g += 3u) {
        sum = post_add_group(sum, a); post_load_group(a, grad, g + 3u, n, lane);
        if (g + 1u < n_groups) { sum = post_add_group(sum, b); post_load_group(b, grad, g + 4u, n, lane); }
        if (g + 2u < n_groups) { sum = post_add_group(sum, c); post_load_group(c, grad, g + 5u, n, lane); }
    }
    if (lane == 0) {                                              // postprocessing.rs:73,87-90
        const float spatial = ceilf(__fmul_rn(0.02f, __fsqrt_rn((float)(w * w + h * h))));
        const float avg = __fdiv_rn(sum, (float)((w - 2u) * (h - 2u)));
        params[0] = spatial;
        params[1] = avg;
        params[2] = __fdiv_rn(0.5f, __fmul_rn(avg, avg));
        params[3] = __fdiv_rn(0.5f, __fmul_rn(spatial, spatial));
    }
}

// postprocessing.rs:108-114; diff enters squared, so its sign is dropped. exp in double, rounded once to f32.
__device__ __forceinline__ float post_weight(float inv_spatial, float inv_range, uint32_t d2, uint32_t adiff) {
    const float t = __fdiv_rn((float)adiff, 255.0f);
    const float arg = __fsub_rn(__fmul_rn(-inv_spatial, (float)d2), __fmul_rn(inv_range, __fmul_rn(t, t)));
    return (float)exp((double)arg);
}

__global__ void post_table_kernel(const float* __restrict__ params, uint32_t size, float* __restrict__ table) {
    const uint32_t n = (size + 1u) * (size + 1u) * 256u;
    const float inv_range = params[2], inv_spatial = params[3];
    for (uint32_t e = blockIdx.x * blockDim.x + threadIdx.x; e < n; e += gridDim.x * blockDim.x) {
        const uint32_t adiff = e & 255u, cell = e >> 8, adx = cell % (size + 1u), ady = cell / (size + 1u);
        table[e] = post_weight(inv_spatial, inv_range, adx * adx + ady * ady, adiff);
    }
}

__device__ __forceinline__ uint8_t post_as_u8(float v) {          // Rust `f32 as u8`: saturating, NaN -> 0
    if (!(v == v) || v <= 0.0f) return 0;
    if (v >= 255.0f) return 255;
    return (uint8_t)v;
}

template <bool kTable>
__global__ void __launch_bounds__(256) post_bilateral_kernel(const uint8_t* __restrict__ img, uint32_t w, uint32_t h, uint32_t size,
                                                             int kind, const float* __restrict__ params,
                                                             const float* __restrict__ table, uint8_t* __restrict__ out) {
    const size_t n = (size_t)w * h * 3u;
    const float inv_range = params[2], inv_spatial = params[3];
    for (size_t tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x; tid < n; tid += (size_t)gridDim.x * blockDim.x) {
        const uint32_t c = (uint32_t)(tid % 3u), pix = (uint32_t)(tid / 3u), x = pix % w, y = pix / w;
        const uint32_t left = x - min(x, size), up = y - min(y, size);
        const uint32_t right = x + min(w - x - 1u, size), down = y + min(h - y - 1u, size);
        const int p = img[tid];
        float col = 0.0f, ws = 0.0f;
        for (uint32_t i = left; i < right; ++i) {
            const uint32_t dx = x < i ? i - x : x - i;
            const uint8_t* colp = img + (size_t)i * 3u + c;
            for (uint32_t j = up; j < down; ++j) {
                const uint32_t dy = y < j ? j - y : y - j;
                if (kind == POST_PROX_EDGES && !(dx + dy < size)) continue;
                const int pi = colp[(size_t)j * w * 3u];
                const uint32_t adiff = (uint32_t)abs(pi - p);
                const float wgt = kTable ? __ldg(table + (((dy * (size + 1u) + dx) << 8) | adiff))
                                         : post_weight(inv_spatial, inv_range, dx * dx + dy * dy, adiff);
                col = __fadd_rn(col, __fdiv_rn(__fmul_rn((float)pi, wgt), 255.0f));
                ws = __fadd_rn(ws, wgt);
            }
        }
        out[tid] = post_as_u8(__fdiv_rn(__fmul_rn(col, 255.0f), ws));
    }
}

int post_grid(size_t n, int block) {
    size_t g = (n + block - 1) / block;
    if (g < 1) g = 1;
    if (g > 148u * 32u) g = 148u * 32u;
    return (int)g;
}

}  // namespace

#define POST_RELEASE() do { if (d_grad) cudaFreeAsync(d_grad, st); if (d_table) cudaFreeAsync(d_table, st); if (d_own) cudaFreeAsync(d_own, st); } while (0)
#define PC(call) do { cudaError_t e__ = (call); if (e__ != cudaSuccess) { POST_RELEASE(); return e__; } } while (0)

cudaError_t bilateral_filter_enqueue(const uint8_t* d_in, uint32_t width, uint32_t height, uint32_t size, int kind,
                                     uint8_t* d_out, float* d_params, cudaStream_t st, int* n_launches) {
    float *d_grad = nullptr, *d_table = nullptr, *d_own = nullptr;
    const uint32_t n_grad = (width - 2u) * (height - 2u);
    const size_t grad_padded = ((size_t)n_grad + kChunk - 1u) / kChunk * kChunk + kChunk;
    const bool use_table = size <= kTableMaxSize;
    const size_t n_table = use_table ? (size_t)(size + 1u) * (size + 1u) * 256u : 0u;
    PC(cudaMallocAsync(reinterpret_cast<void**>(&d_grad), grad_padded * sizeof(float), st));
    if (n_table) PC(cudaMallocAsync(reinterpret_cast<void**>(&d_table), n_table * sizeof(float), st));
    if (!d_params) { PC(cudaMallocAsync(reinterpret_cast<void**>(&d_own), 4 * sizeof(float), st)); d_params = d_own; }
    int launches = 0;
    if (n_grad) { post_gradient_kernel<<<post_grid(n_grad, 256), 256, 0, st>>>(d_in, width, height, d_grad); ++launches; }
    post_params_kernel<<<1, 32, 0, st>>>(d_grad, n_grad, width, height, d_params); ++launches;
    const size_t n = (size_t)width * height * 3u;
    if (use_table) {
        post_table_kernel<<<post_grid(n_table, 256), 256, 0, st>>>(d_params, size, d_table); ++launches;
        post_bilateral_kernel<true><<<post_grid(n, 256), 256, 0, st>>>(d_in, width, height, size, kind, d_params, d_table, d_out);
    } else {
        post_bilateral_kernel<false><<<post_grid(n, 256), 256, 0, st>>>(d_in, width, height, size, kind, d_params, nullptr, d_out);
    }
    ++launches;
    PC(cudaGetLastError());
    if (n_launches) *n_launches = launches;
    POST_RELEASE();
    return cudaSuccess;
}


// ---------------------------------------------------------------------------------------------------------------
// The same filter for callers that may not allocate (rtw_bilateral_filter_device_noalloc, include/rtw.h). The three
// scratch arrays of bilateral_filter_enqueue (the gradient array, the weight table, the four parameters) live in ONE
// file-scope device array of fixed size instead of coming from cudaMallocAsync; every device that loads the library
// holds its own copy (4.3 MiB) until the process exits. The kernels are the ones above, handed addresses inside it.
//
//   floats [0, 64)                              params {spatial, avg_gradient, inv_range, inv_spatial} (when the caller passes none)
//   floats [64, 64 + kStaticTableFloats)        weight table, (size+1)^2 * 256 entries, size <= RTW_NOALLOC_MAX_WINDOW
//   floats [.., .. + kStaticGradFloats)         gradient array, (w-2)(h-2) values + the ordered sum's read-ahead padding
//
// Nothing in it has to start at zero: the gradient kernel writes every value the sum reads (values past n are never
// loaded, post_load4 substitutes +0), the table kernel writes every entry the filter looks up, and the params kernel
// writes all four parameters, so whatever the previous call left behind is overwritten before it is read. All calls
// on one device share the array: two calls whose kernels overlap on the GPU (different streams, no dependency
// between them) race on it.
constexpr uint32_t kStaticMaxWindow = RTW_NOALLOC_MAX_WINDOW;                             // 16
constexpr uint32_t kStaticMaxGrad = RTW_NOALLOC_MAX_INTERIOR_PIXELS;                      // 1 Mi - 2 Ki values
constexpr size_t kStaticParamFloats = 64;                                                 // 256 B: keeps the table aligned
constexpr size_t kStaticTableFloats = (size_t)(kStaticMaxWindow + 1u) * (kStaticMaxWindow + 1u) * 256u;
constexpr size_t kStaticGradFloats = ((size_t)kStaticMaxGrad + kChunk - 1u) / kChunk * kChunk + kChunk;
constexpr size_t kStaticFloats = kStaticParamFloats + kStaticTableFloats + kStaticGradFloats;
static_assert(kStaticTableFloats % 64u == 0 && kStaticGradFloats % 64u == 0, "sub-arrays stay 256-byte aligned");
static_assert(kStaticFloats * sizeof(float) <= (5u << 20), "the static scratch is meant to stay modest");

static __device__ __align__(256) float g_post_scratch[kStaticFloats];

namespace {

// In host code `g_post_scratch` names a host-side shadow of the symbol, not device memory: the device address comes
// from cudaGetSymbolAddress, asked once per device with that device current.
constexpr int kMaxDevices = 64;
std::mutex g_scratch_mutex;
float* g_scratch_addr[kMaxDevices] = {};

cudaError_t post_scratch_for_current_device(float** out) {
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (dev < 0 || dev >= kMaxDevices) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lk(g_scratch_mutex);
    if (!g_scratch_addr[dev]) {
        void* p = nullptr;
        e = cudaGetSymbolAddress(&p, g_post_scratch);
        if (e != cudaSuccess) return e;
        g_scratch_addr[dev] = static_cast<float*>(p);
    }
    *out = g_scratch_addr[dev];
    return cudaSuccess;
}

}  // namespace

// No cudaMalloc / cudaMallocAsync / cudaFree on this path: kernel launches only.
cudaError_t bilateral_filter_enqueue_static(const uint8_t* d_in, uint32_t width, uint32_t height, uint32_t size, int kind,
                                            uint8_t* d_out, float* d_params, cudaStream_t st, int* n_launches) {
    const uint32_t n_grad = (width - 2u) * (height - 2u);
    if (n_grad > kStaticMaxGrad || size > kStaticMaxWindow) return cudaErrorInvalidValue;
    float* base = nullptr;
    cudaError_t e = post_scratch_for_current_device(&base);
    if (e != cudaSuccess) return e;
    float* d_table = base + kStaticParamFloats;
    float* d_grad = d_table + kStaticTableFloats;
    if (!d_params) d_params = base;
    const size_t n_table = (size_t)(size + 1u) * (size + 1u) * 256u;
    int launches = 0;
    if (n_grad) { post_gradient_kernel<<<post_grid(n_grad, 256), 256, 0, st>>>(d_in, width, height, d_grad); ++launches; }
    post_params_kernel<<<1, 32, 0, st>>>(d_grad, n_grad, width, height, d_params); ++launches;
    post_table_kernel<<<post_grid(n_table, 256), 256, 0, st>>>(d_params, size, d_table); ++launches;
    const size_t n = (size_t)width * height * 3u;
    post_bilateral_kernel<true><<<post_grid(n, 256), 256, 0, st>>>(d_in, width, height, size, kind, d_params, d_table, d_out);
    ++launches;
    if (n_launches) *n_launches = launches;
    return cudaGetLastError();
}

}  // namespace rtw

extern "C" int rtw_bilateral_filter_device_noalloc(const uint8_t* d_rgb8, uint32_t width, uint32_t height, uint32_t size,
                                                   int proximity, uint8_t* d_out_rgb8, float* d_params, int device,
                                                   void* stream) {
    auto fail = [](int code, const std::string& msg) { rtw::set_last_error(msg); return code; };
    if (!d_rgb8 || !d_out_rgb8) return fail(RTW_E_INVALID, "null argument");
    if (d_rgb8 == d_out_rgb8) return fail(RTW_E_INVALID, "bilateral filter cannot run in place");
    if (width < 2u || height < 2u || width > 32768u || height > 32768u)
        return fail(RTW_E_INVALID, "bilateral filter: image must be 2..32768 pixels a side");
    if (proximity != RTW_PROX_SQUARE && proximity != RTW_PROX_EDGES) return fail(RTW_E_INVALID, "unknown proximity type");
    const uint64_t interior = (uint64_t)(width - 2u) * (height - 2u);
    if (interior > RTW_NOALLOC_MAX_INTERIOR_PIXELS)
        return fail(RTW_E_UNSUPPORTED, "rtw_bilateral_filter_device_noalloc: (width-2)*(height-2) = " + std::to_string(interior) +
                                           " exceeds the static scratch (RTW_NOALLOC_MAX_INTERIOR_PIXELS = " +
                                           std::to_string(RTW_NOALLOC_MAX_INTERIOR_PIXELS) + "); use rtw_bilateral_filter_device");
    if (size > RTW_NOALLOC_MAX_WINDOW)
        return fail(RTW_E_UNSUPPORTED, "rtw_bilateral_filter_device_noalloc: window size " + std::to_string(size) +
                                           " exceeds the static weight table (RTW_NOALLOC_MAX_WINDOW = " +
                                           std::to_string(RTW_NOALLOC_MAX_WINDOW) + "); use rtw_bilateral_filter_device");
    int prev = -1;
    cudaError_t e = cudaGetDevice(&prev);
    if (e == cudaSuccess && device >= 0 && device != prev) e = cudaSetDevice(device);
    if (e == cudaSuccess)
        e = rtw::bilateral_filter_enqueue_static(d_rgb8, width, height, size, proximity, d_out_rgb8, d_params,
                                                 static_cast<cudaStream_t>(stream), nullptr);
    if (prev >= 0 && device >= 0 && device != prev) cudaSetDevice(prev);
    if (e != cudaSuccess) {
        cudaGetLastError();
        return fail(RTW_E_CUDA, std::string("rtw_bilateral_filter_device_noalloc: ") + cudaGetErrorString(e));
    }
    return RTW_OK;
}
