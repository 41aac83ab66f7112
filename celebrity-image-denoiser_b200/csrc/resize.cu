// Decode-side data prep: Pillow-compatible bicubic resize (u8) and ToTensor+Normalize (u8 -> f32).
//
// Reference (file:line in /root/reference):
//   Image.resize(size, BICUBIC)      trainingcode/denoise_gan_code/noise_generation.py:61,
//                                    denoise_eavl_iter.py:89, denoisegan_eval.py:89
//   ToTensor + Normalize([.5],[.5])  denoisegan_eval.py:76-79, backend/app.py:401-406
//
// The arithmetic of Image.resize lives in Pillow (unpinned in backend/requirements.txt; 12.2.0
// in this image), src/libImaging/Resample.c: a horizontal then a vertical pass, each a
// normalised bicubic (a = -0.5) kernel whose support is stretched by the down-scaling factor,
// evaluated in 22-bit fixed point on u8 with round-half-up and clipping. The coefficient tables
// are computed on the host in double exactly as Pillow does; the device does the integer MACs,
// so the result is bit-identical to Pillow.
#include <math.h>

#include "cid_internal.h"
#include "ptx.cuh"

namespace cid {

constexpr int kPrecisionBits = 32 - 8 - 2;

struct ResizeCoeffs {
    int ksize = 0;
    int* bounds = nullptr;  // device [out][2] = (xmin, count)
    int* kk = nullptr;      // device [out][ksize] fixed-point weights
};

static double bicubic_filter(double x) {
    const double a = -0.5;
    if (x < 0.0) x = -x;
    if (x < 1.0) return ((a + 2.0) * x - (a + 3.0)) * x * x + 1;
    if (x < 2.0) return (((x - 5) * x + 8) * x - 4) * a;
    return 0.0;
}

// host: Pillow's precompute_coeffs + normalize_coeffs_8bpc for the full-image box (0, in_size)
static void host_coeffs(int in_size, int out_size, int* ksize_out, std::vector<int>& bounds, std::vector<int>& kk) {
    const float in0 = 0.f, in1 = static_cast<float>(in_size);
    double scale = static_cast<double>(in1 - in0) / out_size;
    double filterscale = scale;
    if (filterscale < 1.0) filterscale = 1.0;
    const double support = 2.0 * filterscale;
    const int ksize = static_cast<int>(ceil(support)) * 2 + 1;
    bounds.assign(static_cast<size_t>(out_size) * 2, 0);
    kk.assign(static_cast<size_t>(out_size) * ksize, 0);
    std::vector<double> k(ksize);
    for (int xx = 0; xx < out_size; ++xx) {
        const double center = in0 + (xx + 0.5) * scale;
        double ww = 0.0;
        const double ss = 1.0 / filterscale;
        int xmin = static_cast<int>(center - support + 0.5);
        if (xmin < 0) xmin = 0;
        int xmax = static_cast<int>(center + support + 0.5);
        if (xmax > in_size) xmax = in_size;
        xmax -= xmin;
        int x;
        for (x = 0; x < xmax; ++x) {
            const double w = bicubic_filter((x + xmin - center + 0.5) * ss);
            k[x] = w;
            ww += w;
        }
        for (x = 0; x < xmax; ++x)
            if (ww != 0.0) k[x] /= ww;
        for (; x < ksize; ++x) k[x] = 0;
        for (x = 0; x < ksize; ++x) {
            const double v = k[x] * (1 << kPrecisionBits);
            kk[static_cast<size_t>(xx) * ksize + x] = k[x] < 0 ? static_cast<int>(-0.5 + v) : static_cast<int>(0.5 + v);
        }
        bounds[xx * 2 + 0] = xmin;
        bounds[xx * 2 + 1] = xmax;
    }
    *ksize_out = ksize;
}

static int get_coeffs(cid_ctx* c, int in_size, int out_size, ResizeCoeffs** out) {
    const uint64_t key = (static_cast<uint64_t>(static_cast<uint32_t>(in_size)) << 32) | static_cast<uint32_t>(out_size);
    auto it = c->resize_coeffs.find(key);
    if (it != c->resize_coeffs.end()) {
        *out = static_cast<ResizeCoeffs*>(it->second);
        return 0;
    }
    std::vector<int> bounds, kk;
    ResizeCoeffs* rc = new ResizeCoeffs();
    host_coeffs(in_size, out_size, &rc->ksize, bounds, kk);
    CID_CHECK_CUDA(cudaMalloc(reinterpret_cast<void**>(&rc->bounds), bounds.size() * 4));
    CID_CHECK_CUDA(cudaMalloc(reinterpret_cast<void**>(&rc->kk), kk.size() * 4));
    CID_CHECK_CUDA(cudaMemcpy(rc->bounds, bounds.data(), bounds.size() * 4, cudaMemcpyHostToDevice));
    CID_CHECK_CUDA(cudaMemcpy(rc->kk, kk.data(), kk.size() * 4, cudaMemcpyHostToDevice));
    c->resize_coeffs[key] = rc;
    *out = rc;
    return 0;
}

__device__ __forceinline__ uint8_t clip8(int v) {
    v >>= kPrecisionBits;
    return static_cast<uint8_t>(min(max(v, 0), 255));
}

// One thread per output byte; consecutive threads walk the contiguous (x, c) axis, so both the
// stores and (for the vertical pass) the loads are coalesced.
// HORIZ: out[b][y][xx][c] = sum_x in[b][y][xmin+x][c] * k[xx][x]      (in_w -> out_w, rows kept)
// VERT : out[b][yy][x][c] = sum_y in[b][ymin+y][x][c] * k[yy][y]      (in_h -> out_h, cols kept)
template <bool HORIZ>
__global__ void __launch_bounds__(256) resample_u8_kernel(const uint8_t* __restrict__ in, uint8_t* __restrict__ out,
                                                          int B, int in_h, int in_w, int out_h, int out_w, int C,
                                                          const int* __restrict__ bounds, const int* __restrict__ kk,
                                                          int ksize) {
    const size_t total = static_cast<size_t>(B) * out_h * out_w * C;
    const size_t t = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (t >= total) return;
    const int c = static_cast<int>(t % C);
    const int x = static_cast<int>((t / C) % out_w);
    const int y = static_cast<int>((t / (static_cast<size_t>(C) * out_w)) % out_h);
    const int b = static_cast<int>(t / (static_cast<size_t>(C) * out_w * out_h));
    const int o = HORIZ ? x : y;
    const int lo = bounds[2 * o], cnt = bounds[2 * o + 1];
    const int* k = kk + static_cast<size_t>(o) * ksize;
    int ss = 1 << (kPrecisionBits - 1);
    if (HORIZ) {
        const uint8_t* row = in + ((static_cast<size_t>(b) * in_h + y) * in_w + lo) * C + c;
        for (int i = 0; i < cnt; ++i) ss += static_cast<int>(row[static_cast<size_t>(i) * C]) * k[i];
    } else {
        const uint8_t* col = in + ((static_cast<size_t>(b) * in_h + lo) * in_w + x) * C + c;
        const size_t stride = static_cast<size_t>(in_w) * C;
        for (int i = 0; i < cnt; ++i) ss += static_cast<int>(col[i * stride]) * k[i];
    }
    out[t] = clip8(ss);
}

// ToTensor + Normalize: u8 NHWC -> fp32 NCHW, (v/255 - mean)/std with each op rounded in fp32.
__global__ void __launch_bounds__(256) normalize_kernel(const uint8_t* __restrict__ in, float* __restrict__ out, int B,
                                                        int H, int W, float mean, float stdv) {
    const size_t total = static_cast<size_t>(B) * H * W;
    const size_t pix = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (pix >= total) return;
    const size_t b = pix / (static_cast<size_t>(H) * W);
    const size_t r = pix - b * static_cast<size_t>(H) * W;
    const uint8_t* px = in + pix * 3;
#pragma unroll
    for (int c = 0; c < 3; ++c)
        out[(b * 3 + c) * static_cast<size_t>(H) * W + r] =
            __fdiv_rn(__fsub_rn(__fdiv_rn(static_cast<float>(px[c]), 255.f), mean), stdv);
}

}  // namespace cid

using namespace cid;

extern "C" {

int cid_resize_bicubic_u8(cid_handle h, const uint8_t* in, uint8_t* out, uint8_t* tmp, int B, int in_h, int in_w,
                          int out_h, int out_w, int C, void* stream) {
    CID_REQUIRE(h && in && out && tmp, "cid_resize_bicubic_u8: null argument");
    CID_REQUIRE(B >= 1 && in_h >= 1 && in_w >= 1 && out_h >= 1 && out_w >= 1 && C >= 1, "bad shape");
    DeviceGuard g(h->device);
    CID_REQUIRE(g.ok, "cannot select CUDA device %d", h->device);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    ResizeCoeffs *ch = nullptr, *cv = nullptr;
    if (get_coeffs(h, in_w, out_w, &ch)) return 1;
    if (get_coeffs(h, in_h, out_h, &cv)) return 1;
    {
        const size_t total = static_cast<size_t>(B) * in_h * out_w * C;
        resample_u8_kernel<true><<<static_cast<unsigned>((total + 255) / 256), 256, 0, st>>>(
            in, tmp, B, in_h, in_w, in_h, out_w, C, ch->bounds, ch->kk, ch->ksize);
        CID_CHECK_CUDA(cudaGetLastError());
        h->launches++;
    }
    {
        const size_t total = static_cast<size_t>(B) * out_h * out_w * C;
        resample_u8_kernel<false><<<static_cast<unsigned>((total + 255) / 256), 256, 0, st>>>(
            tmp, out, B, in_h, out_w, out_h, out_w, C, cv->bounds, cv->kk, cv->ksize);
        CID_CHECK_CUDA(cudaGetLastError());
        h->launches++;
    }
    return 0;
}

int cid_normalize_u8_to_f32_nchw(cid_handle h, const uint8_t* in, float* out, int B, int H, int W, float mean,
                                 float stdv, void* stream) {
    CID_REQUIRE(h && in && out, "cid_normalize_u8_to_f32_nchw: null argument");
    CID_REQUIRE(B >= 1 && H >= 1 && W >= 1, "bad shape");
    DeviceGuard g(h->device);
    CID_REQUIRE(g.ok, "cannot select CUDA device %d", h->device);
    const size_t total = static_cast<size_t>(B) * H * W;
    normalize_kernel<<<static_cast<unsigned>((total + 255) / 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
        in, out, B, H, W, mean, stdv);
    CID_CHECK_CUDA(cudaGetLastError());
    h->launches++;
    return 0;
}

}  // extern "C"
