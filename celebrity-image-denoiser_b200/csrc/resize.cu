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

#include <algorithm>
#include <mutex>

#include "cid_internal.h"
#include "ptx.cuh"

namespace cid {

constexpr int kPrecisionBits = 32 - 8 - 2;

struct ResizeCoeffs {
    int ksize = 0;
    int out_size = 0;
    int* bounds = nullptr;  // device [out][2] = (xmin, count)
    int* kk = nullptr;      // device [out][ksize] fixed-point weights
    std::vector<int> host_bounds;
    // largest input span (last start + count - first start) over aligned groups of `tile` consecutive outputs
    int max_span(int tile) const {
        int best = 0;
        for (int o0 = 0; o0 < out_size; o0 += tile) {
            const int o1 = std::min(out_size, o0 + tile) - 1;
            best = std::max(best, host_bounds[2 * o1] + host_bounds[2 * o1 + 1] - host_bounds[2 * o0]);
        }
        return best;
    }
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
    rc->out_size = out_size;
    rc->host_bounds = bounds;
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

// ---------------------------------------------------------------------------------------------------------------------
// Both passes in ONE kernel for 3-channel images: a block owns a TX x TY tile of the output, stages the input window it
// needs in shared memory with 16-byte loads, runs the horizontal pass into a second shared buffer (rounded and clipped to
// u8 exactly as Pillow's intermediate image is) and the vertical pass from there to global memory with 4-byte stores.
// The intermediate image never touches HBM and no thread does a 64-bit division. Same integer arithmetic as the
// two-pass kernels above (which stay as the fallback for other channel counts / extreme scale factors): bit-identical.
//
// The kernel is bound by instruction issue, not by HBM: a 2x down-scale needs 9 + 9 taps per output byte and the input
// is 4x the output, i.e. ~31 integer MACs per output byte = 6 MACs per byte of HBM traffic - at 6.5 TB/s that would be
// 39 T MAC/s, more than the 36 T lane-instructions/s the whole GPU issues. So the work per MAC is what counts:
//   * KS (compile-time tap count >= the table's ksize; taps past an output's own count have zero coefficients, so every
//     output runs the same fully unrolled KS taps - reads past its window hit padding bytes times zero);
//   * horizontal pass: one thread = one output pixel (3 channels). Its 3*KS input bytes are contiguous: KS*3/4+1 aligned
//     32-bit shared loads, funnel-shifted to byte 0, then one PRMT + one IMAD per MAC; the coefficients come as 16-byte
//     loads shared by the 3 channels;
//   * vertical pass: one thread = 4 consecutive bytes of an output row: one 32-bit shared load per tap, PRMT + IMAD per MAC.
// Shared memory: [row offsets R ints][kH TX*KS ints][kV TY*KS ints][bH TX][bV TY][s_in R x pitch][s_mid R x TX*3][slack]
// ---------------------------------------------------------------------------------------------------------------------
struct ResizeFusedParams {
    const uint8_t* in;
    uint8_t* out;
    int B, in_h, in_w, out_h, out_w;
    const int *boundsH, *kkH, *boundsV, *kkV;
    int ksH, ksV;
    int tiles_x, tiles_y;
    int max_rows;   // input rows a tile can need
    int pitch;      // bytes per staged input row (multiple of 16, includes the alignment slack and 3*KS bytes of padding)
};

__device__ __forceinline__ int byte_of(uint32_t w, int k) {  // byte k of w, zero-extended: one PRMT
    return static_cast<int>(__byte_perm(w, 0u, 0x4440u + static_cast<uint32_t>(k)));
}

template <int TX, int TY, int KS>
__global__ void __launch_bounds__(256) resize_fused_kernel(const ResizeFusedParams p) {
    constexpr int C = 3, ROWB = TX * C;  // bytes of one tile row of the output / intermediate
    constexpr int KP = (KS + 3) & ~3;     // coefficient row pitch (ints): 16-byte loads
    constexpr int NW = (C * KS + 3) / 4;  // normalised words that hold one output pixel's 3*KS input bytes
    static_assert(ROWB % 4 == 0, "4-byte stores");
    extern __shared__ __align__(16) uint8_t rs_smem[];
    int* kH = reinterpret_cast<int*>(rs_smem);
    int* kV = kH + TX * KP;
    int* row_off = kV + TY * KP;
    int* bH = row_off + ((p.max_rows + 3) & ~3);
    int* bV = bH + TX;
    uint8_t* s_in = reinterpret_cast<uint8_t*>(bV + TY);
    s_in += (16 - (reinterpret_cast<uintptr_t>(s_in) & 15)) & 15;
    uint8_t* s_mid = s_in + static_cast<size_t>(p.max_rows) * p.pitch;

    const int tile = blockIdx.x;
    const int tx = tile % p.tiles_x;
    const int ty = (tile / p.tiles_x) % p.tiles_y;
    const int b = tile / (p.tiles_x * p.tiles_y);
    const int x0 = tx * TX, y0 = ty * TY;
    const int nx = min(TX, p.out_w - x0), ny = min(TY, p.out_h - y0);
    const int tid = threadIdx.x;

    // input window: the bounds are monotone, so the first / last output column (row) of the tile give the range
    const int xlo = __ldg(p.boundsH + 2 * x0);
    const int xhi = __ldg(p.boundsH + 2 * (x0 + nx - 1)) + __ldg(p.boundsH + 2 * (x0 + nx - 1) + 1);
    const int ylo = __ldg(p.boundsV + 2 * y0);
    const int yhi = __ldg(p.boundsV + 2 * (y0 + ny - 1)) + __ldg(p.boundsV + 2 * (y0 + ny - 1) + 1);
    const int rows = yhi - ylo, row_bytes = (xhi - xlo) * C;
    // coefficient tables (zero beyond the table's own ksize and beyond the tile) and window-relative first taps
    for (int i = tid; i < TX * KP; i += 256) {
        const int xx = i / KP, t = i - xx * KP;
        kH[i] = (xx < nx && t < p.ksH) ? __ldg(p.kkH + static_cast<size_t>(x0 + xx) * p.ksH + t) : 0;
    }
    for (int i = tid; i < TY * KP; i += 256) {
        const int yy = i / KP, t = i - yy * KP;
        kV[i] = (yy < ny && t < p.ksV) ? __ldg(p.kkV + static_cast<size_t>(y0 + yy) * p.ksV + t) : 0;
    }
    for (int i = tid; i < TX; i += 256) bH[i] = i < nx ? __ldg(p.boundsH + 2 * (x0 + i)) - xlo : 0;
    for (int i = tid; i < TY; i += 256) bV[i] = i < ny ? __ldg(p.boundsV + 2 * (y0 + i)) - ylo : 0;
    // stage the window: per row, the 16-byte chunks that cover [start, start + row_bytes); the row keeps its misalignment
    const size_t img_end = static_cast<size_t>(p.B) * p.in_h * p.in_w * C;
    const int chunks = (row_bytes + 15 + 15) / 16;  // upper bound incl. the misalignment
    const bool aligned_base = (reinterpret_cast<uintptr_t>(p.in) & 15) == 0;
    for (int i = tid; i < rows * chunks; i += 256) {
        const int r = i / chunks, ck = i - r * chunks;
        const size_t start = ((static_cast<size_t>(b) * p.in_h + ylo + r) * p.in_w + xlo) * C;
        const size_t a0 = start & ~static_cast<size_t>(15);
        const size_t a = a0 + static_cast<size_t>(ck) * 16;
        if (ck == 0) row_off[r] = static_cast<int>(start - a0);
        if (a < start + row_bytes) {
            uint4 v;
            if (a + 16 <= img_end && aligned_base) {
                v = __ldg(reinterpret_cast<const uint4*>(p.in + a));
            } else {  // the tensor's tail (or an unaligned base pointer): byte loads
                uint8_t t[16];
                for (int e = 0; e < 16; ++e) t[e] = a + e < img_end ? p.in[a + e] : 0;
                v = *reinterpret_cast<const uint4*>(t);
            }
            *reinterpret_cast<uint4*>(s_in + static_cast<size_t>(r) * p.pitch + ck * 16) = v;
        }
    }
    __syncthreads();
    // horizontal pass: one output pixel (3 channels) of one staged row per thread-iteration
    for (int i = tid; i < rows * TX; i += 256) {
        const int r = i / TX, xx = i - r * TX;
        const uint32_t addr = static_cast<uint32_t>(r) * p.pitch + row_off[r] + bH[xx] * C;  // byte offset into s_in
        const uint32_t* wp = reinterpret_cast<const uint32_t*>(s_in + (addr & ~3u));
        const uint32_t sh = (addr & 3u) * 8u;
        uint32_t w[NW + 1];
#pragma unroll
        for (int j = 0; j <= NW; ++j) w[j] = wp[j];
#pragma unroll
        for (int j = 0; j < NW; ++j) w[j] = __funnelshift_r(w[j], w[j + 1], sh);
        const int4* k4 = reinterpret_cast<const int4*>(kH + xx * KP);
        int s0 = 1 << (kPrecisionBits - 1), s1 = s0, s2 = s0;
#pragma unroll
        for (int q = 0; q < KP / 4; ++q) {
            const int4 kk = k4[q];
            const int kq[4] = {kk.x, kk.y, kk.z, kk.w};
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const int t = q * 4 + e;
                if (t < KS) {
                    s0 += byte_of(w[(3 * t) >> 2], (3 * t) & 3) * kq[e];
                    s1 += byte_of(w[(3 * t + 1) >> 2], (3 * t + 1) & 3) * kq[e];
                    s2 += byte_of(w[(3 * t + 2) >> 2], (3 * t + 2) & 3) * kq[e];
                }
            }
        }
        uint8_t* m = s_mid + r * ROWB + xx * C;
        m[0] = clip8(s0);
        m[1] = clip8(s1);
        m[2] = clip8(s2);
    }
    __syncthreads();
    // vertical pass: 4 consecutive bytes of one output row per thread-iteration
    const bool wide = (p.out_w * C) % 4 == 0 && (reinterpret_cast<uintptr_t>(p.out) & 3) == 0;
    for (int i = tid; i < ny * (ROWB / 4); i += 256) {
        const int yy = i / (ROWB / 4), j4 = (i - yy * (ROWB / 4)) * 4;
        if (j4 >= nx * C) continue;
        const uint8_t* col = s_mid + bV[yy] * ROWB + j4;
        const int4* k4 = reinterpret_cast<const int4*>(kV + yy * KP);
        int s0 = 1 << (kPrecisionBits - 1), s1 = s0, s2 = s0, s3 = s0;
#pragma unroll
        for (int q = 0; q < KP / 4; ++q) {
            const int4 kk = k4[q];
            const int kq[4] = {kk.x, kk.y, kk.z, kk.w};
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const int t = q * 4 + e;
                if (t < KS) {
                    const uint32_t v = *reinterpret_cast<const uint32_t*>(col + t * ROWB);
                    s0 += byte_of(v, 0) * kq[e];
                    s1 += byte_of(v, 1) * kq[e];
                    s2 += byte_of(v, 2) * kq[e];
                    s3 += byte_of(v, 3) * kq[e];
                }
            }
        }
        uint8_t* dst = p.out + ((static_cast<size_t>(b) * p.out_h + y0 + yy) * p.out_w + x0) * C + j4;
        const uint32_t o = static_cast<uint32_t>(clip8(s0)) | (static_cast<uint32_t>(clip8(s1)) << 8) |
                           (static_cast<uint32_t>(clip8(s2)) << 16) | (static_cast<uint32_t>(clip8(s3)) << 24);
        if (wide && j4 + 4 <= nx * C) {
            *reinterpret_cast<uint32_t*>(dst) = o;
        } else {
            for (int e = 0; e < 4; ++e)
                if (j4 + e < nx * C) dst[e] = static_cast<uint8_t>(o >> (8 * e));
        }
    }
}

// ToTensor + Normalize: u8 NHWC -> fp32 NCHW, (v/255 - mean)/std with each op rounded in fp32.
// Vector form (H*W a multiple of 4): a thread converts 4 consecutive pixels - three 4-byte loads, a 256-entry table in
// shared memory built with the exact fp32 operation sequence, three 16-byte stores (one per channel plane).
__global__ void __launch_bounds__(256) normalize4_kernel(const uint8_t* __restrict__ in, float* __restrict__ out, size_t quads,
                                                         size_t plane, float mean, float stdv) {
    __shared__ float lut[256];
    lut[threadIdx.x] = __fdiv_rn(__fsub_rn(__fdiv_rn(static_cast<float>(threadIdx.x), 255.f), mean), stdv);
    __syncthreads();
    const size_t q = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (q >= quads) return;
    const size_t pix = q * 4;
    const size_t b = pix / plane, r = pix - b * plane;
    const uint32_t* src = reinterpret_cast<const uint32_t*>(in + pix * 3);
    const uint32_t w0 = __ldg(src), w1 = __ldg(src + 1), w2 = __ldg(src + 2);  // r0 g0 b0 r1 | g1 b1 r2 g2 | b2 r3 g3 b3
    float* o = out + b * 3 * plane + r;
    const float4 cr = make_float4(lut[w0 & 0xff], lut[w0 >> 24], lut[(w1 >> 16) & 0xff], lut[(w2 >> 8) & 0xff]);
    const float4 cg = make_float4(lut[(w0 >> 8) & 0xff], lut[w1 & 0xff], lut[w1 >> 24], lut[(w2 >> 16) & 0xff]);
    const float4 cb = make_float4(lut[(w0 >> 16) & 0xff], lut[(w1 >> 8) & 0xff], lut[w2 & 0xff], lut[w2 >> 24]);
    *reinterpret_cast<float4*>(o) = cr;
    *reinterpret_cast<float4*>(o + plane) = cg;
    *reinterpret_cast<float4*>(o + 2 * plane) = cb;
}

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

// ---------------------------------------------------------------------------------------------------------------------
// Allocation-free variant of the two-pass resize (cid_resize_bicubic_u8_static, below). Everything the two passes need
// besides the caller's input and output lives in ONE file-scope device array: the four coefficient tables (computed on
// the device, so there is no host vector and no copy either) and the intermediate image of the horizontal pass.
// Every GPU that loads the library holds the array (kStaticScratchBytes) until the process exits.
//   [0, kStaticTableBytes)                    boundsH | kkH | boundsV | kkV   (ints)
//   [kStaticTableBytes, kStaticScratchBytes)  intermediate image, B x in_h x out_w x C bytes
// ---------------------------------------------------------------------------------------------------------------------
constexpr size_t kStaticTableBytes = CID_RESIZE_STATIC_TABLE_BYTES;
constexpr size_t kStaticTmpBytes = CID_RESIZE_STATIC_TMP_BYTES;
constexpr size_t kStaticScratchBytes = kStaticTableBytes + kStaticTmpBytes;
static __device__ __align__(256) uint8_t g_resize_static_scratch[kStaticScratchBytes];

// host_coeffs()'s arithmetic on the device: the same double operations in the same order, each one rounded on its own
// (no fused multiply-add), so the fixed-point tables - and with them the image - are the ones the host builds.
__device__ __forceinline__ double bicubic_filter_dev(double x) {
    const double a = -0.5;
    if (x < 0.0) x = -x;
    if (x < 1.0) {
        const double t = __dsub_rn(__dmul_rn(a + 2.0, x), a + 3.0);
        return __dadd_rn(__dmul_rn(__dmul_rn(t, x), x), 1.0);
    }
    if (x < 2.0) {
        const double t = __dadd_rn(__dmul_rn(__dsub_rn(x, 5.0), x), 8.0);
        return __dmul_rn(__dsub_rn(__dmul_rn(t, x), 4.0), a);
    }
    return 0.0;
}

__device__ __forceinline__ int static_ksize(int in_size, int out_size) {
    double fs = __ddiv_rn(static_cast<double>(static_cast<float>(in_size)), static_cast<double>(out_size));
    if (fs < 1.0) fs = 1.0;
    return static_cast<int>(ceil(__dmul_rn(2.0, fs))) * 2 + 1;
}

// One thread per output column (first out_w threads) or output row (the next out_h): its (first tap, count) pair and
// its ksize fixed-point weights, written into the table region of the static array (named directly: device code).
__global__ void __launch_bounds__(128) static_coeffs_kernel(int in_w, int out_w, int in_h, int out_h) {
    int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= out_w + out_h) return;
    int* tab = reinterpret_cast<int*>(g_resize_static_scratch);
    const int ksH = static_ksize(in_w, out_w);
    int in_size = in_w, out_size = out_w;
    int* bounds = tab;
    int* kk = tab + 2 * out_w;
    if (t >= out_w) {
        t -= out_w;
        in_size = in_h;
        out_size = out_h;
        bounds = tab + out_w * (2 + ksH);
        kk = bounds + 2 * out_h;
    }
    const double scale = __ddiv_rn(static_cast<double>(static_cast<float>(in_size)), static_cast<double>(out_size));
    const double filterscale = scale < 1.0 ? 1.0 : scale;
    const double support = __dmul_rn(2.0, filterscale);
    const int ksize = static_cast<int>(ceil(support)) * 2 + 1;
    const double center = __dmul_rn(__dadd_rn(static_cast<double>(t), 0.5), scale);
    const double ss = __ddiv_rn(1.0, filterscale);
    int xmin = static_cast<int>(__dadd_rn(__dsub_rn(center, support), 0.5));
    if (xmin < 0) xmin = 0;
    int xmax = static_cast<int>(__dadd_rn(__dadd_rn(center, support), 0.5));
    if (xmax > in_size) xmax = in_size;
    xmax -= xmin;
    double ww = 0.0;
    for (int x = 0; x < xmax; ++x)
        ww = __dadd_rn(ww, bicubic_filter_dev(__dmul_rn(__dadd_rn(__dsub_rn(static_cast<double>(x + xmin), center), 0.5), ss)));
    int* k = kk + static_cast<size_t>(t) * ksize;
    for (int x = 0; x < ksize; ++x) {
        double w = 0.0;
        if (x < xmax) {
            w = bicubic_filter_dev(__dmul_rn(__dadd_rn(__dsub_rn(static_cast<double>(x + xmin), center), 0.5), ss));
            if (ww != 0.0) w = __ddiv_rn(w, ww);
        }
        const double v = __dmul_rn(w, static_cast<double>(1 << kPrecisionBits));
        k[x] = w < 0 ? static_cast<int>(__dadd_rn(-0.5, v)) : static_cast<int>(__dadd_rn(0.5, v));
    }
    bounds[2 * t] = xmin;
    bounds[2 * t + 1] = xmax;
}

// The array's device address, looked up once per device with that device current (in host code the array's name is only
// a host-side stand-in for the symbol, not an address a kernel can use).
static int static_scratch_ptr(int device, uint8_t** out) {
    constexpr int kMaxDevices = 64;
    static std::mutex mu;
    static uint8_t* ptrs[kMaxDevices] = {};
    CID_REQUIRE(device >= 0 && device < kMaxDevices, "cid_resize_bicubic_u8_static: device %d out of range", device);
    std::lock_guard<std::mutex> lock(mu);
    if (!ptrs[device]) {
        void* p = nullptr;
        CID_CHECK_CUDA(cudaGetSymbolAddress(&p, g_resize_static_scratch));
        CID_REQUIRE(p != nullptr, "cid_resize_bicubic_u8_static: no device address for the scratch array");
        ptrs[device] = static_cast<uint8_t*>(p);
    }
    *out = ptrs[device];
    return 0;
}

}  // namespace cid

using namespace cid;

extern "C" {

int cid_resize_bicubic_u8_static(int device, const uint8_t* in, uint8_t* out, int B, int in_h, int in_w, int out_h,
                                 int out_w, int C, void* stream) {
    CID_REQUIRE(in && out, "cid_resize_bicubic_u8_static: null argument");
    CID_REQUIRE(B >= 1 && in_h >= 1 && in_w >= 1 && out_h >= 1 && out_w >= 1 && C >= 1, "bad shape");
    // the same ksize host_coeffs() derives (float box, double scale)
    auto ksize_of = [](int in_size, int out_size) {
        double fs = static_cast<double>(static_cast<float>(in_size)) / out_size;
        if (fs < 1.0) fs = 1.0;
        return static_cast<int>(ceil(2.0 * fs)) * 2 + 1;
    };
    const int ksH = ksize_of(in_w, out_w), ksV = ksize_of(in_h, out_h);
    const size_t table_bytes = (static_cast<size_t>(out_w) * (2 + ksH) + static_cast<size_t>(out_h) * (2 + ksV)) * 4;
    const size_t tmp_bytes = static_cast<size_t>(B) * in_h * out_w * C;
    CID_REQUIRE(table_bytes <= kStaticTableBytes,
                "cid_resize_bicubic_u8_static: coefficient tables of %zu bytes exceed the static limit of %zu "
                "(use cid_resize_bicubic_u8)", table_bytes, kStaticTableBytes);
    CID_REQUIRE(tmp_bytes <= kStaticTmpBytes,
                "cid_resize_bicubic_u8_static: B*in_h*out_w*C = %zu bytes exceeds the static limit of %zu "
                "(use cid_resize_bicubic_u8)", tmp_bytes, kStaticTmpBytes);
    DeviceGuard g(device);
    CID_REQUIRE(g.ok, "cannot select CUDA device %d", device);
    uint8_t* scratch = nullptr;
    if (static_scratch_ptr(device, &scratch)) return 1;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int* boundsH = reinterpret_cast<const int*>(scratch);
    const int* kkH = boundsH + 2 * out_w;
    const int* boundsV = boundsH + static_cast<size_t>(out_w) * (2 + ksH);
    const int* kkV = boundsV + 2 * out_h;
    uint8_t* tmp = scratch + kStaticTableBytes;
    // nothing in the array has to start at zero: the first kernel rewrites every table entry the passes read, the
    // horizontal pass every intermediate byte the vertical pass reads
    static_coeffs_kernel<<<(out_w + out_h + 127) / 128, 128, 0, st>>>(in_w, out_w, in_h, out_h);
    CID_CHECK_CUDA(cudaGetLastError());
    {
        const size_t total = tmp_bytes;
        resample_u8_kernel<true><<<static_cast<unsigned>((total + 255) / 256), 256, 0, st>>>(
            in, tmp, B, in_h, in_w, in_h, out_w, C, boundsH, kkH, ksH);
        CID_CHECK_CUDA(cudaGetLastError());
    }
    {
        const size_t total = static_cast<size_t>(B) * out_h * out_w * C;
        CID_REQUIRE(total + 255 < (1ull << 39), "resize: output too large");
        resample_u8_kernel<false><<<static_cast<unsigned>((total + 255) / 256), 256, 0, st>>>(
            tmp, out, B, in_h, out_w, out_h, out_w, C, boundsV, kkV, ksV);
        CID_CHECK_CUDA(cudaGetLastError());
    }
    return 0;
}

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
    if (C == 3 && std::max(ch->ksize, cv->ksize) <= 17) {
        // fused two-pass kernel: the largest tile whose input window fits in shared memory (two blocks per SM at least);
        // KS = the compile-time tap count that covers both tables
        const int ks = std::max(ch->ksize, cv->ksize);
        const int KS = ks <= 5 ? 5 : ks <= 9 ? 9 : 17;
        const int tiles[3][2] = {{64, 32}, {64, 16}, {32, 8}};
        for (int ti = 0; ti < 3; ++ti) {
            const int TX = tiles[ti][0], TY = tiles[ti][1];
            const int KP = (KS + 3) & ~3;
            const int max_cols = std::min(in_w, ch->max_span(TX)), max_rows = std::min(in_h, cv->max_span(TY)) ;
            ResizeFusedParams p;
            p.in = in; p.out = out; p.B = B; p.in_h = in_h; p.in_w = in_w; p.out_h = out_h; p.out_w = out_w;
            p.boundsH = ch->bounds; p.kkH = ch->kk; p.boundsV = cv->bounds; p.kkV = cv->kk;
            p.ksH = ch->ksize; p.ksV = cv->ksize;
            p.tiles_x = (out_w + TX - 1) / TX; p.tiles_y = (out_h + TY - 1) / TY;
            // the vertical pass reads KS rows from an output's first row: rows past the window are padding (zero taps)
            p.max_rows = max_rows + KS;
            p.pitch = (max_cols * 3 + 15 + 15 + 3 * KS + 8) / 16 * 16 + 16;
            const size_t smem = static_cast<size_t>(TX + TY) * KP * 4 + static_cast<size_t>((p.max_rows + 3) & ~3) * 4 + (TX + TY) * 4 + 16 +
                                static_cast<size_t>(p.max_rows) * p.pitch + static_cast<size_t>(p.max_rows) * TX * 3 + 64;
            if (smem > 100 * 1024) continue;
            const long long nblocks = static_cast<long long>(B) * p.tiles_x * p.tiles_y;
            CID_REQUIRE(nblocks < (1ll << 31), "resize: too many tiles");
            auto launch = [&](auto kern) -> int {
                CID_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
                kern<<<static_cast<unsigned>(nblocks), 256, smem, st>>>(p);
                CID_CHECK_CUDA(cudaGetLastError());
                return 0;
            };
            int rc;
            if (ti == 0) rc = KS == 5 ? launch(resize_fused_kernel<64, 32, 5>) : KS == 9 ? launch(resize_fused_kernel<64, 32, 9>) : launch(resize_fused_kernel<64, 32, 17>);
            else if (ti == 1) rc = KS == 5 ? launch(resize_fused_kernel<64, 16, 5>) : KS == 9 ? launch(resize_fused_kernel<64, 16, 9>) : launch(resize_fused_kernel<64, 16, 17>);
            else rc = KS == 5 ? launch(resize_fused_kernel<32, 8, 5>) : KS == 9 ? launch(resize_fused_kernel<32, 8, 9>) : launch(resize_fused_kernel<32, 8, 17>);
            if (rc) return 1;
            h->launches++;
            return 0;
        }
    }
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
    const size_t plane = static_cast<size_t>(H) * W;
    if (plane % 4 == 0 && ((reinterpret_cast<uintptr_t>(in) & 3) | (reinterpret_cast<uintptr_t>(out) & 15)) == 0)
        normalize4_kernel<<<static_cast<unsigned>((total / 4 + 255) / 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
            in, out, total / 4, plane, mean, stdv);
    else
        normalize_kernel<<<static_cast<unsigned>((total + 255) / 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
            in, out, B, H, W, mean, stdv);
    CID_CHECK_CUDA(cudaGetLastError());
    h->launches++;
    return 0;
}

}  // extern "C"
