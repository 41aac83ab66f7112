// Synthetic noiser: Gaussian and salt-and-pepper corruption of u8 images, HBM-bound.
//
// Reference (file:line in /root/reference):
//   add_gaussian_noise     backend/trainingcode/denoise_gan_code/noise_generation.py:6-10
//       noisy = np.clip(image + N(mean, sigma), 0, 255).astype(np.uint8)   (float64, truncating)
//   add_salt_pepper_noise  backend/trainingcode/denoise_gan_code/noise_generation.py:12-22
//       n = int(H*W*C*p) coordinates from randint(0, dim-1): rows in [0,H-2], cols in [0,W-2];
//       salt (255, all channels) first, pepper (0) second.
//
// numpy's MT19937 stream cannot be reproduced on the device, so each kernel has two modes:
// "supplied" (the caller passes the reference's own noise / coordinates -> bit-exact parity) and
// "philox" (Philox4x32-10 counter RNG, Box-Muller for normals -> statistical parity).
#include "cid_internal.h"
#include "ptx.cuh"

namespace cid {

// ---- Philox4x32-10 (Salmon et al. 2011), counter = (ctr_lo, ctr_hi, stream, 0), key = seed ----
__device__ __forceinline__ uint4 philox4x32_10(uint4 ctr, uint2 key) {
    constexpr uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        const uint32_t hi0 = __umulhi(M0, ctr.x), lo0 = M0 * ctr.x;
        const uint32_t hi1 = __umulhi(M1, ctr.z), lo1 = M1 * ctr.z;
        ctr = make_uint4(hi1 ^ ctr.y ^ key.x, lo1, hi0 ^ ctr.w ^ key.y, lo0);
        key.x += W0;
        key.y += W1;
    }
    return ctr;
}

__device__ __forceinline__ float u32_to_unit_open(uint32_t u) {
    // (0,1]: never 0, so log() is finite
    return (static_cast<float>(u >> 8) + 1.0f) * (1.0f / 16777216.0f);
}

__device__ __forceinline__ void box_muller(uint32_t a, uint32_t b, float& z0, float& z1) {
    const float r = sqrtf(-2.0f * logf(u32_to_unit_open(a)));
    float s, c;
    sincospif(2.0f * u32_to_unit_open(b), &s, &c);
    z0 = r * c;
    z1 = r * s;
}

__device__ __forceinline__ uint8_t clip_trunc_u8(double v) {
    v = fmin(fmax(v, 0.0), 255.0);
    return static_cast<uint8_t>(v);  // truncation, like ndarray.astype(np.uint8) on [0,255]
}

// 16 elements per thread: one 16-byte load and one 16-byte store of pixels.
__global__ void __launch_bounds__(256) gaussian_u8_kernel(const uint8_t* __restrict__ in, uint8_t* __restrict__ out,
                                                          size_t n, float mean, float sigma,
                                                          const double* __restrict__ noise, uint64_t seed,
                                                          uint64_t offset) {
    const size_t t = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    const size_t base = t * 16;
    if (base >= n) return;
    const bool full = base + 16 <= n && ((reinterpret_cast<uintptr_t>(in + base) | reinterpret_cast<uintptr_t>(out + base)) & 15) == 0;
    uint8_t px[16];
    if (full) {
        *reinterpret_cast<uint4*>(px) = __ldg(reinterpret_cast<const uint4*>(in + base));
    } else {
        for (int i = 0; i < 16; ++i) px[i] = base + i < n ? in[base + i] : 0;
    }
    double nz[16];
    if (noise) {
        for (int i = 0; i < 16; ++i) nz[i] = base + i < n ? noise[base + i] : 0.0;
    } else {
        const uint2 key = make_uint2(static_cast<uint32_t>(seed), static_cast<uint32_t>(seed >> 32));
#pragma unroll
        for (int g = 0; g < 4; ++g) {
            const uint64_t ctr = offset + t * 4 + g;
            const uint4 r = philox4x32_10(make_uint4(static_cast<uint32_t>(ctr), static_cast<uint32_t>(ctr >> 32), 0u, 0u), key);
            float z0, z1, z2, z3;
            box_muller(r.x, r.y, z0, z1);
            box_muller(r.z, r.w, z2, z3);
            nz[4 * g + 0] = static_cast<double>(fmaf(sigma, z0, mean));
            nz[4 * g + 1] = static_cast<double>(fmaf(sigma, z1, mean));
            nz[4 * g + 2] = static_cast<double>(fmaf(sigma, z2, mean));
            nz[4 * g + 3] = static_cast<double>(fmaf(sigma, z3, mean));
        }
    }
    uint8_t o[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) o[i] = clip_trunc_u8(static_cast<double>(px[i]) + nz[i]);
    if (full) {
        *reinterpret_cast<uint4*>(out + base) = *reinterpret_cast<const uint4*>(o);
    } else {
        for (int i = 0; i < 16; ++i)
            if (base + i < n) out[base + i] = o[i];
    }
}

// The other element-wise noise types of the reference's generator (noise_generation.py:24-39), same skeleton:
//   KIND 1 speckle  clip(x + x * N(mean, sigma), 0, 255)       p0 = mean, p1 = sigma; draws = the normal array
//   KIND 2 uniform  clip(x + U(low, high), 0, 255)             p0 = low,  p1 = high;  draws = the uniform array
//   KIND 3 poisson  np.random.poisson(x).astype(np.uint8)      draws = the Poisson counts (int64 -> u8 WRAPS mod 256)
// Philox mode (draws == NULL) gives statistical parity: Poisson counts by sequential inversion for lambda < 24 and by
// the rounded normal approximation above.
template <int KIND>
__global__ void __launch_bounds__(256) elementwise_noise_kernel(const uint8_t* __restrict__ in, uint8_t* __restrict__ out,
                                                                size_t n, float p0, float p1,
                                                                const double* __restrict__ draws, uint64_t seed,
                                                                uint64_t offset) {
    const size_t t = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    const size_t base = t * 16;
    if (base >= n) return;
    const bool full = base + 16 <= n && ((reinterpret_cast<uintptr_t>(in + base) | reinterpret_cast<uintptr_t>(out + base)) & 15) == 0;
    uint8_t px[16];
    if (full) {
        *reinterpret_cast<uint4*>(px) = __ldg(reinterpret_cast<const uint4*>(in + base));
    } else {
        for (int i = 0; i < 16; ++i) px[i] = base + i < n ? in[base + i] : 0;
    }
    double d[16];
    if (draws) {
        for (int i = 0; i < 16; ++i) d[i] = base + i < n ? draws[base + i] : 0.0;
    } else {
        const uint2 key = make_uint2(static_cast<uint32_t>(seed), static_cast<uint32_t>(seed >> 32));
#pragma unroll
        for (int g = 0; g < 4; ++g) {
            const uint64_t ctr = offset + t * 4 + g;
            const uint4 r = philox4x32_10(make_uint4(static_cast<uint32_t>(ctr), static_cast<uint32_t>(ctr >> 32), 2u + KIND, 0u), key);
            const uint32_t w[4] = {r.x, r.y, r.z, r.w};
            if (KIND == 1) {
                float z0, z1, z2, z3;
                box_muller(r.x, r.y, z0, z1);
                box_muller(r.z, r.w, z2, z3);
                d[4 * g + 0] = static_cast<double>(fmaf(p1, z0, p0));
                d[4 * g + 1] = static_cast<double>(fmaf(p1, z1, p0));
                d[4 * g + 2] = static_cast<double>(fmaf(p1, z2, p0));
                d[4 * g + 3] = static_cast<double>(fmaf(p1, z3, p0));
            } else if (KIND == 2) {
#pragma unroll
                for (int e = 0; e < 4; ++e)
                    d[4 * g + e] = static_cast<double>(p0) + static_cast<double>(p1 - p0) * (static_cast<double>(w[e]) * (1.0 / 4294967296.0));
            } else {
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    const float lam = static_cast<float>(px[4 * g + e]);
                    const float u = u32_to_unit_open(w[e]);
                    float k;
                    if (lam < 24.f) {  // inversion: smallest k with CDF(k) >= u
                        float pk = expf(-lam), cdf = pk;
                        k = 0.f;
                        while (cdf < u && k < 96.f) {
                            k += 1.f;
                            pk *= lam / k;
                            cdf += pk;
                        }
                    } else {           // N(lam, lam) rounded; the second uniform comes from a scrambled copy of the word
                        float z0, z1;
                        box_muller(w[e], w[e] * 2654435761u + 0x9E3779B9u, z0, z1);
                        k = fmaxf(floorf(lam + sqrtf(lam) * z0 + 0.5f), 0.f);
                    }
                    d[4 * g + e] = static_cast<double>(k);
                }
            }
        }
    }
    uint8_t o[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) {
        const double x = static_cast<double>(px[i]);
        if (KIND == 1) o[i] = clip_trunc_u8(x + x * d[i]);
        else if (KIND == 2) o[i] = clip_trunc_u8(x + d[i]);
        else o[i] = static_cast<uint8_t>(static_cast<long long>(d[i]) & 0xff);  // int64 -> uint8 wraps
    }
    if (full) {
        *reinterpret_cast<uint4*>(out + base) = *reinterpret_cast<const uint4*>(o);
    } else {
        for (int i = 0; i < 16; ++i)
            if (base + i < n) out[base + i] = o[i];
    }
}

__global__ void __launch_bounds__(256) copy_u8_kernel(const uint8_t* __restrict__ in, uint8_t* __restrict__ out, size_t n) {
    const size_t base = (static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x) * 16;
    if (base >= n) return;
    if (base + 16 <= n && ((reinterpret_cast<uintptr_t>(in + base) | reinterpret_cast<uintptr_t>(out + base)) & 15) == 0) {
        *reinterpret_cast<uint4*>(out + base) = __ldg(reinterpret_cast<const uint4*>(in + base));
    } else {
        for (int i = 0; i < 16 && base + i < n; ++i) out[base + i] = in[base + i];
    }
}

// one thread per (image, draw): sets all C channels of pixel (row, col) to `value`.
__global__ void __launch_bounds__(256) impulse_kernel(uint8_t* __restrict__ out, int B, int H, int W, int C, int count,
                                                      int first, int stride, uint8_t value,
                                                      const int32_t* __restrict__ coords, uint64_t seed,
                                                      uint64_t offset) {
    const size_t t = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (t >= static_cast<size_t>(B) * count) return;
    const int b = static_cast<int>(t / count);
    const int i = static_cast<int>(t % count);
    int row, col;
    if (coords) {
        const int32_t* cc = coords + (static_cast<size_t>(b) * stride + first + i) * 2;
        row = cc[0];
        col = cc[1];
    } else {
        const uint2 key = make_uint2(static_cast<uint32_t>(seed), static_cast<uint32_t>(seed >> 32));
        const uint64_t ctr = offset + static_cast<uint64_t>(b) * stride + first + i;
        const uint4 r = philox4x32_10(make_uint4(static_cast<uint32_t>(ctr), static_cast<uint32_t>(ctr >> 32), 1u, 0u), key);
        // randint(0, dim-1): high is exclusive, so the last row / column is never hit
        row = static_cast<int>(__umulhi(r.x, static_cast<uint32_t>(H - 1)));
        col = static_cast<int>(__umulhi(r.y, static_cast<uint32_t>(W - 1)));
    }
    if (row < 0 || row >= H || col < 0 || col >= W) return;
    uint8_t* px = out + ((static_cast<size_t>(b) * H + row) * W + col) * C;
    for (int c = 0; c < C; ++c) px[c] = value;
}

}  // namespace cid

using namespace cid;

extern "C" {

int cid_noise_gaussian_u8(cid_handle h, const uint8_t* in, uint8_t* out, size_t n, float mean, float sigma,
                          const double* noise, uint64_t seed, uint64_t offset, void* stream) {
    CID_REQUIRE(h && in && out, "cid_noise_gaussian_u8: null argument");
    if (n == 0) return 0;
    DeviceGuard g(h->device);
    CID_REQUIRE(g.ok, "cannot select CUDA device %d", h->device);
    const size_t threads = (n + 15) / 16;
    const unsigned grid = static_cast<unsigned>((threads + 255) / 256);
    gaussian_u8_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(in, out, n, mean, sigma, noise, seed, offset);
    CID_CHECK_CUDA(cudaGetLastError());
    h->launches++;
    return 0;
}

int cid_noise_elementwise_u8(cid_handle h, const uint8_t* in, uint8_t* out, size_t n, int kind, float p0, float p1,
                             const double* draws, uint64_t seed, uint64_t offset, void* stream) {
    CID_REQUIRE(h && in && out, "cid_noise_elementwise_u8: null argument");
    CID_REQUIRE(kind >= 0 && kind <= 3, "unknown noise kind %d (0 gaussian, 1 speckle, 2 uniform, 3 poisson)", kind);
    if (kind == 0) return cid_noise_gaussian_u8(h, in, out, n, p0, p1, draws, seed, offset, stream);
    if (n == 0) return 0;
    DeviceGuard g(h->device);
    CID_REQUIRE(g.ok, "cannot select CUDA device %d", h->device);
    const size_t threads = (n + 15) / 16;
    const unsigned grid = static_cast<unsigned>((threads + 255) / 256);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (kind == 1)
        elementwise_noise_kernel<1><<<grid, 256, 0, st>>>(in, out, n, p0, p1, draws, seed, offset);
    else if (kind == 2)
        elementwise_noise_kernel<2><<<grid, 256, 0, st>>>(in, out, n, p0, p1, draws, seed, offset);
    else
        elementwise_noise_kernel<3><<<grid, 256, 0, st>>>(in, out, n, p0, p1, draws, seed, offset);
    CID_CHECK_CUDA(cudaGetLastError());
    h->launches++;
    return 0;
}

int cid_noise_salt_pepper_u8(cid_handle h, const uint8_t* in, uint8_t* out, int B, int H, int W, int C,
                             double salt_prob, double pepper_prob, const int32_t* coords, int64_t coords_per_image,
                             uint64_t seed, uint64_t offset, void* stream) {
    CID_REQUIRE(h && in && out, "cid_noise_salt_pepper_u8: null argument");
    CID_REQUIRE(B >= 0 && H >= 2 && W >= 2 && C >= 1, "bad shape B=%d H=%d W=%d C=%d", B, H, W, C);
    if (B == 0) return 0;
    DeviceGuard g(h->device);
    CID_REQUIRE(g.ok, "cannot select CUDA device %d", h->device);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const size_t n = static_cast<size_t>(B) * H * W * C;
    // int(total_pixels * prob) with total_pixels = image_array.size = H*W*C: a float64 product of the caller's float64
    // probabilities, as the reference computes it (a float32 0.02 gives 599 instead of 600 draws for 100x100x3)
    const int n_salt = static_cast<int>(static_cast<double>(static_cast<size_t>(H) * W * C) * salt_prob);
    const int n_pepper = static_cast<int>(static_cast<double>(static_cast<size_t>(H) * W * C) * pepper_prob);
    CID_REQUIRE(n_salt >= 0 && n_pepper >= 0, "negative salt / pepper probability");
    CID_REQUIRE(!coords || coords_per_image == static_cast<int64_t>(n_salt) + n_pepper,
                "coords holds %lld draws per image; this image size and these probabilities need %d salt + %d pepper",
                static_cast<long long>(coords_per_image), n_salt, n_pepper);
    if (in != out) {
        const size_t threads = (n + 15) / 16;
        copy_u8_kernel<<<static_cast<unsigned>((threads + 255) / 256), 256, 0, st>>>(in, out, n);
        CID_CHECK_CUDA(cudaGetLastError());
        h->launches++;
    }
    const int stride = n_salt + n_pepper;
    if (n_salt > 0) {
        const size_t t = static_cast<size_t>(B) * n_salt;
        impulse_kernel<<<static_cast<unsigned>((t + 255) / 256), 256, 0, st>>>(out, B, H, W, C, n_salt, 0, stride, 255,
                                                                               coords, seed, offset);
        CID_CHECK_CUDA(cudaGetLastError());
        h->launches++;
    }
    if (n_pepper > 0) {
        const size_t t = static_cast<size_t>(B) * n_pepper;
        impulse_kernel<<<static_cast<unsigned>((t + 255) / 256), 256, 0, st>>>(out, B, H, W, C, n_pepper, n_salt, stride,
                                                                               0, coords, seed, offset);
        CID_CHECK_CUDA(cudaGetLastError());
        h->launches++;
    }
    return 0;
}

}  // extern "C"
