// Synthetic noiser: Gaussian and salt-and-pepper corruption of u8 images (plus the reference's other noise types), in the
// u8 domain of noise_generation.py and in the [0,1] domain of esrgan_addNoise.py.
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
#include <string.h>

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

// ---- fast normals for the Philox modes (statistical parity only; the supplied-draw modes keep float64) ----
// One 32-bit Philox word gives TWO normals: the high half is the radius uniform, the low half the angle (Box-Muller on
// 16-bit uniforms (k + 0.5) / 65536: tails reach 4.86 sigma). Everything runs on the special-function unit's approx
// instructions (lg2, sqrt, sin, cos: 2 MUFU per normal) and full-rate ALU ops - the kernels were ALU-bound, not HBM-bound,
// with libdevice's logf / sincospif and float64 clipping (round 1: no GB/s figure; see profiles/r02*_membound*).
__device__ __forceinline__ float u16_to_float(uint32_t word, bool high) {
    // (0x4B000000 | k) is the float 8388608 + k: integer -> float without the conversion pipe
    const uint32_t m = __byte_perm(word, 0x4B000000u, high ? 0x7632u : 0x7610u);
    return __uint_as_float(m);
}
__device__ __forceinline__ void normals_from_word(uint32_t word, float& z0, float& z1) {
    constexpr float kBias = 8388608.0f - 0.5f;
    const float u = (u16_to_float(word, true) - kBias) * (1.0f / 65536.0f);             // (0, 1)
    const float a = (u16_to_float(word, false) - kBias) * (6.283185307179586f / 65536.0f);  // (0, 2 pi)
    float l, r, sn, cs;
    asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(l) : "f"(u));
    asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(l * -1.3862943611198906f));      // sqrt(-2 ln u)
    asm("sin.approx.ftz.f32 %0, %1;" : "=f"(sn) : "f"(a));
    asm("cos.approx.ftz.f32 %0, %1;" : "=f"(cs) : "f"(a));
    z0 = r * cs;
    z1 = r * sn;
}
// 16 standard normals for thread-counter t (two Philox4x32-10 calls)
__device__ __forceinline__ void normals16(uint64_t seed, uint64_t ctr0, uint32_t stream, float (&z)[16]) {
    const uint2 key = make_uint2(static_cast<uint32_t>(seed), static_cast<uint32_t>(seed >> 32));
#pragma unroll
    for (int g = 0; g < 2; ++g) {
        const uint64_t ctr = ctr0 + g;
        const uint4 r = philox4x32_10(make_uint4(static_cast<uint32_t>(ctr), static_cast<uint32_t>(ctr >> 32), stream, 0u), key);
        normals_from_word(r.x, z[8 * g + 0], z[8 * g + 1]);
        normals_from_word(r.y, z[8 * g + 2], z[8 * g + 3]);
        normals_from_word(r.z, z[8 * g + 4], z[8 * g + 5]);
        normals_from_word(r.w, z[8 * g + 6], z[8 * g + 7]);
    }
}
// 16 uniforms in [0, 1) with 16 bits each
__device__ __forceinline__ void uniforms16(uint64_t seed, uint64_t ctr0, uint32_t stream, float (&u)[16]) {
    const uint2 key = make_uint2(static_cast<uint32_t>(seed), static_cast<uint32_t>(seed >> 32));
#pragma unroll
    for (int g = 0; g < 2; ++g) {
        const uint64_t ctr = ctr0 + g;
        const uint4 r = philox4x32_10(make_uint4(static_cast<uint32_t>(ctr), static_cast<uint32_t>(ctr >> 32), stream, 0u), key);
        const uint32_t w[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            u[8 * g + 2 * e + 0] = (u16_to_float(w[e], true) - 8388608.0f) * (1.0f / 65536.0f);
            u[8 * g + 2 * e + 1] = (u16_to_float(w[e], false) - 8388608.0f) * (1.0f / 65536.0f);
        }
    }
}
// pixel byte i of a packed word as a float, without the conversion pipe
__device__ __forceinline__ float byte_to_float(uint32_t word, int i) {
    return __uint_as_float(__byte_perm(word, 0x4B000000u, 0x7440u | static_cast<uint32_t>(i))) - 8388608.0f;
}
// clip to [0, hi] and truncate: the float path of u8(np.clip(v, 0, 255)); returns the integer in the low byte
__device__ __forceinline__ uint32_t clip_trunc_bits(float v, float hi) {
    v = fminf(fmaxf(v, 0.f), hi);
    return __float_as_uint(__fadd_rz(v, 8388608.0f));  // 8388608 + floor(v): the low byte is the result
}
__device__ __forceinline__ uint32_t pack4_low_bytes(uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
    return __byte_perm(__byte_perm(a, b, 0x0040u), __byte_perm(c, d, 0x0040u), 0x5410u);
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
    __align__(16) uint8_t px[16];
    if (full) {
        *reinterpret_cast<uint4*>(px) = __ldg(reinterpret_cast<const uint4*>(in + base));
    } else {
        for (int i = 0; i < 16; ++i) px[i] = base + i < n ? in[base + i] : 0;
    }
    __align__(16) uint8_t o[16];
    if (noise) {
        double nz[16];
        for (int i = 0; i < 16; ++i) nz[i] = base + i < n ? noise[base + i] : 0.0;
#pragma unroll
        for (int i = 0; i < 16; ++i) o[i] = clip_trunc_u8(static_cast<double>(px[i]) + nz[i]);
    } else {
        float z[16];
        normals16(seed, offset + t * 2, 0u, z);
        const uint32_t* pw = reinterpret_cast<const uint32_t*>(px);
        uint32_t* ow = reinterpret_cast<uint32_t*>(o);
#pragma unroll
        for (int w = 0; w < 4; ++w) {
            uint32_t b[4];
#pragma unroll
            for (int e = 0; e < 4; ++e) b[e] = clip_trunc_bits(fmaf(sigma, z[4 * w + e], byte_to_float(pw[w], e) + mean), 255.f);
            ow[w] = pack4_low_bytes(b[0], b[1], b[2], b[3]);
        }
    }
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
    __align__(16) uint8_t px[16];
    if (full) {
        *reinterpret_cast<uint4*>(px) = __ldg(reinterpret_cast<const uint4*>(in + base));
    } else {
        for (int i = 0; i < 16; ++i) px[i] = base + i < n ? in[base + i] : 0;
    }
    if (draws == nullptr && (KIND == 1 || KIND == 2)) {
        // Philox mode (statistical parity): fp32 arithmetic, 16-bit-uniform Box-Muller (see normals_from_word)
        float r[16];
        if (KIND == 1) normals16(seed, offset + t * 2, 2u + KIND, r);
        else uniforms16(seed, offset + t * 2, 2u + KIND, r);
        const uint32_t* pw = reinterpret_cast<const uint32_t*>(px);
        __align__(16) uint32_t ow[4];
#pragma unroll
        for (int w = 0; w < 4; ++w) {
            uint32_t b[4];
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const float x = byte_to_float(pw[w], e);
                const float d = KIND == 1 ? fmaf(p1, r[4 * w + e], p0) : fmaf(p1 - p0, r[4 * w + e], p0);
                b[e] = clip_trunc_bits(KIND == 1 ? fmaf(x, d, x) : x + d, 255.f);
            }
            ow[w] = pack4_low_bytes(b[0], b[1], b[2], b[3]);
        }
        if (full) {
            *reinterpret_cast<uint4*>(out + base) = *reinterpret_cast<const uint4*>(ow);
        } else {
            const uint8_t* ob = reinterpret_cast<const uint8_t*>(ow);
            for (int i = 0; i < 16; ++i)
                if (base + i < n) out[base + i] = ob[i];
        }
        return;
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

// ---------------------------------------------------------------------------------------------------------------------
// [0,1]-domain noiser of the ESRGAN data-prep script (backend/trainingcode/esrgan_code/esrgan_addNoise.py), u8 in -> u8 out
// including the script's own conversions: img = u8 / 255.0 (:66), noise function, (noisy * 255).astype(np.uint8) (:79).
//   KIND 0 add_gaussian (:7-13)   clip(img + N(p0, p1), 0, 1)      reference: mean 0, sigma = sqrt(0.01) = 0.1 - the sigma
//                                                                  BASELINE.json configs[0] names for the synthetic faces
//   KIND 1 add_speckle  (:27-30)  clip(img + img * N(p0, p1), 0, 1)  reference: randn = N(0, 1)
//   KIND 2 add_uniform  (:38-41)  clip(img + U(p0, p1), 0, 1)        reference: U(-0.05, 0.05)
// draws (float64, the reference's own random array) -> float64 arithmetic as numpy, bit-exact; NULL -> Philox, fp32.
// ---------------------------------------------------------------------------------------------------------------------
template <int KIND>
__global__ void __launch_bounds__(256) unit_noise_kernel(const uint8_t* __restrict__ in, uint8_t* __restrict__ out, size_t n,
                                                         float p0, float p1, const double* __restrict__ draws, uint64_t seed,
                                                         uint64_t offset) {
    const size_t t = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    const size_t base = t * 16;
    if (base >= n) return;
    const bool full = base + 16 <= n && ((reinterpret_cast<uintptr_t>(in + base) | reinterpret_cast<uintptr_t>(out + base)) & 15) == 0;
    __align__(16) uint8_t px[16];
    if (full) {
        *reinterpret_cast<uint4*>(px) = __ldg(reinterpret_cast<const uint4*>(in + base));
    } else {
        for (int i = 0; i < 16; ++i) px[i] = base + i < n ? in[base + i] : 0;
    }
    __align__(16) uint8_t o[16];
    if (draws) {
#pragma unroll
        for (int i = 0; i < 16; ++i) {
            const double x = static_cast<double>(px[i]) / 255.0;
            const double d = base + i < n ? draws[base + i] : 0.0;
            double v = KIND == 1 ? x + x * d : x + d;
            v = fmin(fmax(v, 0.0), 1.0);
            o[i] = static_cast<uint8_t>(v * 255.0);  // (noisy * 255).astype(np.uint8): truncation
        }
    } else {
        float r[16];
        if (KIND == 2) uniforms16(seed, offset + t * 2, 8u + KIND, r);
        else normals16(seed, offset + t * 2, 8u + KIND, r);
        const uint32_t* pw = reinterpret_cast<const uint32_t*>(px);
        uint32_t* ow = reinterpret_cast<uint32_t*>(o);
#pragma unroll
        for (int w = 0; w < 4; ++w) {
            uint32_t b[4];
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const float x = byte_to_float(pw[w], e) * (1.0f / 255.0f);
                const float d = KIND == 2 ? fmaf(p1 - p0, r[4 * w + e], p0) : fmaf(p1, r[4 * w + e], p0);
                const float v = fminf(fmaxf(KIND == 1 ? fmaf(x, d, x) : x + d, 0.f), 1.f);
                b[e] = clip_trunc_bits(v * 255.0f, 255.f);
            }
            ow[w] = pack4_low_bytes(b[0], b[1], b[2], b[3]);
        }
    }
    if (full) {
        *reinterpret_cast<uint4*>(out + base) = *reinterpret_cast<const uint4*>(o);
    } else {
        for (int i = 0; i < 16; ++i)
            if (base + i < n) out[base + i] = o[i];
    }
}

// add_salt_pepper (esrgan_addNoise.py:15-25): ceil(amount * img.size * 0.5) ELEMENT coordinates (row, col, channel) from
// randint(0, dim) - the full range, unlike the denoise variant - set to 1 (-> 255), then as many set to 0.
// one thread per (image, draw); coords int32 [B][2 * count][3], salt draws first.
__global__ void __launch_bounds__(256) unit_impulse_kernel(uint8_t* __restrict__ out, int B, int H, int W, int C, int count,
                                                           int first, int stride, uint8_t value,
                                                           const int32_t* __restrict__ coords, uint64_t seed, uint64_t offset) {
    const size_t t = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (t >= static_cast<size_t>(B) * count) return;
    const int b = static_cast<int>(t / count);
    const int i = static_cast<int>(t % count);
    int row, col, ch;
    if (coords) {
        const int32_t* cc = coords + (static_cast<size_t>(b) * stride + first + i) * 3;
        row = cc[0];
        col = cc[1];
        ch = cc[2];
    } else {
        const uint2 key = make_uint2(static_cast<uint32_t>(seed), static_cast<uint32_t>(seed >> 32));
        const uint64_t ctr = offset + static_cast<uint64_t>(b) * stride + first + i;
        const uint4 r = philox4x32_10(make_uint4(static_cast<uint32_t>(ctr), static_cast<uint32_t>(ctr >> 32), 7u, 0u), key);
        row = static_cast<int>(__umulhi(r.x, static_cast<uint32_t>(H)));
        col = static_cast<int>(__umulhi(r.y, static_cast<uint32_t>(W)));
        ch = static_cast<int>(__umulhi(r.z, static_cast<uint32_t>(C)));
    }
    if (row < 0 || row >= H || col < 0 || col >= W || ch < 0 || ch >= C) return;
    out[((static_cast<size_t>(b) * H + row) * W + col) * C + ch] = value;
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
        // draw j = b * stride + first + i of the batch: Philox counter offset + j / 2, words (x, y) or (z, w)
        const uint64_t j = static_cast<uint64_t>(b) * stride + first + i;
        const uint64_t ctr = offset + (j >> 1);
        const uint4 r = philox4x32_10(make_uint4(static_cast<uint32_t>(ctr), static_cast<uint32_t>(ctr >> 32), 1u, 0u), key);
        // randint(0, dim-1): high is exclusive, so the last row / column is never hit
        row = static_cast<int>(__umulhi((j & 1) ? r.z : r.x, static_cast<uint32_t>(H - 1)));
        col = static_cast<int>(__umulhi((j & 1) ? r.w : r.y, static_cast<uint32_t>(W - 1)));
    }
    if (row < 0 || row >= H || col < 0 || col >= W) return;
    uint8_t* px = out + ((static_cast<size_t>(b) * H + row) * W + col) * C;
    for (int c = 0; c < C; ++c) px[c] = value;
}

// ---------------------------------------------------------------------------------------------------------------------
// Salt & pepper in ONE pass. The three-launch form above (copy, salt scatter, pepper scatter) moves the batch through HBM
// once for the copy and then pays a read-modify-write of a 32-byte sector for every one of the millions of scattered
// byte stores (ncu, 256 x 512^2: copy 59 us at 6.8 TB/s, each scatter 130 us with 92 MB read + 100 MB written).
// Here a CTA owns a slab of `rows` image rows in shared memory: bulk load (cp.async.bulk, no thread touches the bytes)
// -> every thread walks the image's draw list and applies the draws that land in the slab (salt first, barrier, then
// pepper: the reference's order, noise_generation.py:17-21) -> bulk store. Every CTA of an image regenerates the image's
// whole draw list (Philox is a pure function of the counter), so the output is bit-identical to the scatter kernels; the
// price is `slabs` times the RNG work - with 31 k draws per 512^2 image that work, not the copy, sets the time, so the
// slab is as large as shared memory allows (one 1024-thread CTA per SM) and one Philox call serves two draws.
// Tried and dropped: the image in the distributed shared memory of an 8-CTA cluster with every draw computed once and
// stored into its owner's slab (st.shared::cluster.u8): 94 k remote byte stores per image made it slower (1.11 vs
// 0.94 ms for 1024 images).
// UNIT = the [0,1]-domain variant (esrgan_addNoise.py:15-25): element coordinates (row, col, channel), full randint range.
// ---------------------------------------------------------------------------------------------------------------------
constexpr int kSpThreads = 1024;
constexpr uint32_t kSpChunk = 32768;  // bytes per bulk copy

template <bool UNIT>
__global__ void __launch_bounds__(kSpThreads, 1) salt_pepper_slab_kernel(const uint8_t* __restrict__ in, uint8_t* __restrict__ out,
                                                                         int H, int W, int C, int rows, int slabs, int n_salt,
                                                                         int n_pepper, const int32_t* __restrict__ coords,
                                                                         uint64_t seed, uint64_t offset) {
    extern __shared__ __align__(128) uint8_t sp_smem[];
    __shared__ __align__(8) uint64_t sp_bar;
    const int b = blockIdx.x / slabs, sl = blockIdx.x - b * slabs;
    const int r0 = sl * rows, nr = min(rows, H - r0);
    const size_t row_bytes = static_cast<size_t>(W) * C;
    const size_t g0 = (static_cast<size_t>(b) * H + r0) * row_bytes;  // first byte of the slab in the batch
    const int nbytes = static_cast<int>(nr * row_bytes);
    const uint8_t* src = in + g0;
    uint8_t* dst = out + g0;
    const int tid = threadIdx.x;
    // The slab keeps its misalignment (shared-memory offset = global address mod 16), so its 16-byte-aligned body moves
    // with bulk async copies and only the <= 15-byte head / tail with ordinary loads and stores.
    const int mis = static_cast<int>(reinterpret_cast<uintptr_t>(src) & 15);
    const bool same_mis = (reinterpret_cast<uintptr_t>(dst) & 15) == static_cast<uintptr_t>(mis);
    uint8_t* slab = sp_smem + mis;
    const int head = mis ? min(16 - mis, nbytes) : 0;
    const int body = (nbytes - head) & ~15;
    const int tail0 = head + body;
    if (tid == 0) {
        mbar_init(&sp_bar, 1);
        fence_mbar_init();
        if (body > 0) {
            mbar_arrive_expect_tx(&sp_bar, static_cast<uint32_t>(body));
            for (int o = 0; o < body; o += kSpChunk)
                bulk_load_1d(slab + head + o, src + head + o, min(static_cast<uint32_t>(body - o), kSpChunk), &sp_bar);
        }
    }
    for (int i = tid; i < head; i += kSpThreads) slab[i] = src[i];
    for (int i = tail0 + tid; i < nbytes; i += kSpThreads) slab[i] = src[i];
    __syncthreads();  // barrier initialised, head / tail in place
    if (body > 0) mbar_wait(&sp_bar, 0);
    const int stride = n_salt + n_pepper;
    const uint2 key = make_uint2(static_cast<uint32_t>(seed), static_cast<uint32_t>(seed >> 32));
    auto apply = [&](int row, int col, int ch, uint8_t value) {
        if (row < r0 || row >= r0 + nr || col < 0 || col >= W) return;
        uint8_t* px = slab + (static_cast<size_t>(row - r0) * W + col) * C;
        if (UNIT) {
            if (ch >= 0 && ch < C) px[ch] = value;
        } else {
            for (int c = 0; c < C; ++c) px[c] = value;
        }
    };
#pragma unroll 1
    for (int pass = 0; pass < 2; ++pass) {
        const int first = pass ? n_salt : 0, count = pass ? n_pepper : n_salt;
        const uint8_t value = pass ? 0 : 255;
        if (coords) {
            for (int i = tid; i < count; i += kSpThreads) {
                const int32_t* cc = coords + (static_cast<size_t>(b) * stride + first + i) * (UNIT ? 3 : 2);
                apply(__ldg(cc), __ldg(cc + 1), UNIT ? __ldg(cc + 2) : 0, value);
            }
        } else {
            // draw j of the image (j = first + i) comes from Philox counter offset + (b * stride + j) / DPC, lane j % DPC
            constexpr int DPC = UNIT ? 1 : 2;  // draws per Philox call
            const uint64_t d0 = static_cast<uint64_t>(b) * stride + first;  // global index of the pass's first draw
            const uint64_t c0 = d0 / DPC, c1 = (d0 + count + DPC - 1) / DPC;
            for (uint64_t cidx = c0 + tid; cidx < c1; cidx += kSpThreads) {
                const uint64_t ctr = offset + cidx;
                const uint4 r = philox4x32_10(make_uint4(static_cast<uint32_t>(ctr), static_cast<uint32_t>(ctr >> 32), UNIT ? 7u : 1u, 0u), key);
                if (UNIT) {
                    apply(static_cast<int>(__umulhi(r.x, static_cast<uint32_t>(H))), static_cast<int>(__umulhi(r.y, static_cast<uint32_t>(W))),
                          static_cast<int>(__umulhi(r.z, static_cast<uint32_t>(C))), value);
                } else {  // randint(0, dim-1): high is exclusive, the last row / column is never hit
                    const uint64_t j0 = cidx * 2;
                    if (j0 >= d0 && j0 < d0 + count)
                        apply(static_cast<int>(__umulhi(r.x, static_cast<uint32_t>(H - 1))),
                              static_cast<int>(__umulhi(r.y, static_cast<uint32_t>(W - 1))), 0, value);
                    if (j0 + 1 >= d0 && j0 + 1 < d0 + count)
                        apply(static_cast<int>(__umulhi(r.z, static_cast<uint32_t>(H - 1))),
                              static_cast<int>(__umulhi(r.w, static_cast<uint32_t>(W - 1))), 0, value);
                }
            }
        }
        __syncthreads();
    }
    if (same_mis) {
        fence_proxy_async();  // the byte stores above must be visible to the bulk-copy engine
        __syncthreads();
        if (tid == 0 && body > 0) {
            for (int o = 0; o < body; o += kSpChunk)
                bulk_store_1d(dst + head + o, slab + head + o, min(static_cast<uint32_t>(body - o), kSpChunk));
            bulk_commit();
        }
        for (int i = tid; i < head; i += kSpThreads) dst[i] = slab[i];
        for (int i = tail0 + tid; i < nbytes; i += kSpThreads) dst[i] = slab[i];
        if (tid == 0 && body > 0) bulk_wait_read<0>();  // shared memory stays allocated until the engine has read it
    } else {
        for (int i = tid; i < nbytes; i += kSpThreads) dst[i] = slab[i];
    }
}

// Launch the slab kernel if an image row group of useful size fits in shared memory; returns -1 when it does not (the
// caller then runs the three-launch form), 0 on success, 1 on error.
template <bool UNIT>
static int launch_salt_pepper_slab(cid_ctx* h, const uint8_t* in, uint8_t* out, int B, int H, int W, int C, int n_salt,
                                   int n_pepper, const int32_t* coords, uint64_t seed, uint64_t offset, cudaStream_t st) {
    const size_t row_bytes = static_cast<size_t>(W) * C;
    const size_t cap = static_cast<size_t>(h->max_smem) - 2048;
    if (row_bytes == 0 || row_bytes * 8 > cap) return -1;
    int slabs = static_cast<int>((static_cast<size_t>(H) * row_bytes + cap - 1) / cap);
    // small batches: more, smaller slabs so that the grid covers the SMs (the RNG work per image grows with `slabs`)
    while (static_cast<long long>(B) * slabs < h->num_sms && (H + slabs) / (slabs + 1) >= 8 &&
           static_cast<long long>(slabs + 1) * (n_salt + n_pepper) <= 4ll * H * W)
        ++slabs;
    const int rows = (H + slabs - 1) / slabs;
    slabs = (H + rows - 1) / rows;
    const size_t smem = static_cast<size_t>(rows) * row_bytes + 32;
    const long long grid = static_cast<long long>(B) * slabs;
    if (grid >= (1ll << 31)) return -1;
    auto kern = salt_pepper_slab_kernel<UNIT>;
    CID_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, h->max_smem - 1024));  // (+ the static mbarrier)
    kern<<<static_cast<unsigned>(grid), kSpThreads, smem, st>>>(in, out, H, W, C, rows, slabs, n_salt, n_pepper, coords, seed, offset);
    CID_CHECK_CUDA(cudaGetLastError());
    h->launches++;
    return 0;
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
    if (!h->sp_scatter) {
        const int rc = launch_salt_pepper_slab<false>(h, in, out, B, H, W, C, n_salt, n_pepper, coords, seed, offset, st);
        if (rc >= 0) return rc;
    }
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

int cid_noise_unit_u8(cid_handle h, const uint8_t* in, uint8_t* out, size_t n, int kind, float p0, float p1,
                      const double* draws, uint64_t seed, uint64_t offset, void* stream) {
    CID_REQUIRE(h && in && out, "cid_noise_unit_u8: null argument");
    CID_REQUIRE(kind >= 0 && kind <= 2, "unknown [0,1]-domain noise kind %d (0 gaussian, 1 speckle, 2 uniform)", kind);
    if (n == 0) return 0;
    DeviceGuard g(h->device);
    CID_REQUIRE(g.ok, "cannot select CUDA device %d", h->device);
    const size_t threads = (n + 15) / 16;
    const unsigned grid = static_cast<unsigned>((threads + 255) / 256);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (kind == 0)
        unit_noise_kernel<0><<<grid, 256, 0, st>>>(in, out, n, p0, p1, draws, seed, offset);
    else if (kind == 1)
        unit_noise_kernel<1><<<grid, 256, 0, st>>>(in, out, n, p0, p1, draws, seed, offset);
    else
        unit_noise_kernel<2><<<grid, 256, 0, st>>>(in, out, n, p0, p1, draws, seed, offset);
    CID_CHECK_CUDA(cudaGetLastError());
    h->launches++;
    return 0;
}

int cid_noise_salt_pepper_unit_u8(cid_handle h, const uint8_t* in, uint8_t* out, int B, int H, int W, int C, double amount,
                                  const int32_t* coords, int64_t coords_per_image, uint64_t seed, uint64_t offset,
                                  void* stream) {
    CID_REQUIRE(h && in && out, "cid_noise_salt_pepper_unit_u8: null argument");
    CID_REQUIRE(B >= 0 && H >= 1 && W >= 1 && C >= 1, "bad shape B=%d H=%d W=%d C=%d", B, H, W, C);
    CID_REQUIRE(amount >= 0.0, "negative amount");
    if (B == 0) return 0;
    DeviceGuard g(h->device);
    CID_REQUIRE(g.ok, "cannot select CUDA device %d", h->device);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const size_t n = static_cast<size_t>(B) * H * W * C;
    // np.ceil(amount * img.size * 0.5), float64, for salt and again for pepper (esrgan_addNoise.py:17, 21)
    const int count = static_cast<int>(ceil(amount * static_cast<double>(static_cast<size_t>(H) * W * C) * 0.5));
    CID_REQUIRE(!coords || coords_per_image == 2 * static_cast<int64_t>(count),
                "coords holds %lld draws per image; this image size and amount need %d salt + %d pepper",
                static_cast<long long>(coords_per_image), count, count);
    if (!h->sp_scatter) {
        const int rc = launch_salt_pepper_slab<true>(h, in, out, B, H, W, C, count, count, coords, seed, offset, st);
        if (rc >= 0) return rc;
    }
    if (in != out) {
        const size_t threads = (n + 15) / 16;
        copy_u8_kernel<<<static_cast<unsigned>((threads + 255) / 256), 256, 0, st>>>(in, out, n);
        CID_CHECK_CUDA(cudaGetLastError());
        h->launches++;
    }
    if (count > 0) {
        const size_t t = static_cast<size_t>(B) * count;
        const unsigned grid = static_cast<unsigned>((t + 255) / 256);
        unit_impulse_kernel<<<grid, 256, 0, st>>>(out, B, H, W, C, count, 0, 2 * count, 255, coords, seed, offset);
        unit_impulse_kernel<<<grid, 256, 0, st>>>(out, B, H, W, C, count, count, 2 * count, 0, coords, seed, offset);
        CID_CHECK_CUDA(cudaGetLastError());
        h->launches += 2;
    }
    return 0;
}

}  // extern "C"
