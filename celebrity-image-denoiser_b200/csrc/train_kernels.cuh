// Memory-bound and small kernels of the training step: MaxPool / ReLU / tanh backward, train-mode BatchNorm apply and backward,
// the discriminator head with BCELoss, column sums, Adam and the fp32 -> bf16 weight re-pack.
//
// Reference ops (file:line in /root/reference/backend/trainingcode/denoise_gan_code/training.py):
//   DenoiseDiscriminator                      :77-98   (Conv k3, LeakyReLU(0.2), Conv k3 s2, BatchNorm2d, AdaptiveAvgPool2d,
//                                                       Conv k1, Sigmoid)
//   nn.MSELoss / nn.BCELoss                   :234-235, 413-423
//   optim.Adam(lr=1e-4, betas=(0.9,0.999))    :237-238, 416, 425
//   autograd of DenoiseGenerator.forward      :19-74 (MaxPool2d, ReLU, tanh, torch.cat backward)
#pragma once
#include <cuda_bf16.h>

#include "ptx.cuh"

namespace cid {

__device__ __forceinline__ float bf16_lo(uint32_t w) { return __uint_as_float(w << 16); }
__device__ __forceinline__ float bf16_hi(uint32_t w) { return __uint_as_float(w & 0xffff0000u); }

// two / eight consecutive channels of a bf16 or fp32 NHWC tensor as floats
__device__ __forceinline__ void load2(const __nv_bfloat16* p, float (&f)[2]) {
    const uint32_t w = *reinterpret_cast<const uint32_t*>(p);
    f[0] = bf16_lo(w); f[1] = bf16_hi(w);
}
__device__ __forceinline__ void load2(const float* p, float (&f)[2]) {
    const float2 v = *reinterpret_cast<const float2*>(p);
    f[0] = v.x; f[1] = v.y;
}
__device__ __forceinline__ void load8(const __nv_bfloat16* p, float (&f)[8]) {
    const uint4 v = __ldg(reinterpret_cast<const uint4*>(p));
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) { f[2 * j] = bf16_lo(w[j]); f[2 * j + 1] = bf16_hi(w[j]); }
}
__device__ __forceinline__ void load8(const float* p, float (&f)[8]) {
    const float4 a = __ldg(reinterpret_cast<const float4*>(p)), b = __ldg(reinterpret_cast<const float4*>(p) + 1);
    f[0] = a.x; f[1] = a.y; f[2] = a.z; f[3] = a.w; f[4] = b.x; f[5] = b.y; f[6] = b.z; f[7] = b.w;
}

__device__ __forceinline__ float block_sum(float v, float* red /* >= 32 floats */) {
    __syncthreads();
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
    __syncthreads();
    const int nw = (blockDim.x + 31) >> 5;
    float s = (threadIdx.x < nw) ? red[threadIdx.x] : 0.f;
    if (threadIdx.x < 32) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    }
    return s;  // valid in thread 0
}

// ---------------------------------------------------------------------------------------------------------------------
// Generator loss head: content = MSE(y, clean); gz = (2 (y - clean) / n + gx) * (1 - y^2) = gradient w.r.t. the last
// conv's pre-tanh output (gx = the discriminator's data gradient, already scaled by 0.001 / B through the BCE head).
// Also the last conv's bias gradient (sum of gz per channel).
// ---------------------------------------------------------------------------------------------------------------------
struct GLossParams {
    int B, H, W;
    const float* y;      // fp32 NCHW generator output
    const float* clean;  // fp32 NCHW
    const float* gx;     // fp32 NHWC-3 or null
    float* gz;           // fp32 NHWC-3 ...
    float* gz_nchw;      // ... and the same as NCHW (input layout of the first-conv kernel that computes the data gradient)
    double* loss_sum;    // += sum (y - clean)^2 / n
    float* bias_grad;    // [3] +=
    float inv_n;
};

__global__ void __launch_bounds__(256) gloss_kernel(const GLossParams p) {
    // programmatic dependent launch: the next kernel may start its set-up now; this one orders itself after its predecessor
    pdl_launch_dependents();
    pdl_wait();
    __shared__ float red[32];
    const size_t total = static_cast<size_t>(p.B) * p.H * p.W;
    const size_t pix = static_cast<size_t>(blockIdx.x) * 256 + threadIdx.x;
    float sq = 0.f, g[3] = {0.f, 0.f, 0.f};
    if (pix < total) {
        const size_t hw = static_cast<size_t>(p.H) * p.W;
        const size_t b = pix / hw, yx = pix % hw;
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            const float yv = p.y[(b * 3 + c) * hw + yx];
            const float d = yv - p.clean[(b * 3 + c) * hw + yx];
            sq += d * d;
            float gr = 2.f * d * p.inv_n;
            if (p.gx) gr += p.gx[pix * 3 + c];
            g[c] = gr * (1.f - yv * yv);
            p.gz[pix * 3 + c] = g[c];
            p.gz_nchw[(b * 3 + c) * hw + yx] = g[c];
        }
    }
    const float s = block_sum(sq, red);
    if (threadIdx.x == 0) atomicAdd(p.loss_sum, static_cast<double>(s) * p.inv_n);
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        const float sg = block_sum(g[c], red);
        if (threadIdx.x == 0) atomicAdd(p.bias_grad + c, sg);
    }
}

// ---------------------------------------------------------------------------------------------------------------------
// Column sums of an NHWC bf16 tensor slice: out[c] += sum_p v[p][c0 + c] (bias gradients), optionally also the sums of
// squares (BatchNorm batch statistics, accumulated in double).
// ---------------------------------------------------------------------------------------------------------------------
template <bool SQ, typename ZT>
__global__ void __launch_bounds__(256) colsum_kernel(const ZT* v, size_t P, int cs, int c0, int C, float* out,
                                                     double* dsum, double* dsq, int ppb) {
    // programmatic dependent launch: the next kernel may start its set-up now; this one orders itself after its predecessor
    pdl_launch_dependents();
    pdl_wait();
    // thread = (8-channel chunk, pixel row group); C/8 chunks (8, 16 or 32) -> 32, 16 or 8 pixels per iteration
    __shared__ float red[256 * 8];
    __shared__ float red2[SQ ? 256 * 8 : 8];
    const int chunks = C >> 3;
    const int rpi = 256 / chunks;
    const int ck = threadIdx.x % chunks, rg = threadIdx.x / chunks;
    const size_t p0 = static_cast<size_t>(blockIdx.x) * ppb, p1 = min(P, p0 + ppb);
    float s[8], q[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) { s[j] = 0.f; q[j] = 0.f; }
#pragma unroll 4
    for (size_t pix = p0 + rg; pix < p1; pix += rpi) {
        float f[8];
        load8(v + pix * cs + c0 + 8 * ck, f);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            s[j] += f[j];
            if (SQ) q[j] = fmaf(f[j], f[j], q[j]);
        }
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        red[(rg * chunks + ck) * 8 + j] = s[j];
        if (SQ) red2[(rg * chunks + ck) * 8 + j] = q[j];
    }
    __syncthreads();
    if (threadIdx.x < C) {
        const int c = threadIdx.x;
        float a = 0.f, b = 0.f;
        for (int r = 0; r < rpi; ++r) {
            a += red[r * C + c];
            if (SQ) b += red2[r * C + c];
        }
        if (SQ) {
            atomicAdd(dsum + c, static_cast<double>(a));
            atomicAdd(dsq + c, static_cast<double>(b));
        } else {
            atomicAdd(out + c, a);
        }
    }
}

// ---------------------------------------------------------------------------------------------------------------------
// MaxPool2d(2,2) backward fused with the skip-connection add and the ReLU mask of the pooled activation e:
//   dz[p][c] = (e[p][c] > 0) ? gskip[p][c] + (p is the first arg-max of its 2x2 window ? gpool[p/2][c] : 0) : 0
// e and gskip live in channel slices of the concat buffers (stride cs, first channel c0).
// ---------------------------------------------------------------------------------------------------------------------
// One thread = one 2x2 window x 4 channels: 4 e + 4 gskip + 1 gpool 8-byte loads, 4 stores (H, W even). Four channels
// (not eight) per thread keep the kernel at ~40 registers, i.e. 5+ blocks per SM instead of 2: the kernel is a pure
// stream and was waiting on its own loads with a quarter of the warps resident. 32-bit index arithmetic (the host checks
// that the item count fits).
__global__ void __launch_bounds__(256) poolbwd_kernel(const __nv_bfloat16* e, const __nv_bfloat16* gskip, int cs, int c0,
                                                      const __nv_bfloat16* gpool, __nv_bfloat16* dz, int B, int H, int W, int C) {
    // programmatic dependent launch: the next kernel may start its set-up now; this one orders itself after its predecessor
    pdl_launch_dependents();
    pdl_wait();
    const uint32_t chunks = static_cast<uint32_t>(C) >> 2;
    const uint32_t Hp = static_cast<uint32_t>(H) >> 1, Wp = static_cast<uint32_t>(W) >> 1;
    const uint32_t total = static_cast<uint32_t>(B) * Hp * Wp * chunks;
    const uint32_t i = blockIdx.x * 256u + threadIdx.x;
    if (i >= total) return;
    const uint32_t ck = i % chunks, pp = i / chunks;  // pooled pixel
    const uint32_t xp = pp % Wp, t = pp / Wp;
    const uint32_t yp = t % Hp, b = t / Hp;
    const size_t q0 = (static_cast<size_t>(b) * H + 2 * yp) * W + 2 * xp;
    size_t q[4] = {q0, q0 + 1, q0 + W, q0 + W + 1};
    uint2 ev[4], gv[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        ev[j] = __ldg(reinterpret_cast<const uint2*>(e + q[j] * cs + c0 + ck * 4));
        gv[j] = __ldg(reinterpret_cast<const uint2*>(gskip + q[j] * cs + c0 + ck * 4));
    }
    const uint2 gp = __ldg(reinterpret_cast<const uint2*>(gpool + static_cast<size_t>(pp) * C + ck * 4));
    uint32_t o[4][2];
#pragma unroll
    for (int wi = 0; wi < 2; ++wi) {
        float r[4][2];
        const uint32_t gpw = wi ? gp.y : gp.x;
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            float ej[4], gj[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const uint32_t w = wi ? ev[j].y : ev[j].x;
                const uint32_t g = wi ? gv[j].y : gv[j].x;
                ej[j] = h ? bf16_hi(w) : bf16_lo(w);
                gj[j] = h ? bf16_hi(g) : bf16_lo(g);
            }
            const float mx = fmaxf(fmaxf(ej[0], ej[1]), fmaxf(ej[2], ej[3]));
            int first = 3;  // torch's MaxPool2d keeps the first maximum in scan order
#pragma unroll
            for (int j = 3; j >= 0; --j)
                if (ej[j] == mx) first = j;
            const float gpv = h ? bf16_hi(gpw) : bf16_lo(gpw);
#pragma unroll
            for (int j = 0; j < 4; ++j) r[j][h] = ej[j] > 0.f ? gj[j] + (first == j ? gpv : 0.f) : 0.f;
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) o[j][wi] = pack_bf16x2(r[j][0], r[j][1]);
    }
#pragma unroll
    for (int j = 0; j < 4; ++j)
        *reinterpret_cast<uint2*>(dz + q[j] * C + ck * 4) = make_uint2(o[j][0], o[j][1]);
}

// ---------------------------------------------------------------------------------------------------------------------
// Train-mode BatchNorm2d (eps 1e-5, momentum 0.1) + LeakyReLU(0.2)
// ---------------------------------------------------------------------------------------------------------------------
// Batch statistics -> the affine table of the apply kernel. Every block of bn_apply_kernel derives the table for itself
// from the double sums the producing conv's epilogue left (C <= 128 channels: one thread each, ~200 cycles - cheaper
// than a one-block kernel and its launch boundary in the dependency chain); block 0 also publishes stats[c] =
// {mean, invstd} for the backward kernels and updates the running statistics as torch does (unbiased variance).
struct BnFwd {
    const double *dsum, *dsq;  // per-channel sum and sum of squares over the n = B*H*W pixels
    double n;
    float eps, momentum;
    float *stats, *running_mean, *running_var, *nbt;
    const float *gamma, *beta;
};

// a = lrelu(z * a2 + b2) with {a2 = gamma * invstd, b2 = beta - gamma * mean * invstd}; 8 channels per thread
constexpr int kBnIter = 8;  // items (8 channels of one pixel) per thread of the BatchNorm apply kernels

template <typename ZT>
__global__ void __launch_bounds__(256) bn_apply_kernel(const ZT* z, __nv_bfloat16* a, size_t P, int C, const BnFwd f,
                                                       float slope) {
    // programmatic dependent launch: the next kernel may start its set-up now; this one orders itself after its predecessor
    pdl_launch_dependents();
    pdl_wait();
    __shared__ float tab[2 * 128];
    if (static_cast<int>(threadIdx.x) < C) {
        const int c = threadIdx.x;
        const double mean = f.dsum[c] / f.n;
        double var = f.dsq[c] / f.n - mean * mean;
        if (var < 0) var = 0;
        const float istd = static_cast<float>(1.0 / sqrt(var + static_cast<double>(f.eps)));
        tab[2 * c] = f.gamma[c] * istd;
        tab[2 * c + 1] = f.beta[c] - f.gamma[c] * istd * static_cast<float>(mean);
        if (blockIdx.x == 0) {
            f.stats[2 * c] = static_cast<float>(mean);
            f.stats[2 * c + 1] = istd;
            f.running_mean[c] = (1.f - f.momentum) * f.running_mean[c] + f.momentum * static_cast<float>(mean);
            f.running_var[c] = (1.f - f.momentum) * f.running_var[c] + f.momentum * static_cast<float>(var * f.n / (f.n - 1.0));
            if (c == 0) f.nbt[0] += 1.f;
        }
    }
    __syncthreads();
    // 256 is a multiple of C/8: a thread keeps the same 8 channels over its items, their table entries in registers
    const int chunks = C >> 3;
    const int c8 = static_cast<int>(threadIdx.x % chunks) * 8;
    float a2[8], b2[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) { a2[j] = tab[2 * (c8 + j)]; b2[j] = tab[2 * (c8 + j) + 1]; }
    const size_t n = P * chunks;
    size_t i = static_cast<size_t>(blockIdx.x) * (256 * kBnIter) + threadIdx.x;
#pragma unroll 4
    for (int it = 0; it < kBnIter; ++it, i += 256) {
        if (i >= n) break;
        float zf[8], r[8];
        load8(z + i * 8, zf);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const float yv = fmaf(zf[j], a2[j], b2[j]);
            r[j] = yv > 0.f ? yv : yv * slope;
        }
        st_global_v4(a + i * 8, pack_bf16x2(r[0], r[1]), pack_bf16x2(r[2], r[3]), pack_bf16x2(r[4], r[5]), pack_bf16x2(r[6], r[7]));
    }
}

// four consecutive channels of a bf16 or fp32 NHWC tensor as floats
__device__ __forceinline__ void load4(const __nv_bfloat16* p, float (&f)[4]) {
    const uint2 v = __ldg(reinterpret_cast<const uint2*>(p));
    f[0] = bf16_lo(v.x); f[1] = bf16_hi(v.x); f[2] = bf16_lo(v.y); f[3] = bf16_hi(v.y);
}
__device__ __forceinline__ void load4(const float* p, float (&f)[4]) {
    const float4 a = __ldg(reinterpret_cast<const float4*>(p));
    f[0] = a.x; f[1] = a.y; f[2] = a.z; f[3] = a.w;
}

// dY = dA * lrelu'(y); sums[c] += {sum dY, sum dY * xhat}. dA comes from a tensor (da) or, for the layer under the
// average pool, is broadcast from dpool[b][c] (already divided by H*W). 4 channels per thread and four pixels in flight:
// with 8 channels the per-channel constants alone took 48 registers and two blocks per SM could not cover the load
// latency of this pure stream (0.43 of HBM peak, `profiles/r02t_train_mem`).
template <typename ZT>
__global__ void __launch_bounds__(256, 5) bn_bwd_reduce_kernel(const ZT* z, const __nv_bfloat16* da,
                                                            const float* dpool, size_t P, size_t hw, int C,
                                                            const float* stats, const float* gamma, const float* beta,
                                                            float slope, double* sums, int ppb, int reps) {
    // programmatic dependent launch: the next kernel may start its set-up now; this one orders itself after its predecessor
    pdl_launch_dependents();
    pdl_wait();
    __shared__ float red[256 * 4];
    __shared__ float red2[256 * 4];
    const int chunks = C >> 2, rpi = 256 / chunks;
    const int ck = threadIdx.x % chunks, rg = threadIdx.x / chunks;
    const size_t p0 = static_cast<size_t>(blockIdx.x) * ppb, p1 = min(P, p0 + ppb);
    const int c4 = ck * 4;
    float mean[4], istd[4], ga[4], be[4], s[4], q[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        mean[j] = stats[2 * (c4 + j)]; istd[j] = stats[2 * (c4 + j) + 1];
        ga[j] = gamma[c4 + j]; be[j] = beta[c4 + j];
        s[j] = 0.f; q[j] = 0.f;
    }
    if (rg < rpi) {  // C / 4 divides 256 for the discriminator's 64 / 128 channels: every thread has a row group
#pragma unroll 4
        for (size_t pix = p0 + rg; pix < p1; pix += rpi) {
            float zz[4], dd[4];
            load4(z + pix * C + c4, zz);
            if (da) {
                load4(da + pix * C + c4, dd);
            } else {
                const size_t b = pix / hw;
                load4(dpool + b * C + c4, dd);
            }
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const float xh = (zz[j] - mean[j]) * istd[j];
                const float yv = ga[j] * xh + be[j];
                const float dy = yv > 0.f ? dd[j] : dd[j] * slope;
                s[j] += dy;
                q[j] = fmaf(dy, xh, q[j]);
            }
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            red[(rg * chunks + ck) * 4 + j] = s[j];
            red2[(rg * chunks + ck) * 4 + j] = q[j];
        }
    }
    __syncthreads();
    if (threadIdx.x < C) {
        const int cc = threadIdx.x;
        float a = 0.f, b2 = 0.f;
        for (int r = 0; r < rpi; ++r) {
            a += red[r * C + cc];
            b2 += red2[r * C + cc];
        }
        // the blocks of the one wave finish together: `reps` replicas of the 2 C sums (summed by the finalize kernel) keep
        // the same-address depth of the double atomics at blocks / reps
        double* dst = sums + static_cast<size_t>(blockIdx.x % reps) * (2 * C);
        atomicAdd(dst + cc, static_cast<double>(a));
        atomicAdd(dst + C + cc, static_cast<double>(b2));
    }
}

// per-channel table of the backward apply: with xh = (z - mean) * istd, y = z * a2 + b2, dy = dA * lrelu'(y):
//   dz = gamma * istd * (dy - sum_dy / n - xh * sum_dyxh / n) = k0 * dy - c1 - z * c2,  table = {a2, b2, k0, c1, c2}
// Every block derives the table for itself from the replicated sums of bn_bwd_reduce_kernel (as the forward apply does);
// block 0 optionally adds the gradients of gamma / beta: += {sum dY xhat, sum dY}.
struct BnBwd {
    const double* sums_rep;  // [reps][2][C]
    int reps;
    double n;
    const float *stats, *gamma, *beta;
    float *ggamma, *gbeta;  // may be null
};

template <typename ZT>
__global__ void __launch_bounds__(256) bn_bwd_apply_kernel(const ZT* z, const __nv_bfloat16* da,
                                                           const float* dpool, __nv_bfloat16* dz, size_t P, size_t hw,
                                                           int C, float slope, const BnBwd f) {
    // programmatic dependent launch: the next kernel may start its set-up now; this one orders itself after its predecessor
    pdl_launch_dependents();
    pdl_wait();
    __shared__ float btab[5 * 128];
    if (static_cast<int>(threadIdx.x) < C) {
        const int c = threadIdx.x;
        double sums[2] = {0.0, 0.0};  // {sum dY, sum dY xhat} of channel c over the replicas
        for (int r = 0; r < f.reps; ++r) {
            sums[0] += f.sums_rep[r * 2 * C + c];
            sums[1] += f.sums_rep[r * 2 * C + C + c];
        }
        const float mean = f.stats[2 * c], istd = f.stats[2 * c + 1], ga = f.gamma[c];
        const float m_dy = static_cast<float>(sums[0] / f.n), m_dyx = static_cast<float>(sums[1] / f.n);
        const float k0 = ga * istd;
        btab[5 * c + 0] = ga * istd;
        btab[5 * c + 1] = f.beta[c] - ga * istd * mean;
        btab[5 * c + 2] = k0;
        btab[5 * c + 3] = k0 * (m_dy - m_dyx * mean * istd);
        btab[5 * c + 4] = k0 * m_dyx * istd;
        if (blockIdx.x == 0 && f.ggamma) {
            f.gbeta[c] += static_cast<float>(sums[0]);
            f.ggamma[c] += static_cast<float>(sums[1]);
        }
    }
    __syncthreads();
    const int chunks = C >> 3;
    const int c8 = static_cast<int>(threadIdx.x % chunks) * 8;
    float t0[8], t1[8], t2[8], t3[8], t4[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        const float* t5 = btab + 5 * (c8 + j);
        t0[j] = t5[0]; t1[j] = t5[1]; t2[j] = t5[2]; t3[j] = t5[3]; t4[j] = t5[4];
    }
    const size_t n = P * chunks;
    size_t i = static_cast<size_t>(blockIdx.x) * (256 * kBnIter) + threadIdx.x;
#pragma unroll 4
    for (int it = 0; it < kBnIter; ++it, i += 256) {
        if (i >= n) break;
        float zf[8], dd[8];
        load8(z + i * 8, zf);
        if (da) {
            load8(da + i * 8, dd);
        } else {
            const size_t b = (i / chunks) / hw;
            const float4 d0 = *reinterpret_cast<const float4*>(dpool + b * C + c8), d1 = *reinterpret_cast<const float4*>(dpool + b * C + c8 + 4);
            dd[0] = d0.x; dd[1] = d0.y; dd[2] = d0.z; dd[3] = d0.w; dd[4] = d1.x; dd[5] = d1.y; dd[6] = d1.z; dd[7] = d1.w;
        }
        float r[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const float yv = fmaf(zf[j], t0[j], t1[j]);
            const float dy = yv > 0.f ? dd[j] : dd[j] * slope;
            r[j] = fmaf(t2[j], dy, -t3[j]) - zf[j] * t4[j];
        }
        st_global_v4(dz + i * 8, pack_bf16x2(r[0], r[1]), pack_bf16x2(r[2], r[3]), pack_bf16x2(r[4], r[5]), pack_bf16x2(r[6], r[7]));
    }
}

// ---------------------------------------------------------------------------------------------------------------------
// Discriminator head: AdaptiveAvgPool2d(1) -> Conv2d(128,1,k1) -> Sigmoid -> BCELoss(mean) and its backward.
// ---------------------------------------------------------------------------------------------------------------------
// pooled[b][c] = mean over hw of a[b][.][c];  grid = B, block = C
// (pooled must be zeroed first: grid.y pixel slices accumulate with atomics)
__global__ void avgpool_kernel(const __nv_bfloat16* a, float* pooled, int hw, int C) {
    // programmatic dependent launch: the next kernel may start its set-up now; this one orders itself after its predecessor
    pdl_launch_dependents();
    pdl_wait();
    const int b = blockIdx.x, c = threadIdx.x;
    const int per = (hw + gridDim.y - 1) / gridDim.y;
    const int i0 = blockIdx.y * per, i1 = min(hw, i0 + per);
    const __nv_bfloat16* src = a + static_cast<size_t>(b) * hw * C + c;
    float s = 0.f;
#pragma unroll 8
    for (int i = i0; i < i1; ++i) s += __bfloat162float(src[static_cast<size_t>(i) * C]);
    atomicAdd(pooled + b * C + c, s / static_cast<float>(hw));
}

// One block of 1024 threads (C = 128). loss_sum += BCE(sigmoid(w.pooled + b), target) (mean over B);
// dlogit = scale * (p - target) / B; dpool[b][c] = dlogit[b] * w[c] / hw; with_param_grads: gw += dlogit^T pooled, gb += sum.
__global__ void __launch_bounds__(1024) dhead_kernel(const float* pooled, const float* w, const float* bias, int B, int hw,
                                                     float target, float scale, double* loss_sum, float* dpool, float* gw,
                                                     float* gb, float* probs, int with_param_grads) {
    // programmatic dependent launch: the next kernel may start its set-up now; this one orders itself after its predecessor
    pdl_launch_dependents();
    pdl_wait();
    extern __shared__ float dl[];  // [B]
    __shared__ float red[32];
    __shared__ float gpart[8][128];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    float lsum = 0.f;
    for (int b = warp; b < B; b += 32) {
        float s = 0.f;
#pragma unroll
        for (int c = lane; c < 128; c += 32) s = fmaf(w[c], pooled[b * 128 + c], s);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
        const float logit = s + bias[0];
        const float pr = 1.f / (1.f + expf(-logit));
        if (lane == 0) {
            const float l1 = fmaxf(logf(pr), -100.f), l0 = fmaxf(log1pf(-pr), -100.f);
            lsum += -(target * l1 + (1.f - target) * l0);
            dl[b] = scale * (pr - target) / static_cast<float>(B);
            if (probs) probs[b] = pr;
        }
    }
    const float tot = block_sum(lsum, red);
    if (threadIdx.x == 0) atomicAdd(loss_sum, static_cast<double>(tot) / B);
    __syncthreads();
    const int c = threadIdx.x & 127, bg = threadIdx.x >> 7;
    float g = 0.f;
    const float wc = w[c], inv_hw = 1.f / static_cast<float>(hw);
    for (int b = bg; b < B; b += 8) {
        g = fmaf(dl[b], pooled[b * 128 + c], g);
        dpool[b * 128 + c] = dl[b] * wc * inv_hw;
    }
    gpart[bg][c] = g;
    __syncthreads();
    if (with_param_grads && threadIdx.x < 128) {
        float t = 0.f;
#pragma unroll
        for (int k = 0; k < 8; ++k) t += gpart[k][c];
        gw[c] += t;
        if (c == 0) {
            float sb = 0.f;
            for (int b = 0; b < B; ++b) sb += dl[b];
            gb[0] += sb;
        }
    }
}

// ---------------------------------------------------------------------------------------------------------------------
// Adam (torch.optim.Adam defaults: eps 1e-8, no weight decay, no amsgrad) over a flat parameter range
// ---------------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) adam_kernel(float* w, const float* g, float* m, float* v, size_t n, float lr,
                                                   float b1, float b2, float eps, float bc1, float bc2_sqrt) {
    // programmatic dependent launch: the next kernel may start its set-up now; this one orders itself after its predecessor
    pdl_launch_dependents();
    pdl_wait();
    const size_t i = static_cast<size_t>(blockIdx.x) * 256 + threadIdx.x;
    if (i >= n) return;
    const float gi = g[i];
    const float mi = b1 * m[i] + (1.f - b1) * gi;
    const float vi = b2 * v[i] + (1.f - b2) * gi * gi;
    m[i] = mi;
    v[i] = vi;
    const float denom = sqrtf(vi) / bc2_sqrt + eps;
    w[i] -= (lr / bc1) * (mi / denom);
}

// ---------------------------------------------------------------------------------------------------------------------
// Data-parallel Adam: gradient all-reduce FUSED into the optimiser step over NVLink peer memory (one launch instead of
// all-reduce + scale + Adam). Every rank's state buffer is symmetric memory mapped into every peer: each rank reads the
// gradient section of all ranks (rank 0 .. world-1 in that fixed order, so all replicas compute bit-identical sums),
// averages, and applies Adam to its own replica of the weights. One-shot is the right shape at this size (7.3 MB / 1 MB
// sections: launch latency, not link bandwidth). Cross-GPU ordering = two flag barriers per launch, per block index:
//   start: "my gradients are complete" (stream order puts this launch after the backward pass) -> peers may read them
//   end:   "I have finished reading yours"                                                    -> peers may overwrite them
// flags[phase][rank][block] live in a symmetric int buffer and hold the epoch of the last launch that passed (monotonic).
// The barriers are per block index, so all kPeerBlocks blocks of every rank must be co-resident (2 per SM on 148 SMs;
// cid_train_set_peers checks the occupancy). A peer that has not arrived after `spin_ns` (default 10 minutes, in the
// range of NCCL's watchdog; cid_train_set_option "peer_timeout_ms") is a fault: the block records it in the sticky
// error words (device + pinned host mirror) and RETURNS WITHOUT TOUCHING weights, exp_avg or exp_avg_sq; every later launch
// sees the sticky word and does nothing either, so its peers time out in turn and the error reaches every rank. The
// host checks the mirror at every cid_train_adam / cid_train_losses call. The replicas may differ after such a fault
// (blocks that had already passed the barrier applied their update): the caller restores a checkpoint.
// ---------------------------------------------------------------------------------------------------------------------
constexpr int kPeerMax = 8;
constexpr int kPeerBlocks = 296;                       // all co-resident: the barriers are per block index
constexpr unsigned long long kPeerSpinNsDefault = 600ull * 1000000000ull;
struct PeerSet {
    const float* grads[kPeerMax];  // peers' gradient sections (this network), rank order
    int* flags[kPeerMax];          // peers' flag buffers [2][kPeerMax][kPeerBlocks]
    int world, rank;
};
__device__ __forceinline__ void st_release_sys(int* p, int v) { asm volatile("st.release.sys.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory"); }
__device__ __forceinline__ int ld_acquire_sys(const int* p) {
    int v;
    asm volatile("ld.acquire.sys.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ float ld_sys_f32(const float* p) {
    float v;
    asm volatile("ld.relaxed.sys.global.f32 %0, [%1];" : "=f"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ float4 ld_sys_v4(const float* p) {
    float4 v;
    asm volatile("ld.relaxed.sys.global.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ unsigned long long global_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
// returns false (block-uniform) if a peer did not arrive in time
__device__ __forceinline__ bool peer_barrier(const PeerSet& ps, int phase, int epoch, unsigned long long spin_ns, int* err_dev,
                                             int* err_host) {
    __shared__ int timed_out;
    if (threadIdx.x == 0) timed_out = 0;
    __syncthreads();
    if (threadIdx.x < ps.world) {
        const int slot = (phase * kPeerMax) * kPeerBlocks + blockIdx.x;
        st_release_sys(ps.flags[threadIdx.x] + slot + ps.rank * kPeerBlocks, epoch);
        const int* mine = ps.flags[ps.rank] + slot + threadIdx.x * kPeerBlocks;
        const unsigned long long t0 = global_ns();
        while (ld_acquire_sys(mine) < epoch) {
            if (global_ns() - t0 > spin_ns) {
                timed_out = 1;
                *err_dev = 1;
                *reinterpret_cast<volatile int*>(err_host) = 1;
                __threadfence_system();
                break;
            }
        }
    }
    __syncthreads();
    return timed_out == 0;
}
__device__ __forceinline__ void adam_update(float& w, float& m, float& v, float g, float lr_bc1, float b1, float b2, float eps,
                                            float bc2_sqrt) {
    m = b1 * m + (1.f - b1) * g;
    v = b2 * v + (1.f - b2) * g * g;
    w -= lr_bc1 * (m / (sqrtf(v) / bc2_sqrt + eps));
}
__global__ void __launch_bounds__(256) adam_allreduce_kernel(float* w, float* m, float* v, size_t n, const __grid_constant__ PeerSet ps,
                                                             int epoch, unsigned long long spin_ns, int* err_dev, int* err_host,
                                                             float lr, float b1, float b2, float eps, float bc1, float bc2_sqrt) {
    if (*reinterpret_cast<volatile int*>(err_dev) != 0) return;        // sticky: an earlier exchange failed
    if (!peer_barrier(ps, 0, epoch, spin_ns, err_dev, err_host)) return;  // no update on a fault (see above)
    const float inv_world = 1.f / static_cast<float>(ps.world), lr_bc1 = lr / bc1;
    // the four sections share their alignment modulo 16 bytes: scalar head up to it, float4 body, scalar tail
    const size_t head = min(n, static_cast<size_t>((4 - ((reinterpret_cast<uintptr_t>(w) >> 2) & 3)) & 3));
    const size_t n4 = (n - head) / 4, tail0 = head + n4 * 4;
    const size_t tid = static_cast<size_t>(blockIdx.x) * 256 + threadIdx.x, nth = static_cast<size_t>(gridDim.x) * 256;
    for (size_t i = tid; i < n4; i += nth) {
        const size_t e = head + i * 4;
        float4 g[kPeerMax];
#pragma unroll
        for (int r = 0; r < kPeerMax; ++r)
            if (r < ps.world) g[r] = ld_sys_v4(ps.grads[r] + e);
        float4 s = g[0];
#pragma unroll
        for (int r = 1; r < kPeerMax; ++r)
            if (r < ps.world) { s.x += g[r].x; s.y += g[r].y; s.z += g[r].z; s.w += g[r].w; }
        float4 wv = *reinterpret_cast<float4*>(w + e), mv = *reinterpret_cast<float4*>(m + e), vv = *reinterpret_cast<float4*>(v + e);
        adam_update(wv.x, mv.x, vv.x, s.x * inv_world, lr_bc1, b1, b2, eps, bc2_sqrt);
        adam_update(wv.y, mv.y, vv.y, s.y * inv_world, lr_bc1, b1, b2, eps, bc2_sqrt);
        adam_update(wv.z, mv.z, vv.z, s.z * inv_world, lr_bc1, b1, b2, eps, bc2_sqrt);
        adam_update(wv.w, mv.w, vv.w, s.w * inv_world, lr_bc1, b1, b2, eps, bc2_sqrt);
        *reinterpret_cast<float4*>(w + e) = wv;
        *reinterpret_cast<float4*>(m + e) = mv;
        *reinterpret_cast<float4*>(v + e) = vv;
    }
    if (blockIdx.x == 0 && threadIdx.x < 8) {  // at most 3 + 3 unaligned elements
        const size_t e = threadIdx.x < 4 ? threadIdx.x : tail0 + (threadIdx.x - 4);
        if (threadIdx.x < 4 ? e < head : e < n) {
            float s = 0.f;
            for (int r = 0; r < ps.world; ++r) s += ld_sys_f32(ps.grads[r] + e);
            adam_update(w[e], m[e], v[e], s * inv_world, lr_bc1, b1, b2, eps, bc2_sqrt);
        }
    }
    peer_barrier(ps, 1, epoch, spin_ns, err_dev, err_host);
}

// ---------------------------------------------------------------------------------------------------------------------
// fp32 master weights (reference layout) -> the bf16 operand layouts of the tensor-core kernels
// ---------------------------------------------------------------------------------------------------------------------
enum RepackMode {
    RP_CONV9_FWD = 0,    // dst[(t*npad + co)*cin + ci]        = w[(co*cin + ci)*9 + t]
    RP_CONV9_DGRAD = 1,  // dst[(t*cin + ci)*cout + co]        = w[(co*cin + ci)*9 + 8 - t]
    RP_CONVT_FWD = 2,    // dst[(q*cout + co)*cin + ci]        = w[(ci*cout + co)*4 + q]
    RP_CONVT_DGRAD = 3,  // dst[(q*cin + ci)*cout + co]        = w[(ci*cout + co)*4 + q]
    RP_K4_FWD = 4,       // dst[((ky*4+kx)*cout + co)*cin + ci] = ky,kx < 3 ? w[(co*cin + ci)*9 + ky*3 + kx] : 0
    RP_T4_DGRAD = 5,     // dst[((ky*4+kx)*cin + ci)*cout + co] = same value
    RP_FIRST = 6,        // first_conv_umma.cuh operand: [tap][k chunk][cout 64][8] with hi/lo split
    RP_LAST = 7,         // last_conv_umma.cuh operand: row tap*3 + co of 128 rows x 64 channels
    RP_FIRST_DGRAD = 9,  // first_conv_umma.cuh operand for the data gradient of a Conv2d(64,3): the 3 -> 64 conv with
                         // weight(ci3 -> co64, tap) = w[ci3][co64][8 - tap]
    RP_LAST_DGRAD = 8    // the same operand for the data gradient of a Conv2d(3,64): row tap*3 + ci = w[co][ci][8 - tap]
};

__device__ __forceinline__ float repack_value(int mode, const float* w, size_t i, int cin, int cout, int npad,
                                              const float* aux = nullptr) {
    float val = 0.f;
    switch (mode) {
        case RP_CONV9_FWD: {
            const int ci = static_cast<int>(i % cin);
            const int co = static_cast<int>((i / cin) % npad);
            const int t = static_cast<int>(i / (static_cast<size_t>(cin) * npad));
            if (co < cout) val = w[(static_cast<size_t>(co) * cin + ci) * 9 + t];
            break;
        }
        case RP_CONV9_DGRAD: {
            const int co = static_cast<int>(i % cout);
            const int ci = static_cast<int>((i / cout) % cin);
            const int t = static_cast<int>(i / (static_cast<size_t>(cin) * cout));
            val = w[(static_cast<size_t>(co) * cin + ci) * 9 + 8 - t];
            break;
        }
        case RP_CONVT_FWD: {
            const int ci = static_cast<int>(i % cin);
            const int co = static_cast<int>((i / cin) % cout);
            const int q = static_cast<int>(i / (static_cast<size_t>(cin) * cout));
            val = w[(static_cast<size_t>(ci) * cout + co) * 4 + q];
            break;
        }
        case RP_CONVT_DGRAD: {
            const int co = static_cast<int>(i % cout);
            const int ci = static_cast<int>((i / cout) % cin);
            const int q = static_cast<int>(i / (static_cast<size_t>(cin) * cout));
            val = w[(static_cast<size_t>(ci) * cout + co) * 4 + q];
            break;
        }
        case RP_K4_FWD: {
            const int ci = static_cast<int>(i % cin);
            const int co = static_cast<int>((i / cin) % cout);
            const int t = static_cast<int>(i / (static_cast<size_t>(cin) * cout));
            const int ky = t >> 2, kx = t & 3;
            if (ky < 3 && kx < 3) val = w[(static_cast<size_t>(co) * cin + ci) * 9 + ky * 3 + kx];
            break;
        }
        case RP_T4_DGRAD: {
            const int co = static_cast<int>(i % cout);
            const int ci = static_cast<int>((i / cout) % cin);
            const int t = static_cast<int>(i / (static_cast<size_t>(cin) * cout));
            const int ky = t >> 2, kx = t & 3;
            if (ky < 3 && kx < 3) val = w[(static_cast<size_t>(co) * cin + ci) * 9 + ky * 3 + kx];
            break;
        }
        case RP_FIRST: {
            const int e = static_cast<int>(i & 7);
            const int co = static_cast<int>((i >> 3) & 63);
            const int chunk = static_cast<int>((i >> 9) & 1);
            const int tap = static_cast<int>(i >> 10);
            if (e == 6) {  // bias rows of the centre tap (pixel slot 6 = 1.0): hi in chunk 0, lo in chunk 1
                if (tap == 4 && aux) val = chunk == 0 ? aux[co] : aux[co] - __bfloat162float(__float2bfloat16_rn(aux[co]));
            } else if (chunk == 0) {
                if (e < 6) val = w[(co * 3 + (e % 3)) * 9 + tap];
            } else if (e < 3) {
                const float wf = w[(co * 3 + e) * 9 + tap];
                val = wf - __bfloat162float(__float2bfloat16_rn(wf));
            }
            break;
        }
        case RP_FIRST_DGRAD: {
            const int e = static_cast<int>(i & 7);
            const int co = static_cast<int>((i >> 3) & 63);
            const int chunk = static_cast<int>((i >> 9) & 1);
            const int tap = static_cast<int>(i >> 10);
            if (chunk == 0) {
                if (e < 6) val = w[((e % 3) * 64 + co) * 9 + 8 - tap];
            } else if (e < 3) {
                const float wf = w[(e * 64 + co) * 9 + 8 - tap];
                val = wf - __bfloat162float(__float2bfloat16_rn(wf));
            }
            break;
        }
        case RP_LAST: {
            const int ci = static_cast<int>(i & 63);
            const int row = static_cast<int>(i >> 6);
            const int tap = row / 3, co = row % 3;
            if (tap < 9) val = w[(static_cast<size_t>(co) * 64 + ci) * 9 + tap];
            break;
        }
        case RP_LAST_DGRAD: {
            const int co = static_cast<int>(i & 63);
            const int row = static_cast<int>(i >> 6);
            const int tap = row / 3, ci = row % 3;
            if (tap < 9) val = w[(static_cast<size_t>(co) * 3 + ci) * 9 + 8 - tap];
            break;
        }
    }
    return val;
}

// One launch re-packs a whole network: job j covers global elements [start_j, start_j + n_j). RP_COPY_F32 jobs copy
// fp32 biases into their padded device arrays.
constexpr int kRepackMaxJobs = 40;
enum { RP_COPY_F32 = 100 };
struct RepackJob {
    const float* w;
    const float* aux;  // RP_FIRST: the layer's bias (folded into the centre tap's K slot 6), else unused
    void* dst;
    int mode, cin, cout, npad;
    unsigned n, start;
};
struct RepackJobs {
    int count;
    unsigned total;
    RepackJob job[kRepackMaxJobs];
};

__global__ void __launch_bounds__(256) repack_multi_kernel(const __grid_constant__ RepackJobs jobs) {
    // programmatic dependent launch: the next kernel may start its set-up now; this one orders itself after its predecessor
    pdl_launch_dependents();
    pdl_wait();
    const unsigned gi = blockIdx.x * 256u + threadIdx.x;
    if (gi >= jobs.total) return;
    int j = 0;
    while (j + 1 < jobs.count && gi >= jobs.job[j + 1].start) ++j;
    const RepackJob& J = jobs.job[j];
    const size_t i = gi - J.start;
    if (J.mode == RP_COPY_F32) {
        static_cast<float*>(J.dst)[i] = J.w[i];
        return;
    }
    static_cast<__nv_bfloat16*>(J.dst)[i] = __float2bfloat16_rn(repack_value(J.mode, J.w, i, J.cin, J.cout, J.npad, J.aux));
}

}  // namespace cid
