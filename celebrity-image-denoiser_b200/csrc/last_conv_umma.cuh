// K3: the last generator layer (64 -> 3, 3x3, pad 1) + tanh (+ de-normalise, clamp, truncating
// quantise and crop on the u8 route path), HBM-bound: 128 B in, 3..12 B out per pixel.
//
// Reference ops (file:line in /root/reference):
//   upconv1[2] = Conv2d(64,3,k3,p1)                  backend/app.py:77
//   torch.tanh                                       backend/app.py:103
//   (y*0.5+0.5).clamp(0,1)                           backend/app.py:435
//   ToPILImage: mul(255) + truncating byte cast      backend/app.py:472
//   crop to the original size                        backend/app.py:474-480
//
// With pixels on the GEMM M side (conv_umma_kernel<16,...>) this layer costs 36 MMAs of N=16 per
// 128-pixel tile, each paced by the 4 KB shared-memory read of its A operand (39 cycles measured,
// tools/micro/mma_rate.cu) = 1404 cycles per tile, 3x slower than HBM can feed it. Here the GEMM is
// transposed: A = the weights, one row per (tap, output channel) = 27 rows (padded to M = 128),
// B = the tile's 10 x 18 halo patch (180 pixels, N = 192), K = 64 channels:
//     P[(tap,c), q] = sum_k W[c,k,tap] * X[q,k]            4 MMAs (K = 16 each) per tile
// and the 3x3 window sum moves to the epilogue, where it is a 27-term gather per output pixel:
//     y[c, (y,x)] = b[c] + sum_{dy,dx} P[(dy*3+dx, c), (y+dy)*10 + (x+dx)]
// The patch arrives by one TMA box load per tile exactly as in conv_umma_kernel (128-byte swizzle,
// rows = patch pixels), which is already the canonical K-major UMMA layout for a B operand.
//
// Warps (384 threads): 0 = TMA producer, 1 = TMEM owner + MMA issuer, 4 and 8 = transposers, 2/3/5/6 and
// 7/9/10/11 = pixel warps: two (transposer, 4 pixel warps) groups take alternate tiles. The 27 used rows sit in TMEM lanes 0..26, i.e. in lane quarter 0, which only a warp
// with (warp & 3) == 0 may read: reading one quarter instead of all 128 lanes cuts the TMEM
// traffic 4x (TMEM reads are what bounded the four-quarter version of this kernel). The
// transposer moves P (TMEM) to shared memory for tile i+1 while the pixel warps gather tile i.
#pragma once
#include "conv_umma.cuh"

namespace cid {

constexpr int kLastThreads = 384;  // warp 0 TMA, 1 MMA, 4 and 8 transposers, {2,3,5,6} and {7,9,10,11} pixel warps
constexpr int kLastN = 192;                    // patch pixels per MMA (180 used)
constexpr int kLastPatchTx = 10 * 18 * 128;    // bytes one patch box delivers
constexpr int kLastStageBytes = kLastN * 128;  // stage pitch: the MMA reads 192 rows
constexpr int kLastWBytes = 128 * 128;         // [128 rows][64 ch] bf16
constexpr int kLastSP = 196;                   // fp32 pitch of one P row in shared memory
constexpr int kLastSBytes = 27 * kLastSP * 4;  // one P tile
constexpr int kLastMaxStages = 7;
constexpr int kLastGroups = 2;                 // tiles alternate between two (transposer, 4 pixel warps) groups

__host__ __device__ constexpr int last_conv_smem(int stages) {
    return 1024 + kLastWBytes + stages * kLastStageBytes + 2 * kLastGroups * kLastSBytes + 320;
}

// row of the packed weight matrix (= TMEM lane) that holds (tap, c)
__host__ __device__ constexpr int last_conv_row(int tap, int c) { return tap * 3 + c; }

struct LastParams {
    int B, H, W;  // activation dims (network-input size)
    int tiles_x, tiles_y, m_tiles;
    int stages;
    const float* bias;  // fp32 [3]
    void* out;          // fp32 NCHW [B,3,H,W], fp32 NHWC [B,H,W,3]   or   u8 NHWC [B,out_h,out_w,3]
    int crop_l, crop_t, out_h, out_w;  // u8 mode
};

template <int EPI>
__global__ void __launch_bounds__(kLastThreads, 1)
last_conv_umma_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmW,
                      const LastParams p) {
    static_assert(EPI == EPI_TANH_F32 || EPI == EPI_TANH_U8 || EPI == EPI_TANH_F32_NHWC || EPI == EPI_LINEAR_F32_NHWC,
                  "final layer epilogues only");
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* w_smem = smem;
    uint8_t* a_smem = smem + kLastWBytes;
    float* s_smem = reinterpret_cast<float*>(a_smem + p.stages * kLastStageBytes);  // [group][2][27][kLastSP]
    uint64_t* bars = reinterpret_cast<uint64_t*>(reinterpret_cast<uint8_t*>(s_smem) + 2 * kLastGroups * kLastSBytes);
    uint64_t* a_full = bars;
    uint64_t* a_empty = bars + kLastMaxStages;
    uint64_t* w_full = bars + 2 * kLastMaxStages;
    uint64_t* acc_full = w_full + 1;
    uint64_t* acc_empty = acc_full + 2;
    uint64_t* s_full = acc_empty + 2;   // [group][2]
    uint64_t* s_empty = s_full + 4;     // [group][2]
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(s_empty + 4);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;

    // rows 180..191 of every stage are read by the MMA but never written by TMA: keep them finite
    for (int i = threadIdx.x; i < p.stages * (kLastStageBytes - kLastPatchTx) / 16; i += kLastThreads) {
        const int per = (kLastStageBytes - kLastPatchTx) / 16;
        reinterpret_cast<uint4*>(a_smem + (i / per) * kLastStageBytes + kLastPatchTx)[i % per] = make_uint4(0u, 0u, 0u, 0u);
    }
    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmA);
        tma_prefetch_desc(&tmW);
        for (int i = 0; i < kLastMaxStages; ++i) {
            mbar_init(&a_full[i], 1);
            mbar_init(&a_empty[i], 1);
        }
        mbar_init(w_full, 1);
        for (int i = 0; i < 2; ++i) {
            mbar_init(&acc_full[i], 1);
            mbar_init(&acc_empty[i], 1);
        }
        for (int i = 0; i < 4; ++i) {
            mbar_init(&s_full[i], 1);
            mbar_init(&s_empty[i], 4);
        }
        fence_mbar_init();
    }
    if (warp == 1) tmem_alloc(tmem_slot, 512);
    fence_proxy_async();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    pdl_launch_dependents();
    // Each CTA walks a CONTIGUOUS range of tiles (x fastest): neighbouring tiles share halo rows in L2 and the pixel
    // warps advance their tile coordinates by increments instead of divisions.
    const int per_cta = (p.m_tiles + gridDim.x - 1) / gridDim.x;
    const int mt_begin = min(static_cast<int>(blockIdx.x) * per_cta, p.m_tiles);
    const int mt_end = min(mt_begin + per_cta, p.m_tiles);

    if (warp == 0) {
        // ======================= TMA producer =======================
        if (lane == 0) {
            mbar_arrive_expect_tx(w_full, kLastWBytes);
            tma_load_2d(w_smem, &tmW, w_full, 0, 0, kEvictLast);
            pdl_wait();  // the activations are the previous layer's output
            uint32_t stage = 0, phase = 0;
            TileWalk tw(p.tiles_x, p.tiles_y, mt_begin);
            for (int mt = mt_begin; mt < mt_end; ++mt) {
                const int tx0 = tw.tx * 8, ty0 = tw.ty * 16, tb = tw.b;
                tw.next(p.tiles_x, p.tiles_y);
                mbar_wait(&a_empty[stage], phase ^ 1);
                mbar_arrive_expect_tx(&a_full[stage], kLastPatchTx);
                tma_load_4d(a_smem + stage * kLastStageBytes, &tmA, &a_full[stage], 0, tx0 - 1, ty0 - 1, tb,
                            kEvictFirst);
                if (++stage == static_cast<uint32_t>(p.stages)) {
                    stage = 0;
                    phase ^= 1;
                }
            }
        }
    } else if (warp == 1) {
        // ======================= MMA issuer =======================
        constexpr uint32_t idesc = umma_idesc_bf16(128, kLastN);
        const uint32_t w_lo = umma_desc_lo(smem_u32(w_smem));
        const uint32_t a_lo0 = umma_desc_lo(smem_u32(a_smem));
        uint32_t stage = 0, phase = 0, acc_it = 0;
        mbar_wait(w_full, 0);
        for (int mt = mt_begin; mt < mt_end; ++mt) {
            const uint32_t buf = acc_it & 1;
            mbar_wait(&acc_empty[buf], ((acc_it >> 1) & 1) ^ 1);
            mbar_wait(&a_full[stage], phase);
            tc_fence_after();
            if (elect_one()) {
                const uint32_t b_lo = a_lo0 + stage * (kLastStageBytes >> 4);
#pragma unroll
                for (int k = 0; k < 4; ++k)
                    umma_bf16_ss(tmem_base + buf * 256, umma_desc_join<1024>(w_lo + 2 * k),
                                 umma_desc_join<1024>(b_lo + 2 * k), idesc, k == 0 ? 0u : 1u);
                umma_commit(&a_empty[stage]);
                umma_commit(&acc_full[buf]);
            }
            __syncwarp();
            if (++stage == static_cast<uint32_t>(p.stages)) {
                stage = 0;
                phase ^= 1;
            }
            ++acc_it;
        }
    } else if (warp == 4 || warp == 8) {
        // ======================= transposers: P rows (TMEM lanes 0..26) -> shared [27][kLastSP] ==========
        // Group g takes the tiles whose per-CTA index has parity g (= accumulator buffer g) and double-buffers
        // its own P tiles, so the TMEM drain of tile i+2 overlaps the gather of tile i.
        const uint32_t g = warp == 4 ? 0u : 1u;
        const bool row_used = lane < 27;
        uint32_t it = 0;
        for (int mt = mt_begin; mt < mt_end; ++mt, ++it) {
            if ((it & 1) != g) continue;
            const uint32_t k = it >> 1, sb = k & 1;
            mbar_wait(&s_empty[g * 2 + sb], ((k >> 1) & 1) ^ 1);
            mbar_wait(&acc_full[g], k & 1);
            tc_fence_after();
            float* srow = s_smem + (g * 2 + sb) * (27 * kLastSP) + lane * kLastSP;
            for_each_chunk<kLastN / 32>(tmem_base + g * 256, [&](int ch, uint32_t (&v)[32]) {
                if (row_used) {
#pragma unroll
                    for (int j = 0; j < 8; ++j)
                        st_shared_v4(srow + ch * 32 + 4 * j, v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
                }
            });
            tc_fence_before();
            __syncwarp();
            if (lane == 0) {
                mbar_arrive(&acc_empty[g]);          // TMEM buffer is free again
                mbar_arrive(&s_full[g * 2 + sb]);    // release: the stores above are visible to the waiters
            }
        }
    } else {
        // ======================= pixel warps: one output pixel per thread =======================
        // group 0 = warps 2,3,5,6; group 1 = warps 7,9,10,11
        const uint32_t g = warp < 7 ? 0u : 1u;
        const int pw = g == 0 ? (warp < 4 ? warp - 2 : warp - 3) : (warp == 7 ? 0 : warp - 8);  // 0..3
        const int et = pw * 32 + lane;                                                           // 0..127
        const int yl = et >> 3, xl = et & 7;
        const float b0 = __ldg(p.bias), b1 = __ldg(p.bias + 1), b2 = __ldg(p.bias + 2);
        uint32_t it = 0;
        TileWalk tw(p.tiles_x, p.tiles_y, mt_begin);
        for (int mt = mt_begin; mt < mt_end; ++mt, ++it) {
            struct { int x0, y0, b; } tc = {tw.tx * 8, tw.ty * 16, tw.b};
            tw.next(p.tiles_x, p.tiles_y);
            if ((it & 1) != g) continue;
            const uint32_t k = it >> 1, sb = k & 1;
            mbar_wait(&s_full[g * 2 + sb], (k >> 1) & 1);
            const float* S = s_smem + (g * 2 + sb) * (27 * kLastSP);
            // 27-term window gather
            float a0 = b0, a1 = b1, a2 = b2;
#pragma unroll
            for (int dy = 0; dy < 3; ++dy)
#pragma unroll
                for (int dx = 0; dx < 3; ++dx) {
                    const float* sp = S + ((dy * 3 + dx) * 3) * kLastSP + (yl + dy) * 10 + (xl + dx);
                    a0 += sp[0];
                    a1 += sp[kLastSP];
                    a2 += sp[2 * kLastSP];
                }
            __syncwarp();
            if (lane == 0) mbar_arrive(&s_empty[g * 2 + sb]);
            const float t[3] = {EPI == EPI_LINEAR_F32_NHWC ? a0 : tanh_fast(a0), EPI == EPI_LINEAR_F32_NHWC ? a1 : tanh_fast(a1),
                                EPI == EPI_LINEAR_F32_NHWC ? a2 : tanh_fast(a2)};
            const int y = tc.y0 + yl, x = tc.x0 + xl;
            if (y < p.H && x < p.W) {
                if constexpr (EPI == EPI_TANH_F32) {
                    float* out = static_cast<float*>(p.out);
#pragma unroll
                    for (int c = 0; c < 3; ++c) out[((static_cast<size_t>(tc.b) * 3 + c) * p.H + y) * p.W + x] = t[c];
                } else if constexpr (EPI == EPI_TANH_F32_NHWC || EPI == EPI_LINEAR_F32_NHWC) {
                    // Keras `predict` surface (backend/app.py:458): fp32 NHWC
                    float* out = static_cast<float*>(p.out) + ((static_cast<size_t>(tc.b) * p.H + y) * p.W + x) * 3;
                    out[0] = t[0];
                    out[1] = t[1];
                    out[2] = t[2];
                } else {
                    // (y*0.5+0.5).clamp(0,1) then ToPILImage: mul(255) and a TRUNCATING byte cast
                    const int oy = y - p.crop_t, ox = x - p.crop_l;
                    if (oy >= 0 && oy < p.out_h && ox >= 0 && ox < p.out_w) {
                        uint8_t* out = static_cast<uint8_t*>(p.out) +
                                       ((static_cast<size_t>(tc.b) * p.out_h + oy) * p.out_w + ox) * 3;
#pragma unroll
                        for (int c = 0; c < 3; ++c) {
                            float u = __fmaf_rn(t[c], 0.5f, 0.5f);
                            u = fminf(fmaxf(u, 0.f), 1.f);
                            out[c] = static_cast<uint8_t>(__fmul_rn(u, 255.f));
                        }
                    }
                }
            }
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 512);
    }
}

}  // namespace cid
