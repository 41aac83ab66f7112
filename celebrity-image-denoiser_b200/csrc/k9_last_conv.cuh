// Last layer of the ESRGAN / SRGAN generators: Conv2d(64 -> 3, 9x9, pad 4) (+ tanh for SRGAN) with the
// route's post-processing (clamp / de-normalise / truncating x255 / crop) in the epilogue.
//
// Reference ops (file:line in /root/reference):
//   ESRGANGenerator.final = Conv2d(64,3,9,1,4), no activation        backend/app.py:212, 214-217
//   postprocess_tensor: clamp(0,1) -> ToPILImage (truncating x255)   backend/app.py:251-254
//   SRGANGenerator.final = Conv2d(64,3,k9,p4) + tanh                 backend/app.py:177, 183-184
//   (y*0.5+0.5).clamp(0,1), ToPILImage                               backend/app.py:435, 472
//
// With 3 output channels a pixels-on-M GEMM needs 81 taps x 4 K steps = 324 MMAs of the minimum width
// N = 16 per 128 pixels, each paced by its 4 KB A-operand read. Here the 9 HORIZONTAL taps move into the
// N dimension instead: for one vertical tap dy
//     D[(yy,xx), (dx,c)] += sum_k X[(y0-4+yy+dy, x0-4+xx), k] * W[c, k, dy, dx]       N = 27 (padded to 32)
// over the tile's 16-column window xx = 0..15, accumulated over dy = 0..8 inside the MMA (the dy shift is a
// descriptor offset of dy window rows), so a tile of 8 x 16 outputs costs 2 x 9 x 4 = 72 MMAs. The horizontal
// window sum is a 27-term gather in the epilogue: y[c,(yy,x)] = b[c] + sum_dx D[(yy, x+dx), (dx,c)].
// The A operand is ONE TMA box {64 ch, 16 px, 24 rows} per tile (128-byte swizzle; out-of-image = zero
// padding), whose 8-pixel groups are already the canonical K-major layout with SBO = 1024 B.
//
// Warps (192 threads): 0 = TMA producer, 1 = TMEM owner + MMA issuer, 2..5 = epilogue (TMEM lane quarter =
// warp & 3): D -> shared [256 window px][28] fp32, barrier, one output pixel per thread gathers its 27 terms.
#pragma once
#include "conv_umma.cuh"

namespace cid {

constexpr int kL9Threads = 192;
constexpr int kL9StageBytes = 24 * 16 * 128;   // one window: 24 rows x 16 px x 64 ch bf16
constexpr int kL9WTile = 32 * 128;             // one dy: 32 rows (dx*3+c, 27 used) x 64 ch
constexpr int kL9WBytes = 9 * kL9WTile;
constexpr int kL9SP = 28;                      // fp32 pitch of one window pixel's 27 partial sums
constexpr int kL9SBytes = 256 * kL9SP * 4;
constexpr int kL9MaxStages = 3;

__host__ __device__ constexpr int k9_last_smem(int stages) {
    return 1024 + kL9WBytes + stages * kL9StageBytes + kL9SBytes + 256;
}

enum : int { L9_NONE_F32 = 0, L9_CLAMP_U8 = 1, L9_TANH_F32 = 2, L9_TANH_U8 = 3 };

struct L9Params {
    int B, H, W;
    int tiles_x, tiles_y, m_tiles;
    int stages;
    const float* bias;  // fp32 [3]
    void* out;          // fp32 NCHW [B,3,H,W]   or   u8 NHWC [B,out_h,out_w,3]
    int crop_l, crop_t, out_h, out_w;  // u8 modes: window of the network output that is kept
};

template <int MODE>
__global__ void __launch_bounds__(kL9Threads, 1)
k9_last_conv_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmW, const L9Params p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* w_smem = smem;
    uint8_t* a_smem = smem + kL9WBytes;
    float* s_smem = reinterpret_cast<float*>(a_smem + p.stages * kL9StageBytes);
    uint64_t* bars = reinterpret_cast<uint64_t*>(reinterpret_cast<uint8_t*>(s_smem) + kL9SBytes);
    uint64_t* a_full = bars;
    uint64_t* a_empty = bars + kL9MaxStages;
    uint64_t* w_full = bars + 2 * kL9MaxStages;
    uint64_t* acc_full = w_full + 1;
    uint64_t* acc_empty = acc_full + 2;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_empty + 2);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmA);
        tma_prefetch_desc(&tmW);
        for (int i = 0; i < kL9MaxStages; ++i) {
            mbar_init(&a_full[i], 1);
            mbar_init(&a_empty[i], 1);
        }
        mbar_init(w_full, 1);
        for (int i = 0; i < 2; ++i) {
            mbar_init(&acc_full[i], 1);
            mbar_init(&acc_empty[i], 4);
        }
        fence_mbar_init();
    }
    if (warp == 1) tmem_alloc(tmem_slot, 128);  // 2 buffers x (2 halves x 32 columns)
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    pdl_launch_dependents();

    if (warp == 0) {
        // ======================= TMA producer =======================
        if (lane == 0) {
            mbar_arrive_expect_tx(w_full, kL9WBytes);
            for (int dy = 0; dy < 9; ++dy) tma_load_2d(w_smem + dy * kL9WTile, &tmW, w_full, 0, dy * 32, kEvictLast);
            pdl_wait();  // the activations are the previous layer's output
            uint32_t stage = 0, phase = 0;
            for (int mt = blockIdx.x; mt < p.m_tiles; mt += gridDim.x) {
                const TileCoord tc = decode_tile(p.tiles_x, p.tiles_y, mt);
                mbar_wait(&a_empty[stage], phase ^ 1);
                mbar_arrive_expect_tx(&a_full[stage], kL9StageBytes);
                tma_load_4d(a_smem + stage * kL9StageBytes, &tmA, &a_full[stage], 0, tc.x0 - 4, tc.y0 - 4, tc.b,
                            kEvictNormal);
                if (++stage == static_cast<uint32_t>(p.stages)) {
                    stage = 0;
                    phase ^= 1;
                }
            }
        }
    } else if (warp == 1) {
        // ======================= MMA issuer =======================
        constexpr uint32_t idesc = umma_idesc_bf16(128, 32);
        const uint32_t w_lo0 = umma_desc_lo(smem_u32(w_smem));
        const uint32_t a_lo0 = umma_desc_lo(smem_u32(a_smem));
        uint32_t stage = 0, phase = 0, acc_it = 0;
        mbar_wait(w_full, 0);
        for (int mt = blockIdx.x; mt < p.m_tiles; mt += gridDim.x) {
            const uint32_t buf = acc_it & 1;
            mbar_wait(&acc_empty[buf], ((acc_it >> 1) & 1) ^ 1);
            mbar_wait(&a_full[stage], phase);
            tc_fence_after();
            if (elect_one()) {
                const uint32_t a_st = a_lo0 + stage * (kL9StageBytes >> 4);
#pragma unroll
                for (int h = 0; h < 2; ++h)
#pragma unroll
                    for (int dy = 0; dy < 9; ++dy) {
                        // output rows 8h..8h+7 of the tile <- window rows 8h+dy..; one window row = 2048 B
                        const uint32_t a_lo = a_st + static_cast<uint32_t>((dy + 8 * h) * (2048 >> 4));
                        const uint32_t w_lo = w_lo0 + static_cast<uint32_t>(dy * (kL9WTile >> 4));
#pragma unroll
                        for (int k = 0; k < 4; ++k)
                            umma_bf16_ss(tmem_base + buf * 64 + h * 32, umma_desc_join<1024>(a_lo + 2 * k),
                                         umma_desc_join<1024>(w_lo + 2 * k), idesc, (dy | k) == 0 ? 0u : 1u);
                    }
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
    } else {
        // ======================= epilogue =======================
        const int q = warp & 3;
        const int et = (warp - 2) * 32 + lane;  // output pixel of this thread, 0..127
        const int yl = et >> 3, xl = et & 7;
        const float b0 = __ldg(p.bias), b1 = __ldg(p.bias + 1), b2 = __ldg(p.bias + 2);
        uint32_t acc_it = 0;
        for (int mt = blockIdx.x; mt < p.m_tiles; mt += gridDim.x) {
            const TileCoord tc = decode_tile(p.tiles_x, p.tiles_y, mt);
            const uint32_t buf = acc_it & 1;
            mbar_wait(&acc_full[buf], (acc_it >> 1) & 1);
            tc_fence_after();
            // D (TMEM lane = window pixel within the half) -> shared [256][28]
            uint32_t v0[32], v1[32];
            const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + buf * 64;
            tmem_ld_32x32(taddr, v0);
            tmem_ld_32x32(taddr + 32, v1);
            tmem_ld_wait();
            float* r0 = s_smem + (q * 32 + lane) * kL9SP;
            float* r1 = r0 + 128 * kL9SP;
#pragma unroll
            for (int j = 0; j < 7; ++j) {
                st_shared_v4(r0 + 4 * j, v0[4 * j], v0[4 * j + 1], v0[4 * j + 2], v0[4 * j + 3]);
                st_shared_v4(r1 + 4 * j, v1[4 * j], v1[4 * j + 1], v1[4 * j + 2], v1[4 * j + 3]);
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&acc_empty[buf]);
            named_bar_sync(1, 128);
            float a0 = b0, a1 = b1, a2 = b2;
            const float* sp = s_smem + (yl * 16 + xl) * kL9SP;
#pragma unroll
            for (int dx = 0; dx < 9; ++dx) {
                a0 += sp[dx * kL9SP + dx * 3 + 0];
                a1 += sp[dx * kL9SP + dx * 3 + 1];
                a2 += sp[dx * kL9SP + dx * 3 + 2];
            }
            named_bar_sync(1, 128);  // every gather is done before the next tile overwrites the buffer
            float t[3] = {a0, a1, a2};
            if constexpr (MODE == L9_TANH_F32 || MODE == L9_TANH_U8) {
                t[0] = tanh_fast(a0);
                t[1] = tanh_fast(a1);
                t[2] = tanh_fast(a2);
            }
            const int y = tc.y0 + yl, x = tc.x0 + xl;
            if (y < p.H && x < p.W) {
                if constexpr (MODE == L9_NONE_F32 || MODE == L9_TANH_F32) {
                    float* out = static_cast<float*>(p.out);
#pragma unroll
                    for (int c = 0; c < 3; ++c) out[((static_cast<size_t>(tc.b) * 3 + c) * p.H + y) * p.W + x] = t[c];
                } else {
                    const int oy = y - p.crop_t, ox = x - p.crop_l;
                    if (oy >= 0 && oy < p.out_h && ox >= 0 && ox < p.out_w) {
                        uint8_t* out = static_cast<uint8_t*>(p.out) +
                                       ((static_cast<size_t>(tc.b) * p.out_h + oy) * p.out_w + ox) * 3;
#pragma unroll
                        for (int c = 0; c < 3; ++c) {
                            float u = MODE == L9_TANH_U8 ? __fmaf_rn(t[c], 0.5f, 0.5f) : t[c];
                            u = fminf(fmaxf(u, 0.f), 1.f);
                            out[c] = static_cast<uint8_t>(__fmul_rn(u, 255.f));  // ToPILImage: truncating cast
                        }
                    }
                }
            }
            ++acc_it;
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 128);
    }
}

}  // namespace cid
