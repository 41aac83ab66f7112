// K0: the first generator layer (3 -> 64, 3x3, pad 1, +ReLU) on tcgen05 tensor cores, with the
// reference's Pad / ToTensor / Normalize folded into the loader.
//
// Reference ops (file:line in /root/reference):
//   get_padding + transforms.Pad(fill=0)          backend/app.py:276-281, 383-385
//   ToTensor -> Normalize([.5]*3,[.5]*3)          backend/app.py:401-406
//   down1[0] = Conv2d(3,64,k3,p1) + ReLU          backend/app.py:43-44
//
// GEMM view per tap: M = 128 pixels (8 x 16 tile), N = 64, K = 16 of which 6 are used: every input
// value x is split x = hi + lo (two bf16) so the INPUT keeps ~16 mantissa bits on a bf16 tensor
// core; K row of a pixel = [hi_r hi_g hi_b lo_r lo_g lo_b 0 0 | 0 x8]. The layer is HBM-bound
// (3..12 B in, 128 B out per pixel), so the padded K is free.
//
// The A operand is the tile's 10 x 18 halo patch, one 16-byte K chunk per pixel, stored in the
// un-swizzled K-major UMMA layout (8-row core matrices: rows 16 B apart, 8-row groups SBO = 160 B
// apart = one patch row). The weights are split w = w_hi + w_lo too: the MMA's second K chunk re-reads
// the SAME pixel chunk (LBO = 0) against [wlo_r wlo_g wlo_b 0 ...], i.e. x_hi*w_lo. Tap (dy,dx) is
// the same patch read at start + (dy*10+dx)*16 B, so each input pixel is converted and written
// to shared memory once per tile (180 writes) instead of 9 times.
//
// The bias rides in the MMA too: K slot 6 of every pixel chunk holds 1.0 and the CENTRE tap's weight tile holds
// bias_hi in slot 6 of K chunk 0 and bias_lo in slot 6 of K chunk 1 (the centre tap reads the output pixel's own
// chunk, which always lies inside the frame), so the epilogue is ReLU on packed bf16 pairs: one cvt + one max per two
// outputs instead of bias loads, adds and a min/max/fma per output - the epilogue warps' issue slots were what kept
// this layer at 44 % of DRAM bandwidth.
//
// Warps (288 threads): 0..3 build patches (u8 mode: 3 byte loads + 3 LUT lookups + one 16-B
// store per pixel); warp 4 owns TMEM and issues 9 MMAs per tile; warps 5..8 run the ReLU
// epilogue (TMEM lane quarter = warp & 3) and store bf16 NHWC.
#pragma once
#include <cuda_bf16.h>

#include "conv_umma.cuh"

namespace cid {

constexpr int kFirstThreads = 288;
constexpr int kFirstStages = 8;
constexpr int kFirstCtasPerSm = 2;               // latency-bound roles: two resident CTAs per SM (<= 113 registers)
constexpr int kFirstPlane = 180 * 16;          // one K chunk of the 10x18 patch
constexpr int kFirstStageBytes = kFirstPlane;
constexpr int kFirstWBytes = 9 * 2 * 64 * 16;  // [tap][k chunk][cout][8 bf16]
constexpr int kFirstOutBufs = 2;               // staged output slabs ([128 px][64 ch] bf16, SW128) in flight
constexpr int kFirstSmem = 1024 + kFirstOutBufs * 16384 + kFirstWBytes + kFirstStages * kFirstStageBytes + 1024 + 32 + 256;
// training (FirstParams::mask): the ReLU mask tiles ([128 px][64 ch] bf16, two in flight) sit behind everything else
constexpr int kFirstMaskOff = (kFirstOutBufs * 16384 + kFirstWBytes + kFirstStages * kFirstStageBytes + 264 * 4 + 256 + 1023) / 1024 * 1024;
constexpr int kFirstSmemMask = 1024 + kFirstMaskOff + 2 * 16384;
static_assert(kFirstMaskOff + 1024 >= kFirstSmem - 1024 && kFirstCtasPerSm * (kFirstSmemMask + 1024) <= 227 * 1024, "first conv smem");

struct FirstParams {
    int B, H, W;          // network-input dims (after padding to a multiple of 4)
    const void* in;       // fp32 NCHW [B,3,H,W] in [-1,1]   or   u8 NHWC [B,in_h,in_w,3]
    int in_h, in_w;       // u8 mode: un-padded image size
    int pad_l, pad_t;     // u8 mode: where the image sits inside the padded frame
    const uint4* w_packed;  // kFirstWBytes of bf16 in the shared-memory layout above
    __nv_bfloat16* out;   // NHWC [B,H,W,64]
    int tiles_x, tiles_y, m_tiles;
    float slope;          // negative slope of the activation: 0 = ReLU (generator), 0.2 = the discriminator's LeakyReLU
    const __nv_bfloat16* mask;  // training (data gradient of the generator's LAST conv run as a 3 -> 64 conv, slope 1):
                                // NHWC-64 activation; the output is zeroed where it is not positive (ReLU backward)
};

// un-swizzled K-major descriptor: start[0,14) LBO[16,30) SBO[32,46) version[46,48)=1 layout[61,64)=0
__device__ __forceinline__ uint64_t umma_desc_nosw(uint32_t smem_addr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
    const uint32_t lo = ((smem_addr & 0x3FFFF) >> 4) | ((lbo_bytes >> 4) << 16);
    const uint32_t hi = (sbo_bytes >> 4) | (1u << 14);
    return (static_cast<uint64_t>(hi) << 32) | lo;
}

__device__ __forceinline__ uint32_t split_hi_lo(float x) {
    const __nv_bfloat16 h = __float2bfloat16_rn(x);
    const __nv_bfloat16 l = __float2bfloat16_rn(x - __bfloat162float(h));
    return static_cast<uint32_t>(__bfloat16_as_ushort(h)) | (static_cast<uint32_t>(__bfloat16_as_ushort(l)) << 16);
}

// tile coordinates of a linear tile index (x fastest, then y, then image), advanced incrementally
template <bool U8>
__global__ void __launch_bounds__(kFirstThreads, kFirstCtasPerSm)
first_conv_umma_kernel(const __grid_constant__ CUtensorMap tmO, const FirstParams p) {
    constexpr int N = 64;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* o_smem = smem;  // 1024-aligned staging slabs first
    uint8_t* w_smem = smem + kFirstOutBufs * 16384;
    uint8_t* a_smem = w_smem + kFirstWBytes;
    uint32_t* lut = reinterpret_cast<uint32_t*>(a_smem + kFirstStages * kFirstStageBytes);  // 256 x {hi | lo<<16}
    uint64_t* bars = reinterpret_cast<uint64_t*>(lut + 264);
    uint64_t* a_full = bars;
    uint64_t* a_empty = bars + kFirstStages;
    uint64_t* acc_full = bars + 2 * kFirstStages;
    uint64_t* acc_empty = acc_full + 2;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_empty + 2);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;

    // one-time: weights -> smem, zero K-chunk-1 planes, u8 LUT, barriers, TMEM
    for (int i = threadIdx.x; i < kFirstWBytes / 16; i += kFirstThreads)
        reinterpret_cast<uint4*>(w_smem)[i] = __ldg(p.w_packed + i);
    for (int i = threadIdx.x; i < kFirstStages * kFirstStageBytes / 16; i += kFirstThreads)
        reinterpret_cast<uint4*>(a_smem)[i] = make_uint4(0u, 0u, 0u, 0u);
    if (U8 && threadIdx.x < 256) {
        // ToTensor: u8 -> f32 / 255 ; Normalize: (v - 0.5) / 0.5, each op rounded (backend/app.py:401-406)
        const float v = __fdiv_rn(__fsub_rn(__fdiv_rn(static_cast<float>(threadIdx.x), 255.f), 0.5f), 0.5f);
        lut[threadIdx.x] = split_hi_lo(v);
        if (threadIdx.x == 0) {
            lut[256] = 0u;           // conv zero padding
            lut[257] = 0x0000BF80u;  // Pad(fill=0) before Normalize: (0-0.5)/0.5 = -1, exact in bf16 (hi=-1, lo=0)
        }
    }
    if (threadIdx.x == 0) {
        tma_prefetch_desc(&tmO);
        for (int i = 0; i < kFirstStages; ++i) {
            mbar_init(&a_full[i], 128);
            mbar_init(&a_empty[i], 1);
        }
        for (int i = 0; i < 2; ++i) {
            mbar_init(&acc_full[i], 1);
            mbar_init(&acc_empty[i], 4);
        }
        fence_mbar_init();
    }
    if (warp == 4) tmem_alloc(tmem_slot, 128);
    fence_proxy_async();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    pdl_launch_dependents();

    // Each CTA walks a CONTIGUOUS range of tiles (x fastest): neighbouring tiles share halo pixels
    // in L1/L2 and the tile coordinates advance by increments instead of divisions.
    const int per_cta = (p.m_tiles + gridDim.x - 1) / gridDim.x;
    const int mt_begin = min(static_cast<int>(blockIdx.x) * per_cta, p.m_tiles);
    const int mt_end = min(mt_begin + per_cta, p.m_tiles);

    if (warp < 4) {
        // ======================= patch producers =======================
        // The raw input values of the next kPF tiles are kept in flight in registers (the loop is
        // unrolled by kPF so every prefetch slot is a statically named register set): without this
        // each tile pays a full DRAM latency between its loads and its LUT lookups.
        constexpr int kPF = 4;
        constexpr uint32_t kZero = U8 ? 256u : 0u;   // conv zero padding (outside the padded frame)
        constexpr uint32_t kFrame = 257u;            // u8 mode: Pad(fill=0) BEFORE Normalize -> -1
        // this thread's (up to) two pixels of the 10 x 18 patch
        const int t1 = threadIdx.x + 128;
        const int py0 = threadIdx.x / 10, px0 = threadIdx.x - py0 * 10;
        const int py1 = t1 / 10, px1 = t1 - py1 * 10;
        const bool has1 = t1 < 180;
        auto load_px = [&](int b, int yy, int xx, bool live, uint32_t& l0, uint32_t& l1, uint32_t& l2) {
            l0 = l1 = l2 = kZero;
            if (live && yy >= 0 && yy < p.H && xx >= 0 && xx < p.W) {
                if constexpr (U8) {
                    const int oy = yy - p.pad_t, ox = xx - p.pad_l;
                    if (oy >= 0 && oy < p.in_h && ox >= 0 && ox < p.in_w) {
                        const uint8_t* s = static_cast<const uint8_t*>(p.in) +
                                           ((static_cast<size_t>(b) * p.in_h + oy) * p.in_w + ox) * 3;
                        l0 = __ldg(s);
                        l1 = __ldg(s + 1);
                        l2 = __ldg(s + 2);
                    } else {
                        l0 = l1 = l2 = kFrame;
                    }
                } else {
                    const float* s = static_cast<const float*>(p.in) + (static_cast<size_t>(b) * 3 * p.H + yy) * p.W + xx;
                    const size_t plane = static_cast<size_t>(p.H) * p.W;
                    l0 = __float_as_uint(__ldg(s));
                    l1 = __float_as_uint(__ldg(s + plane));
                    l2 = __float_as_uint(__ldg(s + 2 * plane));
                }
            }
        };
        pdl_wait();  // the input image may be the previous kernel's output (noiser, resize, an earlier pass)
        TileWalk pf(p.tiles_x, p.tiles_y, mt_begin);  // prefetch cursor, kPF tiles ahead of the store cursor
        auto load_raw = [&](uint32_t (&raw)[6]) {
            const int y0 = pf.ty * 16 - 1, x0 = pf.tx * 8 - 1;
            load_px(pf.b, y0 + py0, x0 + px0, true, raw[0], raw[1], raw[2]);
            load_px(pf.b, y0 + py1, x0 + px1, has1, raw[3], raw[4], raw[5]);
            pf.next(p.tiles_x, p.tiles_y);
        };
        uint32_t rq[kPF][6];
#pragma unroll
        for (int j = 0; j < kPF; ++j)
            if (mt_begin + j < mt_end) load_raw(rq[j]);
        uint32_t stage = 0, phase = 0;
        for (int base = mt_begin; base < mt_end; base += kPF) {
#pragma unroll
            for (int j = 0; j < kPF; ++j) {
                const int mt = base + j;
                if (mt >= mt_end) break;
                uint4 px[2];
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    uint32_t l0, l1, l2;
                    if constexpr (U8) {
                        l0 = lut[rq[j][3 * h + 0]];
                        l1 = lut[rq[j][3 * h + 1]];
                        l2 = lut[rq[j][3 * h + 2]];
                    } else {
                        l0 = split_hi_lo(__uint_as_float(rq[j][3 * h + 0]));
                        l1 = split_hi_lo(__uint_as_float(rq[j][3 * h + 1]));
                        l2 = split_hi_lo(__uint_as_float(rq[j][3 * h + 2]));
                    }
                    // K chunk 0 = [hi0 hi1 hi2 lo0 lo1 lo2 0 0]
                    // K chunk 0 = [hi0 hi1 hi2 lo0 lo1 lo2 1.0 0]: slot 6 multiplies the bias rows of the centre tap
                    px[h] = make_uint4((l0 & 0xffffu) | (l1 << 16), (l2 & 0xffffu) | (l0 & 0xffff0000u),
                                       (l1 >> 16) | (l2 & 0xffff0000u), 0x00003f80u);
                }
                if (mt + kPF < mt_end) load_raw(rq[j]);
                mbar_wait(&a_empty[stage], phase ^ 1);
                uint4* plane0 = reinterpret_cast<uint4*>(a_smem + stage * kFirstStageBytes);
                plane0[threadIdx.x] = px[0];
                if (has1) plane0[t1] = px[1];
                fence_proxy_async();  // generic-proxy smem writes -> visible to the tensor core's async proxy
                mbar_arrive(&a_full[stage]);
                if (++stage == kFirstStages) {
                    stage = 0;
                    phase ^= 1;
                }
            }
        }
    } else if (warp == 4) {
        // ======================= MMA issuer =======================
        constexpr uint32_t idesc = umma_idesc_bf16(128, N);
        const uint32_t a0 = smem_u32(a_smem), w0 = smem_u32(w_smem);
        uint32_t stage = 0, phase = 0, acc_it = 0;
        for (int mt = mt_begin; mt < mt_end; ++mt) {
            const uint32_t buf = acc_it & 1;
            mbar_wait(&acc_empty[buf], ((acc_it >> 1) & 1) ^ 1);
            mbar_wait(&a_full[stage], phase);
            tc_fence_after();
            if (elect_one()) {
                const uint32_t ab = a0 + stage * kFirstStageBytes;
#pragma unroll
                for (int dy = 0; dy < 3; ++dy)
#pragma unroll
                    for (int dx = 0; dx < 3; ++dx) {
                        const int tap = dy * 3 + dx;
                        umma_bf16_ss(tmem_base + buf * N, umma_desc_nosw(ab + (dy * 10 + dx) * 16, 0, 160),
                                     umma_desc_nosw(w0 + tap * 2048, 1024, 128), idesc, tap == 0 ? 0u : 1u);
                    }
                umma_commit(&a_empty[stage]);
                umma_commit(&acc_full[buf]);
            }
            __syncwarp();
            if (++stage == kFirstStages) {
                stage = 0;
                phase ^= 1;
            }
            ++acc_it;
        }
    } else {
        // ======================= epilogue: (Leaky)ReLU -> bf16 NHWC (the bias came through the MMA) =======================
        const int q = warp & 3;
        const int row = q * 32 + lane;
        const int yl = row >> 3, xl = row & 7;
        uint32_t acc_it = 0;
        const bool issuer = (warp == 5 && lane == 0);
        (void)yl; (void)xl;
        TileWalk tw(p.tiles_x, p.tiles_y, mt_begin);
        // ReLU mask (training): every thread copies its own pixel's 128 bytes of the NEXT tile into shared memory with
        // cp.async while it works on the current one (a synchronous load per tile cost ~1 us of exposed latency each)
        uint8_t* m_smem = smem + kFirstMaskOff;
        auto mask_fetch = [&](int tx, int ty, int b, uint32_t slot) {
            // coalesced: 8 lanes copy the 8 chunks of one pixel row, a warp instruction covers 4 whole 128-byte lines
            const int c = lane & 7;
#pragma unroll
            for (int it = 0; it < 8; ++it) {
                const int r = q * 32 + it * 4 + (lane >> 3);
                const int yy = ty * 16 + (r >> 3), xx = tx * 8 + (r & 7);
                if (b < p.B && yy < p.H && xx < p.W)
                    cp_async_16(smem_u32(m_smem + slot * 16384 + r * 128 + ((c ^ (r & 7)) << 4)),
                                p.mask + ((static_cast<size_t>(b) * p.H + yy) * p.W + xx) * 64 + c * 8);
            }
            cp_async_commit();
        };
        // ... and pulls its line of the tile kMaskAhead tiles ahead into L2 (one tile of cp.async distance does not
        // cover the DRAM latency; a third 16 KB buffer does not fit beside two resident CTAs)
        constexpr int kMaskAhead = 4;
        TileWalk twp(p.tiles_x, p.tiles_y, mt_begin);
        int pf = mt_begin;
        auto mask_prefetch_l2 = [&]() {
            if (pf < mt_end) {
                const int yy = twp.ty * 16 + yl, xx = twp.tx * 8 + xl;
                if (twp.b < p.B && yy < p.H && xx < p.W)
                    prefetch_l2(p.mask + ((static_cast<size_t>(twp.b) * p.H + yy) * p.W + xx) * 64);
                twp.next(p.tiles_x, p.tiles_y);
                ++pf;
            }
        };
        if (p.mask != nullptr) {
            for (int k = 0; k < kMaskAhead; ++k) mask_prefetch_l2();
            if (mt_begin < mt_end) mask_fetch(tw.tx, tw.ty, tw.b, 0);
        }
        for (int mt = mt_begin; mt < mt_end; ++mt) {
            const int tx = tw.tx, ty = tw.ty, b = tw.b;
            tw.next(p.tiles_x, p.tiles_y);
            if (p.mask != nullptr) {
                mask_prefetch_l2();
                if (mt + 1 < mt_end) mask_fetch(tw.tx, tw.ty, tw.b, (acc_it + 1) & 1);
                else cp_async_commit();
            }
            const uint32_t buf = acc_it & 1;
            mbar_wait(&acc_full[buf], (acc_it >> 1) & 1);
            tc_fence_after();
            if (p.mask != nullptr) {  // this tile's mask has landed (copied by the lanes of this warp)
                cp_async_wait<1>();
                __syncwarp();
            }
            const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + buf * N;
            const uint32_t ob = acc_it % kFirstOutBufs;
            // the TMA store that used this slab kFirstOutBufs tiles ago must have finished reading it
            if (issuer) bulk_wait_read<kFirstOutBufs - 1>();
            named_bar_sync(1, 128);
            uint8_t* rowp = o_smem + ob * 16384 + row * 128;
            const bool plain_relu = p.slope == 0.f && p.mask == nullptr;
            for_each_chunk<N / 32>(taddr, [&](int ch, uint32_t (&v)[32]) {
                uint32_t pk[16];
                if (plain_relu) {
                    // ReLU after rounding == rounding after ReLU: max on packed bf16 pairs
                    const __nv_bfloat162 zero2 = __float2bfloat162_rn(0.f);
#pragma unroll
                    for (int j = 0; j < 16; ++j) {
                        const uint32_t w2 = pack_bf16x2(__uint_as_float(v[2 * j]), __uint_as_float(v[2 * j + 1]));
                        const __nv_bfloat162 r2 = __hmax2(*reinterpret_cast<const __nv_bfloat162*>(&w2), zero2);
                        pk[j] = *reinterpret_cast<const uint32_t*>(&r2);
                    }
                } else {
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        // max(x,0) + slope * min(x,0): LeakyReLU, exactly ReLU for slope 0 (+0 + -0 = +0)
                        const float x0 = __uint_as_float(v[4 * j + 0]), x1 = __uint_as_float(v[4 * j + 1]);
                        const float x2 = __uint_as_float(v[4 * j + 2]), x3 = __uint_as_float(v[4 * j + 3]);
                        pk[2 * j + 0] = pack_bf16x2(fmaxf(x0, 0.f) + p.slope * fminf(x0, 0.f), fmaxf(x1, 0.f) + p.slope * fminf(x1, 0.f));
                        pk[2 * j + 1] = pack_bf16x2(fmaxf(x2, 0.f) + p.slope * fminf(x2, 0.f), fmaxf(x3, 0.f) + p.slope * fminf(x3, 0.f));
                    }
                }
                if (p.mask != nullptr) {
                    const int yy = ty * 16 + yl, xx = tx * 8 + xl;
                    if (b < p.B && yy < p.H && xx < p.W) {
                        const uint8_t* mrow = m_smem + (acc_it & 1) * 16384 + row * 128;
#pragma unroll
                        for (int i = 0; i < 4; ++i) {
                            const uint4 m = *reinterpret_cast<const uint4*>(mrow + (((ch * 4 + i) ^ (row & 7)) << 4));
                            const uint32_t w[4] = {m.x, m.y, m.z, m.w};
#pragma unroll
                            for (int e = 0; e < 4; ++e)  // one packed compare (0xffff per positive half) + one AND per pair
                                pk[i * 4 + e] &= __hgt2_mask(*reinterpret_cast<const __nv_bfloat162*>(&w[e]), __float2bfloat162_rn(0.f));
                        }
                    }
                }
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const int cj = ch * 4 + j;
                    st_shared_v4(rowp + ((cj ^ (row & 7)) << 4), pk[4 * j], pk[4 * j + 1], pk[4 * j + 2], pk[4 * j + 3]);
                }
            });
            tc_fence_before();
            fence_proxy_async();
            named_bar_sync(1, 128);
            if (issuer) {
                tma_store_4d(&tmO, o_smem + ob * 16384, 0, tx * 8, ty * 16, b);  // partial tiles are clipped
                bulk_commit();
            }
            if (lane == 0) mbar_arrive(&acc_empty[buf]);
            ++acc_it;
        }
        if (issuer) bulk_wait_read<0>();  // shared memory may go; the global writes complete with the grid
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 4) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 128);
    }
}

}  // namespace cid
