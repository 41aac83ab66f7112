// G0: first layer of the Keras cGAN generator — Conv2D(3 -> 64, 4x4, stride 2, 'same') + LeakyReLU(0.2)
// on tcgen05 tensor cores, with the route's Pad / "/255" / "(x-0.5)/0.5" folded into the loader.
//
// Reference ops (file:line in /root/reference):
//   layers.Conv2D(64,(4,4),strides=(2,2),padding='same') + LeakyReLU(0.2)
//                                                   backend/trainingcode/cgan_code/training5barrev.py:27-28
//   pad_tf (Pad fill=0), /255, (x-0.5)/0.5          backend/app.py:383-385, 441-444
//
// Output tile = 8 x 16 pixels of the half-resolution output; it reads a 18 x 34 window of the input
// starting at (2*y0-1, 2*x0-1). The window is stored in shared memory as FOUR phase patches
// (phy,phx) of 9 x 17 pixels (input pixel (2j-1+phy, 2i-1+phx) -> patch (phy,phx) entry (j,i)), one
// 16-byte K chunk per pixel = [hi_r hi_g hi_b lo_r lo_g lo_b 0 0] (x = hi + lo, two bf16, so the
// input keeps ~16 mantissa bits), un-swizzled K-major UMMA layout: rows 16 B apart, 8-row groups
// SBO = 144 B apart (one patch row). The weights are split w = w_hi + w_lo as well: the MMA's second K chunk
// re-reads the SAME pixel chunk (LBO = 0) against [wlo_r wlo_g wlo_b 0 ...], so the product keeps
// x_hi*w_hi + x_lo*w_hi + x_hi*w_lo (~16 mantissa bits on both operands; only x_lo*w_lo is dropped). Tap
// (ky,kx) = (2dy+phy, 2dx+phx) is phase patch (phy,phx) read at + (dy*9+dx)*16 B: 16 MMAs
// (M128 x N64 x K16) per tile and every input pixel is converted and written once per tile.
// The layer is HBM-bound (3..12 B in per input pixel, 128 B out per output pixel).
//
// Warps (288 threads): 0..3 build patches, warp 4 owns TMEM and issues the MMAs, warps 5..8 run the
// bias + LeakyReLU epilogue (TMEM lane quarter = warp & 3) and store bf16 NHWC through TMA.
#pragma once
#include "first_conv_umma.cuh"

namespace cid {

constexpr int kG0Threads = 288;
constexpr int kG0Stages = 4;
constexpr int kG0CtasPerSm = 2;               // latency-bound roles: two resident CTAs per SM
constexpr int kG0Phase = 153 * 16;            // one 9 x 17 phase patch
constexpr int kG0Plane = 4 * kG0Phase;        // 9792 B: the four phase patches (K chunk 0)
constexpr int kG0StageBytes = kG0Plane;
constexpr int kG0WBytes = 16 * 2 * 64 * 16;   // [tap][k chunk][cout][8 bf16]
constexpr int kG0OutBufs = 2;
constexpr int kG0Entries = 18 * 34;           // input pixels per tile window
constexpr int kG0PerThread = (kG0Entries + 127) / 128;  // 5
constexpr int kG0Smem = 1024 + kG0OutBufs * 16384 + kG0WBytes + kG0Stages * kG0StageBytes + 1088 + 256;

struct G0Params {
    int B, H, W;          // network-input dims (padded frame, multiples of 4)
    const void* in;       // fp32 NHWC [B,H,W,3] in [-1,1]   or   u8 NHWC [B,in_h,in_w,3]
    int in_h, in_w;       // u8 mode: un-padded image size
    int pad_l, pad_t;     // u8 mode: where the image sits inside the padded frame
    const uint4* w_packed;  // kG0WBytes of bf16 in the shared-memory layout above
    const float* bias;    // fp32 [64]
    float slope;          // LeakyReLU negative slope
    int tiles_x, tiles_y, m_tiles;  // tiles of the (H/2) x (W/2) output
};

template <bool U8>
__global__ void __launch_bounds__(kG0Threads, kG0CtasPerSm)
cgan_first_conv_kernel(const __grid_constant__ CUtensorMap tmO, const G0Params p) {
    constexpr int N = 64;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* o_smem = smem;
    uint8_t* w_smem = smem + kG0OutBufs * 16384;
    uint8_t* a_smem = w_smem + kG0WBytes;
    uint32_t* lut = reinterpret_cast<uint32_t*>(a_smem + kG0Stages * kG0StageBytes);  // 256 (+2) x {hi | lo<<16}
    uint64_t* bars = reinterpret_cast<uint64_t*>(lut + 264);
    uint64_t* a_full = bars;
    uint64_t* a_empty = bars + kG0Stages;
    uint64_t* acc_full = bars + 2 * kG0Stages;
    uint64_t* acc_empty = acc_full + 2;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_empty + 2);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;

    for (int i = threadIdx.x; i < kG0WBytes / 16; i += kG0Threads)
        reinterpret_cast<uint4*>(w_smem)[i] = __ldg(p.w_packed + i);
    for (int i = threadIdx.x; i < kG0Stages * kG0StageBytes / 16; i += kG0Threads)
        reinterpret_cast<uint4*>(a_smem)[i] = make_uint4(0u, 0u, 0u, 0u);
    if (U8 && threadIdx.x < 256) {
        // x/255 then (x-0.5)/0.5, each op rounded in fp32 as numpy does (backend/app.py:442-443)
        const float v = __fdiv_rn(__fsub_rn(__fdiv_rn(static_cast<float>(threadIdx.x), 255.f), 0.5f), 0.5f);
        lut[threadIdx.x] = split_hi_lo(v);
        if (threadIdx.x == 0) {
            lut[256] = 0u;           // conv zero padding ('same')
            lut[257] = 0x0000BF80u;  // Pad(fill=0) before the normalise: (0-0.5)/0.5 = -1
        }
    }
    if (threadIdx.x == 0) {
        tma_prefetch_desc(&tmO);
        for (int i = 0; i < kG0Stages; ++i) {
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

    const int per_cta = (p.m_tiles + gridDim.x - 1) / gridDim.x;
    const int mt_begin = min(static_cast<int>(blockIdx.x) * per_cta, p.m_tiles);
    const int mt_end = min(mt_begin + per_cta, p.m_tiles);

    if (warp < 4) {
        // ======================= patch producers =======================
        constexpr uint32_t kZero = U8 ? 256u : 0u;
        constexpr uint32_t kFrame = 257u;
        // this thread's (up to) 5 window pixels e = t + 128*i: window row wy = e/18, column wx = e%18
        // (consecutive threads read consecutive input pixels); its shared-memory slot is entry
        // (phase (wy&1, wx&1), row wy>>1, column wx>>1).
        int wy[kG0PerThread], wx[kG0PerThread], slot[kG0PerThread];
#pragma unroll
        for (int i = 0; i < kG0PerThread; ++i) {
            const int e = threadIdx.x + 128 * i;
            wy[i] = e / 18;
            wx[i] = e - wy[i] * 18;
            slot[i] = ((wy[i] & 1) * 2 + (wx[i] & 1)) * 153 + (wy[i] >> 1) * 9 + (wx[i] >> 1);
        }
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
                    const float* s = static_cast<const float*>(p.in) + ((static_cast<size_t>(b) * p.H + yy) * p.W + xx) * 3;
                    l0 = __float_as_uint(__ldg(s));
                    l1 = __float_as_uint(__ldg(s + 1));
                    l2 = __float_as_uint(__ldg(s + 2));
                }
            }
        };
        pdl_wait();  // the input image may be the previous kernel's output (noiser, resize, an earlier pass)
        TileWalk pf(p.tiles_x, p.tiles_y, mt_begin);
        auto load_raw = [&](uint32_t (&raw)[3 * kG0PerThread]) {
            const int y0 = 2 * pf.ty * 16 - 1, x0 = 2 * pf.tx * 8 - 1;
#pragma unroll
            for (int i = 0; i < kG0PerThread; ++i)
                load_px(pf.b, y0 + wy[i], x0 + wx[i], threadIdx.x + 128 * i < kG0Entries, raw[3 * i], raw[3 * i + 1],
                        raw[3 * i + 2]);
            pf.next(p.tiles_x, p.tiles_y);
        };
        constexpr int kPF = 2;
        uint32_t rq[kPF][3 * kG0PerThread];
#pragma unroll
        for (int j = 0; j < kPF; ++j)
            if (mt_begin + j < mt_end) load_raw(rq[j]);
        uint32_t stage = 0, phase = 0;
        for (int base = mt_begin; base < mt_end; base += kPF) {
#pragma unroll
            for (int j = 0; j < kPF; ++j) {
                const int mt = base + j;
                if (mt >= mt_end) break;
                uint4 px[kG0PerThread];
#pragma unroll
                for (int i = 0; i < kG0PerThread; ++i) {
                    uint32_t l0, l1, l2;
                    if constexpr (U8) {
                        l0 = lut[rq[j][3 * i + 0]];
                        l1 = lut[rq[j][3 * i + 1]];
                        l2 = lut[rq[j][3 * i + 2]];
                    } else {
                        l0 = split_hi_lo(__uint_as_float(rq[j][3 * i + 0]));
                        l1 = split_hi_lo(__uint_as_float(rq[j][3 * i + 1]));
                        l2 = split_hi_lo(__uint_as_float(rq[j][3 * i + 2]));
                    }
                    px[i] = make_uint4((l0 & 0xffffu) | (l1 << 16), (l2 & 0xffffu) | (l0 & 0xffff0000u),
                                       (l1 >> 16) | (l2 & 0xffff0000u), 0u);
                }
                if (mt + kPF < mt_end) load_raw(rq[j]);
                mbar_wait(&a_empty[stage], phase ^ 1);
                uint4* plane0 = reinterpret_cast<uint4*>(a_smem + stage * kG0StageBytes);
#pragma unroll
                for (int i = 0; i < kG0PerThread; ++i)
                    if (threadIdx.x + 128 * i < kG0Entries) plane0[slot[i]] = px[i];
                fence_proxy_async();
                mbar_arrive(&a_full[stage]);
                if (++stage == kG0Stages) {
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
                const uint32_t ab = a0 + stage * kG0StageBytes;
#pragma unroll
                for (int ph = 0; ph < 4; ++ph)
#pragma unroll
                    for (int dy = 0; dy < 2; ++dy)
#pragma unroll
                        for (int dx = 0; dx < 2; ++dx) {
                            const int tap = (2 * dy + (ph >> 1)) * 4 + 2 * dx + (ph & 1);
                            umma_bf16_ss(tmem_base + buf * N,
                                         umma_desc_nosw(ab + ph * kG0Phase + (dy * 9 + dx) * 16, 0, 144),
                                         umma_desc_nosw(w0 + tap * 2048, 1024, 128), idesc,
                                         (ph | dy | dx) == 0 ? 0u : 1u);
                        }
                umma_commit(&a_empty[stage]);
                umma_commit(&acc_full[buf]);
            }
            __syncwarp();
            if (++stage == kG0Stages) {
                stage = 0;
                phase ^= 1;
            }
            ++acc_it;
        }
    } else {
        // ======================= epilogue: bias + LeakyReLU -> bf16 NHWC =======================
        const int q = warp & 3;
        const int row = q * 32 + lane;
        uint32_t acc_it = 0;
        const bool issuer = (warp == 5 && lane == 0);
        const float slope = p.slope;
        TileWalk tw(p.tiles_x, p.tiles_y, mt_begin);
        for (int mt = mt_begin; mt < mt_end; ++mt) {
            const int tx = tw.tx, ty = tw.ty, b = tw.b;
            tw.next(p.tiles_x, p.tiles_y);
            const uint32_t buf = acc_it & 1;
            mbar_wait(&acc_full[buf], (acc_it >> 1) & 1);
            tc_fence_after();
            const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + buf * N;
            const uint32_t ob = acc_it % kG0OutBufs;
            if (issuer) bulk_wait_read<kG0OutBufs - 1>();
            named_bar_sync(1, 128);
            uint8_t* rowp = o_smem + ob * 16384 + row * 128;
            for_each_chunk<N / 32>(taddr, [&](int ch, uint32_t (&v)[32]) {
                const float4* b4 = reinterpret_cast<const float4*>(p.bias + ch * 32);
                uint32_t pk[16];
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    const float4 bb = __ldg(b4 + j);
                    float f0 = __uint_as_float(v[4 * j + 0]) + bb.x;
                    float f1 = __uint_as_float(v[4 * j + 1]) + bb.y;
                    float f2 = __uint_as_float(v[4 * j + 2]) + bb.z;
                    float f3 = __uint_as_float(v[4 * j + 3]) + bb.w;
                    f0 = f0 > 0.f ? f0 : f0 * slope;
                    f1 = f1 > 0.f ? f1 : f1 * slope;
                    f2 = f2 > 0.f ? f2 : f2 * slope;
                    f3 = f3 > 0.f ? f3 : f3 * slope;
                    pk[2 * j + 0] = pack_bf16x2(f0, f1);
                    pk[2 * j + 1] = pack_bf16x2(f2, f3);
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
                tma_store_4d(&tmO, o_smem + ob * 16384, 0, tx * 8, ty * 16, b);
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
