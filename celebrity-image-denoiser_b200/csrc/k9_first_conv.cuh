// First layer of the ESRGAN / SRGAN generators: Conv2d(3 -> 64, 9x9, pad 4) + PReLU on tcgen05 tensor cores,
// with ToTensor (and, for the SRGAN route, Pad + Normalize) folded into the loader.
//
// Reference ops (file:line in /root/reference):
//   ESRGANGenerator.initial = Conv2d(3,64,9,1,4) + PReLU()          backend/app.py:205-208
//   SRGANGenerator.initial  = Conv2d(3,64,k9,p4) + PReLU()          backend/app.py:152-155
//   preprocess_image: ToTensor only (ESRGAN)                        backend/app.py:243-249
//   Pad(fill=0) -> ToTensor -> Normalize(.5,.5) (SRGAN)             backend/app.py:383-385, 401-406
//
// Output tile = 8 x 16 pixels; its 16 x 24 input window (x0-4.., y0-4..) is stored once per tile as one 16-byte
// K chunk per pixel = [hi_r hi_g hi_b lo_r lo_g lo_b 0 0] (x = hi + lo, two bf16), un-swizzled K-major UMMA
// layout: rows 16 B apart, 8-row groups SBO = 256 B apart (one window row). Tap (dy,dx) is the window read at
// +(dy*16+dx)*16 B, and two horizontally adjacent taps share one K=16 MMA: the second K chunk is the NEXT pixel
// (LBO = 16 B) against the next tap's weights, so 81 taps cost 45 MMAs (M128 x N64 x K16) per tile.
//
// Warps (288 threads): 0..3 build windows, warp 4 owns TMEM and issues the MMAs, warps 5..8 run the
// bias + PReLU epilogue (TMEM lane quarter = warp & 3) and store bf16 NHWC through TMA.
#pragma once
#include "first_conv_umma.cuh"

namespace cid {

constexpr int kK9Threads = 288;
constexpr int kK9Stages = 6;
constexpr int kK9CtasPerSm = 1;               // 92 KB of resident weights per CTA
constexpr int kK9Row = 16 * 16;               // one window row: 16 pixels x 16 B
constexpr int kK9Window = 24 * kK9Row;        // 6144 B: 24 rows x 16 px
constexpr int kK9StageBytes = kK9Window;
constexpr int kK9WBytes = 45 * 2 * 64 * 16;   // [dy 9][dx pair 5][k chunk 2][cout 64][8 bf16]
constexpr int kK9OutBufs = 2;
constexpr int kK9Entries = 16 * 24;           // input pixels per tile window
constexpr int kK9PerThread = kK9Entries / 128;  // 3
constexpr int kK9Smem = 1024 + kK9OutBufs * 16384 + kK9WBytes + kK9Stages * kK9StageBytes + 64 + 1088 + 256;

struct K9Params {
    int B, H, W;          // network-input dims (padded frame, multiples of 4)
    const void* in;       // fp32 NCHW [B,3,H,W]   or   u8 NHWC [B,in_h,in_w,3]
    int norm;             // u8 mode: 1 = ToTensor + Normalize(.5,.5) (SRGAN route), 0 = ToTensor only (ESRGAN)
    int in_h, in_w;       // u8 mode: un-padded image size
    int pad_l, pad_t;     // u8 mode: where the image sits inside the padded frame
    const uint4* w_packed;  // kK9WBytes of bf16 in the shared-memory layout above
    const float* bias;    // fp32 [64]
    float slope;          // PReLU parameter
    int tiles_x, tiles_y, m_tiles;  // tiles of the H x W output
};

template <bool U8>
__global__ void __launch_bounds__(kK9Threads, kK9CtasPerSm)
k9_first_conv_kernel(const __grid_constant__ CUtensorMap tmO, const K9Params p) {
    constexpr int N = 64;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* o_smem = smem;
    uint8_t* w_smem = smem + kK9OutBufs * 16384;
    uint8_t* a_smem = w_smem + kK9WBytes;
    uint32_t* lut = reinterpret_cast<uint32_t*>(a_smem + kK9Stages * kK9StageBytes + 64);  // +64: the last tap pair reads 16 B past a stage  // 256 (+2) x {hi | lo<<16}
    uint64_t* bars = reinterpret_cast<uint64_t*>(lut + 264);
    uint64_t* a_full = bars;
    uint64_t* a_empty = bars + kK9Stages;
    uint64_t* acc_full = bars + 2 * kK9Stages;
    uint64_t* acc_empty = acc_full + 2;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_empty + 2);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;

    for (int i = threadIdx.x; i < kK9WBytes / 16; i += kK9Threads)
        reinterpret_cast<uint4*>(w_smem)[i] = __ldg(p.w_packed + i);
    for (int i = threadIdx.x; i < (kK9Stages * kK9StageBytes + 64) / 16; i += kK9Threads)
        reinterpret_cast<uint4*>(a_smem)[i] = make_uint4(0u, 0u, 0u, 0u);
    if (U8 && threadIdx.x < 256) {
        // ToTensor: v/255; SRGAN route: then Normalize (v-0.5)/0.5, each op rounded in fp32 (backend/app.py:401-406)
        float v = __fdiv_rn(static_cast<float>(threadIdx.x), 255.f);
        if (p.norm) v = __fdiv_rn(__fsub_rn(v, 0.5f), 0.5f);
        lut[threadIdx.x] = split_hi_lo(v);
        if (threadIdx.x == 0) {
            lut[256] = 0u;           // conv zero padding
            lut[257] = p.norm ? 0x0000BF80u : 0u;  // Pad(fill=0) before Normalize: (0-0.5)/0.5 = -1
        }
    }
    if (threadIdx.x == 0) {
        tma_prefetch_desc(&tmO);
        for (int i = 0; i < kK9Stages; ++i) {
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
        // this thread's 3 window pixels e = t + 128*i: window row wy = e/16, column wx = e%16; slot = e
        int wy[kK9PerThread], wx[kK9PerThread], slot[kK9PerThread];
#pragma unroll
        for (int i = 0; i < kK9PerThread; ++i) {
            const int e = threadIdx.x + 128 * i;
            wy[i] = e >> 4;
            wx[i] = e & 15;
            slot[i] = e;
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
                    const float* s = static_cast<const float*>(p.in) + (static_cast<size_t>(b) * 3 * p.H + yy) * p.W + xx;
                    const size_t plane = static_cast<size_t>(p.H) * p.W;
                    l0 = __float_as_uint(__ldg(s));
                    l1 = __float_as_uint(__ldg(s + plane));
                    l2 = __float_as_uint(__ldg(s + 2 * plane));
                }
            }
        };
        pdl_wait();  // the input image may be the previous kernel's output (noiser, resize, an earlier pass)
        TileWalk pf(p.tiles_x, p.tiles_y, mt_begin);
        auto load_raw = [&](uint32_t (&raw)[3 * kK9PerThread]) {
            const int y0 = pf.ty * 16 - 4, x0 = pf.tx * 8 - 4;
#pragma unroll
            for (int i = 0; i < kK9PerThread; ++i)
                load_px(pf.b, y0 + wy[i], x0 + wx[i], threadIdx.x + 128 * i < kK9Entries, raw[3 * i], raw[3 * i + 1],
                        raw[3 * i + 2]);
            pf.next(p.tiles_x, p.tiles_y);
        };
        constexpr int kPF = 2;
        uint32_t rq[kPF][3 * kK9PerThread];
#pragma unroll
        for (int j = 0; j < kPF; ++j)
            if (mt_begin + j < mt_end) load_raw(rq[j]);
        uint32_t stage = 0, phase = 0;
        for (int base = mt_begin; base < mt_end; base += kPF) {
#pragma unroll
            for (int j = 0; j < kPF; ++j) {
                const int mt = base + j;
                if (mt >= mt_end) break;
                uint4 px[kK9PerThread];
#pragma unroll
                for (int i = 0; i < kK9PerThread; ++i) {
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
                uint4* plane0 = reinterpret_cast<uint4*>(a_smem + stage * kK9StageBytes);
#pragma unroll
                for (int i = 0; i < kK9PerThread; ++i)
                    if (threadIdx.x + 128 * i < kK9Entries) plane0[slot[i]] = px[i];
                fence_proxy_async();
                mbar_arrive(&a_full[stage]);
                if (++stage == kK9Stages) {
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
                const uint32_t ab = a0 + stage * kK9StageBytes;
                // tap pair (dy, dx = 2j, 2j+1): K chunk 0 = window pixel +dx, K chunk 1 = the next pixel (LBO = 16 B)
#pragma unroll
                for (int dy = 0; dy < 9; ++dy)
#pragma unroll
                    for (int j = 0; j < 5; ++j)
                        umma_bf16_ss(tmem_base + buf * N, umma_desc_nosw(ab + (dy * 16 + 2 * j) * 16, 16, kK9Row),
                                     umma_desc_nosw(w0 + (dy * 5 + j) * 2048, 1024, 128), idesc, (dy | j) == 0 ? 0u : 1u);
                umma_commit(&a_empty[stage]);
                umma_commit(&acc_full[buf]);
            }
            __syncwarp();
            if (++stage == kK9Stages) {
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
            const uint32_t ob = acc_it % kK9OutBufs;
            if (issuer) bulk_wait_read<kK9OutBufs - 1>();
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
