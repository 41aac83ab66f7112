// K1/K2: implicit-GEMM 3x3 convolution (and 1x1 "GEMM" mode for the 2x2-stride-2 transposed
// convolutions) on tcgen05 tensor cores, NHWC bf16 activations staged by TMA, fp32 accumulators
// in TMEM, fused epilogues.
//
// Reference ops this replaces (file:line in /root/reference):
//   nn.Conv2d(k3,p1)+nn.ReLU            backend/app.py:43-77
//   nn.MaxPool2d(2,2)                   backend/app.py:48,56   (fused: EPI_RELU_POOL)
//   nn.ConvTranspose2d(k2,s2)           backend/app.py:65,73   (TAPS=1 + EPI_CONVT)
//   torch.cat([d,e],dim=1)              backend/app.py:93,100  (eliminated: producers write
//                                                               channel slices of one buffer)
//   torch.tanh + *0.5+0.5, clamp, u8    backend/app.py:103,435,472 (EPI_TANH_*)
//
// Geometry. One M-tile = 128 output pixels = 8 wide x 16 tall; GEMM row r <-> pixel
// (y0 + r/8, x0 + r%8). The A operand for one 64-channel K block and one horizontal tap dx is ONE
// TMA box {64 ch, 8 px, 18 rows} (rows y0-1 .. y0+16; out-of-image coordinates are zero-filled
// by TMA = the conv's zero padding). In shared memory that box is 18 groups of 8 rows x 128 B
// (128-byte swizzle), i.e. exactly the K-major SW128 UMMA layout with SBO = 1024 B, and the three
// vertical taps dy are the same box read at start address + dy*1024 B. So 9 taps cost 3 box
// loads (3.4x the minimum bytes instead of 9x).
//
// Warp roles (192 threads): warp 0 = TMA producer (one lane), warp 1 = TMEM owner + MMA issuer
// (one lane), warps 2..5 = epilogue (TMEM lane quarter = warp & 3). TMEM holds two accumulator
// buffers of S*N fp32 columns so the epilogue of work item i overlaps the MMAs of item i+1.
// Weights are either resident in shared memory for the whole (persistent) CTA (WRES) or streamed
// through a ring; with S > 1 one streamed weight tile feeds S pixel tiles.
#pragma once
#include "ptx.cuh"

namespace cid {

enum : int {  // TANH_*: last_conv_umma.cuh; LRELU/T4: the Keras cGAN layers (leaky ReLU with ConvParams::slope)
    EPI_RELU = 0, EPI_RELU_POOL = 1, EPI_CONVT = 2, EPI_TANH_F32 = 3, EPI_TANH_U8 = 4, EPI_LRELU = 5, EPI_T4 = 6, EPI_TANH_F32_NHWC = 7,
    EPI_ADD = 8,  // bias + residual add(s), no activation (ESRGAN / SRGAN skip connections)
    EPI_LINEAR_F32_NHWC = 10,  // last_conv_umma.cuh without tanh: the discriminator's 64 -> 3 input gradient
    EPI_LRELU_BN = 11,  // training: EPI_LRELU whose result goes out as fp32 NHWC (ConvParams::out_f32) with the BatchNorm batch
                        // statistics reduced in the epilogue (ConvParams::stats). A separate instantiation: the extra live
                        // registers and branches cost the inference LeakyReLU layers 40 % (ESRGAN block.0 0.25 -> 0.41 ms)
    EPI_DGRAD = 9  // training: data gradient = bias-free linear output, optionally masked by the ReLU of the layer's
                   // own input activation (ConvParams::res2 = that activation, > 0 keeps the gradient)
};

constexpr int kConvThreads = 320;  // warp 0 TMA, warp 1 MMA, warps 2..9 two epilogue groups
constexpr int kMaxStages = 8;    // A-operand ring (halo patches)
constexpr int kMaxWStages = 16;  // streamed-weight ring

struct ConvParams {
    int B, H, W;         // A-operand activation: batch, height, width (pixels)
    int kb_count;        // Cin / 64
    int n_tiles;         // tiles of width N across the GEMM N extent
    int tiles_x, tiles_y;
    int m_tiles;         // B * tiles_x * tiles_y
    int tx_sh, ty_sh;    // log2(tiles_x), log2(tiles_y) when both are powers of two, else -1 (decode_tile)
    int num_work;        // ceil(m_tiles / S) * n_tiles
    int a_stages, w_stages;
    int w_rows_per_tap;  // rows of the packed weight matrix per tap (= padded Cout)
    const float* bias;   // fp32 [cout]
    int cout;            // real output channels (EPI_CONVT: channels per quadrant)
    void* out;           // primary output
    int out_cs, out_c0;  // channel count of the output buffer / first channel we write
    void* out2;          // EPI_RELU_POOL: pooled output buffer
    int out2_cs, out2_c0;
    int crop_l, crop_t, out_h, out_w;  // EPI_TANH_U8: crop window into the u8 NHWC result
    float slope;         // EPI_LRELU / EPI_T4 / EPI_CONVT: negative slope of the leaky ReLU (PReLU with one parameter)
    const void* res;     // EPI_ADD: bf16 NHWC residual(s) with the output's spatial dims, res_cs channels per pixel
    const void* res2;    //          (second one optional)
    int res_cs, res2_cs;
    int bias_per_column; // EPI_CONVT as pixel shuffle: bias indexed by GEMM column instead of by channel
    int linear;          // EPI_T4: bias is all zeros and slope is 1 (the training data-gradient launches): the epilogue packs the
                         // accumulator as it is
    float mask_slope;    // EPI_T4 with res2 set (training): the output gradient is multiplied by the LeakyReLU derivative of
                         // the activation res2 at the OUTPUT pixel (1 where it is positive, mask_slope elsewhere)
    double* stats;       // EPI_LRELU_BN (training): per-channel sum (stats[c]) and sum of squares (stats[128 + c])
                         // of the fp32 outputs over all real pixels = the batch statistics of the BatchNorm that follows,
                         // reduced in the epilogue (warp transpose-sum, per-CTA register accumulators, one double atomic
                         // per channel per warp at the end); may be null
    float* out_f32;      // EPI_LRELU_BN: the fp32 NHWC result (= what tmO maps; staged in shared memory and written by TMA):
                         // the discriminator's pre-BatchNorm outputs, whose mean >> std
};

// Tap geometry (TAPS):
//   9  = 3x3, pad 1: one halo patch {64 ch, PW=10 px, 18 rows} per K block; tap (dy,dx) is the same patch
//        read at start + (dy*PW + dx)*128 B with SBO = PW*128 B (relies on the 128-B swizzle being a function
//        of the absolute shared-memory address, for TMA writes and UMMA reads alike).
//   1  = 1x1 (the 2x2-stride-2 transposed conv as a GEMM): one box {64 ch, 8 px, 16 rows}.
//   16 = 4x4 stride 2 pad 1 (Keras Conv2D k4 s2 'same'): a K block is (64 input channels, input phase
//        (phy,phx)); its patch {64 ch, PW=9 px, 17 rows} is a STRIDE-2 TMA box starting at input pixel
//        (2*y0-1+phy, 2*x0-1+phx) and serves the 4 taps ky = 2*dy+phy, kx = 2*dx+phx (dy,dx in {0,1}).
//   8  = 4x4 stride 2 pad 1 TRANSPOSED (Keras Conv2DTranspose k4 s2 'same'): the work item's n_tile is the
//        output row phase py; the accumulator holds both column phases px (2*N columns). Output
//        (2y+py, 2x+px) sums input (y-1+ty, x-1+tx) * w[ky = py+3-2ty, kx = px+3-2tx] over ty in {py,py+1}
//        and the (tx,px) pairs (0,0) (1,0) (1,1) (2,1): 8 MMA steps of width N over the 3x3 halo patch.
//   4  = 2x2 stride 2 pad 0 (the data gradient of ConvTranspose2d k2 s2, training only): a K block is (64 input
//        channels, input phase (phy,phx)); its box {64 ch, 8 px, 16 rows} is a stride-2 TMA box starting at input pixel
//        (2*y0+phy, 2*x0+phx) and serves the single tap (phy,phx).
// CG = 2: CTA pairs (cluster of 2, tcgen05 cta_group::2). Each CTA owns its own S pixel tiles (A operand,
//         accumulators, epilogue) but only HALF of every weight tile (N/2 rows); the leader CTA's MMAs
//         read both halves, so the per-SM shared-memory traffic of the B operand is halved.
template <int N, int TAPS, int S, bool WRES, int EPI, int PW = 0, int TS = 0, int CG = 1>
struct ConvCfg {
    static_assert(CG == 1 || CG == 2, "single CTA or CTA pair");
    static_assert(CG == 1 || N % 32 == 0, "pair mode: each CTA holds N/2 weight rows, N multiple of 16 for M=256");
    static_assert(TAPS == 9 || TAPS == 1 || TAPS == 16 || TAPS == 8 || TAPS == 4, "3x3, 1x1, 4x4 s2, transposed 4x4 s2 or 2x2 s2");
    static_assert(N % 16 == 0 && N >= 16 && N <= 256, "UMMA N for M=128");
    static constexpr bool K4 = TAPS == 16, T4 = TAPS == 8, K2 = TAPS == 4;
    static_assert(!(K4 || T4 || K2) || !WRES, "stride-2 layers stream their weights");
    static constexpr int DX = (TAPS == 9 || T4) ? 3 : K4 ? 2 : 1;
    static constexpr int DY = DX;
    static_assert(PW == (DX == 1 ? 0 : 8 + DX - 1), "patch width = 8 + DX - 1 (0 = plain box for 1x1)");
    static constexpr int STEPS = TAPS == 9 ? 9 : K4 ? 4 : T4 ? 8 : 1;  // weight tiles (MMA steps) per K block
    static constexpr int KSUB = (K4 || K2) ? 4 : 1;                    // K blocks per 64 input channels
    static constexpr int A_ROWS = 16 + DY - 1;
    static constexpr int A_TX = (PW ? PW : 8) * A_ROWS * 128;   // bytes one box delivers
    static constexpr int A_BYTES = (A_TX + 1023) / 1024 * 1024; // stage pitch (1024-B aligned)
    static constexpr int SBO = (PW ? PW : 8) * 128;
    static constexpr int W_BYTES = (N / CG) * 128;  // this CTA's share of one weight tile
    static constexpr int NACC = T4 ? 2 * N : N;     // accumulator columns per pixel tile
    static constexpr int ACC_COLS = S * NACC;
    static_assert(2 * ACC_COLS <= 512, "two accumulator buffers must fit TMEM");
    static constexpr int TMEM_COLS = (2 * ACC_COLS <= 32)    ? 32
                                     : (2 * ACC_COLS <= 64)  ? 64
                                     : (2 * ACC_COLS <= 128) ? 128
                                     : (2 * ACC_COLS <= 256) ? 256
                                                             : 512;
    // The epilogue stages each 64-channel slab of a tile in shared memory (SW128 rows, same geometry as
    // the A boxes) and writes it with one TMA store. Each of the 2 epilogue groups has NOB buffers.
    // TS = 1: the epilogue stages each output slab in shared memory and writes it with one TMA store (the only form; a
    // register-direct 16-byte-store epilogue for the N = 64 layers measured 60 % SLOWER on down1.2 - every warp store
    // touches 32 lines with 16 bytes each - profiles/r02b_direct_store_*.json - and was removed).
    static_assert(TS == 1, "staged epilogue");
    static constexpr int NOB = 1;  // staging buffers per group (the other group's buffer gives the overlap)
    static constexpr int O_SLAB = 128 * 128;
    static constexpr int O2_SLAB = 32 * 128;
    static constexpr int OT_SLAB = 256 * 128;  // transposed convs: [16 rows][16 px][64 ch] of the 2x-upsampled output
    // (EPI_LRELU_BN stages fp32: a 64-channel slab = two 32-channel half-slabs of [128 px][128 B], stored by two TMA stores)
    static constexpr int O_GROUP = EPI == EPI_LRELU_BN ? 2 * O_SLAB : NOB * (O_SLAB + (EPI == EPI_RELU_POOL ? O2_SLAB : 0));
    // EPI_ADD: each group also owns a residual buffer (one slab) that the block-skip tile is TMA-loaded into one work item
    // ahead, so the load overlaps the previous slab's TMA store and the wait for the next accumulator
    static constexpr int R_BYTES = EPI == EPI_ADD ? 2 * O_SLAB : 0;
    static constexpr int O_BYTES = !TS ? 0 : (EPI == EPI_CONVT || EPI == EPI_T4) ? 2 * OT_SLAB : 2 * O_GROUP + R_BYTES;
    // dynamic shared memory needed (bytes), including 1 KB of alignment slack
    static int smem_bytes(int kb_count, int a_stages, int w_stages) {
        int w_items = WRES ? kb_count * TAPS : w_stages;
        return 1024 + w_items * W_BYTES + a_stages * A_BYTES + O_BYTES + 1024;
    }
};

// One MMA step of a K block: which patch offset it reads, which packed weight rows it uses, which accumulator
// columns it feeds. `sub` = the K block's input phase (K4) ; `nt` = the work item's n_tile (T4: row phase py).
template <int N, int TAPS, int PW>
struct StepMap {
    int a_px;      // patch pixel offset of the tap (x 128 B)
    int w_tap;     // tap index into the packed weights (rows w_tap * w_rows_per_tap ...)
    int d_col;     // accumulator column offset inside the tile
    bool first;    // first MMA step into its accumulator columns within a K block 0
    __device__ __forceinline__ StepMap(int step, int sub, int nt) {
        if constexpr (TAPS == 9) {
            const int dx = step / 3, dy = step - dx * 3;
            a_px = dy * PW + dx; w_tap = dy * 3 + dx; d_col = 0; first = step == 0;
        } else if constexpr (TAPS == 16) {
            const int dy = step >> 1, dx = step & 1;
            a_px = dy * PW + dx; w_tap = (2 * dy + (sub >> 1)) * 4 + 2 * dx + (sub & 1); d_col = 0; first = step == 0;
        } else if constexpr (TAPS == 8) {
            const int ty = nt + (step >> 2), q = step & 3;
            const int tx = (q + 1) >> 1, px = q >> 1;
            a_px = ty * PW + tx; w_tap = (nt + 3 - 2 * ty) * 4 + (px + 3 - 2 * tx); d_col = px * N;
            first = (step == 0 || step == 2);
        } else if constexpr (TAPS == 4) {
            a_px = 0; w_tap = sub; d_col = 0; first = true;
        } else {
            a_px = 0; w_tap = 0; d_col = 0; first = true;
        }
    }
};

struct TileCoord {
    int b, y0, x0;
};

__device__ __forceinline__ TileCoord decode_tile(int tiles_x, int tiles_y, int mt) {
    TileCoord t;
    int tx = mt % tiles_x;
    int r = mt / tiles_x;
    int ty = r % tiles_y;
    t.b = r / tiles_y;  // == B for padding tiles past m_tiles: TMA zero-fills, stores are masked
    t.y0 = ty * 16;
    t.x0 = tx * 8;
    return t;
}
__device__ __forceinline__ TileCoord decode_tile(const ConvParams& p, int mt) {
    if (p.tx_sh >= 0) {  // tiles_x and tiles_y are powers of two (128 x 128 images: 16 x 8): shifts instead of two divisions
        TileCoord t;
        const int r = mt >> p.tx_sh;
        t.x0 = (mt & (p.tiles_x - 1)) * 8;
        t.y0 = (r & (p.tiles_y - 1)) * 16;
        t.b = r >> p.ty_sh;
        return t;
    }
    return decode_tile(p.tiles_x, p.tiles_y, mt);
}

// Tile cursor for kernels whose CTAs walk a contiguous range of tiles (x fastest): increments instead of divisions.
struct TileWalk {
    int tx, ty, b;
    __device__ TileWalk(int tiles_x, int tiles_y, int mt) {
        tx = mt % tiles_x;
        const int r = mt / tiles_x;
        ty = r % tiles_y;
        b = r / tiles_y;
    }
    __device__ __forceinline__ void next(int tiles_x, int tiles_y) {
        if (++tx == tiles_x) {
            tx = 0;
            if (++ty == tiles_y) {
                ty = 0;
                ++b;
            }
        }
    }
};


// Walk an accumulator row in 32-column chunks with the next chunk's tcgen05.ld already in
// flight while the current one is processed (two register buffers, fully unrolled).
template <int NCH, class F>
__device__ __forceinline__ void for_each_chunk(uint32_t taddr, F&& f) {
    uint32_t va[32], vb[32];
    tmem_ld_32x32(taddr, va);
#pragma unroll
    for (int ch = 0; ch < NCH; ch += 2) {
        tmem_ld_wait();
        if (ch + 1 < NCH) tmem_ld_32x32(taddr + (ch + 1) * 32, vb);
        f(ch, va);
        if (ch + 1 < NCH) {
            tmem_ld_wait();
            if (ch + 2 < NCH) tmem_ld_32x32(taddr + (ch + 2) * 32, va);
            f(ch + 1, vb);
        }
    }
}

// Same, over NCH chunks whose 32-column chunk index is given by map(i), i = 0..NCH-1.
template <int NCH, class M, class F>
__device__ __forceinline__ void for_each_chunk_mapped(uint32_t taddr, M&& map, F&& f) {
    uint32_t va[32], vb[32];
    tmem_ld_32x32(taddr + map(0) * 32, va);
#pragma unroll
    for (int i = 0; i < NCH; i += 2) {
        tmem_ld_wait();
        if (i + 1 < NCH) tmem_ld_32x32(taddr + map(i + 1) * 32, vb);
        f(map(i), va);
        if (i + 1 < NCH) {
            tmem_ld_wait();
            if (i + 2 < NCH) tmem_ld_32x32(taddr + map(i + 2) * 32, va);
            f(map(i + 1), vb);
        }
    }
}

template <int N, int TAPS, int S, bool WRES, int EPI, int PW = 0, int TS = 0, int CG = 1>
__global__ void __launch_bounds__(kConvThreads, 1)
conv_umma_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmW,
                 const __grid_constant__ CUtensorMap tmO, const __grid_constant__ CUtensorMap tmO2,
                 const ConvParams p) {
    using Cfg = ConvCfg<N, TAPS, S, WRES, EPI, PW, TS, CG>;
    constexpr bool PAIR = CG == 2;
    constexpr int STEPS = Cfg::STEPS, KSUB = Cfg::KSUB;
    constexpr int A_BYTES = Cfg::A_BYTES, W_BYTES = Cfg::W_BYTES;
    using Step = StepMap<N, TAPS, PW>;

    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));

    const int w_items = WRES ? p.kb_count * TAPS : p.w_stages;
    uint8_t* w_smem = smem;
    uint8_t* a_smem = smem + w_items * W_BYTES;
    uint8_t* o_smem = a_smem + p.a_stages * A_BYTES;               // epilogue staging, Cfg::O_BYTES
    uint64_t* bars = reinterpret_cast<uint64_t*>(o_smem + Cfg::O_BYTES);
    uint64_t* a_full = bars;
    uint64_t* a_empty = bars + kMaxStages;
    uint64_t* w_full = bars + 2 * kMaxStages;
    uint64_t* w_empty = w_full + kMaxWStages;
    uint64_t* acc_full = w_empty + kMaxWStages;
    uint64_t* acc_empty = acc_full + 2;
    uint64_t* res_full = acc_empty + 2;  // EPI_ADD: residual tile of each epilogue group has landed in its slab
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(res_full + 2);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int AST = p.a_stages;
    const int WST = p.w_stages;
    const uint32_t rank = PAIR ? cluster_ctarank() : 0u;           // 0 = leader (issues the pair's MMAs)
    const int unit = PAIR ? (blockIdx.x >> 1) : blockIdx.x;        // CTA (or CTA pair) index
    const int units = PAIR ? (gridDim.x >> 1) : gridDim.x;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmA);
        tma_prefetch_desc(&tmW);
        if (TS) {
            tma_prefetch_desc(&tmO);
            if (EPI == EPI_RELU_POOL || EPI == EPI_ADD) tma_prefetch_desc(&tmO2);
        }
        // pair mode: the leader's "full" barriers collect one expect_tx arrival per CTA, its acc_empty
        // one arrival per epilogue warp of both CTAs; "empty"/acc_full get one multicast commit each.
        for (int i = 0; i < kMaxStages; ++i) {
            mbar_init(&a_full[i], CG);
            mbar_init(&a_empty[i], 1);
        }
        for (int i = 0; i < kMaxWStages; ++i) {
            mbar_init(&w_full[i], CG);
            mbar_init(&w_empty[i], 1);
        }
        for (int i = 0; i < 2; ++i) {
            mbar_init(&acc_full[i], 1);
            mbar_init(&acc_empty[i], 8 * CG);
            mbar_init(&res_full[i], 1);
        }
        fence_mbar_init();
    }
    if (warp == 1) {
        if (PAIR)
            tmem_alloc_cg2(tmem_slot, Cfg::TMEM_COLS);
        else
            tmem_alloc(tmem_slot, Cfg::TMEM_COLS);
    }
    tc_fence_before();
    if (PAIR)
        cluster_sync_all();  // the peer's barriers are initialised before anything arrives on them
    else
        __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    pdl_launch_dependents();  // the next layer's CTAs may start their own set-up on SMs that are (or become) free

    if (warp == 0) {
        // ======================= TMA producer =======================
        if (lane == 0) {
            // completion barrier of a load: this CTA's own, or (pair mode) the LEADER's at the same offset
            auto full_bar = [&](uint64_t* bar) { return PAIR ? mapa_u32(smem_u32(bar), 0) : smem_u32(bar); };
            auto expect = [&](uint64_t* bar, uint32_t bytes) {
                if (PAIR)
                    mbar_arrive_expect_tx_cluster(full_bar(bar), bytes);
                else
                    mbar_arrive_expect_tx(bar, bytes);
            };
            auto load_w = [&](void* dst, uint64_t* bar, int c0, int row) {
                if (PAIR)
                    tma_load_2d_cg2(dst, &tmW, full_bar(bar), c0, row + static_cast<int>(rank) * (N / CG), kEvictLast);
                else
                    tma_load_2d(dst, &tmW, bar, c0, row, kEvictLast);
            };
            if (WRES) {
                // n_tiles > 1 (host guarantees units % n_tiles == 0): this CTA only ever sees column block unit % n_tiles
                const int wrow0 = (unit % p.n_tiles) * N;
                expect(&w_full[0], static_cast<uint32_t>(w_items) * W_BYTES);
                for (int kb = 0; kb < p.kb_count; ++kb)
                    for (int tap = 0; tap < TAPS; ++tap)
                        load_w(w_smem + (kb * TAPS + tap) * W_BYTES, &w_full[0], kb * 64, tap * p.w_rows_per_tap + wrow0);
            }
            uint32_t a_stage = 0, a_phase = 0, w_stage = 0, w_phase = 0;
            const int kb_total = p.kb_count * KSUB;
            // Streamed weights are constants too: the first ring-full of weight tiles of this CTA's first work item is
            // requested BEFORE waiting for the previous layer, so on a launch-bound (small-batch) pass they are in flight
            // while that layer finishes. The main loop below skips exactly these tiles.
            int w_pre = 0;
            if (!WRES && unit < p.num_work) {
                const int n_tile0 = unit % p.n_tiles;
                for (int kb = 0; kb < kb_total && w_pre < WST; ++kb) {
                    const int cb = kb / KSUB, sub = kb % KSUB;
                    for (int step = 0; step < STEPS && w_pre < WST; ++step) {
                        const Step sm(step, sub, n_tile0);
                        expect(&w_full[w_pre], W_BYTES);   // first use of every stage: its "empty" phase is trivially free
                        load_w(w_smem + w_pre * W_BYTES, &w_full[w_pre], cb * 64,
                               sm.w_tap * p.w_rows_per_tap + (Cfg::T4 ? 0 : n_tile0 * N));
                        ++w_pre;
                    }
                }
            }
            pdl_wait();  // activations are the previous layer's output (the weights above are constants)
            for (int work = unit; work < p.num_work; work += units) {
                const int n_tile = work % p.n_tiles;
                const int st = work / p.n_tiles;
                TileCoord tc[S];
#pragma unroll
                for (int s = 0; s < S; ++s) tc[s] = decode_tile(p, (st * CG + static_cast<int>(rank)) * S + s);
                for (int kb = 0; kb < kb_total; ++kb) {
                    const int cb = kb / KSUB, sub = kb % KSUB;  // 64-channel block, input phase (4x4 s2 only)
#pragma unroll
                    for (int s = 0; s < S; ++s) {
                        // patch origin in the A tensor: 3x3 / transposed 4x4 halo starts one pixel up-left;
                        // 4x4 s2 reads input phase (phy,phx) of the 2x larger input through a stride-2 box
                        const int ax = Cfg::K4   ? 2 * tc[s].x0 - 1 + (sub & 1)
                                       : Cfg::K2 ? 2 * tc[s].x0 + (sub & 1)
                                                 : tc[s].x0 - (Cfg::DX == 3 ? 1 : 0);
                        const int ay = Cfg::K4   ? 2 * tc[s].y0 - 1 + (sub >> 1)
                                       : Cfg::K2 ? 2 * tc[s].y0 + (sub >> 1)
                                                 : tc[s].y0 - (Cfg::DY == 3 ? 1 : 0);
                        mbar_wait(&a_empty[a_stage], a_phase ^ 1);
                        expect(&a_full[a_stage], Cfg::A_TX);
                        if (PAIR)
                            tma_load_4d_cg2(a_smem + a_stage * A_BYTES, &tmA, full_bar(&a_full[a_stage]), cb * 64, ax, ay,
                                            tc[s].b, kEvictNormal);
                        else
                            tma_load_4d(a_smem + a_stage * A_BYTES, &tmA, &a_full[a_stage], cb * 64, ax, ay, tc[s].b,
                                        kEvictNormal);
                        if (++a_stage == static_cast<uint32_t>(AST)) {
                            a_stage = 0;
                            a_phase ^= 1;
                        }
                    }
                    if (!WRES) {
                        for (int step = 0; step < STEPS; ++step) {
                            const Step sm(step, sub, n_tile);
                            if (w_pre > 0) {
                                --w_pre;  // already requested above
                            } else {
                                mbar_wait(&w_empty[w_stage], w_phase ^ 1);
                                expect(&w_full[w_stage], W_BYTES);
                                load_w(w_smem + w_stage * W_BYTES, &w_full[w_stage], cb * 64,
                                       sm.w_tap * p.w_rows_per_tap + (Cfg::T4 ? 0 : n_tile * N));
                            }
                            if (++w_stage == static_cast<uint32_t>(WST)) {
                                w_stage = 0;
                                w_phase ^= 1;
                            }
                        }
                    }
                }
            }
        }
    } else if (warp == 1 && rank == 0) {
        // ======================= MMA issuer (pair mode: the leader CTA only) =======================
        // The whole warp walks the loop (warp-uniform control flow keeps descriptors in uniform
        // registers); one elected lane issues the MMAs and the commits.
        constexpr uint32_t idesc = umma_idesc_bf16(128 * CG, N);
        auto wait_full = [&](uint64_t* bar, uint32_t parity) {
            if (PAIR)
                mbar_wait_cluster(bar, parity);
            else
                mbar_wait(bar, parity);
        };
        auto mma = [&](uint32_t d, uint64_t ad, uint64_t bd, uint32_t acc) {
            if (PAIR)
                umma_bf16_ss_cg2(d, ad, bd, idesc, acc);
            else
                umma_bf16_ss(d, ad, bd, idesc, acc);
        };
        auto commit = [&](uint64_t* bar) {
            if (PAIR)
                umma_commit_cg2(bar);
            else
                umma_commit(bar);
        };
        const uint32_t a_lo0 = umma_desc_lo(smem_u32(a_smem));
        const uint32_t w_lo0 = umma_desc_lo(smem_u32(w_smem));
        uint32_t a_stage = 0, a_phase = 0, w_stage = 0, w_phase = 0, acc_it = 0;
        if (WRES) {
            wait_full(&w_full[0], 0);
            tc_fence_after();
        }
        for (int work = unit; work < p.num_work; work += units) {
            const uint32_t buf = acc_it & 1;
            const uint32_t aph = (acc_it >> 1) & 1;
            wait_full(&acc_empty[buf], aph ^ 1);
            tc_fence_after();
            const uint32_t d_base = tmem_base + buf * Cfg::ACC_COLS;
            const int n_tile = work % p.n_tiles;
            const int kb_total = p.kb_count * KSUB;
            for (int kb = 0; kb < kb_total; ++kb) {
                const int sub = kb % KSUB;
                uint32_t stg[S];
#pragma unroll
                for (int s = 0; s < S; ++s) {
                    wait_full(&a_full[a_stage], a_phase);
                    stg[s] = a_stage;
                    if (++a_stage == static_cast<uint32_t>(AST)) {
                        a_stage = 0;
                        a_phase ^= 1;
                    }
                }
                tc_fence_after();
#pragma unroll
                for (int step = 0; step < STEPS; ++step) {
                    const Step sm(step, sub, n_tile);
                    uint32_t w_lo;
                    if (WRES) {
                        w_lo = w_lo0 + static_cast<uint32_t>(kb * TAPS + sm.w_tap) * (W_BYTES >> 4);
                    } else {
                        wait_full(&w_full[w_stage], w_phase);
                        tc_fence_after();
                        w_lo = w_lo0 + w_stage * (W_BYTES >> 4);
                    }
                    const uint32_t first = (kb == 0 && sm.first) ? 1u : 0u;
                    if (elect_one()) {
#pragma unroll
                        for (int s = 0; s < S; ++s) {
                            const uint32_t a_lo = a_lo0 + stg[s] * (A_BYTES >> 4) + sm.a_px * (128 >> 4);
#pragma unroll
                            for (int k = 0; k < 4; ++k) {
                                mma(d_base + s * Cfg::NACC + sm.d_col, umma_desc_join<Cfg::SBO>(a_lo + 2 * k),
                                    umma_desc_join<1024>(w_lo + 2 * k), (first && k == 0) ? 0u : 1u);
                            }
                        }
                        if (!WRES) commit(&w_empty[w_stage]);
                    }
                    __syncwarp();
                    if (!WRES) {
                        if (++w_stage == static_cast<uint32_t>(WST)) {
                            w_stage = 0;
                            w_phase ^= 1;
                        }
                    }
                }
                if (elect_one()) {
#pragma unroll
                    for (int s = 0; s < S; ++s) commit(&a_empty[stg[s]]);
                }
                __syncwarp();
            }
            if (elect_one()) commit(&acc_full[buf]);
            __syncwarp();
            ++acc_it;
        }
    } else if (warp >= 2) {
        // ======================= epilogue: 2 groups x 4 warps =======================
        // Group g (warps 2+4g .. 5+4g; TMEM lane quarter = warp & 3) owns the 64-channel output slabs
        // of the work item whose index has parity g, with its own staging buffers, named barrier and
        // TMA-store issuer, so the two groups drain one accumulator buffer concurrently.
        const int g = (warp - 2) >> 2;
        const int q = warp & 3;
        const int row = q * 32 + lane;
        const int yl = row >> 3;
        const int xl = row & 7;
        const bool issuer = (threadIdx.x == 64 + 128 * g);  // lane 0 of the group's first warp
        const int bar_id = 1 + g;
        uint8_t* og = o_smem + g * Cfg::O_GROUP;
        uint8_t* rg = o_smem + 2 * Cfg::O_GROUP + g * Cfg::O_SLAB;  // EPI_ADD: this group's residual buffer
        (void)rg;
        uint32_t acc_it = 0;
        uint32_t o_count = 0;  // slabs this group has staged so far
        // BatchNorm statistics (EPI_LRELU_BN): lane = channel inside a 32-channel chunk
        // (group g stages the slabs j = g, g + 2: always the SAME 64-channel column block cs = j % SL - two accumulators)
        static_assert(EPI != EPI_LRELU_BN || N <= 128, "EPI_LRELU_BN: one column block per epilogue group");
        float st_sum[2] = {0.f, 0.f}, st_sq[2] = {0.f, 0.f};
        // EPI_ADD: the first residual tensor (tmO2 = its tile map, same geometry as the output's) is TMA-loaded into
        // the group's residual buffer one work item ahead: as soon as the previous item's add has read the buffer.
        // Group g owns exactly slab j = g of every work item (S * N/64 == 2). Out-of-image pixels load as zeros.
        const bool res_tma = EPI == EPI_ADD && p.res != nullptr;
        auto issue_residual = [&](int w) {
            if constexpr (EPI == EPI_ADD) {
                static_assert(EPI != EPI_ADD || S * (N / 64) == 2, "one slab per group per work item");
                const int nt = w % p.n_tiles, stw = w / p.n_tiles;
                const int sj = g / (N / 64), csj = g % (N / 64);
                const TileCoord tcr = decode_tile(p, (stw * CG + static_cast<int>(rank)) * S + sj);
                mbar_arrive_expect_tx(&res_full[g], Cfg::O_SLAB);
                tma_load_4d(rg, &tmO2, &res_full[g], nt * N + csj * 64, tcr.x0, tcr.y0, tcr.b, kEvictFirst);
            }
        };
        if (res_tma && issuer && unit < p.num_work) {
            pdl_wait();  // the residual is an earlier layer's output: under programmatic dependent launch this warp
                         // may run before those layers have finished (the TMA producer's wait does not cover it)
            issue_residual(unit);
        }
        // Training (EPI_DGRAD / EPI_T4 with res2): the derivative mask is one 128-byte line of the layer's input activation
        // per output pixel, read by the thread that owns the pixel. The epilogue is the critical path of these kernels and
        // never waits for an accumulator, so each of those reads used to cost a full DRAM latency: every thread now pulls
        // the lines of its NEXT work item into L2 while it works on the current one.
        auto prefetch_mask = [&](int w) {
            if constexpr (EPI == EPI_DGRAD || EPI == EPI_T4) {
                if (p.res2 == nullptr || w >= p.num_work) return;
                constexpr int SLp = N / 64;
                const int nt = w % p.n_tiles, stw = w / p.n_tiles;
#pragma unroll
                for (int j = 0; j < S * SLp; ++j) {
                    if ((j & 1) != g) continue;
                    const TileCoord tcw = decode_tile(p, (stw * CG + static_cast<int>(rank)) * S + j / SLp);
                    const uint16_t* base = static_cast<const uint16_t*>(p.res2);
                    if constexpr (EPI == EPI_DGRAD) {
                        const int yy = tcw.y0 + yl, xx = tcw.x0 + xl;
                        if (tcw.b < p.B && yy < p.H && xx < p.W)
                            prefetch_l2(base + ((static_cast<size_t>(tcw.b) * p.H + yy) * p.W + xx) * p.res2_cs + nt * N + (j % SLp) * 64);
                    } else {
                        const int oy = 2 * (tcw.y0 + yl) + nt;
#pragma unroll
                        for (int px = 0; px < 2; ++px) {
                            const int ox = 2 * (tcw.x0 + xl) + px;
                            if (tcw.b < p.B && oy < 2 * p.H && ox < 2 * p.W)
                                prefetch_l2(base + ((static_cast<size_t>(tcw.b) * (2 * p.H) + oy) * (2 * p.W) + ox) * p.res2_cs + (j % SLp) * 64);
                        }
                    }
                }
            }
        };
        prefetch_mask(unit);
        // EPI_T4 with res2 (training): the LeakyReLU mask of chunk i = 2 px + half of a slab = 64 bytes of the activation
        // at this thread's output pixel (2y+py, 2x+px), held in t4_mk[i & 1]. Chunks 0 and 1 of the group's NEXT slab are
        // requested as soon as the current slab has consumed its last chunk (across work items when a group has one slab
        // per work item), chunks 2 and 3 when 0 and 1 have been used: waiting for the loads in place, four times per slab,
        // was a quarter of this kernel's stall samples.
        uint4 t4_mk[2][4];
        bool t4_live[2] = {false, false};
        (void)t4_mk;
        auto t4_mask_load = [&](int i, const TileCoord& tcw, int nt, int csw) {  // i = 2 px + half
            if constexpr (EPI == EPI_T4) {
                const int px = i >> 1;
                if (t4_live[px]) {
                    const size_t pix = (static_cast<size_t>(tcw.b) * (2 * p.H) + 2 * (tcw.y0 + yl) + nt) * (2 * p.W) + 2 * (tcw.x0 + xl) + px;
                    const uint4* m4 = reinterpret_cast<const uint4*>(static_cast<const uint16_t*>(p.res2) + pix * p.res2_cs + csw * 64 + (i & 1) * 32);
#pragma unroll
                    for (int k = 0; k < 4; ++k) t4_mk[i & 1][k] = __ldg(m4 + k);
                }
            }
        };
        auto t4_mask_begin = [&](const TileCoord& tcw, int nt, int csw) {  // first two chunks of a slab
            if constexpr (EPI == EPI_T4) {
#pragma unroll
                for (int px = 0; px < 2; ++px)
                    t4_live[px] = p.res2 != nullptr && tcw.b < p.B && 2 * (tcw.y0 + yl) + nt < 2 * p.H && 2 * (tcw.x0 + xl) + px < 2 * p.W;
                t4_mask_load(0, tcw, nt, csw);
                t4_mask_load(1, tcw, nt, csw);
            }
        };
        constexpr bool T4_ONE_SLAB = EPI == EPI_T4 && S * (N / 64) == 2;  // one slab per group per work item: j = g
        if constexpr (T4_ONE_SLAB) {
            if (unit < p.num_work)
                t4_mask_begin(decode_tile(p, ((unit / p.n_tiles) * CG + static_cast<int>(rank)) * S + g / (N / 64)), unit % p.n_tiles, g % (N / 64));
        }
        for (int work = unit; work < p.num_work; work += units) {
            const int n_tile = work % p.n_tiles;
            const int st = work / p.n_tiles;
            const uint32_t buf = acc_it & 1;
            const uint32_t aph = (acc_it >> 1) & 1;
            prefetch_mask(work + units);
            mbar_wait(&acc_full[buf], aph);
            tc_fence_after();
            const uint32_t tbase = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + buf * Cfg::ACC_COLS;
            {
                if constexpr (EPI == EPI_RELU || EPI == EPI_RELU_POOL || EPI == EPI_LRELU || EPI == EPI_LRELU_BN || EPI == EPI_ADD || EPI == EPI_DGRAD) {
                    constexpr int SL = N / 64;  // 64-channel slabs per pixel tile
                    const int sub = (xl & 1) | ((yl & 1) << 1);
                    const int prow_i = (yl >> 1) * 4 + (xl >> 1);
                    const float slope = p.slope;
                    auto act = [&](float x) {  // ReLU, or LeakyReLU(slope) for the Keras layers
                        if constexpr (EPI == EPI_LRELU) return x > 0.f ? x : x * slope;
                        if constexpr (EPI == EPI_ADD || EPI == EPI_DGRAD || EPI == EPI_LRELU_BN) return x;  // (BN: z = conv + bias)
                        return fmaxf(x, 0.f);
                    };
                    (void)slope;
#pragma unroll((EPI == EPI_LRELU_BN || EPI == EPI_DGRAD) ? 1 : S * SL)  // training-only epilogues: rolled (they stalled on instruction fetch)
                    for (int j = 0; j < S * SL; ++j) {
                        if ((j & 1) != g) continue;
                        const int s = j / SL, cs = j % SL;
                        const TileCoord tc = decode_tile(p, (st * CG + static_cast<int>(rank)) * S + s);
                        const uint32_t ob = o_count % Cfg::NOB;
                        uint8_t* slab = og + ob * Cfg::O_SLAB;
                        uint8_t* slab2 = og + Cfg::NOB * Cfg::O_SLAB + ob * Cfg::O2_SLAB;
                        // the store that used this buffer NOB slabs ago must have finished READING it
                        if constexpr (TS) {
                            if (issuer) bulk_wait_read<Cfg::NOB - 1>();
                            named_bar_sync(bar_id, 128);
                        }
                        // EPI_ADD: the optional SECOND residual (ESRGAN's long skip, last block only) comes straight from
                        // global memory, this thread's pixel = one 128-byte line
                        uint4 rv2[(EPI == EPI_ADD || EPI == EPI_DGRAD) ? 8 : 1];
                        if constexpr (EPI == EPI_DGRAD) {
                            // ReLU mask: this thread's pixel of the layer's input activation, one 128-byte line
                            const int yy = tc.y0 + yl, xx = tc.x0 + xl;
                            const bool live = tc.b < p.B && yy < p.H && xx < p.W;
#pragma unroll
                            for (int i = 0; i < 8; ++i) rv2[i] = make_uint4(0x3f803f80u, 0x3f803f80u, 0x3f803f80u, 0x3f803f80u);
                            if (live && p.res2 != nullptr) {
                                const size_t pix = (static_cast<size_t>(tc.b) * p.H + yy) * p.W + xx;
                                const uint4* r2 = reinterpret_cast<const uint4*>(
                                    static_cast<const uint16_t*>(p.res2) + pix * p.res2_cs + n_tile * N + cs * 64);
#pragma unroll
                                for (int i = 0; i < 8; ++i) rv2[i] = __ldg(r2 + i);
                            }
                        }
                        if constexpr (EPI == EPI_ADD) {
                            const int yy = tc.y0 + yl, xx = tc.x0 + xl;
                            const bool live = tc.b < p.B && yy < p.H && xx < p.W;
#pragma unroll
                            for (int i = 0; i < 8; ++i) rv2[i] = make_uint4(0u, 0u, 0u, 0u);
                            if (live && p.res2 != nullptr) {
                                const size_t pix = (static_cast<size_t>(tc.b) * p.H + yy) * p.W + xx;
                                const uint4* r2 = reinterpret_cast<const uint4*>(
                                    static_cast<const uint16_t*>(p.res2) + pix * p.res2_cs + n_tile * N + cs * 64);
#pragma unroll
                                for (int i = 0; i < 8; ++i) rv2[i] = __ldg(r2 + i);
                            }
                            if (res_tma) mbar_wait(&res_full[g], acc_it & 1);  // the residual tile is in `rg`
                        }
                        for_each_chunk<2>(tbase + s * Cfg::NACC + cs * 64, [&](int ch, uint32_t (&v)[32]) {
                            const float4* b4 = reinterpret_cast<const float4*>(p.bias + n_tile * N + cs * 64 + ch * 32);
                            uint32_t pk[16];
                            if constexpr (EPI == EPI_ADD) {
                                // v[2i], v[2i+1] += the bf16 pair i of this 32-channel half
#pragma unroll
                                for (int i = 0; i < 4; ++i) {
                                    const int cjr = ch * 4 + i;
                                    const uint4 a = res_tma ? ld_shared_v4(rg + row * 128 + ((cjr ^ (row & 7)) << 4))
                                                            : make_uint4(0u, 0u, 0u, 0u);
                                    const uint4 b = rv2[ch * 4 + i];
                                    const uint32_t w[4] = {a.x, a.y, a.z, a.w}, w2[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
                                    for (int e = 0; e < 4; ++e) {
                                        const int col = (i * 4 + e) * 2;
                                        float lo = __uint_as_float(v[col]) + __uint_as_float(w[e] << 16);
                                        float hi = __uint_as_float(v[col + 1]) + __uint_as_float(w[e] & 0xffff0000u);
                                        lo += __uint_as_float(w2[e] << 16);
                                        hi += __uint_as_float(w2[e] & 0xffff0000u);
                                        v[col] = __float_as_uint(lo);
                                        v[col + 1] = __float_as_uint(hi);
                                    }
                                }
                            }
#pragma unroll
                            for (int jj = 0; jj < 8; ++jj) {
                                const float4 bb = EPI == EPI_DGRAD ? make_float4(0.f, 0.f, 0.f, 0.f) : __ldg(b4 + jj);
                                const float f0 = act(__uint_as_float(v[4 * jj + 0]) + bb.x);
                                const float f1 = act(__uint_as_float(v[4 * jj + 1]) + bb.y);
                                const float f2 = act(__uint_as_float(v[4 * jj + 2]) + bb.z);
                                const float f3 = act(__uint_as_float(v[4 * jj + 3]) + bb.w);
                                pk[2 * jj + 0] = pack_bf16x2(f0, f1);
                                pk[2 * jj + 1] = pack_bf16x2(f2, f3);
                                if constexpr (EPI == EPI_LRELU_BN) {
                                    // fp32 result: this pixel's 32 channels = one 128-byte row of half-slab `ch` (SW128),
                                    // written out by a TMA store below (per-thread 16-byte global stores touch 32 lines
                                    // with 16 bytes each per warp instruction: the layer ran at 21 % of the tensor peak).
                                    // Pixels outside the image are staged as zeros (the store clips them; the batch
                                    // statistics below then need no mask): a branch, uniform for every full tile.
                                    const bool live = tc.b < p.B && tc.y0 + yl < p.H && tc.x0 + xl < p.W;
                                    uint8_t* dst = slab + ch * Cfg::O_SLAB + row * 128 + ((jj ^ (row & 7)) << 4);
                                    if (live)
                                        st_shared_v4(dst, __float_as_uint(f0), __float_as_uint(f1), __float_as_uint(f2), __float_as_uint(f3));
                                    else
                                        st_shared_v4(dst, 0u, 0u, 0u, 0u);
                                }
                            }
                            if constexpr (EPI == EPI_DGRAD) {
                                // zero the gradient where the activation (bf16, post-ReLU: >= 0) is not positive: on the
                                // packed pairs, one packed compare (0xffff per true half) and one AND per two values
                                const __nv_bfloat162 zero2 = __float2bfloat162_rn(0.f);
#pragma unroll
                                for (int i = 0; i < 4; ++i) {
                                    const uint4 m = rv2[ch * 4 + i];
                                    const uint32_t w[4] = {m.x, m.y, m.z, m.w};
#pragma unroll
                                    for (int e = 0; e < 4; ++e)
                                        pk[i * 4 + e] &= __hgt2_mask(*reinterpret_cast<const __nv_bfloat162*>(&w[e]), zero2);
                                }
                            }
                            if constexpr (TS && EPI != EPI_LRELU_BN) {
                                uint8_t* rowp = slab + row * 128;
#pragma unroll
                                for (int jj = 0; jj < 4; ++jj) {
                                    const int cj = ch * 4 + jj;  // 16-byte chunk inside the 128-byte row
                                    st_shared_v4(rowp + ((cj ^ (row & 7)) << 4), pk[4 * jj], pk[4 * jj + 1], pk[4 * jj + 2],
                                                 pk[4 * jj + 3]);
                                }
                            }
                            if constexpr (EPI == EPI_RELU_POOL) {
                                // 2x2 max: x partner is lane^1, y partner is lane^8 (row = y*8 + x).
#pragma unroll
                                for (int jj = 0; jj < 16; ++jj) {
                                    uint32_t m = max_bf16x2(pk[jj], __shfl_xor_sync(0xffffffffu, pk[jj], 1));
                                    pk[jj] = max_bf16x2(m, __shfl_xor_sync(0xffffffffu, m, 8));
                                }
                                // each of the 4 lanes of a 2x2 block writes a different 8-channel quarter
                                const uint32_t o0 = sub == 0 ? pk[0] : sub == 1 ? pk[4] : sub == 2 ? pk[8] : pk[12];
                                const uint32_t o1 = sub == 0 ? pk[1] : sub == 1 ? pk[5] : sub == 2 ? pk[9] : pk[13];
                                const uint32_t o2 = sub == 0 ? pk[2] : sub == 1 ? pk[6] : sub == 2 ? pk[10] : pk[14];
                                const uint32_t o3 = sub == 0 ? pk[3] : sub == 1 ? pk[7] : sub == 2 ? pk[11] : pk[15];
                                if constexpr (TS) {
                                    const int cj = ch * 4 + sub;
                                    st_shared_v4(slab2 + prow_i * 128 + ((cj ^ (prow_i & 7)) << 4), o0, o1, o2, o3);
                                }
                            }
                        });
                        if constexpr (TS) {
                            fence_proxy_async();
                            named_bar_sync(bar_id, 128);
                        }
                        if constexpr (EPI == EPI_LRELU_BN) {
                            if (issuer) {  // tmO = the fp32 map of the output (box 32 ch x 8 px x 16 rows)
                                const int c64 = n_tile * N + cs * 64;
                                tma_store_4d(&tmO, slab, p.out_c0 + c64, tc.x0, tc.y0, tc.b);
                                tma_store_4d(&tmO, slab + Cfg::O_SLAB, p.out_c0 + c64 + 32, tc.x0, tc.y0, tc.b);
                                bulk_commit();
                            }
                            // BatchNorm batch statistics from the staged tile: lane = channel, this warp's 32 pixel rows
                            // (one row = 32 banks, the swizzle keeps the lanes on distinct banks: conflict-free), summed
                            // into per-CTA register accumulators - no shuffles
                            if (p.stats != nullptr) {
#pragma unroll
                                for (int hh = 0; hh < 2; ++hh) {
                                    const uint8_t* hs = slab + hh * Cfg::O_SLAB + ((lane & 3) << 2);
                                    float a1 = 0.f, a2 = 0.f;
#pragma unroll 8
                                    for (int r = q * 32; r < q * 32 + 32; ++r) {
                                        const float v = *reinterpret_cast<const float*>(hs + r * 128 + (((lane >> 2) ^ (r & 7)) << 4));
                                        a1 += v;
                                        a2 = fmaf(v, v, a2);
                                    }
                                    st_sum[hh] += a1;
                                    st_sq[hh] += a2;
                                }
                            }
                        } else if (TS && issuer) {
                            const int c64 = n_tile * N + cs * 64;
                            tma_store_4d(&tmO, slab, p.out_c0 + c64, tc.x0, tc.y0, tc.b);
                            if constexpr (EPI == EPI_RELU_POOL)
                                tma_store_4d(&tmO2, slab2, p.out2_c0 + c64, tc.x0 >> 1, tc.y0 >> 1, tc.b);
                            bulk_commit();
                        }
                        // (every thread of the group passed the barrier above after its last read of `rg`)
                        if (res_tma && issuer && work + units < p.num_work) issue_residual(work + units);
                        ++o_count;
                    }
                } else if constexpr (EPI == EPI_T4) {
                    // ---- transposed 4x4 s2: the tile's accumulator is [px][N] for output row phase py = n_tile.
                    // Slab (s, cs) = channels [64cs, 64cs+64) of tile s, both px: [16 y][16 px][64 ch] of output
                    // rows 2y+py, one 5-D TMA store over (C, 2W, py, H, B). Slabs alternate between the groups.
                    constexpr int SL = N / 64;
                    const float slope = p.slope;
#pragma unroll
                    for (int j = 0; j < S * SL; ++j) {
                        if ((j & 1) != g) continue;
                        const int s = j / SL, cs = j % SL;
                        const TileCoord tc = decode_tile(p, (st * CG + static_cast<int>(rank)) * S + s);
                        uint8_t* slab = o_smem + g * Cfg::OT_SLAB;
                        if constexpr (!T4_ONE_SLAB) t4_mask_begin(tc, n_tile, cs);  // (else requested one slab ago)
                        if (issuer) bulk_wait_read<0>();
                        named_bar_sync(bar_id, 128);
                        // chunk i: px = i >> 1, 32-channel half = i & 1
                        auto chunk_of = [&](int i) { return ((i >> 1) * N + cs * 64 + (i & 1) * 32) >> 5; };
                        for_each_chunk_mapped<4>(tbase + s * Cfg::NACC, chunk_of, [&](int ch, uint32_t (&v)[32]) {
                            const int col = ch * 32;
                            const int px = col >= N ? 1 : 0;
                            const int c = col - px * N;  // channel of v[0]
                            const float4* b4 = reinterpret_cast<const float4*>(p.bias + c);
                            uint32_t pk[16];
                            if (p.res2 != nullptr) {
                                // LeakyReLU backward of the layer below, fused (mask loaded two chunks ahead, see above)
                                const int ci = px * 2 + (((c - cs * 64) >> 5) & 1);  // this chunk's index in the slab
                                if (t4_live[px]) {
#pragma unroll
                                    for (int i = 0; i < 4; ++i) {
                                        const uint4 m = t4_mk[ci & 1][i];
                                        const uint32_t w[4] = {m.x, m.y, m.z, m.w};
#pragma unroll
                                        for (int e = 0; e < 4; ++e) {
                                            const int k = (i * 4 + e) * 2;
                                            if (!(__uint_as_float(w[e] << 16) > 0.f)) v[k] = __float_as_uint(__uint_as_float(v[k]) * p.mask_slope);
                                            if (!(__uint_as_float(w[e] & 0xffff0000u) > 0.f))
                                                v[k + 1] = __float_as_uint(__uint_as_float(v[k + 1]) * p.mask_slope);
                                        }
                                    }
                                }
                                if (ci + 2 < 4) t4_mask_load(ci + 2, tc, n_tile, cs);
                            }
                            if (p.linear) {  // zero bias, slope 1 (training data gradients): nothing to add or select
#pragma unroll
                                for (int jj = 0; jj < 16; ++jj)
                                    pk[jj] = pack_bf16x2(__uint_as_float(v[2 * jj]), __uint_as_float(v[2 * jj + 1]));
                            } else {
#pragma unroll
                                for (int jj = 0; jj < 8; ++jj) {
                                    const float4 bb = __ldg(b4 + jj);
                                    float f0 = __uint_as_float(v[4 * jj + 0]) + bb.x;
                                    float f1 = __uint_as_float(v[4 * jj + 1]) + bb.y;
                                    float f2 = __uint_as_float(v[4 * jj + 2]) + bb.z;
                                    float f3 = __uint_as_float(v[4 * jj + 3]) + bb.w;
                                    f0 = f0 > 0.f ? f0 : f0 * slope;
                                    f1 = f1 > 0.f ? f1 : f1 * slope;
                                    f2 = f2 > 0.f ? f2 : f2 * slope;
                                    f3 = f3 > 0.f ? f3 : f3 * slope;
                                    pk[2 * jj + 0] = pack_bf16x2(f0, f1);
                                    pk[2 * jj + 1] = pack_bf16x2(f2, f3);
                                }
                            }
                            const int orow_i = yl * 16 + 2 * xl + px;
                            uint8_t* rowp = slab + orow_i * 128;
                            const int half = (c >> 5) & 1;
#pragma unroll
                            for (int jj = 0; jj < 4; ++jj) {
                                const int cj = half * 4 + jj;
                                st_shared_v4(rowp + ((cj ^ (orow_i & 7)) << 4), pk[4 * jj], pk[4 * jj + 1], pk[4 * jj + 2],
                                             pk[4 * jj + 3]);
                            }
                        });
                        if constexpr (T4_ONE_SLAB) {  // the mask registers are free: request the next work item's first chunks
                            const int wn = work + units;
                            if (wn < p.num_work)
                                t4_mask_begin(decode_tile(p, ((wn / p.n_tiles) * CG + static_cast<int>(rank)) * S + s), wn % p.n_tiles, cs);
                        }
                        fence_proxy_async();
                        named_bar_sync(bar_id, 128);
                        if (issuer) {
                            tma_store_5d(&tmO, slab, p.out_c0 + cs * 64, 2 * tc.x0, n_tile, tc.y0, tc.b);
                            bulk_commit();
                        }
                    }
                } else {
                    // ---- transposed conv: GEMM column n = quadrant*cout + c, quadrant = dy*2+dx. The tile's 8
                    // chunks fill two slabs [16 y][16 px][64 ch] of output rows 2y+dy (cout 64: slab = dy;
                    // cout 128: slab = 64-channel half, dy = n_tile), each written by one 5-D TMA store over
                    // (C, 2W, dy, H, B). Group g fills and stores slab g.
                    static_assert(EPI == EPI_CONVT && N == 256 && S == 1, "convT / pixel-shuffle epilogue: 8 chunks per tile");
                    const TileCoord tc = decode_tile(p, st * CG + static_cast<int>(rank));
                    const bool c64mode = (p.cout == 64);
                    const float ct_slope = p.slope;
                    const bool act_on = ct_slope != 1.f;
                    const int slab_i = g;
                    uint8_t* slab = o_smem + slab_i * Cfg::OT_SLAB;
                    if (issuer) bulk_wait_read<0>();  // this group's previous store has drained the slab
                    named_bar_sync(bar_id, 128);
                    // chunks of this slab: cout 64 -> 4 consecutive (quadrants 2*slab, 2*slab+1);
                    // cout 128 -> channels [64*slab, 64*slab+64) of both quadrants of this n_tile
                    auto chunk_of = [&](int i) { return c64mode ? slab_i * 4 + i : (i >> 1) * 4 + slab_i * 2 + (i & 1); };
                    for_each_chunk_mapped<4>(tbase, chunk_of, [&](int ch, uint32_t (&v)[32]) {
                        const int n = n_tile * N + ch * 32;
                        const int quad = c64mode ? (n >> 6) : (n >> 7);  // p.cout is 64 or 128
                        const int c = n - quad * p.cout;
                        // transposed conv: one bias per output channel; pixel shuffle: one per GEMM column
                        const float4* b4 = reinterpret_cast<const float4*>(p.bias + (p.bias_per_column ? n : c));
                        uint32_t pk[16];
#pragma unroll
                        for (int jj = 0; jj < 8; ++jj) {
                            const float4 bb = __ldg(b4 + jj);
                            float f0 = __uint_as_float(v[4 * jj + 0]) + bb.x, f1 = __uint_as_float(v[4 * jj + 1]) + bb.y;
                            float f2 = __uint_as_float(v[4 * jj + 2]) + bb.z, f3 = __uint_as_float(v[4 * jj + 3]) + bb.w;
                            if (act_on) {  // PixelShuffle + PReLU (SRGAN); the plain transposed convs have no activation
                                f0 = f0 > 0.f ? f0 : f0 * ct_slope;
                                f1 = f1 > 0.f ? f1 : f1 * ct_slope;
                                f2 = f2 > 0.f ? f2 : f2 * ct_slope;
                                f3 = f3 > 0.f ? f3 : f3 * ct_slope;
                            }
                            pk[2 * jj + 0] = pack_bf16x2(f0, f1);
                            pk[2 * jj + 1] = pack_bf16x2(f2, f3);
                        }
                        const int orow_i = yl * 16 + 2 * xl + (quad & 1);
                        uint8_t* rowp = slab + orow_i * 128;
#pragma unroll
                        for (int jj = 0; jj < 4; ++jj) {
                            const int cj = (ch & 1) * 4 + jj;
                            st_shared_v4(rowp + ((cj ^ (orow_i & 7)) << 4), pk[4 * jj], pk[4 * jj + 1], pk[4 * jj + 2],
                                         pk[4 * jj + 3]);
                        }
                    });
                    fence_proxy_async();
                    named_bar_sync(bar_id, 128);
                    if (issuer) {
                        const int dy = c64mode ? slab_i : n_tile;
                        const int c0 = p.out_c0 + (c64mode ? 0 : slab_i * 64);
                        tma_store_5d(&tmO, slab, c0, 2 * tc.x0, dy, tc.y0, tc.b);
                        bulk_commit();
                    }
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) {
                if (PAIR)
                    mbar_arrive_cluster(mapa_u32(smem_u32(&acc_empty[buf]), 0));  // the leader's barrier
                else
                    mbar_arrive(&acc_empty[buf]);
            }
            ++acc_it;
        }
        if constexpr (EPI == EPI_LRELU_BN) {
            if (p.stats != nullptr) {
                // (n_tiles == 1 on this path) this group's column block: channels 64 cs + 32 hh + lane
                const int c0 = (N / 64 == 1 ? 0 : g) * 64 + lane;
#pragma unroll
                for (int hh = 0; hh < 2; ++hh) {
                    if (st_sum[hh] != 0.f || st_sq[hh] != 0.f) {
                        atomicAdd(p.stats + c0 + hh * 32, static_cast<double>(st_sum[hh]));
                        atomicAdd(p.stats + 128 + c0 + hh * 32, static_cast<double>(st_sq[hh]));
                    }
                }
            }
        }
        if (issuer) bulk_wait_read<0>();  // the staged slabs have been read; the writes complete with the grid
    }

    tc_fence_before();
    if (PAIR)
        cluster_sync_all();  // neither CTA leaves while the other may still touch its smem / TMEM / barriers
    else
        __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        if (PAIR)
            tmem_dealloc_cg2(tmem_base, Cfg::TMEM_COLS);
        else
            tmem_dealloc(tmem_base, Cfg::TMEM_COLS);
    }
}

}  // namespace cid
