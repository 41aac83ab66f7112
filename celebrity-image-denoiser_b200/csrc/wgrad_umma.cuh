// K3: weight gradients of the convolutions on tcgen05 tensor cores (training step).
//
// Reference ops this replaces (autograd's conv weight-gradient under `g_loss.backward()` / `d_loss.backward()`,
// backend/trainingcode/denoise_gan_code/training.py:410-425):
//   dW[co][ci][ky][kx] = sum over (b,y,x) of dZ[b,y,x,co] * X[b, s*y-1+ky, s*x-1+kx, ci]      Conv2d k3 p1, stride s
//   dW[ci][co][dy][dx] = sum over (b,y,x) of X[b,y,x,ci]  * dY[b, 2y+dy, 2x+dx, co]            ConvTranspose2d k2 s2
//
// Both are GEMMs whose K dimension is the PIXEL index, with NHWC bf16 operands whose contiguous dimension is the
// channel = the GEMM's M / N dimension: both operands are "MN-major" UMMA operands (instruction-descriptor bits 15/16),
// read straight from the same TMA boxes the forward kernels use:
//   * operand A (M = 128 rows): a plain 8 px x 16 row pixel tile {64 ch, 8 px, 16 rows} per 64-channel atom. With only 64
//     channels the second atom re-reads the first (LBO = 0) and accumulator rows 64..127 are ignored.
//   * operand B (N = 64 or 128 columns): a patch {64 ch, PW px, 16 rows} (PW = 8 + halo) per 64-channel atom; the T taps of
//     one role are the same patch read at start + t * 128 B (one pixel to the right) - the absolute-address 128-byte
//     swizzle makes the shifted reads legal, exactly as in conv_umma.cuh. For stride-2 layers the patch is a stride-2
//     TMA box (the operand with twice the resolution).
// One MMA covers K = 16 pixels = two tile rows: 8 consecutive pixels are 8 consecutive 128-byte lines (one swizzle atom
// of the canonical MN-major SW128 layout ((8,n),(8,k)):((16B,LBO),(128B,SBO))), the second row sits SBO = PW*128 B further.
//
// Work split. A ROLE fixes (M channel block, N channel block, tap row): its T accumulators [128 x N] stay in TMEM for
// the CTA's whole life (split-K over pixel tiles: CTA (role, slice) walks its contiguous share of the tiles), and the
// epilogue adds them into the fp32 gradient tensor (reference layout) with red.global.add.f32.
// Warp roles (192 threads): warp 0 = TMA producer, warp 1 = TMEM owner + MMA issuer, warps 2..5 = epilogue; while the
// main loop runs they also sum the dZ tiles over the pixels (the layer's bias gradient) straight from shared memory.
#pragma once
#include "ptx.cuh"

namespace cid {

constexpr int kWgThreads = 192;
constexpr int kWgMaxRoles = 48;
constexpr int kWgMaxStages = 6;

struct WgRole {
    int a_c0, b_c0;  // first channel of this role's M / N block (coordinates in the TMA maps' channel dimension)
    int m0, n0;      // the same blocks as indices of the gradient tensor
    int oy, ox;      // patch origin = scale * (tile y0, x0) + (oy, ox)
    int tap[3];      // gradient tap index of MMA tap t (patch pixel offset t); -1 = computed but dropped
    int bias;        // this role also reduces its A tiles over the pixels: bias_out[m0 + c] += sum_p A[p][c]
};

struct WgParams {
    int B, H, W;  // pixel grid of operand A (tiles of 8 x 16)
    int tiles_x, tiles_y, m_tiles;
    int n_roles, slices;
    int scale;    // 1, or 2: operand B has twice the resolution and is read through a stride-2 box
    int pw;       // patch pixels per row (B operand)
    int stages;
    float* out;   // out[((m0 + row) * s_m + n0 + col) * kk + tap]
    int s_m, kk;
    int m_valid;  // rows of the 128-row accumulator that are real channels (64 or 128)
    int fuse_taps;  // T = 3 only: the three taps of a row are ONE MMA of N = 192 per 64-channel B atom (the atoms of the
                    // MN-major operand are the same patch one pixel = 128 B apart: LBO = 128), so operand A is read from
                    // shared memory once per K step instead of three times. Accumulator column = q*192 + t*64 + c.
    float* bias_out;  // bias gradient (column sums of operand A = dZ), reduced by the otherwise idle epilogue warps of
                      // the roles flagged WgRole::bias while the MMAs run; may be null
    WgRole roles[kWgMaxRoles];
};

// MN-major SW128 shared-memory descriptor (same bit layout as the K-major one; majorness lives in the instruction
// descriptor): start[0,14) lbo[16,30) sbo[32,46) version[46,48)=1 layout[61,64)=2.
__device__ __forceinline__ uint64_t umma_desc_mn_sw128(uint32_t smem_addr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
    uint64_t d = 0;
    d |= static_cast<uint64_t>((smem_addr & 0x3FFFF) >> 4);
    d |= static_cast<uint64_t>((lbo_bytes >> 4) & 0x3FFF) << 16;
    d |= static_cast<uint64_t>((sbo_bytes >> 4) & 0x3FFF) << 32;
    d |= static_cast<uint64_t>(1) << 46;
    d |= static_cast<uint64_t>(2) << 61;
    return d;
}

// kind::f16 instruction descriptor, bf16 x bf16 -> fp32, BOTH operands MN-major.
__host__ __device__ constexpr uint32_t umma_idesc_bf16_mn(int m, int n) {
    return (1u << 4) | (1u << 7) | (1u << 10) | (1u << 15) | (1u << 16) | (static_cast<uint32_t>(n >> 3) << 17) |
           (static_cast<uint32_t>(m >> 4) << 24);
}

__device__ __forceinline__ void red_add_f32(float* p, float v) {
    asm volatile("red.global.add.f32 [%0], %1;" ::"l"(p), "f"(v) : "memory");
}

// NA: 64-channel atoms of operand A that are real (1: rows 64..127 duplicate rows 0..63), NB: atoms of operand B
// (N = 64 * NB), T: taps per role (accumulators).
// CL = 3 (3x3 stride-1 layers): the three tap-row roles of one (M block, N block) run as a CLUSTER of 3 CTAs on the same
// pixel tiles. Every operand tile is fetched from L2 ONCE per cluster and TMA-multicast into all three CTAs (rank 0 loads
// the A atoms, ranks 1 / 2 the B atoms: the full 18-row halo patch, of which CTA `rank` reads rows rank .. rank+15), so
// the L2 -> SM traffic per FLOP drops 2.8x; a stage is re-filled only after all three CTAs' MMAs (and the bias-reducing
// warps) have released it (multicast commits onto every CTA's "empty" barrier).
template <int NA, int NB, int T, int CL = 1>
__global__ void __launch_bounds__(kWgThreads, 1)
wgrad_umma_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                  const __grid_constant__ WgParams p) {
    static_assert(NA == 1 || NA == 2, "M = 128");
    static_assert(NB == 1 || NB == 2, "N = 64 or 128");
    static_assert(T >= 1 && T <= 3 && T * NB * 64 <= 512, "accumulators must fit TMEM");
    constexpr int N = 64 * NB;
    constexpr int A_ATOM = 128 * 128;  // one 64-channel atom of the pixel tile
    constexpr uint32_t TMEM_COLS = (T * N <= 64) ? 64 : (T * N <= 128) ? 128 : (T * N <= 256) ? 256 : 512;

    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    static_assert(CL == 1 || (CL == 3 && T == 3), "cluster mode = the three tap rows of a 3x3 layer");
    constexpr int PROWS = CL == 3 ? 18 : 16;           // patch rows per B atom
    const int pb = (p.pw * PROWS * 128 + 1023) & ~1023;  // bytes of one B atom (patch), 1024-aligned pitch
    const int stage_bytes = NA * A_ATOM + NB * pb;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + p.stages * stage_bytes);
    uint64_t* full = bars;
    uint64_t* empty = bars + kWgMaxStages;
    uint64_t* acc_full = bars + 2 * kWgMaxStages;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_full + 1);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const uint32_t rank = CL == 3 ? cluster_ctarank() : 0u;
    const int unit = blockIdx.x / CL;                  // CTA (or cluster) index
    const int n_units = p.n_roles / CL;                // role blocks
    const int role_i = (unit % n_units) * CL + static_cast<int>(rank);
    const int slice = unit / n_units;
    const WgRole& role = p.roles[role_i];
    const int per = (p.m_tiles + p.slices - 1) / p.slices;
    const int t_begin = slice * per;
    const int t_end = min(p.m_tiles, t_begin + per);
    const bool has_work = t_begin < t_end;
    const bool do_bias = role.bias != 0 && p.bias_out != nullptr;
    // cluster mode: the bias role is rank 0; its four warps release a stage in ALL three CTAs
    const bool blk_bias = CL == 3 ? (p.roles[(unit % n_units) * CL].bias != 0 && p.bias_out != nullptr) : do_bias;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmA);
        tma_prefetch_desc(&tmB);
        for (int i = 0; i < kWgMaxStages; ++i) {
            mbar_init(&full[i], 1);
            mbar_init(&empty[i], CL + (blk_bias ? 4 : 0));  // MMA commit(s) (+ one arrival per bias-reducing warp)
        }
        mbar_init(acc_full, 1);
        fence_mbar_init();
    }
    if (warp == 1) tmem_alloc(tmem_slot, TMEM_COLS);
    tc_fence_before();
    if (CL == 3)
        cluster_sync_all();  // the peers' barriers are initialised before anything arrives on them
    else
        __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    pdl_launch_dependents();  // the next kernel may begin its set-up; it orders itself with griddepcontrol.wait

    if (warp == 0) {
        if (lane == 0 && has_work) {
            pdl_wait();
            uint32_t stage = 0, phase = 0;
            for (int mt = t_begin; mt < t_end; ++mt) {
                const int tx = mt % p.tiles_x;
                const int r = mt / p.tiles_x;
                const int ty = r % p.tiles_y;
                const int b = r / p.tiles_y;
                const int y0 = ty * 16, x0 = tx * 8;
                uint8_t* sa = smem + stage * stage_bytes;
                uint8_t* sb = sa + NA * A_ATOM;
                mbar_wait(&empty[stage], phase ^ 1);
                mbar_arrive_expect_tx(&full[stage], NA * A_ATOM + NB * p.pw * PROWS * 128);
                if constexpr (CL == 3) {
                    // each rank fetches its share once and multicasts it to the whole cluster
                    if (rank == 0) {
#pragma unroll
                        for (int a = 0; a < NA; ++a)
                            tma_load_4d_mc(sa + a * A_ATOM, &tmA, &full[stage], role.a_c0 + a * 64, x0, y0, b, 7, kEvictNormal);
                    } else if (static_cast<int>(rank) <= NB) {
                        const int q = static_cast<int>(rank) - 1;
                        tma_load_4d_mc(sb + q * pb, &tmB, &full[stage], role.b_c0 + q * 64, x0 - 1, y0 - 1, b, 7, kEvictNormal);
                    }
                } else {
#pragma unroll
                    for (int a = 0; a < NA; ++a)
                        tma_load_4d(sa + a * A_ATOM, &tmA, &full[stage], role.a_c0 + a * 64, x0, y0, b, kEvictNormal);
#pragma unroll
                    for (int q = 0; q < NB; ++q)
                        tma_load_4d(sb + q * pb, &tmB, &full[stage], role.b_c0 + q * 64, p.scale * x0 + role.ox,
                                    p.scale * y0 + role.oy, b, kEvictNormal);
                }
                if (++stage == static_cast<uint32_t>(p.stages)) {
                    stage = 0;
                    phase ^= 1;
                }
            }
        }
    } else if (warp == 1) {
        if (has_work) {
            constexpr uint32_t idesc = umma_idesc_bf16_mn(128, N);
            const uint32_t sbo_b = static_cast<uint32_t>(p.pw) * 128u;
            uint32_t stage = 0, phase = 0;
            for (int mt = t_begin; mt < t_end; ++mt) {
                mbar_wait(&full[stage], phase);
                tc_fence_after();
                const uint32_t sa = smem_u32(smem + stage * stage_bytes);
                const uint32_t sb = sa + NA * A_ATOM;
                if (elect_one()) {
#pragma unroll
                    for (int k8 = 0; k8 < 8; ++k8) {
                        // K = 16 pixels = tile rows 2*k8 and 2*k8+1
                        const uint64_t ad = umma_desc_mn_sw128(sa + k8 * 2048, NA == 2 ? A_ATOM : 0, 1024);
                        if (T == 3 && p.fuse_taps) {
                            constexpr uint32_t idesc3 = umma_idesc_bf16_mn(128, 192);
#pragma unroll
                            for (int q = 0; q < NB; ++q) {
                                const uint64_t bd = umma_desc_mn_sw128(
                                    sb + q * pb + ((k8 * 2 + (CL == 3 ? static_cast<int>(rank) : 0)) * p.pw) * 128, 128, sbo_b);
                                umma_bf16_ss(tmem_base + q * 192, ad, bd, idesc3, (mt == t_begin && k8 == 0) ? 0u : 1u);
                            }
                        } else {
#pragma unroll
                            for (int t = 0; t < T; ++t) {
                                const uint64_t bd = umma_desc_mn_sw128(
                                    sb + ((k8 * 2 + (CL == 3 ? static_cast<int>(rank) : 0)) * p.pw + t) * 128,
                                    static_cast<uint32_t>(pb), sbo_b);
                                umma_bf16_ss(tmem_base + t * N, ad, bd, idesc, (mt == t_begin && k8 == 0) ? 0u : 1u);
                            }
                        }
                    }
                    if (CL == 3)
                        umma_commit_mc(&empty[stage], 7);
                    else
                        umma_commit(&empty[stage]);
                }
                __syncwarp();
                if (++stage == static_cast<uint32_t>(p.stages)) {
                    stage = 0;
                    phase ^= 1;
                }
            }
            if (elect_one()) umma_commit(acc_full);
            __syncwarp();
        }
    } else if (has_work) {
        // epilogue: TMEM lane quarter = warp & 3; lane = accumulator row = M channel
        const int q = warp & 3;
        const int row = q * 32 + lane;
        if (do_bias) {
            // column sums of the A tiles: thread = (16-byte channel chunk j, pixel-row group rg of 16)
            const int et = (warp - 2) * 32 + lane;
            const int j = et & 7, rg = et >> 3;
            float acc[NA][8];
#pragma unroll
            for (int a = 0; a < NA; ++a)
#pragma unroll
                for (int e = 0; e < 8; ++e) acc[a][e] = 0.f;
            uint32_t stage = 0, phase = 0;
            for (int mt = t_begin; mt < t_end; ++mt) {
                mbar_wait(&full[stage], phase);
                const uint8_t* sa = smem + stage * stage_bytes;
#pragma unroll
                for (int a = 0; a < NA; ++a)
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        const int r = rg + 16 * i;
                        const uint4 v = ld_shared_v4(sa + a * A_ATOM + r * 128 + ((j ^ (r & 7)) << 4));
                        const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
                        for (int e = 0; e < 4; ++e) {
                            acc[a][2 * e] += __uint_as_float(w[e] << 16);
                            acc[a][2 * e + 1] += __uint_as_float(w[e] & 0xffff0000u);
                        }
                    }
                __syncwarp();
                if (lane == 0) {
                    if (CL == 3) {
#pragma unroll
                        for (uint32_t r = 0; r < 3; ++r) mbar_arrive_cluster(mapa_u32(smem_u32(&empty[stage]), r));
                    } else {
                        mbar_arrive(&empty[stage]);
                    }
                }
                if (++stage == static_cast<uint32_t>(p.stages)) {
                    stage = 0;
                    phase ^= 1;
                }
            }
#pragma unroll
            for (int a = 0; a < NA; ++a)
#pragma unroll
                for (int e = 0; e < 8; ++e) {
                    float v = acc[a][e];
                    v += __shfl_xor_sync(0xffffffffu, v, 8);
                    v += __shfl_xor_sync(0xffffffffu, v, 16);
                    if (lane < 8) red_add_f32(p.bias_out + role.m0 + a * 64 + j * 8 + e, v);
                }
        }
        mbar_wait(acc_full, 0);
        tc_fence_after();
        if (q * 32 < p.m_valid) {  // warp-uniform: quarters that hold duplicate rows do nothing
            float* orow = p.out + (static_cast<size_t>(role.m0 + row) * p.s_m + role.n0) * p.kk;
#pragma unroll
            for (int t = 0; t < T; ++t) {
                const int tap = role.tap[t];
                if (tap < 0) continue;
#pragma unroll
                for (int ch = 0; ch < N / 32; ++ch) {
                    uint32_t v[32];
                    // column of (tap t, channel chunk ch): t*N + 32 ch, or q*192 + t*64 + ... in the fused-tap layout
                    const int col = (T == 3 && p.fuse_taps) ? (ch >> 1) * 192 + t * 64 + (ch & 1) * 32 : t * N + ch * 32;
                    tmem_ld_32x32(tmem_base + (static_cast<uint32_t>(q * 32) << 16) + col, v);
                    tmem_ld_wait();
#pragma unroll
                    for (int j = 0; j < 32; ++j) red_add_f32(orow + static_cast<size_t>(ch * 32 + j) * p.kk + tap, __uint_as_float(v[j]));
                }
            }
        }
    }

    tc_fence_before();
    if (CL == 3)
        cluster_sync_all();  // no CTA leaves while a peer may still multicast into its smem or arrive on its barriers
    else
        __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, TMEM_COLS);
    }
}

// ---------------------------------------------------------------------------------------------------------------------
// Weight gradients of the two 3-channel layers (Conv2d(3,64): V64 = dZ, V3 = the layer input, sign +1; Conv2d(64,3):
// V64 = the layer input, V3 = dZ, sign -1) with the im2col operand built IN the kernel:
//   R[c64][ci*9 + tap] = sum_p V64[p][c64] * V3[p + sign*(tap - centre)][ci]
// Operand A = the 8 x 16 pixel tile of V64 (TMA, as above). Operand B = U[p][ci*9 + tap] (27 of 64 channels, the rest
// zero): the four epilogue warps - idle during the main loop otherwise - gather it, one tile per warp, from a shared-memory
// window of the fp32 3-channel source (10 x 18 pixels, loaded one tile ahead) and write it straight into the stage's
// SW128 tile, the layout TMA would have produced. Saves the im2col kernel and its 64-channel bf16 tensor (a write and a read of 128 B per pixel
// for 54 B of information) and the scatter kernel: the epilogue adds into the gradient tensor directly.
// Channel 27 of U is all ones: column 27 of the result is the layer's bias gradient (Conv2d(3,64): sum of dZ), no reduction
// pass. full[stage] = the TMA transaction + the builder warp's arrival; empty[stage] = the MMA commit.
// ---------------------------------------------------------------------------------------------------------------------
// one pixel's 27 im2col values from a source window [18 rows][32 slots] (slot = x*3 + ci for an NHWC source, ci*10 + x for
// NCHW); wp points at the pixel's own slot (ci = 0). Compile-time sign / layout: every read is one LDS with an immediate.
template <int SIGN, bool NCHW>
__device__ __forceinline__ void wg_edge_gather(const float* wp, float (&f)[32]) {
    constexpr int SX = NCHW ? 1 : 3, SC = NCHW ? 10 : 1;
#pragma unroll
    for (int ci = 0; ci < 3; ++ci)
#pragma unroll
        for (int ky = 0; ky < 3; ++ky)
#pragma unroll
            for (int kx = 0; kx < 3; ++kx) f[ci * 9 + ky * 3 + kx] = wp[ci * SC + SIGN * ((ky - 1) * 32 + (kx - 1) * SX)];
}

struct WgEdgeParams {
    int B, H, W;
    int tiles_x, tiles_y, m_tiles, stages;
    const float* v3;  // fp32 NHWC-3 or NCHW-3
    int v3_nchw, sign;
    float* out;       // mode 0: out[(c64*3 + ci)*9 + tap], mode 1: out[(ci*64 + c64)*9 + tap]
    int out_mode;
    float* bias_out;  // += column sums of V64 (may be null)
};

constexpr int kWgEdgeStage = 2 * 128 * 128;  // A tile + U tile
// builder warps (warps 2..): each owns every 6th tile of the CTA. Must not exceed the number of stages: a warp that has
// finished tile k (so tile k - stages has been consumed) goes on to tile k + builders, whose parity wait on the stage's
// "empty" barrier is only unambiguous if tile k + builders - 2*stages has been consumed as well.
constexpr int kWgEdgeBuilders = 6;
constexpr int kWgEdgeThreads = 64 + 32 * kWgEdgeBuilders;
constexpr int kWgEdgeWinW = 18 * 32;         // floats of one builder warp's source window: [18 rows][32 slots, 30 used]

__global__ void __launch_bounds__(kWgEdgeThreads, 1)
wgrad_edge_umma_kernel(const __grid_constant__ CUtensorMap tmA, const WgEdgeParams p) {
    constexpr int A_ATOM = 128 * 128;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    float* win = reinterpret_cast<float*>(smem + p.stages * kWgEdgeStage);  // [builder warp][2][18][32]
    uint64_t* bars = reinterpret_cast<uint64_t*>(win + kWgEdgeBuilders * 2 * kWgEdgeWinW + 8);
    uint64_t* full = bars;
    uint64_t* empty = bars + kWgMaxStages;
    uint64_t* acc_full = bars + 2 * kWgMaxStages;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_full + 1);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int per = (p.m_tiles + static_cast<int>(gridDim.x) - 1) / static_cast<int>(gridDim.x);
    const int t_begin = blockIdx.x * per;
    const int t_end = min(p.m_tiles, t_begin + per);
    const bool has_work = t_begin < t_end;
    const bool do_bias = p.bias_out != nullptr;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmA);
        for (int i = 0; i < kWgMaxStages; ++i) {
            mbar_init(&full[i], 1 + 1);                   // the TMA transaction + the tile's builder warp
            mbar_init(&empty[i], 1);                      // the MMA commit
        }
        mbar_init(acc_full, 1);
        fence_mbar_init();
    }
    // channels 32..63 of every U tile stay zero for the whole kernel (chunks 4..7 of each 128-byte row, in any swizzle)
    for (int i = threadIdx.x; i < p.stages * 128 * 4; i += kWgEdgeThreads) {
        const int st = i / 512, r = (i >> 2) & 127, c = 4 + (i & 3);
        st_shared_v4(smem + st * kWgEdgeStage + A_ATOM + r * 128 + ((c ^ (r & 7)) << 4), 0u, 0u, 0u, 0u);
    }
    fence_proxy_async();
    if (warp == 1) tmem_alloc(tmem_slot, 64);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    pdl_launch_dependents();

    if (warp == 0) {
        if (lane == 0 && has_work) {
            pdl_wait();
            uint32_t stage = 0, phase = 0;
            for (int mt = t_begin; mt < t_end; ++mt) {
                const int tx = mt % p.tiles_x;
                const int r = mt / p.tiles_x;
                const int ty = r % p.tiles_y;
                const int b = r / p.tiles_y;
                mbar_wait(&empty[stage], phase ^ 1);
                mbar_arrive_expect_tx(&full[stage], A_ATOM);
                tma_load_4d(smem + stage * kWgEdgeStage, &tmA, &full[stage], 0, tx * 8, ty * 16, b, kEvictNormal);
                if (++stage == static_cast<uint32_t>(p.stages)) {
                    stage = 0;
                    phase ^= 1;
                }
            }
        }
    } else if (warp == 1) {
        if (has_work) {
            constexpr uint32_t idesc = umma_idesc_bf16_mn(128, 64);
            uint32_t stage = 0, phase = 0;
            for (int mt = t_begin; mt < t_end; ++mt) {
                mbar_wait(&full[stage], phase);
                tc_fence_after();
                const uint32_t sa = smem_u32(smem + stage * kWgEdgeStage);
                if (elect_one()) {
#pragma unroll
                    for (int k8 = 0; k8 < 8; ++k8)  // K = 16 pixels = tile rows 2*k8, 2*k8+1
                        umma_bf16_ss(tmem_base, umma_desc_mn_sw128(sa + k8 * 2048, 0, 1024),
                                     umma_desc_mn_sw128(sa + A_ATOM + k8 * 2048, 0, 1024), idesc, (mt == t_begin && k8 == 0) ? 0u : 1u);
                    umma_commit(&empty[stage]);
                }
                __syncwarp();
                if (++stage == static_cast<uint32_t>(p.stages)) {
                    stage = 0;
                    phase ^= 1;
                }
            }
            if (elect_one()) umma_commit(acc_full);
            __syncwarp();
        }
    } else if (has_work) {
        // ======================= U-tile builders (then bias reducers, then the epilogue) =======================
        // Every builder warp owns every 6th tile of the CTA and its own pair of source windows, so six tiles are being
        // built at any time (one warp's chain window -> gather -> fence -> arrive is ~1.5 us of latencies; 128 threads on
        // one tile at a time made the kernel builder-bound at 1.9 us per tile).
        const int bw = warp - 2;
        const int q = warp & 3;
        float* wwin = win + bw * (2 * kWgEdgeWinW);  // two windows [18 rows][32 slots]
        // window slot of this lane: NHWC source: slot = wx*3 + ci (30 contiguous floats per row); NCHW: slot = ci*10 + wx
        const int wx_l = p.v3_nchw ? lane % 10 : lane / 3, ci_l = p.v3_nchw ? lane / 10 : lane % 3;
        const int sx = p.v3_nchw ? 1 : 3;
        auto load_window = [&](int mt, float (&w)[18]) {
            const int tx = mt % p.tiles_x;
            const int r0 = mt / p.tiles_x;
            const int ty = r0 % p.tiles_y;
            const int b = r0 / p.tiles_y;
            const int xx = tx * 8 - 1 + wx_l;
            const bool col_ok = lane < 30 && mt < t_end && b < p.B && xx >= 0 && xx < p.W;
#pragma unroll
            for (int k = 0; k < 18; ++k) {
                const int yy = ty * 16 - 1 + k;
                float v = 0.f;
                if (col_ok && yy >= 0 && yy < p.H)
                    v = p.v3_nchw ? __ldg(p.v3 + ((static_cast<size_t>(b) * 3 + ci_l) * p.H + yy) * p.W + xx)
                                  : __ldg(p.v3 + ((static_cast<size_t>(b) * p.H + yy) * p.W + xx) * 3 + ci_l);
                w[k] = v;
            }
        };
        pdl_wait();  // the 3-channel source is an earlier kernel's output
        float wn[18];
        load_window(t_begin + bw, wn);
        int buf = 0;
        for (int kt = bw; t_begin + kt < t_end; kt += kWgEdgeBuilders, buf ^= 1) {
            const int mt = t_begin + kt;
            const uint32_t stage = static_cast<uint32_t>(kt % p.stages), phase = static_cast<uint32_t>(kt / p.stages) & 1u;
            float* wb = wwin + buf * kWgEdgeWinW;
#pragma unroll
            for (int k = 0; k < 18; ++k) wb[k * 32 + lane] = wn[k];
            load_window(mt + kWgEdgeBuilders, wn);  // in flight while this tile is gathered
            __syncwarp();
            mbar_wait(&empty[stage], phase ^ 1);
            uint8_t* ut = smem + stage * kWgEdgeStage + A_ATOM;
#pragma unroll 1
            for (int pass = 0; pass < 4; ++pass) {
                const int et = pass * 32 + lane, yl = et >> 3, xl = et & 7;
                // channel ci*9 + ky*3 + kx = source pixel + sign*(ky-1, kx-1)
                const float* wp = wb + (yl + 1) * 32 + (xl + 1) * sx;
                float f[32];
                f[27] = 1.f;  // a channel of ones: column 27 of the result = sum_p V64[p][c64] = the bias gradient, for free
#pragma unroll
                for (int i = 28; i < 32; ++i) f[i] = 0.f;
                if (p.v3_nchw) {
                    if (p.sign > 0) wg_edge_gather<1, true>(wp, f); else wg_edge_gather<-1, true>(wp, f);
                } else {
                    if (p.sign > 0) wg_edge_gather<1, false>(wp, f); else wg_edge_gather<-1, false>(wp, f);
                }
                uint8_t* ub = ut + et * 128;
#pragma unroll
                for (int c = 0; c < 4; ++c)
                    st_shared_v4(ub + ((c ^ (et & 7)) << 4), pack_bf16x2(f[8 * c], f[8 * c + 1]), pack_bf16x2(f[8 * c + 2], f[8 * c + 3]),
                                 pack_bf16x2(f[8 * c + 4], f[8 * c + 5]), pack_bf16x2(f[8 * c + 6], f[8 * c + 7]));
            }
            fence_proxy_async();  // generic-proxy writes -> visible to the tensor core
            __syncwarp();
            if (lane == 0) mbar_arrive(&full[stage]);
        }
        mbar_wait(acc_full, 0);
        tc_fence_after();
        if (q < 2 && bw < 4) {  // (warps 4, 5) accumulator rows 0..63 = the 64 channels of V64; rows 64..127 repeat them
            const int c64 = q * 32 + lane;
            uint32_t v[32];
            tmem_ld_32x32(tmem_base + (static_cast<uint32_t>(q * 32) << 16), v);
            tmem_ld_wait();
#pragma unroll
            for (int j = 0; j < 27; ++j) {
                const int ci = j / 9, tap = j - ci * 9;
                float* dst = p.out_mode ? p.out + (ci * 64 + c64) * 9 + tap : p.out + (c64 * 3 + ci) * 9 + tap;
                red_add_f32(dst, __uint_as_float(v[j]));
            }
            if (do_bias) red_add_f32(p.bias_out + c64, __uint_as_float(v[27]));
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 64);
    }
}


}  // namespace cid
