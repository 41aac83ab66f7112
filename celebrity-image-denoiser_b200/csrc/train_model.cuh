// The denoise-GAN training step (BASELINE config #5): generator + discriminator forward, both
// backward passes, Adam - the work of one iteration of DenoiseGANTrainer.train
// (/root/reference/backend/trainingcode/denoise_gan_code/training.py:403-427), split into the phases a data-parallel
// caller needs (gradient all-reduce between backward and Adam):
//
//   cid_train_d_backward : denoised = G(noisy); d_loss = BCE(D(clean),1) + BCE(D(denoised.detach()),0); D gradients
//   cid_train_adam(D)    : d_optimizer.step()  (+ re-pack of the bf16 tensor-core operands)
//   cid_train_g_backward : g_loss = MSE(denoised, clean) + 1e-3 * BCE(D(denoised),1); G gradients
//   cid_train_adam(G)    : g_optimizer.step()
//
// Precision: activations and activation gradients are bf16 NHWC, accumulation fp32 (TMEM), master weights, weight
// gradients, Adam moments and BatchNorm statistics fp32 (the state buffer the caller owns).
// Kernels: forward = the inference kernels (conv_umma.cuh ...); data gradients = the same conv kernel on tap-flipped,
// transposed weight packs (EPI_DGRAD), stride-2 layers through the 4x4-stride-2 modes with zero-padded taps; weight
// gradients = wgrad_umma.cuh; everything memory-bound = train_kernels.cuh.
#pragma once
#include "train_kernels.cuh"
#include "wgrad_umma.cuh"

namespace cid {

// ---------------------------------------------------------------------------------------------------------------------
// parameter table
// ---------------------------------------------------------------------------------------------------------------------
struct TrainTensor {
    std::string key;  // "generator.<state_dict key>" / "discriminator.<state_dict key>"
    size_t off, n;    // offset / element count inside one section of the state buffer
    bool buffer;      // BatchNorm running statistics: live in the buffer section, not trained
};

enum { DL_CONV0 = 0, DL_CONV1 = 1, DL_CONV2 = 2, DL_CONV3 = 3 };
static const int kDCin[4] = {3, 64, 64, 128};
static const int kDCout[4] = {64, 64, 128, 128};
static const int kDIdx[4] = {0, 2, 5, 8};      // nn.Sequential indices of the convs
static const int kDBnIdx[4] = {-1, 3, 6, 9};   // ... and of the BatchNorms that follow them

// training workspace buffers that follow the generator's activation buffers
enum TBuf {
    TB_Y, TB_GZ, TB_GX,                                                             // fp32, 3 channels, level 0
    TB_DZ10, TB_GCAT1, TB_DZ8, TB_DZ7, TB_GCAT2, TB_DZ5, TB_DZ4, TB_GP2, TB_DZ3, TB_DZ2, TB_GP1, TB_DZ1, TB_DZ0,
    TB_DA0, TB_DZ1D, TB_DA1, TB_DZ2D, TB_DA2, TB_DZ3D, TB_DA3, TB_DG3, TB_DG2, TB_DG1, TB_DG0,
    TB_GZC,  // gz as fp32 NCHW
    TB_COUNT
};
static const int kTBufC[TB_COUNT] = {3, 3, 3, 64, 128, 128, 128, 256, 256, 256, 128, 128, 128, 64, 64, 64,
                                     64, 64, 64, 128, 128, 128, 128, 128, 128, 64, 64, 3};
static const int kTBufLevel[TB_COUNT] = {0, 0, 0, 0, 0, 1, 1, 1, 2, 2, 2, 1, 1, 1, 0, 0,
                                         0, 1, 1, 1, 1, 2, 2, 2, 1, 1, 0, 0};

}  // namespace cid

struct cid_train {
    cid_ctx* g = nullptr;  // generator inference context: forward weight packs + the forward launch path
    int device = 0, num_sms = 0, max_smem = 0;
    float* state = nullptr;
    size_t state_floats = 0;
    std::vector<cid::TrainTensor> tensors;
    size_t n_g = 0, n_d = 0, n_buf = 0;
    bool g_ready = false;
    // bf16 operand packs owned by the library
    void* gd_w[12] = {nullptr};       // generator data-gradient packs (layers 1..10; 11 = first-conv operand of upconv1.2's)
    CUtensorMap gd_tm[12];
    CUtensorMap gd_tmh[12];           // same packs, box of N/2 rows (CTA-pair kernels load half a weight tile each)
    int gd_n[12] = {0};               // tile width N of the pack's tensor map
    void* df_w[4] = {nullptr};        // discriminator forward packs (conv1..3)
    void* dd_w[4] = {nullptr};        // discriminator data-gradient packs (conv1..3)
    CUtensorMap df_tm[4], df_tmh[4], dd_tm[4], dd_tmh[4];  // (tmh: half-height weight boxes of the CTA-pair launches)
    float* zero_bias = nullptr;       // 512 zeros
    float* d_bias_pad[4] = {nullptr}; // fp32 conv biases padded for float4 epilogue loads
    double* dsc = nullptr;            // double scratch: BN sums, backward sums, losses
    float* fsc = nullptr;             // float scratch: BN stats, pooled, dpool, probs
    std::map<std::string, CUtensorMap> maps;
    int64_t step_g = 0, step_d = 0;
    // data-parallel peers (cid_train_set_peers): symmetric state / flag buffers of every rank, mapped into this process
    int peer_world = 0, peer_rank = 0, peer_epoch = 0;
    const float* peer_state[8] = {nullptr};
    int* peer_flags[8] = {nullptr};
    int* peer_err = nullptr;          // device int, sticky: a peer barrier timed out (the fused optimiser then does nothing)
    int* peer_err_host = nullptr;     // its pinned, device-mapped host mirror: read by the host without synchronising
    int* peer_err_host_dev = nullptr; // the mirror's device address
    unsigned long long peer_spin_ns = 600ull * 1000000000ull;  // = kPeerSpinNsDefault (train_kernels.cuh)
    int64_t launches = 0;
    int B = 0, H = 0, W = 0;          // shape of the last d_backward (g_backward continues from it)
    const void* ws = nullptr;
};

namespace cid {

// double scratch layout
constexpr int kBwdReps = 8;  // replicas of the BatchNorm-backward sums (see bn_bwd_reduce_kernel)
enum { DS_BN = 0 /* 3 layers x 2 x 128 */, DS_BWD = 768 /* 3 x kBwdReps x 2 x 128 */, DS_LOSS = 768 + 3 * kBwdReps * 256 /* 8 */, DS_COUNT = DS_LOSS + 8 };
// float scratch layout
enum { FS_STATS = 0 /* 3 x 256 */, FS_POOLED = 768 /* B*128 */, FS_DPOOL = 768 + 128 * 1024, FS_PROBS = 768 + 2 * 128 * 1024,
       FS_COUNT = FS_PROBS + 1024 };
constexpr int kTrainMaxBatch = 1024;

static void train_build_table(cid_train* t) {
    t->tensors.clear();
    size_t off = 0;
    auto add = [&](const std::string& key, size_t n) {
        t->tensors.push_back({key, off, n, false});
        off += n;
    };
    for (int l = 0; l < 12; ++l) {
        const LayerDesc& L = kLayers[l];
        const size_t taps = L.kind == L_CONVT ? 4 : 9;
        add(std::string("generator.") + L.name + ".weight", static_cast<size_t>(L.cin) * L.cout * taps);
        add(std::string("generator.") + L.name + ".bias", L.cout);
    }
    t->n_g = off;
    for (int l = 0; l < 4; ++l) {
        const std::string base = "discriminator.model." + std::to_string(kDIdx[l]);
        add(base + ".weight", static_cast<size_t>(kDCin[l]) * kDCout[l] * 9);
        add(base + ".bias", kDCout[l]);
        if (kDBnIdx[l] >= 0) {
            const std::string bn = "discriminator.model." + std::to_string(kDBnIdx[l]);
            add(bn + ".weight", kDCout[l]);
            add(bn + ".bias", kDCout[l]);
        }
    }
    add("discriminator.model.12.weight", 128);
    add("discriminator.model.12.bias", 1);
    t->n_d = off - t->n_g;
    size_t boff = 0;
    for (int l = 1; l < 4; ++l) {
        const std::string bn = "discriminator.model." + std::to_string(kDBnIdx[l]);
        t->tensors.push_back({bn + ".running_mean", boff, static_cast<size_t>(kDCout[l]), true});
        boff += kDCout[l];
        t->tensors.push_back({bn + ".running_var", boff, static_cast<size_t>(kDCout[l]), true});
        boff += kDCout[l];
        t->tensors.push_back({bn + ".num_batches_tracked", boff, 1, true});
        boff += 1;
    }
    t->n_buf = (boff + 3) & ~static_cast<size_t>(3);
}

static const TrainTensor* train_find(const cid_train* t, const std::string& key) {
    for (const auto& x : t->tensors)
        if (x.key == key) return &x;
    return nullptr;
}
static size_t train_P(const cid_train* t) { return t->n_g + t->n_d; }
// section: 0 params, 1 grads, 2 Adam m, 3 Adam v; buffers follow the four sections
static float* train_ptr(const cid_train* t, const std::string& key, int section) {
    const TrainTensor* x = train_find(t, key);
    if (!x) return nullptr;
    if (x->buffer) return t->state + 4 * train_P(t) + x->off;
    return t->state + section * train_P(t) + x->off;
}

// ---------------------------------------------------------------------------------------------------------------------
// workspace
// ---------------------------------------------------------------------------------------------------------------------
// fp32: the 3-channel image-side tensors and the discriminator's pre-BatchNorm conv outputs (mean >> std on smooth
// images: bf16 storage would cost ~3 % of a standard deviation per layer in the normalised value)
static size_t tbuf_elem(int b) { return (b <= TB_GX || b == TB_DZ1D || b == TB_DZ2D || b == TB_DZ3D || b == TB_GZC) ? 4 : 2; }
static size_t tbuf_bytes(int b, int B, int H, int W) {
    const int lv = kTBufLevel[b];
    const size_t es = tbuf_elem(b);
    const size_t n = static_cast<size_t>(B) * (H >> lv) * (W >> lv) * kTBufC[b] * es;
    return (n + 1023) & ~static_cast<size_t>(1023);
}
static size_t train_layout(int B, int H, int W, size_t* toff) {
    size_t cur = layout_offsets(B, H, W, nullptr);
    for (int i = 0; i < TB_COUNT; ++i) {
        if (toff) toff[i] = cur;
        cur += tbuf_bytes(i, B, H, W);
    }
    return cur;
}

// ---------------------------------------------------------------------------------------------------------------------
// tensor maps (cached by geometry)
// ---------------------------------------------------------------------------------------------------------------------
static int make_act_map_s2(CUtensorMap* tm, const void* base, int B, int H, int W, int C, int box_rows, int box_px) {
    EncodeTiledFn fn = get_encode_fn();
    CID_REQUIRE(fn, "cuTensorMapEncodeTiled entry point not available");
    cuuint64_t dims[4] = {static_cast<cuuint64_t>(C), static_cast<cuuint64_t>(W), static_cast<cuuint64_t>(H),
                          static_cast<cuuint64_t>(B)};
    cuuint64_t strides[3] = {static_cast<cuuint64_t>(C) * 2, static_cast<cuuint64_t>(W) * C * 2,
                             static_cast<cuuint64_t>(H) * W * C * 2};
    cuuint32_t box[4] = {64, static_cast<cuuint32_t>(2 * box_px), static_cast<cuuint32_t>(2 * box_rows), 1};
    cuuint32_t estr[4] = {1, 2, 2, 1};
    CUresult r = fn(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(base), dims, strides, box, estr,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    CID_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled(stride-2 %dx%dx%dx%d box %dx%d) failed: %d", B, H, W, C, box_rows,
                box_px, static_cast<int>(r));
    return 0;
}

// fp32 NHWC [B,H,W,C]: box = 32 channels (one 128-byte swizzle row) x 8 pixels x 16 rows - the staged output tile of
// the discriminator's pre-BatchNorm convolutions (EPI_LRELU_BN)
static int make_act_map_f32(CUtensorMap* tm, const void* base, int B, int H, int W, int C) {
    EncodeTiledFn fn = get_encode_fn();
    CID_REQUIRE(fn, "cuTensorMapEncodeTiled entry point not available");
    cuuint64_t dims[4] = {static_cast<cuuint64_t>(C), static_cast<cuuint64_t>(W), static_cast<cuuint64_t>(H),
                          static_cast<cuuint64_t>(B)};
    cuuint64_t strides[3] = {static_cast<cuuint64_t>(C) * 4, static_cast<cuuint64_t>(W) * C * 4,
                             static_cast<cuuint64_t>(H) * W * C * 4};
    cuuint32_t box[4] = {32, 8, 16, 1};
    cuuint32_t estr[4] = {1, 1, 1, 1};
    CUresult r = fn(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, const_cast<void*>(base), dims, strides, box, estr,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    CID_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled(fp32 activation %dx%dx%dx%d) failed: %d", B, H, W, C, static_cast<int>(r));
    return 0;
}

// kind: 0 plain box, 1 stride-2 box, 2 the 5-D 2x-upsampling scatter map (H, W = INPUT resolution), 3 fp32 output tile
static const CUtensorMap* tmap(cid_train* t, const void* base, int B, int H, int W, int C, int rows, int px, int kind) {
    char key[160];
    snprintf(key, sizeof(key), "%p/%d/%d/%d/%d/%d/%d/%d", base, B, H, W, C, rows, px, kind);
    auto it = t->maps.find(key);
    if (it != t->maps.end()) return &it->second;
    CUtensorMap tm;  // (the cache is trimmed only between steps, train_make_run: pointers handed out stay valid)
    int rc = kind == 0   ? make_act_map(&tm, base, B, H, W, C, rows, px)
             : kind == 1 ? make_act_map_s2(&tm, base, B, H, W, C, rows, px)
             : kind == 3 ? make_act_map_f32(&tm, base, B, H, W, C)
                         : make_convt_out_map(&tm, base, B, H, W, C);
    if (rc) return nullptr;
    return &(t->maps[key] = tm);
}

// ---------------------------------------------------------------------------------------------------------------------
// launches
// ---------------------------------------------------------------------------------------------------------------------
template <int N, int TAPS, int S, bool WRES, int EPI, int PW = 0, int TS = 0, int CG = 1>
static int launch_conv_t(cid_train* t, const CUtensorMap& tmA, const CUtensorMap& tmW, const CUtensorMap& tmO, ConvParams p,
                         cudaStream_t st) {
    using Cfg = ConvCfg<N, TAPS, S, WRES, EPI, PW, TS, CG>;
    cid_ctx* c = t->g;
    const int budget = c->max_smem - 1024 - 1024 - Cfg::O_BYTES;
    int w_stages = 1, a_cap = kMaxStages;
    if (!WRES) stage_split<Cfg, N, S>(budget, &a_cap, &w_stages);  // (cid_api.cu: deep weight ring, A ring of ~2 K blocks)
    const int w_bytes = (WRES ? p.kb_count * TAPS : w_stages) * Cfg::W_BYTES;
    int a_stages = std::min(a_cap, (budget - w_bytes) / Cfg::A_BYTES);
    CID_REQUIRE(a_stages >= S + 1, "train conv<%d,%d,%d>: shared memory too small (%d stages)", N, TAPS, S, a_stages);
    p.a_stages = a_stages;
    p.w_stages = w_stages;
    p.tiles_x = (p.W + 7) / 8;
    p.tiles_y = (p.H + 15) / 16;
    p.m_tiles = p.B * p.tiles_x * p.tiles_y;
    {
        auto lg = [](int v) { int k = 0; while ((1 << k) < v) ++k; return (1 << k) == v ? k : -1; };
        p.tx_sh = lg(p.tiles_x);
        p.ty_sh = lg(p.tiles_y);
        if (p.tx_sh < 0 || p.ty_sh < 0) p.tx_sh = p.ty_sh = -1;
    }
    p.num_work = ((p.m_tiles + S * CG - 1) / (S * CG)) * p.n_tiles;
    const int smem = Cfg::smem_bytes(p.kb_count, a_stages, w_stages);
    auto kern = conv_umma_kernel<N, TAPS, S, WRES, EPI, PW, TS, CG>;
    static uint64_t attr_set = 0;  // bit per device: function attributes are per device
    if (!(attr_set >> (c->device & 63) & 1)) {
        CID_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, c->max_smem));
        attr_set |= 1ull << (c->device & 63);
    }
    const int grid = std::min(p.num_work * CG, c->num_sms / CG * CG);
    CID_CHECK_CUDA(launch_ex(c, kern, grid, kConvThreads, smem, st, CG, tmA, tmW, tmO, tmO, p));
    CID_CHECK_CUDA(cudaGetLastError());
    t->launches++;
    return 0;
}

template <int NA, int NB, int T, int CL = 1>
static int launch_wgrad(cid_train* t, const CUtensorMap& tmA, const CUtensorMap& tmB, WgParams p, cudaStream_t st) {
    p.tiles_x = (p.W + 7) / 8;
    p.tiles_y = (p.H + 15) / 16;
    p.m_tiles = p.B * p.tiles_x * p.tiles_y;
    CID_REQUIRE(p.n_roles >= 1 && p.n_roles <= kWgMaxRoles && p.n_roles % CL == 0, "wgrad: %d roles", p.n_roles);
    p.m_valid = 64 * NA;
    const int pb = (p.pw * (CL == 3 ? 18 : 16) * 128 + 1023) & ~1023;
    const int stage_bytes = NA * 16384 + NB * pb;
    p.stages = std::min(kWgMaxStages, (t->max_smem - 2048) / stage_bytes);
    CID_REQUIRE(p.stages >= 2, "wgrad: shared memory too small");
    // > half of the SM's shared memory: one CTA per SM (each CTA allocates up to all 512 TMEM columns)
    const int smem = std::max(p.stages * stage_bytes + 1024 + 256, 120 * 1024);
    auto kern = wgrad_umma_kernel<NA, NB, T, CL>;
    static uint64_t attr_set = 0;  // bit per device
    static int max_units = 0;      // co-resident CTAs (CL = 1) or clusters (CL = 3); same on every B200 of the box
    if (!(attr_set >> (t->device & 63) & 1)) {
        CID_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, t->max_smem));
        max_units = t->num_sms;
        if (CL > 1) {
            cudaLaunchConfig_t cfg;
            memset(&cfg, 0, sizeof(cfg));
            cfg.gridDim = dim3(t->num_sms / CL * CL);
            cfg.blockDim = dim3(kWgThreads);
            cfg.dynamicSmemBytes = t->max_smem;
            cudaLaunchAttribute at;
            at.id = cudaLaunchAttributeClusterDimension;
            at.val.clusterDim.x = CL; at.val.clusterDim.y = 1; at.val.clusterDim.z = 1;
            cfg.attrs = &at;
            cfg.numAttrs = 1;
            int n = 0;
            CID_CHECK_CUDA(cudaOccupancyMaxActiveClusters(&n, kern, &cfg));
            CID_REQUIRE(n >= 1, "wgrad: no cluster of %d CTAs fits", CL);
            max_units = n;
        }
        attr_set |= 1ull << (t->device & 63);
    }
    const int n_units = p.n_roles / CL;
    p.slices = std::max(1, std::min(max_units / n_units, p.m_tiles));
    const int grid = n_units * p.slices * CL;
    if (CL > 1) {
        cudaLaunchConfig_t cfg;
        memset(&cfg, 0, sizeof(cfg));
        cfg.gridDim = dim3(grid);
        cfg.blockDim = dim3(kWgThreads);
        cfg.dynamicSmemBytes = smem;
        cfg.stream = st;
        cudaLaunchAttribute at;
        at.id = cudaLaunchAttributeClusterDimension;
        at.val.clusterDim.x = CL; at.val.clusterDim.y = 1; at.val.clusterDim.z = 1;
        cfg.attrs = &at;
        cfg.numAttrs = 1;
        CID_CHECK_CUDA(cudaLaunchKernelEx(&cfg, kern, tmA, tmB, p));
    } else {
        launch_k(t->g->pdl, kern, grid, kWgThreads, smem, st, tmA, tmB, p);
    }
    CID_CHECK_CUDA(cudaGetLastError());
    t->launches++;
    return 0;
}

// geometry of a weight gradient: 0 = 3x3 stride 1, 1 = 3x3 stride 2 (operand B = the 2x larger layer input),
// 2 = ConvTranspose k2 s2 (operand A = the layer input, operand B = the 2x larger output gradient)
static void wgrad_roles(WgParams& p, int geom, int m_ch, int n_ch, int NA, int NB) {
    p.n_roles = 0;
    for (int mb = 0; mb < m_ch / (64 * NA); ++mb)
        for (int nb = 0; nb < n_ch / (64 * NB); ++nb) {
            bool first_of_block = true;
            auto push = [&](int oy, int ox, int t0, int t1, int t2) {
                WgRole& r = p.roles[p.n_roles++];
                r.a_c0 = r.m0 = mb * 64 * NA;
                r.b_c0 = r.n0 = nb * 64 * NB;
                r.oy = oy; r.ox = ox;
                r.tap[0] = t0; r.tap[1] = t1; r.tap[2] = t2;
                r.bias = (nb == 0 && first_of_block) ? 1 : 0;
                first_of_block = false;
            };
            if (geom == 0) {
                for (int ky = 0; ky < 3; ++ky) push(ky - 1, -1, ky * 3, ky * 3 + 1, ky * 3 + 2);
            } else if (geom == 1) {
                for (int ky = 0; ky < 3; ++ky) {
                    push(ky - 1, -1, ky * 3 + 0, ky * 3 + 2, -1);  // even columns of the 2x input: kx = 0 (px 0), kx = 2 (px 1)
                    push(ky - 1, 0, ky * 3 + 1, -1, -1);           // odd columns: kx = 1
                }
            } else {
                for (int q = 0; q < 4; ++q) push(q >> 1, q & 1, q, -1, -1);
            }
        }
    p.scale = geom == 0 ? 1 : 2;
    p.pw = geom == 0 ? 10 : geom == 1 ? 9 : 8;
    p.kk = geom == 2 ? 4 : 9;
    p.s_m = n_ch;
}

// dW += wgrad. A: [B,H,W,a_C] (tile operand, m_ch channels used from channel 0), Bt: operand with the taps
// ([B,H*scale,W*scale,b_C], n_ch channels used from channel 0).
static int wgrad(cid_train* t, int geom, const void* A, int a_C, int m_ch, const void* Bt, int b_C, int n_ch, int B, int H,
                 int W, float* out, cudaStream_t st, float* bias_out = nullptr) {
    WgParams p;
    memset(&p, 0, sizeof(p));
    p.B = B; p.H = H; p.W = W;
    p.out = out;
    p.bias_out = bias_out;
    p.fuse_taps = geom == 0 ? 1 : 0;  // 3x3 stride-1 layers: the three taps of a patch row as ONE N = 192 MMA (6.34 -> 6.23 ms)
    const int NA = m_ch >= 128 ? 2 : 1, NB = n_ch >= 128 ? 2 : 1;
    wgrad_roles(p, geom, m_ch, n_ch, NA, NB);
    const CUtensorMap* tmA = tmap(t, A, B, H, W, a_C, 16, 8, 0);
    const CUtensorMap* tmB = geom == 0 ? tmap(t, Bt, B, H, W, b_C, 16, 10, 0) : tmap(t, Bt, B, 2 * H, 2 * W, b_C, 16, p.pw, 1);
    if (!tmA || !tmB) return 1;
    const int T = geom == 0 ? 3 : geom == 1 ? 2 : 1;
#define CID_WG(na, nb, tt) \
    if (NA == na && NB == nb && T == tt) return launch_wgrad<na, nb, tt>(t, *tmA, *tmB, p, st);
    CID_WG(1, 1, 3) CID_WG(1, 2, 3) CID_WG(2, 1, 3) CID_WG(2, 2, 3)
    CID_WG(1, 1, 2) CID_WG(2, 2, 2)
    CID_WG(2, 1, 1) CID_WG(2, 2, 1)
#undef CID_WG
    set_error("wgrad: no kernel for NA=%d NB=%d T=%d", NA, NB, T);
    return 1;
}

// pixels per block of the column reductions: ONE wave of `per_sm` resident blocks per SM (per_sm = what the kernel's
// registers allow: a second, partly filled wave costs a whole block time; small tensors were latency-bound with few blocks)
static int reduce_ppb(const cid_train* t, size_t P, int per_sm = 8) {
    const size_t target = static_cast<size_t>(t->num_sms) * per_sm;
    size_t ppb = (P + target - 1) / target;
    ppb = (ppb + 31) / 32 * 32;
    return static_cast<int>(std::min<size_t>(std::max<size_t>(ppb, 64), 4096));
}

static int colsum(cid_train* t, const void* v, size_t P, int cs, int c0, int C, float* out, cudaStream_t st) {
    const int ppb = reduce_ppb(t, P);
    launch_k(t->g->pdl, colsum_kernel<false, __nv_bfloat16>, static_cast<unsigned>((P + ppb - 1) / ppb), 256, 0, st, 
        static_cast<const __nv_bfloat16*>(v), P, cs, c0, C, out, nullptr, nullptr, ppb);
    CID_CHECK_CUDA(cudaGetLastError());
    t->launches++;
    return 0;
}

// weight gradient of a 3-channel layer: out += R[c64][ci*9+tap] = sum_p v64[p][c64] * v3[p + sign*(tap - centre)][ci]
// (wgrad_edge_umma_kernel: the im2col operand is built in shared memory, the result goes straight into the gradient
// tensor; bias_out += column sums of v64)
static int wgrad_edge(cid_train* t, const void* v64, const float* v3, int v3_nchw, int sign, float* out, int out_mode,
                      int B, int H, int W, cudaStream_t st, float* bias_out = nullptr) {
    WgEdgeParams p;
    memset(&p, 0, sizeof(p));
    p.B = B; p.H = H; p.W = W;
    p.tiles_x = (W + 7) / 8;
    p.tiles_y = (H + 15) / 16;
    p.m_tiles = B * p.tiles_x * p.tiles_y;
    p.v3 = v3; p.v3_nchw = v3_nchw; p.sign = sign;
    p.out = out; p.out_mode = out_mode; p.bias_out = bias_out;
    const int extra = kWgEdgeBuilders * 2 * kWgEdgeWinW * 4 + 32 + 256;
    p.stages = std::min(kWgMaxStages, (t->max_smem - 2048 - extra) / kWgEdgeStage);
    CID_REQUIRE(p.stages >= kWgEdgeBuilders, "wgrad_edge: shared memory too small (%d stages for %d builder warps)", p.stages, kWgEdgeBuilders);
    const CUtensorMap* tmA = tmap(t, v64, B, H, W, 64, 16, 8, 0);
    if (!tmA) return 1;
    const int smem = std::max(p.stages * kWgEdgeStage + extra + 1024, 120 * 1024);  // one CTA per SM
    static uint64_t attr_set = 0;  // bit per device
    if (!(attr_set >> (t->device & 63) & 1)) {
        CID_CHECK_CUDA(cudaFuncSetAttribute(wgrad_edge_umma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, t->max_smem));
        attr_set |= 1ull << (t->device & 63);
    }
    const int grid = std::max(1, std::min(t->num_sms, p.m_tiles));
    launch_k(t->g->pdl, wgrad_edge_umma_kernel, grid, kWgEdgeThreads, smem, st, *tmA, p);
    CID_CHECK_CUDA(cudaGetLastError());
    t->launches++;
    return 0;
}

// job list of one repack_multi_kernel launch
struct RepackList {
    RepackJobs jobs;
    RepackList() { jobs.count = 0; jobs.total = 0; }
    int add(int mode, const float* w, void* dst, int cin, int cout, int npad, size_t n, const float* aux = nullptr) {
        CID_REQUIRE(jobs.count < kRepackMaxJobs, "repack: too many jobs");
        RepackJob& J = jobs.job[jobs.count++];
        J.w = w; J.aux = aux; J.dst = dst; J.mode = mode; J.cin = cin; J.cout = cout; J.npad = npad;
        J.n = static_cast<unsigned>(n);
        J.start = jobs.total;
        jobs.total += static_cast<unsigned>(n);
        return 0;
    }
    int launch(cid_train* t, cudaStream_t st) {
        if (!jobs.count) return 0;
        launch_k(t->g->pdl, repack_multi_kernel, (jobs.total + 255) / 256, 256, 0, st, jobs);
        CID_CHECK_CUDA(cudaGetLastError());
        t->launches++;
        return 0;
    }
};

// ---------------------------------------------------------------------------------------------------------------------
// weight packs: fp32 masters (state buffer) -> bf16 operands
// ---------------------------------------------------------------------------------------------------------------------
static int train_alloc_packs(cid_train* t) {
    for (int l = 1; l <= 10; ++l) {
        const LayerDesc& L = kLayers[l];
        const size_t taps = L.kind == L_CONVT ? 4 : 9;
        const size_t n = taps * L.cin * L.cout;
        CID_CHECK_CUDA(cudaMalloc(&t->gd_w[l], n * 2));
        // data-gradient operand: rows = taps * Cin (the gradient's output channels), row length Cout
        t->gd_n[l] = std::min(L.cin, 256);
        if (make_weight_map(&t->gd_tm[l], t->gd_w[l], static_cast<int>(taps) * L.cin, L.cout, t->gd_n[l])) return 1;
        if (make_weight_map(&t->gd_tmh[l], t->gd_w[l], static_cast<int>(taps) * L.cin, L.cout, t->gd_n[l] / 2)) return 1;
    }
    for (int l = 1; l < 4; ++l) {
        const int cin = kDCin[l], cout = kDCout[l];
        const bool s2 = l != DL_CONV2;
        const size_t n = static_cast<size_t>(s2 ? 16 : 9) * cin * cout;
        CID_CHECK_CUDA(cudaMalloc(&t->df_w[l], n * 2));
        CID_CHECK_CUDA(cudaMalloc(&t->dd_w[l], n * 2));
        if (make_weight_map(&t->df_tm[l], t->df_w[l], (s2 ? 16 : 9) * cout, cin, cout)) return 1;
        if (make_weight_map(&t->df_tmh[l], t->df_w[l], (s2 ? 16 : 9) * cout, cin, cout / 2)) return 1;
        if (make_weight_map(&t->dd_tm[l], t->dd_w[l], (s2 ? 16 : 9) * cin, cout, cin)) return 1;
        if (make_weight_map(&t->dd_tmh[l], t->dd_w[l], (s2 ? 16 : 9) * cin, cout, cin / 2)) return 1;
    }
    CID_CHECK_CUDA(cudaMalloc(&t->gd_w[11], kFirstWBytes));     // first_conv_umma operand of upconv1.2's data gradient
    CID_CHECK_CUDA(cudaMalloc(&t->df_w[0], kFirstWBytes));      // first_conv_umma operand of model.0
    CID_CHECK_CUDA(cudaMalloc(&t->dd_w[0], 128 * 64 * 2));      // last_conv_umma operand of model.0's data gradient
    if (make_weight_map(&t->dd_tm[0], t->dd_w[0], 128, 64, 128)) return 1;
    for (int l = 0; l < 4; ++l) {
        CID_CHECK_CUDA(cudaMalloc(reinterpret_cast<void**>(&t->d_bias_pad[l]), 256 * 4));
        CID_CHECK_CUDA(cudaMemset(t->d_bias_pad[l], 0, 256 * 4));
    }
    CID_CHECK_CUDA(cudaMalloc(reinterpret_cast<void**>(&t->zero_bias), 512 * 4));
    CID_CHECK_CUDA(cudaMemset(t->zero_bias, 0, 512 * 4));
    CID_CHECK_CUDA(cudaMalloc(reinterpret_cast<void**>(&t->dsc), DS_COUNT * 8));
    CID_CHECK_CUDA(cudaMemset(t->dsc, 0, DS_COUNT * 8));
    CID_CHECK_CUDA(cudaMalloc(reinterpret_cast<void**>(&t->fsc), FS_COUNT * 4));
    CID_CHECK_CUDA(cudaMemset(t->fsc, 0, FS_COUNT * 4));
    return 0;
}

// which: bit 0 = generator, bit 1 = discriminator
static int train_repack(cid_train* t, int which, cudaStream_t st) {
    cid_ctx* c = t->g;
    if (which & 1) {
        const bool first_time = !t->g_ready;
        if (first_time) {
            // first time: build the forward packs, biases and TMA maps through the inference path (host pack)
            CID_CHECK_CUDA(cudaStreamSynchronize(st));
            for (int l = 0; l < 12; ++l) {
                const LayerDesc& L = kLayers[l];
                for (int wb = 0; wb < 2; ++wb) {
                    const std::string k = std::string(L.name) + (wb ? ".bias" : ".weight");
                    const TrainTensor* x = train_find(t, "generator." + k);
                    int64_t shape[4];
                    int nd = 1;
                    if (wb) {
                        shape[0] = L.cout;
                    } else if (L.kind == L_CONVT) {
                        shape[0] = L.cin; shape[1] = L.cout; shape[2] = 2; shape[3] = 2; nd = 4;
                    } else {
                        shape[0] = L.cout; shape[1] = L.cin; shape[2] = 3; shape[3] = 3; nd = 4;
                    }
                    if (cid_set_param(c, k.c_str(), t->state + x->off, shape, nd)) return 1;
                }
            }
            if (cid_finalize(c)) return 1;
            t->g_ready = true;
        }
        RepackList rl;
        if (!first_time) {
            for (int l = 0; l < 12; ++l) {
                const LayerDesc& L = kLayers[l];
                const float* w = train_ptr(t, std::string("generator.") + L.name + ".weight", 0);
                const float* b = train_ptr(t, std::string("generator.") + L.name + ".bias", 0);
                if (l == 0) {
                    if (rl.add(RP_FIRST, w, c->w_dev[0], 3, 64, 64, kFirstWBytes / 2, b)) return 1;
                } else if (l == 11) {
                    if (rl.add(RP_LAST, w, c->w_dev[11], 64, 3, 128, 128 * 64)) return 1;
                } else if (L.kind == L_CONV) {
                    if (rl.add(RP_CONV9_FWD, w, c->w_dev[l], L.cin, L.cout, L.cout, static_cast<size_t>(9) * L.cin * L.cout)) return 1;
                } else {
                    if (rl.add(RP_CONVT_FWD, w, c->w_dev[l], L.cin, L.cout, L.cout, static_cast<size_t>(4) * L.cin * L.cout)) return 1;
                }
                if (rl.add(RP_COPY_F32, b, c->bias_dev[l], 0, 0, 0, L.cout)) return 1;
            }
        }
        for (int l = 1; l <= 10; ++l) {
            const LayerDesc& L = kLayers[l];
            const float* w = train_ptr(t, std::string("generator.") + L.name + ".weight", 0);
            const size_t taps = L.kind == L_CONVT ? 4 : 9;
            if (rl.add(L.kind == L_CONVT ? RP_CONVT_DGRAD : RP_CONV9_DGRAD, w, t->gd_w[l], L.cin, L.cout, L.cout,
                       taps * L.cin * L.cout))
                return 1;
        }
        if (rl.add(RP_FIRST_DGRAD, train_ptr(t, "generator.upconv1.2.weight", 0), t->gd_w[11], 64, 3, 64, kFirstWBytes / 2)) return 1;
        if (rl.launch(t, st)) return 1;
    }
    if (which & 2) {
        RepackList rl;
        for (int l = 0; l < 4; ++l) {
            const std::string base = "discriminator.model." + std::to_string(kDIdx[l]);
            const float* w = train_ptr(t, base + ".weight", 0);
            const float* b = train_ptr(t, base + ".bias", 0);
            if (rl.add(RP_COPY_F32, b, t->d_bias_pad[l], 0, 0, 0, kDCout[l])) return 1;
            if (l == 0) {
                if (rl.add(RP_FIRST, w, t->df_w[0], 3, 64, 64, kFirstWBytes / 2, b)) return 1;
                if (rl.add(RP_LAST_DGRAD, w, t->dd_w[0], 3, 64, 64, 128 * 64)) return 1;
                continue;
            }
            const int cin = kDCin[l], cout = kDCout[l];
            const bool s2 = l != DL_CONV2;
            const size_t n = static_cast<size_t>(s2 ? 16 : 9) * cin * cout;
            if (rl.add(s2 ? RP_K4_FWD : RP_CONV9_FWD, w, t->df_w[l], cin, cout, cout, n)) return 1;
            if (rl.add(s2 ? RP_T4_DGRAD : RP_CONV9_DGRAD, w, t->dd_w[l], cin, cout, cout, n)) return 1;
        }
        if (rl.launch(t, st)) return 1;
    }
    return 0;
}

// ---------------------------------------------------------------------------------------------------------------------
// per-step context
// ---------------------------------------------------------------------------------------------------------------------
struct TrainRun {
    cid_train* t;
    int B, H, W;
    uint8_t* ws;
    size_t goff[B_COUNT];
    size_t toff[TB_COUNT];
    cudaStream_t st;
    void* gbuf(int b) const { return ws + goff[b]; }
    void* tbuf(int b) const { return ws + toff[b]; }
    size_t pixels(int lv) const { return static_cast<size_t>(B) * (H >> lv) * (W >> lv); }
};

static ConvParams conv_base(const TrainRun& r, int lv, int k_ch, int n_ch) {
    ConvParams p;
    memset(&p, 0, sizeof(p));
    p.B = r.B;
    p.H = r.H >> lv;
    p.W = r.W >> lv;
    p.kb_count = k_ch / 64;
    p.n_tiles = 1;
    p.bias = r.t->zero_bias;
    p.cout = n_ch;
    p.w_rows_per_tap = n_ch;
    p.slope = 1.f;
    p.linear = 1;  // (zero bias, slope 1: what the data-gradient launches keep; the forward convs set their own bias and do not read it)
    return p;
}

// data gradient of a 3x3 stride-1 conv: dX[B,h,w,n_ch] = conv(dZ[B,h,w,k_ch], flipped W^T), optionally ReLU-masked
static int dgrad3x3(const TrainRun& r, int lv, const void* dz, int k_ch, const CUtensorMap& tmW, void* dx, int n_ch,
                    const void* mask, int mask_cs, const CUtensorMap* tmWh = nullptr) {
    cid_train* t = r.t;
    const int h = r.H >> lv, w = r.W >> lv;
    const CUtensorMap* tmA = tmap(t, dz, r.B, h, w, k_ch, 18, 10, 0);
    const CUtensorMap* tmO = tmap(t, dx, r.B, h, w, n_ch, 16, 8, 0);
    if (!tmA || !tmO) return 1;
    ConvParams p = conv_base(r, lv, k_ch, n_ch);
    p.out = dx; p.out_cs = n_ch; p.out_c0 = 0;
    p.res2 = mask; p.res2_cs = mask_cs;
    // (resident weights / CTA pairs where the forward's layers use them: measured 7.07 -> 6.96 ms per iteration)
    if (n_ch == 64) {
        // resident weights as the forward's N = 64 layers: 74 KB for K = 64; K = 128 only fits halved, as a CTA pair
        if (k_ch == 64) return launch_conv_t<64, 9, 2, true, EPI_DGRAD, 10, 1>(t, *tmA, tmW, *tmO, p, r.st);
        if (k_ch == 128 && tmWh) return launch_conv_t<64, 9, 2, true, EPI_DGRAD, 10, 1, 2>(t, *tmA, *tmWh, *tmO, p, r.st);
        return launch_conv_t<64, 9, 2, false, EPI_DGRAD, 10, 1>(t, *tmA, tmW, *tmO, p, r.st);
    }
    if (n_ch == 128) {
        if (tmWh) return launch_conv_t<128, 9, 2, false, EPI_DGRAD, 10, 1, 2>(t, *tmA, *tmWh, *tmO, p, r.st);
        return launch_conv_t<128, 9, 2, false, EPI_DGRAD, 10, 1>(t, *tmA, tmW, *tmO, p, r.st);
    }
    if (n_ch == 256) {  // CTA pair: each CTA streams half of every 32 KB weight tile (as the forward's bottleneck layers)
        if (tmWh) return launch_conv_t<256, 9, 1, false, EPI_DGRAD, 10, 1, 2>(t, *tmA, *tmWh, *tmO, p, r.st);
        return launch_conv_t<256, 9, 1, false, EPI_DGRAD, 10, 1>(t, *tmA, tmW, *tmO, p, r.st);
    }
    set_error("dgrad3x3: unsupported width %d", n_ch);
    return 1;
}

// data gradient of ConvTranspose2d k2 s2: dX[B,h,w,n_ch] = 2x2-stride-2 conv of dY[B,2h,2w,(k_ch of dy_C)]
static int dgrad_convt(const TrainRun& r, int lv, const void* dy, int dy_C, int k_ch, const CUtensorMap& tmW, void* dx,
                       int n_ch, const void* mask) {
    cid_train* t = r.t;
    const int h = r.H >> lv, w = r.W >> lv;
    const CUtensorMap* tmA = tmap(t, dy, r.B, 2 * h, 2 * w, dy_C, 16, 8, 1);
    const CUtensorMap* tmO = tmap(t, dx, r.B, h, w, n_ch, 16, 8, 0);
    if (!tmA || !tmO) return 1;
    ConvParams p = conv_base(r, lv, k_ch, n_ch);
    p.out = dx; p.out_cs = n_ch; p.out_c0 = 0;
    p.res2 = mask; p.res2_cs = n_ch;
    if (n_ch == 128) return launch_conv_t<128, 4, 2, false, EPI_DGRAD, 0, 1>(t, *tmA, tmW, *tmO, p, r.st);
    if (n_ch == 256) return launch_conv_t<256, 4, 1, false, EPI_DGRAD, 0, 1>(t, *tmA, tmW, *tmO, p, r.st);
    set_error("dgrad_convt: unsupported width %d", n_ch);
    return 1;
}

// ---------------------------------------------------------------------------------------------------------------------
// discriminator forward (train mode) on an fp32 NCHW image batch; leaves a0, z1..z3, a1..a3, stats, pooled in place
// ---------------------------------------------------------------------------------------------------------------------
static int bn_forward(const TrainRun& r, int l, const void* z, void* a, int lv) {
    cid_train* t = r.t;
    const int C = kDCout[l];
    const size_t P = r.pixels(lv);
    double* sums = t->dsc + DS_BN + (l - 1) * 256;
    float* stats = t->fsc + FS_STATS + (l - 1) * 256;
    const std::string bn = "discriminator.model." + std::to_string(kDBnIdx[l]);
    BnFwd f;
    f.dsum = sums; f.dsq = sums + 128; f.n = static_cast<double>(P); f.eps = 1e-5f; f.momentum = 0.1f; f.stats = stats;
    f.running_mean = train_ptr(t, bn + ".running_mean", 0); f.running_var = train_ptr(t, bn + ".running_var", 0);
    f.nbt = train_ptr(t, bn + ".num_batches_tracked", 0);
    f.gamma = train_ptr(t, bn + ".weight", 0); f.beta = train_ptr(t, bn + ".bias", 0);
    const size_t n8 = P * C / 8;
    launch_k(t->g->pdl, bn_apply_kernel<float>, static_cast<unsigned>((n8 + 256 * kBnIter - 1) / (256 * kBnIter)), 256, 0, r.st, 
        static_cast<const float*>(z), static_cast<__nv_bfloat16*>(a), P, C, f, 0.2f);
    CID_CHECK_CUDA(cudaGetLastError());
    t->launches += 1;
    return 0;
}

static int d_forward(const TrainRun& r, const float* img) {
    cid_train* t = r.t;
    const int B = r.B, H = r.H, W = r.W;
    {  // model.0 + LeakyReLU(0.2): the generator's first-conv kernel (hi/lo-split input and weights) with slope 0.2
        const CUtensorMap* tmO = tmap(t, r.tbuf(TB_DA0), B, H, W, 64, 16, 8, 0);
        if (!tmO) return 1;
        FirstParams fp;
        memset(&fp, 0, sizeof(fp));
        fp.B = B; fp.H = H; fp.W = W;
        fp.in = img;
        fp.out = static_cast<__nv_bfloat16*>(r.tbuf(TB_DA0));
        fp.tiles_x = (W + 7) / 8;
        fp.tiles_y = (H + 15) / 16;
        fp.m_tiles = B * fp.tiles_x * fp.tiles_y;
        fp.w_packed = static_cast<const uint4*>(t->df_w[0]);
        fp.slope = 0.2f;
        CID_CHECK_CUDA(cudaFuncSetAttribute(first_conv_umma_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kFirstSmemMask));
        const int grid = std::min(fp.m_tiles, t->num_sms * kFirstCtasPerSm);
        CID_CHECK_CUDA(launch_ex(t->g, first_conv_umma_kernel<false>, grid, kFirstThreads, kFirstSmem, r.st, 1, *tmO, fp));
        CID_CHECK_CUDA(cudaGetLastError());
        t->launches++;
    }
    {  // model.2: 3x3 stride 2 (as 4x4 stride 2 with a zero 4th tap row / column) -> z1
        const CUtensorMap* tmA = tmap(t, r.tbuf(TB_DA0), B, H, W, 64, 17, 9, 1);
        const CUtensorMap* tmO = tmap(t, r.tbuf(TB_DZ1D), B, H >> 1, W >> 1, 64, 16, 8, 3);  // fp32 output tile
        if (!tmA || !tmO) return 1;
        ConvParams p = conv_base(r, 1, 64, 64);
        p.bias = t->d_bias_pad[1];
        p.out = r.tbuf(TB_DZ1D); p.out_cs = 64;
        p.out_f32 = static_cast<float*>(r.tbuf(TB_DZ1D));
        p.stats = t->dsc + DS_BN + 0 * 256;  // batch statistics reduced in the epilogue
        CID_CHECK_CUDA(cudaMemsetAsync(p.stats, 0, 256 * 8, r.st));
        if (launch_conv_t<64, 16, 2, false, EPI_LRELU_BN, 9, 1, 2>(t, *tmA, t->df_tmh[1], *tmO, p, r.st)) return 1;  // CTA pair
    }
    if (bn_forward(r, 1, r.tbuf(TB_DZ1D), r.tbuf(TB_DA1), 1)) return 1;
    {  // model.5: 3x3 64 -> 128 -> z2
        const CUtensorMap* tmA = tmap(t, r.tbuf(TB_DA1), B, H >> 1, W >> 1, 64, 18, 10, 0);
        const CUtensorMap* tmO = tmap(t, r.tbuf(TB_DZ2D), B, H >> 1, W >> 1, 128, 16, 8, 3);  // fp32 output tile
        if (!tmA || !tmO) return 1;
        ConvParams p = conv_base(r, 1, 64, 128);
        p.bias = t->d_bias_pad[2];
        p.out = r.tbuf(TB_DZ2D); p.out_cs = 128;
        p.out_f32 = static_cast<float*>(r.tbuf(TB_DZ2D));
        p.stats = t->dsc + DS_BN + 1 * 256;  // batch statistics reduced in the epilogue
        CID_CHECK_CUDA(cudaMemsetAsync(p.stats, 0, 256 * 8, r.st));
        // CTA pair (as the generator's same-shaped down2.0): with K = 64 a weight tile feeds only 8 MMAs - the single-CTA form
        // waited on weight tiles (tensor pipe 46 % busy)
        if (launch_conv_t<128, 9, 2, false, EPI_LRELU_BN, 10, 1, 2>(t, *tmA, t->df_tmh[2], *tmO, p, r.st)) return 1;
    }
    if (bn_forward(r, 2, r.tbuf(TB_DZ2D), r.tbuf(TB_DA2), 1)) return 1;
    {  // model.8: 3x3 stride 2 128 -> 128 -> z3
        const CUtensorMap* tmA = tmap(t, r.tbuf(TB_DA2), B, H >> 1, W >> 1, 128, 17, 9, 1);
        const CUtensorMap* tmO = tmap(t, r.tbuf(TB_DZ3D), B, H >> 2, W >> 2, 128, 16, 8, 3);  // fp32 output tile
        if (!tmA || !tmO) return 1;
        ConvParams p = conv_base(r, 2, 128, 128);
        p.bias = t->d_bias_pad[3];
        p.out = r.tbuf(TB_DZ3D); p.out_cs = 128;
        p.out_f32 = static_cast<float*>(r.tbuf(TB_DZ3D));
        p.stats = t->dsc + DS_BN + 2 * 256;  // batch statistics reduced in the epilogue
        CID_CHECK_CUDA(cudaMemsetAsync(p.stats, 0, 256 * 8, r.st));
        if (launch_conv_t<128, 16, 2, false, EPI_LRELU_BN, 9, 1, 2>(t, *tmA, t->df_tmh[3], *tmO, p, r.st)) return 1;  // CTA pair
    }
    if (bn_forward(r, 3, r.tbuf(TB_DZ3D), r.tbuf(TB_DA3), 2)) return 1;
    CID_CHECK_CUDA(cudaMemsetAsync(t->fsc + FS_POOLED, 0, static_cast<size_t>(B) * 128 * 4, r.st));
    launch_k(t->g->pdl, avgpool_kernel, dim3(B, 16), 128, 0, r.st, static_cast<const __nv_bfloat16*>(r.tbuf(TB_DA3)), t->fsc + FS_POOLED,
                                                  (H >> 2) * (W >> 2), 128);
    CID_CHECK_CUDA(cudaGetLastError());
    t->launches++;
    return 0;
}

// BatchNorm + LeakyReLU backward of layer l: da (tensor, or null = broadcast from dpool) -> dz
static int bn_backward(const TrainRun& r, int l, const void* z, const void* da, void* dz, int lv, bool param_grads) {
    cid_train* t = r.t;
    const int C = kDCout[l];
    const size_t P = r.pixels(lv);
    const size_t hw = static_cast<size_t>(r.H >> lv) * (r.W >> lv);
    double* sums = t->dsc + DS_BWD + (l - 1) * (kBwdReps * 256);
    const float* stats = t->fsc + FS_STATS + (l - 1) * 256;
    const std::string bn = "discriminator.model." + std::to_string(kDBnIdx[l]);
    const float* gamma = train_ptr(t, bn + ".weight", 0);
    const float* beta = train_ptr(t, bn + ".bias", 0);
    CID_CHECK_CUDA(cudaMemsetAsync(sums, 0, kBwdReps * 256 * 8, r.st));
    static int per_sm = 0;  // resident blocks per SM of the reduce kernel (one wave: see reduce_ppb)
    if (per_sm == 0) {
        int nb = 0;
        CID_CHECK_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, bn_bwd_reduce_kernel<float>, 256, 0));
        per_sm = std::max(nb, 1);
    }
    const int ppb = reduce_ppb(t, P, per_sm);
    launch_k(t->g->pdl, bn_bwd_reduce_kernel<float>, static_cast<unsigned>((P + ppb - 1) / ppb), 256, 0, r.st, 
        static_cast<const float*>(z), static_cast<const __nv_bfloat16*>(da), t->fsc + FS_DPOOL, P, hw, C, stats, gamma,
        beta, 0.2f, sums, ppb, kBwdReps);
    BnBwd f;
    f.sums_rep = sums; f.reps = kBwdReps; f.n = static_cast<double>(P); f.stats = stats; f.gamma = gamma; f.beta = beta;
    f.ggamma = param_grads ? train_ptr(t, bn + ".weight", 1) : nullptr;
    f.gbeta = param_grads ? train_ptr(t, bn + ".bias", 1) : nullptr;
    const size_t n8 = P * C / 8;
    launch_k(t->g->pdl, bn_bwd_apply_kernel<float>, static_cast<unsigned>((n8 + 256 * kBnIter - 1) / (256 * kBnIter)), 256, 0, r.st, 
        static_cast<const float*>(z), static_cast<const __nv_bfloat16*>(da), t->fsc + FS_DPOOL,
        static_cast<__nv_bfloat16*>(dz), P, hw, C, 0.2f, f);
    t->launches += 2;
    CID_CHECK_CUDA(cudaGetLastError());
    return 0;
}

// BCE head + the discriminator's backward. target / scale: see dhead_kernel. param_grads: accumulate D's gradients.
// want_gx: also the gradient w.r.t. the input image (fp32 NHWC-3 in TB_GX).
static int d_backward(const TrainRun& r, const float* img, float target, float scale, int loss_slot, bool param_grads,
                      bool want_gx) {
    cid_train* t = r.t;
    const int B = r.B, H = r.H, W = r.W;
    launch_k(t->g->pdl, dhead_kernel, 1, 1024, B * sizeof(float), r.st, 
        t->fsc + FS_POOLED, train_ptr(t, "discriminator.model.12.weight", 0), train_ptr(t, "discriminator.model.12.bias", 0), B,
        (H >> 2) * (W >> 2), target, scale, t->dsc + DS_LOSS + loss_slot, t->fsc + FS_DPOOL,
        train_ptr(t, "discriminator.model.12.weight", 1), train_ptr(t, "discriminator.model.12.bias", 1), t->fsc + FS_PROBS,
        param_grads ? 1 : 0);
    CID_CHECK_CUDA(cudaGetLastError());
    t->launches++;
    auto gradp = [&](int l, const char* what) {
        return train_ptr(t, "discriminator.model." + std::to_string(kDIdx[l]) + what, 1);
    };
    // layer 3 (model.8 / model.9)
    if (bn_backward(r, 3, r.tbuf(TB_DZ3D), nullptr, r.tbuf(TB_DG3), 2, param_grads)) return 1;
    if (param_grads) {
            if (wgrad(t, 1, r.tbuf(TB_DG3), 128, 128, r.tbuf(TB_DA2), 128, 128, B, H >> 2, W >> 2, gradp(3, ".weight"), r.st, gradp(3, ".bias"))) return 1;
    }
    {  // data gradient: transposed 4x4 stride 2 (zero-padded taps) -> dA2
        const CUtensorMap* tmA = tmap(t, r.tbuf(TB_DG3), B, H >> 2, W >> 2, 128, 18, 10, 0);
        const CUtensorMap* tmO = tmap(t, r.tbuf(TB_DG2), B, H >> 2, W >> 2, 128, 0, 0, 2);
        if (!tmA || !tmO) return 1;
        ConvParams p = conv_base(r, 2, 128, 128);
        p.n_tiles = 2;
        p.out = r.tbuf(TB_DG2); p.out_cs = 128;
        if (launch_conv_t<128, 8, 1, false, EPI_T4, 10, 1, 2>(t, *tmA, t->dd_tmh[3], *tmO, p, r.st)) return 1;  // CTA pair
    }
    // layer 2 (model.5 / model.6)
    if (bn_backward(r, 2, r.tbuf(TB_DZ2D), r.tbuf(TB_DG2), r.tbuf(TB_DG2), 1, param_grads)) return 1;
    if (param_grads) {
            if (wgrad(t, 0, r.tbuf(TB_DG2), 128, 128, r.tbuf(TB_DA1), 64, 64, B, H >> 1, W >> 1, gradp(2, ".weight"), r.st, gradp(2, ".bias"))) return 1;
    }
    if (dgrad3x3(r, 1, r.tbuf(TB_DG2), 128, t->dd_tm[2], r.tbuf(TB_DG1), 64, nullptr, 0, &t->dd_tmh[2])) return 1;
    // layer 1 (model.2 / model.3)
    if (bn_backward(r, 1, r.tbuf(TB_DZ1D), r.tbuf(TB_DG1), r.tbuf(TB_DG1), 1, param_grads)) return 1;
    if (param_grads) {
            if (wgrad(t, 1, r.tbuf(TB_DG1), 64, 64, r.tbuf(TB_DA0), 64, 64, B, H >> 1, W >> 1, gradp(1, ".weight"), r.st, gradp(1, ".bias"))) return 1;
    }
    {  // data gradient -> dA0
        const CUtensorMap* tmA = tmap(t, r.tbuf(TB_DG1), B, H >> 1, W >> 1, 64, 18, 10, 0);
        const CUtensorMap* tmO = tmap(t, r.tbuf(TB_DG0), B, H >> 1, W >> 1, 64, 0, 0, 2);
        if (!tmA || !tmO) return 1;
        ConvParams p = conv_base(r, 1, 64, 64);
        p.n_tiles = 2;
        p.out = r.tbuf(TB_DG0); p.out_cs = 64;
        // layer 0's LeakyReLU backward rides in the epilogue: g *= lrelu'(a0) (a0 is post-activation: same sign)
        p.res2 = r.tbuf(TB_DA0); p.res2_cs = 64; p.mask_slope = 0.2f;
        // CTA pair, as the same-shaped layer of the Keras generator (cgan_model.cuh G3: 1.48 -> 1.28 ms)
        if (launch_conv_t<64, 8, 2, false, EPI_T4, 10, 1, 2>(t, *tmA, t->dd_tmh[1], *tmO, p, r.st)) return 1;
    }
    if (param_grads) {
                if (wgrad_edge(t, r.tbuf(TB_DG0), img, 1, 1, gradp(0, ".weight"), 0, B, H, W, r.st, gradp(0, ".bias"))) return 1;
    }
    if (want_gx) {
        // 64 -> 3 data gradient on the last-conv kernel (transposed GEMM) with a linear fp32 NHWC epilogue
        const CUtensorMap* tmA = tmap(t, r.tbuf(TB_DG0), B, H, W, 64, 18, 10, 0);
        if (!tmA) return 1;
        LastParams lp;
        memset(&lp, 0, sizeof(lp));
        lp.B = B; lp.H = H; lp.W = W;
        lp.tiles_x = (W + 7) / 8;
        lp.tiles_y = (H + 15) / 16;
        lp.m_tiles = B * lp.tiles_x * lp.tiles_y;
        lp.stages = std::min(kLastMaxStages, (t->max_smem - last_conv_smem(0)) / kLastStageBytes);
        CID_REQUIRE(lp.stages >= 2, "last conv: shared memory too small");
        lp.bias = t->zero_bias;
        lp.out = r.tbuf(TB_GX);
        const int smem = last_conv_smem(lp.stages);
        CID_CHECK_CUDA(cudaFuncSetAttribute(last_conv_umma_kernel<EPI_LINEAR_F32_NHWC>, cudaFuncAttributeMaxDynamicSharedMemorySize, t->max_smem));
        const int grid = std::min(lp.m_tiles, t->num_sms);
        CID_CHECK_CUDA(launch_ex(t->g, last_conv_umma_kernel<EPI_LINEAR_F32_NHWC>, grid, kLastThreads, smem, r.st, 1, *tmA, t->dd_tm[0], lp));
        CID_CHECK_CUDA(cudaGetLastError());
        t->launches++;
    }
    return 0;
}

// ---------------------------------------------------------------------------------------------------------------------
// generator backward from gz (TB_GZ, gradient w.r.t. the last conv's pre-tanh output)
// ---------------------------------------------------------------------------------------------------------------------
static int g_backward_layers(const TrainRun& r, const float* noisy) {
    cid_train* t = r.t;
    const int B = r.B, H = r.H, W = r.W;
    auto gw = [&](int l) { return train_ptr(t, std::string("generator.") + kLayers[l].name + ".weight", 1); };
    auto gb = [&](int l) { return train_ptr(t, std::string("generator.") + kLayers[l].name + ".bias", 1); };
    auto pool_bwd = [&](int cat, int C, int lv, int gcat, int gpool, int dz) {
        const size_t total = r.pixels(lv) / 4 * (C / 4);  // one thread per 2x2 window x 4 channels (< 2^32: workspace-bound)
        launch_k(t->g->pdl, poolbwd_kernel, static_cast<unsigned>((total + 255) / 256), 256, 0, r.st, 
            static_cast<const __nv_bfloat16*>(r.gbuf(cat)), static_cast<const __nv_bfloat16*>(r.tbuf(gcat)), 2 * C, C,
            static_cast<const __nv_bfloat16*>(r.tbuf(gpool)), static_cast<__nv_bfloat16*>(r.tbuf(dz)), B, H >> lv, W >> lv, C);
        t->launches++;
    };
    // L11 upconv1.2 (64 -> 3): bias gradient came with gz
    {
        if (wgrad_edge(t, r.gbuf(B_U1A), static_cast<const float*>(r.tbuf(TB_GZ)), 0, -1, gw(11), 1, B, H, W, r.st))
            return 1;
        // data gradient (3 -> 64 conv of gz with flipped weights, masked by u1a > 0) on the tensor-core first-conv kernel
        const CUtensorMap* tmO = tmap(t, r.tbuf(TB_DZ10), B, H, W, 64, 16, 8, 0);
        if (!tmO) return 1;
        FirstParams fp;
        memset(&fp, 0, sizeof(fp));
        fp.B = B; fp.H = H; fp.W = W;
        fp.in = r.tbuf(TB_GZC);
        fp.out = static_cast<__nv_bfloat16*>(r.tbuf(TB_DZ10));
        fp.tiles_x = (W + 7) / 8;
        fp.tiles_y = (H + 15) / 16;
        fp.m_tiles = B * fp.tiles_x * fp.tiles_y;
        fp.w_packed = static_cast<const uint4*>(t->gd_w[11]);
        fp.slope = 1.f;
        fp.mask = static_cast<const __nv_bfloat16*>(r.gbuf(B_U1A));
        const int grid = std::min(fp.m_tiles, t->num_sms * kFirstCtasPerSm);
        CID_CHECK_CUDA(cudaFuncSetAttribute(first_conv_umma_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kFirstSmemMask));
        CID_CHECK_CUDA(launch_ex(t->g, first_conv_umma_kernel<false>, grid, kFirstThreads, kFirstSmemMask, r.st, 1, *tmO, fp));
        CID_CHECK_CUDA(cudaGetLastError());
        t->launches++;
    }
    // L10 upconv1.0 (cat1 128 -> 64)
    if (wgrad(t, 0, r.tbuf(TB_DZ10), 64, 64, r.gbuf(B_CAT1), 128, 128, B, H, W, gw(10), r.st, gb(10))) return 1;
    if (dgrad3x3(r, 0, r.tbuf(TB_DZ10), 64, t->gd_tm[10], r.tbuf(TB_GCAT1), 128, nullptr, 0, &t->gd_tmh[10])) return 1;
    // L9 up1 (ConvT 128 -> 64, u2b -> cat1[:, 0:64])
    if (colsum(t, r.tbuf(TB_GCAT1), r.pixels(0), 128, 0, 64, gb(9), r.st)) return 1;
    if (wgrad(t, 2, r.gbuf(B_U2B), 128, 128, r.tbuf(TB_GCAT1), 128, 64, B, H >> 1, W >> 1, gw(9), r.st)) return 1;
    if (dgrad_convt(r, 1, r.tbuf(TB_GCAT1), 128, 64, t->gd_tm[9], r.tbuf(TB_DZ8), 128, r.gbuf(B_U2B))) return 1;
    // L8 upconv2.2 (u2a 128 -> 128)
    if (wgrad(t, 0, r.tbuf(TB_DZ8), 128, 128, r.gbuf(B_U2A), 128, 128, B, H >> 1, W >> 1, gw(8), r.st, gb(8))) return 1;
    if (dgrad3x3(r, 1, r.tbuf(TB_DZ8), 128, t->gd_tm[8], r.tbuf(TB_DZ7), 128, r.gbuf(B_U2A), 128, &t->gd_tmh[8])) return 1;
    // L7 upconv2.0 (cat2 256 -> 128)
    if (wgrad(t, 0, r.tbuf(TB_DZ7), 128, 128, r.gbuf(B_CAT2), 256, 256, B, H >> 1, W >> 1, gw(7), r.st, gb(7))) return 1;
    if (dgrad3x3(r, 1, r.tbuf(TB_DZ7), 128, t->gd_tm[7], r.tbuf(TB_GCAT2), 256, nullptr, 0, &t->gd_tmh[7])) return 1;
    // L6 up2 (ConvT 256 -> 128, b1 -> cat2[:, 0:128])
    if (colsum(t, r.tbuf(TB_GCAT2), r.pixels(1), 256, 0, 128, gb(6), r.st)) return 1;
    if (wgrad(t, 2, r.gbuf(B_B1), 256, 256, r.tbuf(TB_GCAT2), 256, 128, B, H >> 2, W >> 2, gw(6), r.st)) return 1;
    if (dgrad_convt(r, 2, r.tbuf(TB_GCAT2), 256, 128, t->gd_tm[6], r.tbuf(TB_DZ5), 256, r.gbuf(B_B1))) return 1;
    // L5 bottleneck.2 (b0 256 -> 256)
    if (wgrad(t, 0, r.tbuf(TB_DZ5), 256, 256, r.gbuf(B_B0), 256, 256, B, H >> 2, W >> 2, gw(5), r.st, gb(5))) return 1;
    if (dgrad3x3(r, 2, r.tbuf(TB_DZ5), 256, t->gd_tm[5], r.tbuf(TB_DZ4), 256, r.gbuf(B_B0), 256, &t->gd_tmh[5])) return 1;
    // L4 bottleneck.0 (p2 128 -> 256)
    if (wgrad(t, 0, r.tbuf(TB_DZ4), 256, 256, r.gbuf(B_P2), 128, 128, B, H >> 2, W >> 2, gw(4), r.st, gb(4))) return 1;
    if (dgrad3x3(r, 2, r.tbuf(TB_DZ4), 256, t->gd_tm[4], r.tbuf(TB_GP2), 128, nullptr, 0, &t->gd_tmh[4])) return 1;
    // pool2 + skip: e2 = cat2[:, 128:256]
    pool_bwd(B_CAT2, 128, 1, TB_GCAT2, TB_GP2, TB_DZ3);
    // L3 down2.2 (a2 128 -> 128)
    if (wgrad(t, 0, r.tbuf(TB_DZ3), 128, 128, r.gbuf(B_A2), 128, 128, B, H >> 1, W >> 1, gw(3), r.st, gb(3))) return 1;
    if (dgrad3x3(r, 1, r.tbuf(TB_DZ3), 128, t->gd_tm[3], r.tbuf(TB_DZ2), 128, r.gbuf(B_A2), 128, &t->gd_tmh[3])) return 1;
    // L2 down2.0 (p1 64 -> 128)
    if (wgrad(t, 0, r.tbuf(TB_DZ2), 128, 128, r.gbuf(B_P1), 64, 64, B, H >> 1, W >> 1, gw(2), r.st, gb(2))) return 1;
    if (dgrad3x3(r, 1, r.tbuf(TB_DZ2), 128, t->gd_tm[2], r.tbuf(TB_GP1), 64, nullptr, 0, &t->gd_tmh[2])) return 1;
    // pool1 + skip: e1 = cat1[:, 64:128]
    pool_bwd(B_CAT1, 64, 0, TB_GCAT1, TB_GP1, TB_DZ1);
    // L1 down1.2 (a0 64 -> 64)
    if (wgrad(t, 0, r.tbuf(TB_DZ1), 64, 64, r.gbuf(B_A0), 64, 64, B, H, W, gw(1), r.st, gb(1))) return 1;
    if (dgrad3x3(r, 0, r.tbuf(TB_DZ1), 64, t->gd_tm[1], r.tbuf(TB_DZ0), 64, r.gbuf(B_A0), 64, &t->gd_tmh[1])) return 1;
    // L0 down1.0 (3 -> 64)
        {
        if (wgrad_edge(t, r.tbuf(TB_DZ0), noisy, 1, 1, gw(0), 0, B, H, W, r.st, gb(0))) return 1;
    }
    return 0;
}

static int train_make_run(cid_train* t, int B, int H, int W, void* ws, size_t ws_bytes, cudaStream_t st, TrainRun* r) {
    CID_REQUIRE(t && t->state, "training state not bound (cid_train_bind_state)");
    CID_REQUIRE(t->g_ready, "weights not synchronised (cid_train_sync_weights)");
    CID_REQUIRE(B >= 2 && B <= kTrainMaxBatch, "training batch %d out of range [2,%d] (train-mode BatchNorm needs > 1)", B,
                kTrainMaxBatch);
    CID_REQUIRE(H >= 16 && W >= 16 && H % 4 == 0 && W % 4 == 0, "training image %dx%d must be multiples of 4, >= 16", H, W);
    CID_REQUIRE(ws && (reinterpret_cast<uintptr_t>(ws) & 1023) == 0, "workspace must be 1024-byte aligned");
    if (t->maps.size() > 384) t->maps.clear();  // many (shape, workspace) combinations seen: start over
    r->t = t; r->B = B; r->H = H; r->W = W;
    r->ws = static_cast<uint8_t*>(ws);
    r->st = st;
    layout_offsets(B, H, W, r->goff);
    const size_t need = train_layout(B, H, W, r->toff);
    CID_REQUIRE(ws_bytes >= need, "training workspace too small: %zu < %zu bytes", ws_bytes, need);
    return 0;
}

}  // namespace cid
