// C-ABI entry points (include/cid_b200.h): context, weight packing, execution plans and the
// layer-by-layer launch sequence of the denoise generator.
//
// Layer table (reference: backend/app.py:42-78, forward :80-103):
//   L0  down1.0       conv3x3   3->64   +ReLU            first_conv_umma_kernel -> a0
//   L1  down1.2       conv3x3  64->64   +ReLU +pool      <64 ,9,1,WRES,POOL>    -> cat1[:,64:128], p1
//   L2  down2.0       conv3x3  64->128  +ReLU            <128,9,1,WRES,RELU>    -> a2
//   L3  down2.2       conv3x3 128->128  +ReLU +pool      <128,9,2,ring,POOL>    -> cat2[:,128:256], p2
//   L4  bottleneck.0  conv3x3 128->256  +ReLU            <256,9,1,ring,RELU>    -> b0
//   L5  bottleneck.2  conv3x3 256->256  +ReLU            <256,9,1,ring,RELU>    -> b1
//   L6  up2           convT2x2 256->128                  <256,1,1,ring,CONVT>   -> cat2[:,0:128]
//   L7  upconv2.0     conv3x3 256->128  +ReLU            <128,9,2,ring,RELU>    -> u2a
//   L8  upconv2.2     conv3x3 128->128  +ReLU            <128,9,2,ring,RELU>    -> u2b
//   L9  up1           convT2x2 128->64                   <256,1,1,ring,CONVT>   -> cat1[:,0:64]
//   L10 upconv1.0     conv3x3 128->64   +ReLU            <64 ,9,1,WRES,RELU>    -> u1a
//   L11 upconv1.2     conv3x3  64->3    +tanh            <16 ,9,1,WRES,TANH_*>  -> y
#include <cuda_bf16.h>
#include <string.h>

#include <algorithm>

#include "cid_internal.h"
#include "conv_umma.cuh"
#include "first_conv_umma.cuh"
#include "last_conv_umma.cuh"
#include "cgan_first_conv.cuh"
#include "k9_first_conv.cuh"
#include "k9_last_conv.cuh"

#define TSV(x) x

namespace cid {

static thread_local char g_err[1024] = "";

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

static void graph_release(cid_graph& g) {
    if (g.exec) cudaGraphExecDestroy(static_cast<cudaGraphExec_t>(g.exec));
    if (g.graph) cudaGraphDestroy(static_cast<cudaGraph_t>(g.graph));
    g = cid_graph();
}

void clear_plans(cid_ctx* c) {
    for (auto& pl : c->plans) {
        graph_release(pl.graph[0]);
        graph_release(pl.graph[1]);
    }
    c->plans.clear();
}

// ---------------------------------------------------------------------------------------------
// layer table
// ---------------------------------------------------------------------------------------------
enum LayerKind { L_FIRST, L_CONV, L_CONVT };
struct LayerDesc {
    const char* name;
    LayerKind kind;
    int cin, cout;
};
static const LayerDesc kLayers[12] = {
    {"down1.0", L_FIRST, 3, 64},        {"down1.2", L_CONV, 64, 64},       {"down2.0", L_CONV, 64, 128},
    {"down2.2", L_CONV, 128, 128},      {"bottleneck.0", L_CONV, 128, 256}, {"bottleneck.2", L_CONV, 256, 256},
    {"up2", L_CONVT, 256, 128},         {"upconv2.0", L_CONV, 256, 128},   {"upconv2.2", L_CONV, 128, 128},
    {"up1", L_CONVT, 128, 64},          {"upconv1.0", L_CONV, 128, 64},    {"upconv1.2", L_CONV, 64, 3},
};

// activation buffers inside the workspace
enum Buf { B_A0, B_CAT1, B_P1, B_A2, B_CAT2, B_P2, B_B0, B_B1, B_U2A, B_U2B, B_U1A, B_COUNT };
static const char* kBufNames[B_COUNT] = {"a0", "cat1", "p1", "a2", "cat2", "p2", "b0", "b1", "u2a", "u2b", "u1a"};
static const int kBufC[B_COUNT] = {64, 128, 64, 128, 256, 128, 256, 256, 128, 128, 64};
// Spatial levels of the U-Net for ANY input size H x W >= 4 (backend/app.py:80-103): 0 = input (e1), 1 = after pool1
// (floor(H/2)), 2 = after pool2, 3 = up2's output (2 * level 2), 4 = up1's output (2 * level 3) = the size of the
// network's result. For multiples of 4, level 3 == level 1 and level 4 == level 0; otherwise the decoder is smaller and
// the reference crops the skip tensors to it (:90-92, :97-99) - here the consumer's tensor map simply views the top-left
// (level 3 / 4) window of the concat buffer, whose skip half was written at the encoder's size.
static const int kBufLevel[B_COUNT] = {0, 0, 1, 1, 1, 2, 2, 2, 3, 3, 4};
struct LevelDims {
    int h[5], w[5];
};
static LevelDims level_dims(int H, int W) {
    LevelDims d;
    d.h[0] = H; d.w[0] = W;
    d.h[1] = H / 2; d.w[1] = W / 2;
    d.h[2] = d.h[1] / 2; d.w[2] = d.w[1] / 2;
    d.h[3] = 2 * d.h[2]; d.w[3] = 2 * d.w[2];
    d.h[4] = 2 * d.h[3]; d.w[4] = 2 * d.w[3];
    return d;
}
// the level a layer's GEMM runs at (its A operand's / pixel tiles' dims)
static const int kLayerLevel[12] = {0, 0, 1, 1, 2, 2, 2, 3, 3, 3, 4, 4};

static size_t buf_bytes(int buf, int B, int H, int W) {
    const LevelDims d = level_dims(H, W);
    const int lv = kBufLevel[buf];
    size_t n = static_cast<size_t>(B) * d.h[lv] * d.w[lv] * kBufC[buf] * 2;
    return (n + 1023) & ~static_cast<size_t>(1023);
}

static size_t layout_offsets(int B, int H, int W, size_t* off) {
    size_t cur = 0;
    for (int i = 0; i < B_COUNT; ++i) {
        if (off) off[i] = cur;
        cur += buf_bytes(i, B, H, W);
    }
    return cur;
}

// ---------------------------------------------------------------------------------------------
// TMA descriptors
// ---------------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
    static EncodeTiledFn fn = nullptr;
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
            qres == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(p);
    }
    return fn;
}

// NHWC bf16 activation [B,H,W,C]: box = 64 channels x 8 pixels x box_rows rows x 1 image, 128B swizzle.
// (pitch_h, pitch_w) = the allocated dims when the map views only the top-left H x W window of a larger buffer.
static int make_act_map(CUtensorMap* tm, const void* base, int B, int H, int W, int C, int box_rows, int box_px,
                        int pitch_h = 0, int pitch_w = 0) {
    EncodeTiledFn fn = get_encode_fn();
    CID_REQUIRE(fn, "cuTensorMapEncodeTiled entry point not available");
    const cuuint64_t ph = pitch_h > 0 ? pitch_h : H, pw = pitch_w > 0 ? pitch_w : W;
    cuuint64_t dims[4] = {static_cast<cuuint64_t>(C), static_cast<cuuint64_t>(W), static_cast<cuuint64_t>(H),
                          static_cast<cuuint64_t>(B)};
    cuuint64_t strides[3] = {static_cast<cuuint64_t>(C) * 2, pw * C * 2, ph * pw * C * 2};
    cuuint32_t box[4] = {64, static_cast<cuuint32_t>(box_px), static_cast<cuuint32_t>(box_rows), 1};
    cuuint32_t estr[4] = {1, 1, 1, 1};
    CUresult r = fn(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(base), dims, strides, box, estr,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    CID_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled(activation %dx%dx%dx%d) failed: %d", B, H, W, C,
                static_cast<int>(r));
    return 0;
}

// convT output viewed as (C, 2W, dy, H, B): box = 64 ch x 16 px x 1 x 16 rows x 1 image (rows 2y+dy of a tile).
// (pitch_h, pitch_w) = the output buffer's allocated dims when they are larger than 2H x 2W (odd encoder sizes).
static int make_convt_out_map(CUtensorMap* tm, const void* base, int B, int H, int W, int C, int pitch_h = 0, int pitch_w = 0) {
    EncodeTiledFn fn = get_encode_fn();
    CID_REQUIRE(fn, "cuTensorMapEncodeTiled entry point not available");
    const cuuint64_t ph = pitch_h > 0 ? pitch_h : 2 * H, pw = pitch_w > 0 ? pitch_w : 2 * W;
    const cuuint64_t row = pw * C * 2;  // bytes of one output row
    cuuint64_t dims[5] = {static_cast<cuuint64_t>(C), static_cast<cuuint64_t>(2 * W), 2, static_cast<cuuint64_t>(H),
                          static_cast<cuuint64_t>(B)};
    cuuint64_t strides[4] = {static_cast<cuuint64_t>(C) * 2, row, 2 * row, ph * row};
    cuuint32_t box[5] = {64, 16, 1, 16, 1};
    cuuint32_t estr[5] = {1, 1, 1, 1, 1};
    CUresult r = fn(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 5, const_cast<void*>(base), dims, strides, box, estr,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    CID_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled(convT out %dx%dx%dx%d) failed: %d", B, H, W, C,
                static_cast<int>(r));
    return 0;
}

// packed weights [rows, Cin] bf16 (Cin contiguous): box = 64 x box_rows, 128B swizzle.
static int make_weight_map(CUtensorMap* tm, const void* base, int rows, int cin, int box_rows) {
    EncodeTiledFn fn = get_encode_fn();
    CID_REQUIRE(fn, "cuTensorMapEncodeTiled entry point not available");
    cuuint64_t dims[2] = {static_cast<cuuint64_t>(cin), static_cast<cuuint64_t>(rows)};
    cuuint64_t strides[1] = {static_cast<cuuint64_t>(cin) * 2};
    cuuint32_t box[2] = {64, static_cast<cuuint32_t>(box_rows)};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = fn(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    CID_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled(weights %dx%d) failed: %d", rows, cin, static_cast<int>(r));
    return 0;
}

// ---------------------------------------------------------------------------------------------
// host-side weight packing
// ---------------------------------------------------------------------------------------------
static uint16_t f32_to_bf16_rne(float f) {
    uint32_t u;
    memcpy(&u, &f, 4);
    if ((u & 0x7fffffffu) > 0x7f800000u) return static_cast<uint16_t>((u >> 16) | 0x40);  // NaN
    u += 0x7fffu + ((u >> 16) & 1u);
    return static_cast<uint16_t>(u >> 16);
}

static float bf16_to_f32(uint16_t h) {
    const uint32_t u = static_cast<uint32_t>(h) << 16;
    float f;
    memcpy(&f, &u, 4);
    return f;
}

// per-layer tensor-core tile width N and padded row count per tap
static int layer_n(int l) {
    switch (l) {
        case 1: case 10: return 64;
        case 2: case 3: case 7: case 8: return 128;
        case 4: case 5: case 6: case 9: return 256;
        case 11: return 16;
        default: return 0;
    }
}

static int pack_layer(cid_ctx* c, int l) {
    const LayerDesc& L = kLayers[l];
    const std::string wkey = std::string(L.name) + ".weight";
    const std::string bkey = std::string(L.name) + ".bias";
    auto wi = c->params.find(wkey);
    auto bi = c->params.find(bkey);
    CID_REQUIRE(wi != c->params.end() && bi != c->params.end(), "missing parameter for layer %s", L.name);
    const std::vector<float>& w = wi->second;
    const std::vector<float>& b = bi->second;

    if (L.kind == L_FIRST) {
        // per tap a [cout 64][K 16] tile in the un-swizzled K-major layout
        // [tap][k chunk][cout][8 bf16]; K chunk 0 = [w_r w_g w_b | w_r w_g w_b | 0 0] (the input is
        // split hi+lo along K), K chunk 1 = [wlo_r wlo_g wlo_b | 0 ...]: the MMA reads the same pixel chunk for
        // both K chunks (LBO = 0), so chunk 1 adds x_hi * w_lo.
        std::vector<uint16_t> pk(kFirstWBytes / 2, 0);
        for (int tap = 0; tap < 9; ++tap)
            for (int co = 0; co < 64; ++co)
                for (int ci = 0; ci < 3; ++ci) {
                    const float wf = w[(static_cast<size_t>(co) * 3 + ci) * 9 + tap];
                    const uint16_t v = f32_to_bf16_rne(wf);
                    pk[((tap * 2 + 0) * 64 + co) * 8 + ci] = v;
                    pk[((tap * 2 + 0) * 64 + co) * 8 + 3 + ci] = v;
                    pk[((tap * 2 + 1) * 64 + co) * 8 + ci] = f32_to_bf16_rne(wf - bf16_to_f32(v));  // w_lo (x_hi)
                }
        // the bias as two more K rows of the centre tap: pixel slot 6 is 1.0 (first_conv_umma.cuh)
        for (int co = 0; co < 64; ++co) {
            const uint16_t bh = f32_to_bf16_rne(b[co]);
            pk[((4 * 2 + 0) * 64 + co) * 8 + 6] = bh;
            pk[((4 * 2 + 1) * 64 + co) * 8 + 6] = f32_to_bf16_rne(b[co] - bf16_to_f32(bh));
        }
        if (c->w_dev[0]) cudaFree(c->w_dev[0]);
        if (c->bias_dev[0]) cudaFree(c->bias_dev[0]);
        c->w_dev[0] = nullptr;
        c->bias_dev[0] = nullptr;
        CID_CHECK_CUDA(cudaMalloc(&c->w_dev[0], pk.size() * 2));
        CID_CHECK_CUDA(cudaMemcpy(c->w_dev[0], pk.data(), pk.size() * 2, cudaMemcpyHostToDevice));
        CID_CHECK_CUDA(cudaMalloc(reinterpret_cast<void**>(&c->bias_dev[0]), 64 * 4));
        CID_CHECK_CUDA(cudaMemcpy(c->bias_dev[0], b.data(), 64 * 4, cudaMemcpyHostToDevice));
        return 0;
    }

    std::vector<uint16_t> packed;
    int rows = 0;
    int box_rows = layer_n(l);
    if (l == 11) {
        // transposed GEMM of last_conv_umma_kernel: one row per (tap, cout), 128 rows of 64 channels
        rows = box_rows = 128;
        packed.assign(static_cast<size_t>(rows) * L.cin, 0);
        for (int co = 0; co < L.cout; ++co)
            for (int ci = 0; ci < L.cin; ++ci)
                for (int t = 0; t < 9; ++t)
                    packed[static_cast<size_t>(last_conv_row(t, co)) * L.cin + ci] =
                        f32_to_bf16_rne(w[(static_cast<size_t>(co) * L.cin + ci) * 9 + t]);
    } else if (L.kind == L_CONV) {
        const int npad = std::max(L.cout, 16);
        rows = 9 * npad;
        packed.assign(static_cast<size_t>(rows) * L.cin, 0);
        for (int co = 0; co < L.cout; ++co)
            for (int ci = 0; ci < L.cin; ++ci)
                for (int t = 0; t < 9; ++t)
                    packed[(static_cast<size_t>(t) * npad + co) * L.cin + ci] =
                        f32_to_bf16_rne(w[(static_cast<size_t>(co) * L.cin + ci) * 9 + t]);
    } else {  // ConvTranspose2d weight [Cin, Cout, 2, 2] -> rows (dy*2+dx)*Cout + co
        rows = 4 * L.cout;
        packed.assign(static_cast<size_t>(rows) * L.cin, 0);
        for (int ci = 0; ci < L.cin; ++ci)
            for (int co = 0; co < L.cout; ++co)
                for (int qd = 0; qd < 4; ++qd)
                    packed[(static_cast<size_t>(qd) * L.cout + co) * L.cin + ci] =
                        f32_to_bf16_rne(w[(static_cast<size_t>(ci) * L.cout + co) * 4 + qd]);
    }
    if (c->w_dev[l]) cudaFree(c->w_dev[l]);
    if (c->bias_dev[l]) cudaFree(c->bias_dev[l]);
    c->w_dev[l] = nullptr;
    c->bias_dev[l] = nullptr;
    CID_CHECK_CUDA(cudaMalloc(&c->w_dev[l], packed.size() * 2));
    CID_CHECK_CUDA(cudaMemcpy(c->w_dev[l], packed.data(), packed.size() * 2, cudaMemcpyHostToDevice));
    // bias padded to a multiple of 32 floats so float4 epilogue loads never run off the end
    std::vector<float> bp((b.size() + 31) / 32 * 32, 0.f);
    std::copy(b.begin(), b.end(), bp.begin());
    CID_CHECK_CUDA(cudaMalloc(reinterpret_cast<void**>(&c->bias_dev[l]), bp.size() * 4));
    CID_CHECK_CUDA(cudaMemcpy(c->bias_dev[l], bp.data(), bp.size() * 4, cudaMemcpyHostToDevice));
    if (box_rows % 2 == 0 && make_weight_map(&c->tmWh[l], c->w_dev[l], rows, L.cin, box_rows / 2)) return 1;
    // 64-row boxes: the N-split form of the 128 / 256-wide layers (small batches, see pick_split)
    if (L.kind == L_CONV && l != 11 && make_weight_map(&c->tmW64[l], c->w_dev[l], rows, L.cin, 64)) return 1;
    return make_weight_map(&c->tmW[l], c->w_dev[l], rows, L.cin, box_rows);
}

// ---------------------------------------------------------------------------------------------
// plans
// ---------------------------------------------------------------------------------------------
static int layer_in_buf(int l) {
    static const int in[12] = {-1, B_A0, B_P1, B_A2, B_P2, B_B0, B_B1, B_CAT2, B_U2A, B_U2B, B_CAT1, B_U1A};
    return in[l];
}

static int get_plan(cid_ctx* c, int Bc, int H, int W, const void* ws, cid_plan** out) {
    for (auto& pl : c->plans)
        if (pl.Bc == Bc && pl.H == H && pl.W == W && pl.ws == ws) {
            *out = &pl;
            return 0;
        }
    if (c->plans.size() >= 64) clear_plans(c);
    cid_plan pl;
    pl.Bc = Bc;
    pl.H = H;
    pl.W = W;
    pl.ws = ws;
    layout_offsets(Bc, H, W, pl.off);
    const LevelDims d = level_dims(H, W);
    for (int l = 1; l < 12; ++l) {
        // the A operand: the layer's level-sized view of its input buffer (the concat buffers are allocated at the
        // encoder's size; for sizes that are not multiples of 4 the decoder views their top-left window)
        const int buf = layer_in_buf(l);
        const int lv = kLayerLevel[l], blv = kBufLevel[buf];
        const int rows = kLayers[l].kind == L_CONVT ? 16 : 18;
        const int px = kLayers[l].kind == L_CONV ? 10 : 8;
        if (make_act_map(&pl.tmA[l], static_cast<const uint8_t*>(ws) + pl.off[buf], Bc, d.h[lv], d.w[lv], kBufC[buf], rows, px,
                         d.h[blv], d.w[blv]))
            return 1;
    }
    static const int out_buf[12] = {B_A0, B_CAT1, B_A2, B_CAT2, B_B0, B_B1, -1, B_U2A, B_U2B, -1, B_U1A, -1};
    for (int l = 0; l < 12; ++l) {
        const int ob = out_buf[l];
        if (ob < 0) continue;
        const int lv = kBufLevel[ob];  // conv outputs are written at their buffer's own size
        if (make_act_map(&pl.tmO[l], static_cast<const uint8_t*>(ws) + pl.off[ob], Bc, d.h[lv], d.w[lv], kBufC[ob], 16, 8))
            return 1;
        if (l == 1 || l == 3) {
            const int pb = l == 1 ? B_P1 : B_P2;
            const int plv = kBufLevel[pb];
            if (make_act_map(&pl.tmO2[l], static_cast<const uint8_t*>(ws) + pl.off[pb], Bc, d.h[plv], d.w[plv],
                             kBufC[pb], 8, 4))
                return 1;
        }
    }
    // convT layers: input level 2 / 3, output = the decoder half of cat2 (allocated at level 1) / cat1 (level 0)
    if (make_convt_out_map(&pl.tmO[6], static_cast<const uint8_t*>(ws) + pl.off[B_CAT2], Bc, d.h[2], d.w[2], 256, d.h[1], d.w[1])) return 1;
    if (make_convt_out_map(&pl.tmO[9], static_cast<const uint8_t*>(ws) + pl.off[B_CAT1], Bc, d.h[3], d.w[3], 128, d.h[0], d.w[0])) return 1;
    c->plans.push_back(pl);
    *out = &c->plans.back();
    return 0;
}

// ---------------------------------------------------------------------------------------------
// launches
// ---------------------------------------------------------------------------------------------
// One launch path for every generator kernel: optional cluster dimension (CTA pairs) and the programmatic-
// dependent-launch attribute, so consecutive layers overlap the next kernel's set-up (barriers, TMEM, resident
// weights) with the tail of the previous one. The kernels order themselves with griddepcontrol.wait.
template <class... KArgs, class... Args>
static cudaError_t launch_ex(const cid_ctx* c, void (*kern)(KArgs...), int grid, int block, int smem, cudaStream_t st,
                             int cluster, Args&&... args) {
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(block);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[2];
    int n = 0;
    if (cluster > 1) {
        attr[n].id = cudaLaunchAttributeClusterDimension;
        attr[n].val.clusterDim.x = cluster;
        attr[n].val.clusterDim.y = 1;
        attr[n].val.clusterDim.z = 1;
        ++n;
    }
    if (c->pdl) {
        attr[n].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[n].val.programmaticStreamSerializationAllowed = 1;
        ++n;
    }
    cfg.attrs = attr;
    cfg.numAttrs = n;
    return cudaLaunchKernelEx(&cfg, kern, std::forward<Args>(args)...);
}

// The small kernels of the training step (reductions, BatchNorm, Adam, re-pack ...) launch with the same
// programmatic-dependent-launch attribute: each begins with griddepcontrol.launch_dependents + griddepcontrol.wait, so the
// launch latency and set-up of the NEXT kernel overlap this one's tail while the memory ordering stays that of the stream.
// Launch errors surface through the caller's cudaGetLastError().
template <class... KArgs, class... Args>
static void launch_k(bool pdl, void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args&&... args) {
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl ? 1 : 0;
    (void)cudaLaunchKernelEx(&cfg, kern, std::forward<Args>(args)...);
}

// Shared-memory split of a streamed-weight kernel between the A ring (halo patches) and the weight ring: the A ring covers
// ~4500 tensor cycles of K blocks (at least two), the weight ring gets the rest (at most kMaxWStages, at least 4 if the
// A ring can give way).
template <class Cfg, int N, int S>
static void stage_split(int budget, int* a_stages, int* w_stages) {
    const int kb_cycles = Cfg::STEPS * S * 4 * (N / 2);  // tensor cycles one K block's patches last
    const int kb_in_flight = std::max(2, (4500 + kb_cycles - 1) / kb_cycles + 1);
    int a = std::min(kMaxStages, std::max(S + 1, S * kb_in_flight));
    auto w_for = [&](int aa) { return std::min(kMaxWStages, (budget - aa * Cfg::A_BYTES) / Cfg::W_BYTES); };
    while (a > S + 1 && w_for(a) < 4) --a;
    *a_stages = a;
    *w_stages = std::max(2, w_for(a));
}

// deep_w: small-batch form - a CTA runs one or two work items, so a few A stages suffice and the rest of the shared
// memory becomes weight stages: the kernel is then bound by weight-tile latency x tiles in flight (ncu, batch 1: a K = 2304
// item took 13.8 us for 2.3 us of MMAs with 4 stages of 8 KB in flight).
template <int N, int TAPS, int S, bool WRES, int EPI, int PW = 0, int TS = 0, int CG = 1>
static int launch_conv(cid_ctx* c, int slot, const CUtensorMap& tmA, const CUtensorMap& tmW, const CUtensorMap& tmO,
                       const CUtensorMap& tmO2, ConvParams p, cudaStream_t st, bool deep_w = false, int w_req = 0, int a_req = 0) {
    using Cfg = ConvCfg<N, TAPS, S, WRES, EPI, PW, TS, CG>;
    const int budget = c->max_smem - 1024 - 1024 - Cfg::O_BYTES;
    // Streamed weights: a weight tile lasts S * 4 MMAs (256-1024 tensor cycles) while a halo patch lasts STEPS times as
    // long, and the MMA warp was found waiting on WEIGHT tiles (L2 latency x tiles in flight: every CTA streams the same
    // tiles at the same time). So the A ring gets two K blocks' worth of patches (2 * S) and the weight ring the rest of
    // the shared memory. Sweep at batch 256 (tools/wst_sweep.sh, profiles/r02j_wst_sweep.txt): upconv2.2 0.239 ms at 4
    // weight stages, 0.216 at 8, 0.205 at 12; bottleneck.2 0.222 / 0.213 at 4 / 8; down2.2 0.261 / 0.230 at 4 / 10 -
    // and a cliff as soon as the A ring drops to S + 1.
    int a_cap = kMaxStages;
    int w_stages = 1;
    if (!WRES) {
        stage_split<Cfg, N, S>(budget, &a_cap, &w_stages);
        if (deep_w) {
            a_cap = std::min(kMaxStages, std::max(S + 1, std::min(p.kb_count + 1, 5)));
            w_stages = std::max(2, std::min(kMaxWStages, (budget - a_cap * Cfg::A_BYTES) / Cfg::W_BYTES));
        }
        if (w_req > 0) {
            w_stages = w_req;
            a_cap = kMaxStages;
        }
    }
    const int w_bytes = (WRES ? p.kb_count * TAPS : w_stages) * Cfg::W_BYTES;
    int a_stages = std::min(a_cap, (budget - w_bytes) / Cfg::A_BYTES);
    if (a_req > 0) a_stages = std::min(a_stages, std::max(a_req, S + 1));
    CID_REQUIRE(a_stages >= S + 1, "conv<%d,%d,%d>: shared memory too small (%d stages)", N, TAPS, S, a_stages);
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
    p.num_work = ((p.m_tiles + S * CG - 1) / (S * CG)) * p.n_tiles;  // work items of one CTA (or CTA pair)
    const int smem = Cfg::smem_bytes(p.kb_count, a_stages, w_stages);
    auto kern = conv_umma_kernel<N, TAPS, S, WRES, EPI, PW, TS, CG>;
    (void)slot;
    static uint64_t attr_set = 0;  // bit per device: function attributes are per device and per instantiation
    if (!(attr_set >> (c->device & 63) & 1)) {
        CID_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, c->max_smem));
        attr_set |= 1ull << (c->device & 63);
    }
    const int grid = std::min(p.num_work * CG, c->num_sms / CG * CG);
    CID_CHECK_CUDA(launch_ex(c, kern, grid, kConvThreads, smem, st, CG, tmA, tmW, tmO, tmO2, p));
    CID_CHECK_CUDA(cudaGetLastError());
    c->launches++;
    return 0;
}

struct IoSpec {
    bool u8;
    const void* x;
    void* y;
    int in_h, in_w, pad_l, pad_t;  // u8 mode
};

static int prof_mark(cid_ctx* c, cudaStream_t st) {
    if (!c->profile) return 0;
    if (c->prof_used == c->prof_events.size()) {
        cudaEvent_t e;
        CID_CHECK_CUDA(cudaEventCreate(&e));
        c->prof_events.push_back(e);
    }
    CID_CHECK_CUDA(cudaEventRecord(c->prof_events[c->prof_used++], st));
    return 0;
}

// ---------------------------------------------------------------------------------------------
// Small batches (the upload route runs batch 1, backend/app.py:406): a layer whose pixel tiles do not fill
// the 148 SMs in its batch-256 configuration (S tiles per work item, CTA pairs, N = Cout) runs instead as
// N = 64 column blocks of single tiles (n_tiles = Cout / 64 work items per pixel tile, one CTA each), so up
// to 4x as many SMs share the layer. Both forms accumulate every output in the same order: bit-identical.
// Cost model: waves x work-item time, time ~ S * N columns, N = 64 MMAs pay ~1.25x (they sit on the
// shared-memory roofline). `S`, `CG` describe the large form of the layer.
// ---------------------------------------------------------------------------------------------
static bool pick_split(const cid_ctx* c, int m_tiles, int n, int S, int CG) {
    if (c->regime == 1) return false;
    if (c->regime == 2) return true;
    const int slots = c->num_sms;
    auto waves = [](int items, int sl) { return (items + sl - 1) / sl; };
    const int large_items = (m_tiles + S * CG - 1) / (S * CG);
    const double large = waves(large_items, slots / CG) * static_cast<double>(S * n) * (n == 64 ? 1.25 : 1.0);
    const double split = waves(m_tiles * (n / 64), slots) * 64.0 * 1.25;
    return split < large;
}

static FirstParams make_first_params(const cid_ctx* c, const cid_plan* pl, int nb, const IoSpec& io) {
    uint8_t* ws = static_cast<uint8_t*>(const_cast<void*>(pl->ws));
    FirstParams fp;
    memset(&fp, 0, sizeof(fp));
    fp.B = nb; fp.H = pl->H; fp.W = pl->W;
    fp.in = io.x; fp.in_h = io.in_h; fp.in_w = io.in_w; fp.pad_l = io.pad_l; fp.pad_t = io.pad_t;
    fp.out = reinterpret_cast<__nv_bfloat16*>(ws + pl->off[B_A0]);
    fp.tiles_x = (pl->W + 7) / 8;
    fp.tiles_y = (pl->H + 15) / 16;
    fp.m_tiles = nb * fp.tiles_x * fp.tiles_y;
    fp.w_packed = static_cast<const uint4*>(c->w_dev[0]);
    return fp;
}

static int make_last_params(const cid_ctx* c, const cid_plan* pl, int nb, const IoSpec& io, LastParams* out) {
    LastParams lp;
    memset(&lp, 0, sizeof(lp));
    const LevelDims d = level_dims(pl->H, pl->W);  // the network's result has the decoder's size (level 4)
    lp.B = nb; lp.H = d.h[4]; lp.W = d.w[4];
    lp.tiles_x = (lp.W + 7) / 8;
    lp.tiles_y = (lp.H + 15) / 16;
    lp.m_tiles = nb * lp.tiles_x * lp.tiles_y;
    lp.stages = std::min(kLastMaxStages, (c->max_smem - last_conv_smem(0)) / kLastStageBytes);
    CID_REQUIRE(lp.stages >= 2, "last conv: shared memory too small");
    lp.bias = c->bias_dev[11];
    lp.out = io.y;
    lp.crop_l = io.pad_l; lp.crop_t = io.pad_t; lp.out_h = io.in_h; lp.out_w = io.in_w;
    *out = lp;
    return 0;
}

// one pass of nb (<= plan.Bc) images through the 12 layers
static int run_layers(cid_ctx* c, cid_plan* pl, int nb, const IoSpec& io, cudaStream_t st) {
    const int H = pl->H, W = pl->W;
    uint8_t* ws = static_cast<uint8_t*>(const_cast<void*>(pl->ws));
    auto buf = [&](int b) { return reinterpret_cast<__nv_bfloat16*>(ws + pl->off[b]); };
    const LevelDims dm = level_dims(H, W);
    auto tiles = [&](int lv) { return nb * ((dm.w[lv] + 7) / 8) * ((dm.h[lv] + 15) / 16); };

    if (prof_mark(c, st)) return 1;
    {  // L0 down1.0 (+ Pad / ToTensor / Normalize) -> a0
        const FirstParams fp = make_first_params(c, pl, nb, io);
        if (!(c->attr_done & (1u << 30))) {
            CID_CHECK_CUDA(cudaFuncSetAttribute(first_conv_umma_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kFirstSmemMask));
            CID_CHECK_CUDA(cudaFuncSetAttribute(first_conv_umma_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kFirstSmemMask));
            c->attr_done |= (1u << 30);
        }
        const int grid = std::min(fp.m_tiles, c->num_sms * kFirstCtasPerSm);
        if (io.u8)
            CID_CHECK_CUDA(launch_ex(c, first_conv_umma_kernel<true>, grid, kFirstThreads, kFirstSmem, st, 1, pl->tmO[0], fp));
        else
            CID_CHECK_CUDA(launch_ex(c, first_conv_umma_kernel<false>, grid, kFirstThreads, kFirstSmem, st, 1, pl->tmO[0], fp));
        CID_CHECK_CUDA(cudaGetLastError());
        c->launches++;
    }
    if (prof_mark(c, st)) return 1;

    auto base = [&](int l, int lv) {
        ConvParams p;
        memset(&p, 0, sizeof(p));
        p.B = nb;
        p.H = dm.h[lv];
        p.W = dm.w[lv];
        p.kb_count = kLayers[l].cin / 64;
        p.n_tiles = 1;
        p.bias = c->bias_dev[l];
        p.cout = kLayers[l].cout;
        p.w_rows_per_tap = std::max(kLayers[l].cout, 16);
        p.slope = 1.f;  // EPI_CONVT: no activation after the transposed convs
        return p;
    };
    // N-split form of a 3x3 layer: column blocks of 64 through the streamed-weight single-CTA kernel
    auto split_relu = [&](int l, ConvParams p) {
        p.n_tiles = kLayers[l].cout / 64;
        return launch_conv<64, 9, 1, false, EPI_RELU, 10, 1>(c, 22, pl->tmA[l], c->tmW64[l], pl->tmO[l], pl->tmO[l], p, st, true);
    };
    auto split_pool = [&](int l, ConvParams p) {
        p.n_tiles = kLayers[l].cout / 64;
        return launch_conv<64, 9, 1, false, EPI_RELU_POOL, 10, 1>(c, 23, pl->tmA[l], c->tmW64[l], pl->tmO[l], pl->tmO2[l], p, st, true);
    };

    {  // L1 down1.2 -> cat1[:,64:128] + p1
        ConvParams p = base(1, 0);
        p.out = buf(B_CAT1); p.out_cs = 128; p.out_c0 = 64;
        p.out2 = buf(B_P1); p.out2_cs = 64; p.out2_c0 = 0;
        if (pick_split(c, tiles(0), 64, 2, 2)) {  // one tile per CTA, weights resident
            if (launch_conv<64, 9, 1, true, EPI_RELU_POOL, 10, 1>(c, 24, pl->tmA[1], c->tmW[1], pl->tmO[1], pl->tmO2[1], p, st)) return 1;
        } else {
            if (launch_conv<64, 9, 2, true, EPI_RELU_POOL, 10, 1, 2>(c, 10, pl->tmA[1], c->tmWh[1], pl->tmO[1], pl->tmO2[1], p, st, false, 0, c->ast[1])) return 1;
        }
    }
    if (prof_mark(c, st)) return 1;
    {  // L2 down2.0 -> a2
        ConvParams p = base(2, 1);
        p.out = buf(B_A2); p.out_cs = 128; p.out_c0 = 0;
        if (pick_split(c, tiles(1), 128, 1, 2)) {
            if (split_relu(2, p)) return 1;
        } else {  // CTA pair: the resident weights (147 KB) only fit next to the staging buffers when halved
            if (launch_conv<128, 9, 1, true, EPI_RELU, 10, 1, 2>(c, 11, pl->tmA[2], c->tmWh[2], pl->tmO[2], pl->tmO[2], p, st, false, 0, c->ast[2])) return 1;
        }
    }
    if (prof_mark(c, st)) return 1;
    {  // L3 down2.2 -> cat2[:,128:256] + p2
        ConvParams p = base(3, 1);
        p.out = buf(B_CAT2); p.out_cs = 256; p.out_c0 = 128;
        p.out2 = buf(B_P2); p.out2_cs = 128; p.out2_c0 = 0;
        if (pick_split(c, tiles(1), 128, 2, 2)) {
            if (split_pool(3, p)) return 1;
        } else {
            if (launch_conv<128, 9, 2, false, EPI_RELU_POOL, 10, 1, 2>(c, 12, pl->tmA[3], c->tmWh[3], pl->tmO[3], pl->tmO2[3], p, st, false, c->wst[3])) return 1;
        }
    }
    if (prof_mark(c, st)) return 1;
    for (int l = 4; l <= 5; ++l) {  // L4 bottleneck.0 -> b0, L5 bottleneck.2 -> b1
        ConvParams p = base(l, 2);
        p.out = buf(l == 4 ? B_B0 : B_B1); p.out_cs = 256; p.out_c0 = 0;
        if (pick_split(c, tiles(2), 256, 1, 1)) {
            if (split_relu(l, p)) return 1;
        } else {  // CTA pair: each CTA streams half of every 32 KB weight tile (single CTAs: bottleneck.2 0.239 ms, pair 0.210)
            if (launch_conv<256, 9, 1, false, EPI_RELU, 10, 1, 2>(c, 3, pl->tmA[l], c->tmWh[l], pl->tmO[l], pl->tmO[l], p, st, false, c->wst[l])) return 1;
        }
        if (prof_mark(c, st)) return 1;
    }
    {  // L6 up2 (convT) -> cat2[:,0:128]; 4*128 columns in two blocks of 256. CTA pair; a pair keeps ITS column block's half of
       // the weight matrix (2 x 64 KB) resident and walks every pixel tile (the A tiles are read twice, the second time from
       // L2): 0.105 -> 0.093 ms against streaming half of the 256 KB matrix per tile pair
        ConvParams p = base(6, 2);
        p.n_tiles = 2;
        p.out = buf(B_CAT2); p.out_cs = 256; p.out_c0 = 0;
        if (launch_conv<256, 1, 1, true, EPI_CONVT, 0, 1, 2>(c, 20, pl->tmA[6], c->tmWh[6], pl->tmO[6], pl->tmO[6], p, st, false, 0, c->ast[6])) return 1;
    }
    if (prof_mark(c, st)) return 1;
    for (int l = 7; l <= 8; ++l) {  // L7 upconv2.0 (reads cat2) -> u2a, L8 upconv2.2 -> u2b
        ConvParams p = base(l, 3);
        p.out = buf(l == 7 ? B_U2A : B_U2B); p.out_cs = 128; p.out_c0 = 0;
        if (pick_split(c, tiles(3), 128, 2, 2)) {
            if (split_relu(l, p)) return 1;
        } else {
            if (launch_conv<128, 9, 2, false, EPI_RELU, 10, 1, 2>(c, 14, pl->tmA[l], c->tmWh[l], pl->tmO[l], pl->tmO[l], p, st, false, c->wst[l])) return 1;
        }
        if (prof_mark(c, st)) return 1;
    }
    {  // L9 up1 (convT) -> cat1[:,0:64]; the whole 128 x 256 weight matrix (64 KB) stays in shared memory
        ConvParams p = base(9, 3);
        p.n_tiles = 1;  // 4*64 columns in one tile of 256
        p.out = buf(B_CAT1); p.out_cs = 128; p.out_c0 = 0;
        if (launch_conv<256, 1, 1, true, EPI_CONVT, 0, 1>(c, 21, pl->tmA[9], c->tmW[9], pl->tmO[9], pl->tmO[9], p, st, false, 0, c->ast[9])) return 1;
    }
    if (prof_mark(c, st)) return 1;
    {  // L10 upconv1.0 (reads cat1) -> u1a; CTA pair (resident weights halved: 74 KB per CTA)
        ConvParams p = base(10, 4);
        p.out = buf(B_U1A); p.out_cs = 64; p.out_c0 = 0;
        if (launch_conv<64, 9, 1, true, EPI_RELU, 10, 1, 2>(c, 15, pl->tmA[10], c->tmWh[10], pl->tmO[10], pl->tmO[10], p, st, false, 0, c->ast[10])) return 1;
    }
    if (prof_mark(c, st)) return 1;
    {  // L11 upconv1.2 + tanh -> y   (transposed GEMM: weights on M, the halo patch on N)
        LastParams lp;
        if (make_last_params(c, pl, nb, io, &lp)) return 1;
        const int smem = last_conv_smem(lp.stages);
        if (!(c->attr_done & (1u << 29))) {
            CID_CHECK_CUDA(cudaFuncSetAttribute(last_conv_umma_kernel<EPI_TANH_U8>, cudaFuncAttributeMaxDynamicSharedMemorySize, c->max_smem));
            CID_CHECK_CUDA(cudaFuncSetAttribute(last_conv_umma_kernel<EPI_TANH_F32>, cudaFuncAttributeMaxDynamicSharedMemorySize, c->max_smem));
            c->attr_done |= (1u << 29);
        }
        const int grid = std::min(lp.m_tiles, c->num_sms);
        if (io.u8)
            CID_CHECK_CUDA(launch_ex(c, last_conv_umma_kernel<EPI_TANH_U8>, grid, kLastThreads, smem, st, 1, pl->tmA[11], c->tmW[11], lp));
        else
            CID_CHECK_CUDA(launch_ex(c, last_conv_umma_kernel<EPI_TANH_F32>, grid, kLastThreads, smem, st, 1, pl->tmA[11], c->tmW[11], lp));
        CID_CHECK_CUDA(cudaGetLastError());
        c->launches++;
    }
    if (prof_mark(c, st)) return 1;
    return 0;
}

// ---------------------------------------------------------------------------------------------
// CUDA-graph replay of one pass (small batches only). A batch-1 forward is 12 launches of a few microseconds
// each: the host's launch calls, not the kernels, would set the pace. The first pass of a plan runs eagerly
// (it also sets the kernels' attributes), the second is captured on the context's own stream - with the
// programmatic-dependent-launch edges between the layers - and from then on a pass is ONE cudaGraphLaunch into
// the caller's stream. The caller's input / output pointers and padding live in the parameters of the first and
// last kernel node: when they change, those two nodes are updated in the instantiated graph (no re-capture).
// ---------------------------------------------------------------------------------------------
static bool same_io(const IoSpec& a, const IoSpec& b) {
    return a.u8 == b.u8 && a.x == b.x && a.y == b.y && a.in_h == b.in_h && a.in_w == b.in_w && a.pad_l == b.pad_l &&
           a.pad_t == b.pad_t;
}

static bool graph_wanted(const cid_ctx* c, const cid_plan* pl, int nb) {
    if (!c->use_graph || c->profile || c->graph_broken || nb != pl->Bc) return false;
    return static_cast<int64_t>(nb) * pl->H * pl->W <= c->graph_max_px;  // longer passes are not launch-bound
}

static int graph_capture(cid_ctx* c, cid_plan* pl, int nb, const IoSpec& io, cid_graph* g) {
    if (!c->capture_stream) CID_CHECK_CUDA(cudaStreamCreateWithFlags(reinterpret_cast<cudaStream_t*>(&c->capture_stream), cudaStreamNonBlocking));
    cudaStream_t cs = static_cast<cudaStream_t>(c->capture_stream);
    cudaGraph_t graph = nullptr;
    const int64_t l0 = c->launches;
    CID_CHECK_CUDA(cudaStreamBeginCapture(cs, cudaStreamCaptureModeThreadLocal));
    const int rc = run_layers(c, pl, nb, io, cs);
    const cudaError_t ee = cudaStreamEndCapture(cs, &graph);
    g->launches = static_cast<int>(c->launches - l0);
    c->launches = l0;
    if (rc) {
        if (graph) cudaGraphDestroy(graph);
        return 1;
    }
    CID_CHECK_CUDA(ee);
    cudaGraphExec_t exec = nullptr;
    cudaError_t e = cudaGraphInstantiate(&exec, graph, 0);
    if (e != cudaSuccess) {
        cudaGraphDestroy(graph);
        CID_CHECK_CUDA(e);
    }
    g->graph = graph;
    g->exec = exec;
    // the nodes that carry the caller's pointers: the first and the last kernel of the pass
    size_t n = 0;
    CID_CHECK_CUDA(cudaGraphGetNodes(graph, nullptr, &n));
    std::vector<cudaGraphNode_t> nodes(n);
    CID_CHECK_CUDA(cudaGraphGetNodes(graph, nodes.data(), &n));
    const void* f_first = io.u8 ? reinterpret_cast<const void*>(first_conv_umma_kernel<true>)
                                : reinterpret_cast<const void*>(first_conv_umma_kernel<false>);
    const void* f_last = io.u8 ? reinterpret_cast<const void*>(last_conv_umma_kernel<EPI_TANH_U8>)
                               : reinterpret_cast<const void*>(last_conv_umma_kernel<EPI_TANH_F32>);
    int found = 0;
    for (size_t i = 0; i < n; ++i) {
        cudaGraphNodeType ty;
        if (cudaGraphNodeGetType(nodes[i], &ty) != cudaSuccess || ty != cudaGraphNodeTypeKernel) continue;
        cudaKernelNodeParams kp;
        if (cudaGraphKernelNodeGetParams(nodes[i], &kp) != cudaSuccess) continue;
        if (kp.func == f_first) { g->first = nodes[i]; ++found; }
        if (kp.func == f_last) { g->last = nodes[i]; ++found; }
    }
    CID_REQUIRE(found == 2 && g->first && g->last, "graph: could not identify the first / last kernel node (%d)", found);
    g->io_u8 = io.u8; g->x = io.x; g->y = io.y; g->in_h = io.in_h; g->in_w = io.in_w; g->pad_l = io.pad_l; g->pad_t = io.pad_t;
    return 0;
}

static int graph_rebind(cid_ctx* c, cid_plan* pl, int nb, const IoSpec& io, cid_graph* g) {
    cudaGraphExec_t exec = static_cast<cudaGraphExec_t>(g->exec);
    {
        FirstParams fp = make_first_params(c, pl, nb, io);
        CUtensorMap tmO = pl->tmO[0];
        void* args[2] = {&tmO, &fp};
        cudaKernelNodeParams kp;
        CID_CHECK_CUDA(cudaGraphKernelNodeGetParams(static_cast<cudaGraphNode_t>(g->first), &kp));
        kp.kernelParams = args;
        kp.extra = nullptr;
        CID_CHECK_CUDA(cudaGraphExecKernelNodeSetParams(exec, static_cast<cudaGraphNode_t>(g->first), &kp));
    }
    {
        LastParams lp;
        if (make_last_params(c, pl, nb, io, &lp)) return 1;
        CUtensorMap tmA = pl->tmA[11], tmW = c->tmW[11];
        void* args[3] = {&tmA, &tmW, &lp};
        cudaKernelNodeParams kp;
        CID_CHECK_CUDA(cudaGraphKernelNodeGetParams(static_cast<cudaGraphNode_t>(g->last), &kp));
        kp.kernelParams = args;
        kp.extra = nullptr;
        CID_CHECK_CUDA(cudaGraphExecKernelNodeSetParams(exec, static_cast<cudaGraphNode_t>(g->last), &kp));
    }
    g->x = io.x; g->y = io.y; g->in_h = io.in_h; g->in_w = io.in_w; g->pad_l = io.pad_l; g->pad_t = io.pad_t;
    return 0;
}

// run_layers, or its graph replay when the pass is launch-bound
static int run_pass(cid_ctx* c, cid_plan* pl, int nb, const IoSpec& io, cudaStream_t st) {
    if (!graph_wanted(c, pl, nb)) return run_layers(c, pl, nb, io, st);
    cid_graph& g = pl->graph[io.u8 ? 1 : 0];
    if (!g.exec) {
        if (g.eager_runs++ == 0) return run_layers(c, pl, nb, io, st);  // first pass: eager (sets kernel attributes)
        if (graph_capture(c, pl, nb, io, &g)) {
            // capture is an optimisation: fall back to eager launches for good, keep the reason for cid_last_error readers
            graph_release(g);
            c->graph_broken = true;
            cudaGetLastError();
            return run_layers(c, pl, nb, io, st);
        }
    } else {
        IoSpec baked;
        baked.u8 = g.io_u8; baked.x = g.x; baked.y = g.y; baked.in_h = g.in_h; baked.in_w = g.in_w; baked.pad_l = g.pad_l; baked.pad_t = g.pad_t;
        if (!same_io(baked, io) && graph_rebind(c, pl, nb, io, &g)) {
            graph_release(g);
            c->graph_broken = true;
            cudaGetLastError();
            return run_layers(c, pl, nb, io, st);
        }
    }
    CID_CHECK_CUDA(cudaGraphLaunch(static_cast<cudaGraphExec_t>(g.exec), st));
    c->launches += g.launches;
    c->graph_replays++;
    return 0;
}

}  // namespace cid
#include "cgan_model.cuh"
#include "sr_model.cuh"
#include "train_model.cuh"
namespace cid {

static int chunk_of(const cid_ctx* c, int B) {
    return (c->chunk > 0 && c->chunk < B) ? static_cast<int>(c->chunk) : B;
}

static int forward_common(cid_ctx* c, int B, int H, int W, void* ws, size_t ws_bytes, cudaStream_t st,
                          const uint8_t* x, size_t x_img_bytes, uint8_t* y, size_t y_img_bytes, IoSpec io) {
    CID_REQUIRE(c && c->finalized, "generator not finalized (call cid_set_param for every tensor, then cid_finalize)");
    const bool sr = c->model == CID_MODEL_ESRGAN || c->model == CID_MODEL_SRGAN;
    if (sr) {  // fully convolutional, stride 1: any size
        CID_REQUIRE(B >= 1 && H >= 1 && W >= 1, "bad shape B=%d H=%d W=%d", B, H, W);
    } else {
        CID_REQUIRE(B >= 1 && H >= 4 && W >= 4, "bad shape B=%d H=%d W=%d", B, H, W);
        // the denoise U-Net takes any size >= 4 on its fp32 entry (the result is 4 * floor(H / 4) x 4 * floor(W / 4), as the
        // reference's forward gives, backend/app.py:89-99); the fused u8 route pads to multiples of 4 first (get_padding,
        // :276-281) and the Keras generator's stride-2 'same' convolutions are only laid out for multiples of 4
        CID_REQUIRE((c->model == CID_MODEL_DENOISE && !io.u8) || (H % 4 == 0 && W % 4 == 0),
                    "network input %dx%d must be multiples of 4 (get_padding, app.py:276)", H, W);
    }
    CID_REQUIRE(x && y && ws, "null pointer argument");
    CID_REQUIRE((reinterpret_cast<uintptr_t>(ws) & 1023) == 0, "workspace must be 1024-byte aligned");
    DeviceGuard g(c->device);
    CID_REQUIRE(g.ok, "cannot select CUDA device %d", c->device);
    const int Bc = chunk_of(c, B);
    const bool cgan = c->model == CID_MODEL_KERAS_CGAN;
    const size_t need = sr     ? sr_layout_offsets(c->model, Bc, H, W, nullptr)
                        : cgan ? cgan_layout_offsets(Bc, H, W, nullptr)
                               : layout_offsets(Bc, H, W, nullptr);
    CID_REQUIRE(ws_bytes >= need, "workspace too small: %zu < %zu bytes", ws_bytes, need);
    cid_plan* pl = nullptr;
    if (sr ? sr_get_plan(c, Bc, H, W, ws, &pl) : cgan ? cgan_get_plan(c, Bc, H, W, ws, &pl) : get_plan(c, Bc, H, W, ws, &pl))
        return 1;
    for (int b0 = 0; b0 < B; b0 += Bc) {
        const int nb = std::min(Bc, B - b0);
        io.x = x + static_cast<size_t>(b0) * x_img_bytes;
        io.y = y + static_cast<size_t>(b0) * y_img_bytes;
        if (sr ? sr_run_layers(c, pl, nb, io, st) : cgan ? cgan_run_layers(c, pl, nb, io, st) : run_pass(c, pl, nb, io, st))
            return 1;
    }
    return 0;
}

}  // namespace cid

using namespace cid;

// =============================================================================================
// C ABI
// =============================================================================================
extern "C" {

int cid_version(void) { return 7; }

const char* cid_last_error(void) { return g_err; }

int cid_create(int device, cid_handle* out) { return cid_create_model(device, "denoise", out); }

int cid_create_model(int device, const char* model, cid_handle* out) {
    CID_REQUIRE(out, "cid_create: out is NULL");
    CID_REQUIRE(model, "cid_create_model: model is NULL");
    int kind = -1;
    if (!strcmp(model, "denoise")) kind = CID_MODEL_DENOISE;
    if (!strcmp(model, "keras_cgan")) kind = CID_MODEL_KERAS_CGAN;
    if (!strcmp(model, "esrgan")) kind = CID_MODEL_ESRGAN;
    if (!strcmp(model, "srgan")) kind = CID_MODEL_SRGAN;
    CID_REQUIRE(kind >= 0, "unknown model '%s' (denoise | keras_cgan | esrgan | srgan)", model);
    *out = nullptr;
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    CID_REQUIRE(e == cudaSuccess && n > 0, "no CUDA device available (%s); this library has no CPU fallback",
                e == cudaSuccess ? "device count is 0" : cudaGetErrorString(e));
    CID_REQUIRE(device >= 0 && device < n, "device %d out of range [0,%d)", device, n);
    cudaDeviceProp prop;
    CID_CHECK_CUDA(cudaGetDeviceProperties(&prop, device));
    CID_REQUIRE(prop.major == 10, "device %d is sm_%d%d; this library is built for sm_100a (B200) only", device,
                prop.major, prop.minor);
    cid_ctx* c = new cid_ctx();
    c->device = device;
    c->model = kind;
    c->num_sms = prop.multiProcessorCount;
    c->max_smem = static_cast<int>(prop.sharedMemPerBlockOptin);
    *out = c;
    return 0;
}

int cid_destroy(cid_handle h) {
    if (!h) return 0;
    DeviceGuard g(h->device);
    clear_plans(h);
    if (h->capture_stream) cudaStreamDestroy(static_cast<cudaStream_t>(h->capture_stream));
    for (int l = 0; l < kMaxLayers; ++l) {
        if (h->w_dev[l]) cudaFree(h->w_dev[l]);
        if (h->bias_dev[l]) cudaFree(h->bias_dev[l]);
    }
    delete h;
    return 0;
}

int cid_set_param(cid_handle h, const char* key, const float* data, const int64_t* shape, int ndim) {
    CID_REQUIRE(h && key && data && shape, "cid_set_param: null argument");
    const std::string k(key);
    if (h->model != CID_MODEL_DENOISE) {
        int64_t expect[4];
        const int en = h->model == CID_MODEL_KERAS_CGAN ? cgan_param_shape(k, expect) : sr_param_shape(h->model, k, expect);
        CID_REQUIRE(en > 0, "unknown parameter '%s' for this generator", key);
        CID_REQUIRE(ndim == en, "%s: expected %d dims, got %d", key, en, ndim);
        size_t n = 1;
        for (int i = 0; i < en; ++i) {
            CID_REQUIRE(shape[i] == expect[i], "%s: dim %d is %lld, expected %lld", key, i,
                        static_cast<long long>(shape[i]), static_cast<long long>(expect[i]));
            n *= static_cast<size_t>(shape[i]);
        }
        DeviceGuard g(h->device);
        CID_REQUIRE(g.ok, "cannot select CUDA device %d", h->device);
        std::vector<float> v(n);
        CID_CHECK_CUDA(cudaMemcpy(v.data(), data, n * 4, cudaMemcpyDefault));
        h->params[k] = std::move(v);
        h->finalized = false;
        return 0;
    }
    const size_t dot = k.rfind('.');
    CID_REQUIRE(dot != std::string::npos, "unknown parameter key '%s'", key);
    const std::string lname = k.substr(0, dot), what = k.substr(dot + 1);
    int l = -1;
    for (int i = 0; i < 12; ++i)
        if (lname == kLayers[i].name) l = i;
    CID_REQUIRE(l >= 0 && (what == "weight" || what == "bias"), "unknown parameter key '%s'", key);
    const LayerDesc& L = kLayers[l];
    int64_t expect[4];
    int en;
    if (what == "bias") {
        expect[0] = L.cout;
        en = 1;
    } else if (L.kind == L_CONVT) {
        expect[0] = L.cin; expect[1] = L.cout; expect[2] = 2; expect[3] = 2;
        en = 4;
    } else {
        expect[0] = L.cout; expect[1] = L.cin; expect[2] = 3; expect[3] = 3;
        en = 4;
    }
    CID_REQUIRE(ndim == en, "%s: expected %d dims, got %d", key, en, ndim);
    size_t n = 1;
    for (int i = 0; i < en; ++i) {
        CID_REQUIRE(shape[i] == expect[i], "%s: dim %d is %lld, expected %lld", key, i,
                    static_cast<long long>(shape[i]), static_cast<long long>(expect[i]));
        n *= static_cast<size_t>(shape[i]);
    }
    DeviceGuard g(h->device);
    CID_REQUIRE(g.ok, "cannot select CUDA device %d", h->device);
    std::vector<float> v(n);
    CID_CHECK_CUDA(cudaMemcpy(v.data(), data, n * 4, cudaMemcpyDefault));
    h->params[k] = std::move(v);
    h->finalized = false;
    return 0;
}

int cid_finalize(cid_handle h) {
    CID_REQUIRE(h, "cid_finalize: null handle");
    DeviceGuard g(h->device);
    CID_REQUIRE(g.ok, "cannot select CUDA device %d", h->device);
    CID_CHECK_CUDA(cudaDeviceSynchronize());  // no forward may still be reading the old packed weights
    if (h->model == CID_MODEL_KERAS_CGAN) {
        if (cgan_pack(h)) return 1;
    } else if (h->model != CID_MODEL_DENOISE) {
        if (sr_pack(h)) return 1;
    } else {
        for (int l = 0; l < 12; ++l)
            if (pack_layer(h, l)) return 1;
    }
    clear_plans(h);
    h->finalized = true;
    return 0;
}

int cid_set_option(cid_handle h, const char* name, int64_t value) {
    CID_REQUIRE(h && name, "cid_set_option: null argument");
    if (!strcmp(name, "chunk")) {
        CID_REQUIRE(value >= 0, "chunk must be >= 0");
        h->chunk = value;
        return 0;
    }
    // 0 = choose per layer (pick_split), 1 = always the batch-256 configuration, 2 = always the N-split small-batch form
    // (the two are bit-identical; the tests use 1 / 2 to compare them at any size)
    if (!strcmp(name, "regime")) {
        CID_REQUIRE(value >= 0 && value <= 2, "regime must be 0 (auto), 1 (large) or 2 (split)");
        h->regime = static_cast<int>(value);
        for (auto& pl : h->plans) { graph_release(pl.graph[0]); graph_release(pl.graph[1]); }
        return 0;
    }
    if (!strncmp(name, "wst", 3) && name[3] >= '0' && name[3] <= '9') {  // "wst7" = weight-ring depth of layer 7 (tuning)
        const int l = atoi(name + 3);
        CID_REQUIRE(l >= 0 && l < 16 && value >= 0 && value <= kMaxWStages, "wst<layer>: layer 0..15, depth 0..%d", kMaxWStages);
        h->wst[l] = static_cast<int>(value);
        return 0;
    }
    if (!strncmp(name, "ast", 3) && name[3] >= '0' && name[3] <= '9') {  // "ast10" = cap on the A-ring depth of layer 10 (tuning)
        const int l = atoi(name + 3);
        CID_REQUIRE(l >= 0 && l < 16 && value >= 0 && value <= kMaxStages, "ast<layer>: layer 0..15, depth 0..%d", kMaxStages);
        h->ast[l] = static_cast<int>(value);
        return 0;
    }
    if (!strcmp(name, "sp_scatter")) {  // salt & pepper noiser: 1 = copy + scatter launches instead of the one-pass slab kernel
        h->sp_scatter = value != 0;
        return 0;
    }
    if (!strcmp(name, "graph_max_px")) {
        CID_REQUIRE(value >= 0, "graph_max_px must be >= 0");
        h->graph_max_px = value;
        return 0;
    }
    if (!strcmp(name, "graph")) {  // CUDA-graph replay of launch-bound passes (default on)
        h->use_graph = value != 0;
        return 0;
    }
    if (!strcmp(name, "pdl")) {  // programmatic dependent launch between consecutive layers (default on)
        h->pdl = value != 0;
        return 0;
    }
    if (!strcmp(name, "profile")) {
        h->profile = value != 0;
        h->prof_used = 0;
        return 0;
    }
    set_error("unknown option '%s'", name);
    return 1;
}

int cid_workspace_bytes(cid_handle h, int B, int H, int W, size_t* bytes) {
    CID_REQUIRE(h && bytes, "cid_workspace_bytes: null argument");
    if (h->model == CID_MODEL_ESRGAN || h->model == CID_MODEL_SRGAN) {
        CID_REQUIRE(B >= 1 && H >= 1 && W >= 1, "bad shape B=%d H=%d W=%d", B, H, W);
        *bytes = sr_layout_offsets(h->model, chunk_of(h, B), H, W, nullptr);
        return 0;
    }
    CID_REQUIRE(B >= 1 && H >= 4 && W >= 4, "bad shape B=%d H=%d W=%d", B, H, W);
    CID_REQUIRE(h->model == CID_MODEL_DENOISE || (H % 4 == 0 && W % 4 == 0),
                "network input %dx%d must be multiples of 4 (get_padding, app.py:276)", H, W);
    *bytes = h->model == CID_MODEL_KERAS_CGAN ? cgan_layout_offsets(chunk_of(h, B), H, W, nullptr)
                                              : layout_offsets(chunk_of(h, B), H, W, nullptr);
    return 0;
}

int cid_forward_f32_nhwc(cid_handle h, const float* x, float* y, int B, int H, int W, void* ws, size_t ws_bytes,
                         void* stream) {
    CID_REQUIRE(h && h->model == CID_MODEL_KERAS_CGAN, "cid_forward_f32_nhwc is the keras_cgan generator's entry point");
    IoSpec io;
    memset(&io, 0, sizeof(io));
    io.u8 = false;
    const size_t img = static_cast<size_t>(3) * H * W * 4;
    return forward_common(h, B, H, W, ws, ws_bytes, static_cast<cudaStream_t>(stream),
                          reinterpret_cast<const uint8_t*>(x), img, reinterpret_cast<uint8_t*>(y), img, io);
}

int cid_forward_f32_nchw(cid_handle h, const float* x, float* y, int B, int H, int W, void* ws, size_t ws_bytes,
                         void* stream) {
    CID_REQUIRE(h && h->model != CID_MODEL_KERAS_CGAN, "cid_forward_f32_nchw is the torch generators' (denoise, esrgan, "
                                                       "srgan) entry point (keras_cgan: cid_forward_f32_nhwc)");
    IoSpec io;
    memset(&io, 0, sizeof(io));
    io.u8 = false;
    const size_t img = static_cast<size_t>(3) * H * W * 4;
    size_t out_img = h->model == CID_MODEL_SRGAN ? img * 16 : img;  // SRGAN: [B,3,4H,4W]
    if (h->model == CID_MODEL_DENOISE)  // [B,3,4*floor(H/4),4*floor(W/4)]: == [B,3,H,W] for multiples of 4
        out_img = static_cast<size_t>(3) * (H / 4 * 4) * (W / 4 * 4) * 4;
    return forward_common(h, B, H, W, ws, ws_bytes, static_cast<cudaStream_t>(stream),
                          reinterpret_cast<const uint8_t*>(x), img, reinterpret_cast<uint8_t*>(y), out_img, io);
}

int cid_forward_u8_nhwc(cid_handle h, const uint8_t* x, uint8_t* y, int B, int H, int W, int pad_l, int pad_t,
                        int pad_r, int pad_b, void* ws, size_t ws_bytes, void* stream) {
    CID_REQUIRE(pad_l >= 0 && pad_t >= 0 && pad_r >= 0 && pad_b >= 0, "negative padding");
    CID_REQUIRE(H >= 1 && W >= 1, "bad image size %dx%d", H, W);
    IoSpec io;
    memset(&io, 0, sizeof(io));
    io.u8 = true;
    io.in_h = H;
    io.in_w = W;
    io.pad_l = pad_l;
    io.pad_t = pad_t;
    const size_t img = static_cast<size_t>(3) * H * W;
    if (h && h->model == CID_MODEL_ESRGAN) {
        // the esrgan branch feeds the UN-padded image (preprocess_image, backend/app.py:387-390) and then crops the
        // un-padded output at the padding offsets (:473-479): PIL fills what falls outside with black
        if (pad_l || pad_t) {
            DeviceGuard g(h->device);
            CID_CHECK_CUDA(cudaMemsetAsync(y, 0, img * B, static_cast<cudaStream_t>(stream)));
        }
        return forward_common(h, B, H, W, ws, ws_bytes, static_cast<cudaStream_t>(stream), x, img, y, img, io);
    }
    const size_t out_img = (h && h->model == CID_MODEL_SRGAN)
                               ? static_cast<size_t>(3) * 16 * (H + pad_t + pad_b) * (W + pad_l + pad_r)
                               : img;  // SRGAN: the 4x output of the PADDED frame, not cropped (:480-485)
    return forward_common(h, B, H + pad_t + pad_b, W + pad_l + pad_r, ws, ws_bytes, static_cast<cudaStream_t>(stream),
                          x, img, y, out_img, io);
}

int64_t cid_kernel_launches(cid_handle h) { return h ? h->launches : -1; }

int64_t cid_graph_replays(cid_handle h) { return h ? h->graph_replays : -1; }

int cid_profile_read(cid_handle h, double* layer_ms, int64_t* passes) {
    CID_REQUIRE(h && layer_ms && passes, "cid_profile_read: null argument");
    DeviceGuard g(h->device);
    CID_REQUIRE(g.ok, "cannot select CUDA device %d", h->device);
    CID_CHECK_CUDA(cudaDeviceSynchronize());
    for (int l = 0; l < kMaxLayers; ++l) layer_ms[l] = 0.0;
    const int nl = h->model == CID_MODEL_KERAS_CGAN ? 5 : h->model == CID_MODEL_ESRGAN ? 18 : h->model == CID_MODEL_SRGAN ? 15 : 12;
    const size_t n = h->prof_used / (nl + 1);
    for (size_t i = 0; i < n; ++i)
        for (int l = 0; l < nl; ++l) {
            float ms = 0.f;
            CID_CHECK_CUDA(cudaEventElapsedTime(&ms, h->prof_events[i * (nl + 1) + l], h->prof_events[i * (nl + 1) + l + 1]));
            layer_ms[l] += ms;
        }
    *passes = static_cast<int64_t>(n);
    h->prof_used = 0;
    return 0;
}

int cid_debug_copy_activation(cid_handle h, const char* name, const void* ws, int B, int H, int W, void* dst_host,
                              size_t dst_bytes, int* channels, int* height, int* width) {
    CID_REQUIRE(h && name && ws && dst_host, "cid_debug_copy_activation: null argument");
    if (h->model == CID_MODEL_KERAS_CGAN) {
        static const char* names[CB_COUNT] = {"c0", "c1", "c2", "c3"};
        int b = -1;
        for (int i = 0; i < CB_COUNT; ++i)
            if (!strcmp(name, names[i])) b = i;
        CID_REQUIRE(b >= 0, "unknown activation '%s'", name);
        DeviceGuard g(h->device);
        CID_REQUIRE(g.ok, "cannot select CUDA device %d", h->device);
        const int Bc = chunk_of(h, B);
        size_t off[CB_COUNT];
        cgan_layout_offsets(Bc, H, W, off);
        const int lv = kCganBufLevel[b];
        const size_t n = static_cast<size_t>(Bc) * (H >> lv) * (W >> lv) * kCganBufC[b] * 2;
        CID_REQUIRE(dst_bytes >= n, "destination too small: %zu < %zu", dst_bytes, n);
        CID_CHECK_CUDA(cudaDeviceSynchronize());
        CID_CHECK_CUDA(cudaMemcpy(dst_host, static_cast<const uint8_t*>(ws) + off[b], n, cudaMemcpyDeviceToHost));
        if (channels) *channels = kCganBufC[b];
        if (height) *height = H >> lv;
        if (width) *width = W >> lv;
        return 0;
    }
    int b = -1;
    for (int i = 0; i < B_COUNT; ++i)
        if (!strcmp(name, kBufNames[i])) b = i;
    CID_REQUIRE(b >= 0, "unknown activation '%s'", name);
    DeviceGuard g(h->device);
    CID_REQUIRE(g.ok, "cannot select CUDA device %d", h->device);
    const int Bc = chunk_of(h, B);
    size_t off[B_COUNT];
    layout_offsets(Bc, H, W, off);
    const int lv = kBufLevel[b];
    const LevelDims dm = level_dims(H, W);
    const size_t n = static_cast<size_t>(Bc) * dm.h[lv] * dm.w[lv] * kBufC[b] * 2;
    CID_REQUIRE(dst_bytes >= n, "destination too small: %zu < %zu", dst_bytes, n);
    CID_CHECK_CUDA(cudaDeviceSynchronize());
    CID_CHECK_CUDA(cudaMemcpy(dst_host, static_cast<const uint8_t*>(ws) + off[b], n, cudaMemcpyDeviceToHost));
    if (channels) *channels = kBufC[b];
    if (height) *height = dm.h[lv];
    if (width) *width = dm.w[lv];
    return 0;
}


// =============================================================================================
// training step
// =============================================================================================
static const char* const kPeerFaultMsg =
    "a data-parallel peer did not reach the fused all-reduce barrier in time: that optimiser step and every later one were "
    "skipped (weights untouched from the fault on); the replicas may differ - restore a checkpoint on every rank";

int cid_train_create(int device, cid_train_handle* out) {
    CID_REQUIRE(out, "cid_train_create: out is NULL");
    *out = nullptr;
    cid_handle g = nullptr;
    if (cid_create_model(device, "denoise", &g)) return 1;
    cid_train* t = new cid_train();
    t->g = g;
    g->use_graph = false;  // the training step launches run_layers directly (its buffers are read back by the backward pass)
    t->device = device;
    t->num_sms = g->num_sms;
    t->max_smem = g->max_smem;
    train_build_table(t);
    DeviceGuard dg(device);
    if (!dg.ok || train_alloc_packs(t)) {
        cid_train_destroy(t);
        return 1;
    }
    *out = t;
    return 0;
}

int cid_train_destroy(cid_train_handle t) {
    if (!t) return 0;
    {
        DeviceGuard dg(t->device);
        for (int l = 0; l < 12; ++l)
            if (t->gd_w[l]) cudaFree(t->gd_w[l]);
        for (int l = 0; l < 4; ++l) {
            if (t->df_w[l]) cudaFree(t->df_w[l]);
            if (t->dd_w[l]) cudaFree(t->dd_w[l]);
            if (t->d_bias_pad[l]) cudaFree(t->d_bias_pad[l]);
        }
        if (t->zero_bias) cudaFree(t->zero_bias);
        if (t->peer_err) cudaFree(t->peer_err);
        if (t->peer_err_host) cudaFreeHost(t->peer_err_host);
        if (t->dsc) cudaFree(t->dsc);
        if (t->fsc) cudaFree(t->fsc);
    }
    cid_destroy(t->g);
    delete t;
    return 0;
}

int cid_train_set_option(cid_train_handle t, const char* name, int64_t value) {
    CID_REQUIRE(t && name, "cid_train_set_option: null argument");
    if (!strcmp(name, "peer_timeout_ms")) {  // fused all-reduce: how long a block waits for a peer before it reports a fault
        CID_REQUIRE(value >= 1, "peer_timeout_ms must be >= 1");
        t->peer_spin_ns = static_cast<unsigned long long>(value) * 1000000ull;
        return 0;
    }
    return cid_set_option(t->g, name, value);  // "pdl", ...
}

int cid_train_state_floats(cid_train_handle t, size_t* n) {
    CID_REQUIRE(t && n, "cid_train_state_floats: null argument");
    *n = 4 * train_P(t) + t->n_buf;
    return 0;
}

int cid_train_bind_state(cid_train_handle t, float* state, size_t n) {
    CID_REQUIRE(t && state, "cid_train_bind_state: null argument");
    CID_REQUIRE(n >= 4 * train_P(t) + t->n_buf, "state buffer too small: %zu < %zu floats", n, 4 * train_P(t) + t->n_buf);
    CID_REQUIRE((reinterpret_cast<uintptr_t>(state) & 15) == 0, "state buffer must be 16-byte aligned");
    t->state = state;
    t->state_floats = n;
    return 0;
}

int cid_train_tensor(cid_train_handle t, const char* key, int section, size_t* offset, size_t* numel) {
    CID_REQUIRE(t && key && offset && numel, "cid_train_tensor: null argument");
    CID_REQUIRE(section >= 0 && section < 4, "section %d out of range (0 params, 1 grads, 2 adam m, 3 adam v)", section);
    const TrainTensor* x = train_find(t, key);
    CID_REQUIRE(x, "unknown training tensor '%s'", key);
    CID_REQUIRE(!x->buffer || section == 0, "'%s' is a buffer: it has no gradient or Adam state", key);
    *offset = x->buffer ? 4 * train_P(t) + x->off : section * train_P(t) + x->off;
    *numel = x->n;
    return 0;
}

int cid_train_tensor_count(cid_train_handle t) { return t ? static_cast<int>(t->tensors.size()) : -1; }

const char* cid_train_tensor_name(cid_train_handle t, int index) {
    if (!t || index < 0 || index >= static_cast<int>(t->tensors.size())) return nullptr;
    return t->tensors[index].key.c_str();
}

int cid_train_section(cid_train_handle t, const char* net, int section, size_t* offset, size_t* numel) {
    CID_REQUIRE(t && net && offset && numel, "cid_train_section: null argument");
    CID_REQUIRE(section >= 0 && section < 4, "section %d out of range", section);
    const bool g = !strcmp(net, "generator"), d = !strcmp(net, "discriminator");
    CID_REQUIRE(g || d, "unknown network '%s' (generator | discriminator)", net);
    *offset = section * train_P(t) + (g ? 0 : t->n_g);
    *numel = g ? t->n_g : t->n_d;
    return 0;
}

int cid_train_sync_weights(cid_train_handle t, void* stream) {
    CID_REQUIRE(t && t->state, "cid_train_sync_weights: state not bound");
    DeviceGuard dg(t->device);
    CID_REQUIRE(dg.ok, "cannot select CUDA device %d", t->device);
    return train_repack(t, 3, static_cast<cudaStream_t>(stream));
}

int cid_train_workspace_bytes(cid_train_handle t, int B, int H, int W, size_t* bytes) {
    CID_REQUIRE(t && bytes, "cid_train_workspace_bytes: null argument");
    CID_REQUIRE(B >= 1 && H >= 4 && W >= 4 && H % 4 == 0 && W % 4 == 0, "bad shape B=%d H=%d W=%d", B, H, W);
    *bytes = train_layout(B, H, W, nullptr);
    return 0;
}

int cid_train_d_backward(cid_train_handle t, const float* noisy, const float* clean, int B, int H, int W, void* ws,
                         size_t ws_bytes, void* stream) {
    CID_REQUIRE(t && noisy && clean, "cid_train_d_backward: null argument");
    DeviceGuard dg(t->device);
    CID_REQUIRE(dg.ok, "cannot select CUDA device %d", t->device);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    TrainRun r;
    if (train_make_run(t, B, H, W, ws, ws_bytes, st, &r)) return 1;
    t->B = B; t->H = H; t->W = W; t->ws = ws;
    CID_CHECK_CUDA(cudaMemsetAsync(t->state + train_P(t) + t->n_g, 0, t->n_d * 4, st));
    CID_CHECK_CUDA(cudaMemsetAsync(t->dsc + DS_LOSS, 0, 2 * 8, st));
    // denoised = G(noisy)  (training.py:411)
    cid_plan* pl = nullptr;
    if (get_plan(t->g, B, H, W, ws, &pl)) return 1;
    IoSpec io;
    memset(&io, 0, sizeof(io));
    io.u8 = false;
    io.x = noisy;
    io.y = r.tbuf(TB_Y);
    const int64_t l0 = t->g->launches;
    if (run_layers(t->g, pl, B, io, st)) return 1;
    t->launches += t->g->launches - l0;
    // d_loss = BCE(D(clean), 1) + BCE(D(denoised.detach()), 0)  (:412-415)
    if (d_forward(r, clean)) return 1;
    if (d_backward(r, clean, 1.f, 1.f, 0, true, false)) return 1;
    const float* fake = static_cast<const float*>(r.tbuf(TB_Y));
    if (d_forward(r, fake)) return 1;
    if (d_backward(r, fake, 0.f, 1.f, 1, true, false)) return 1;
    return 0;
}

int cid_train_g_backward(cid_train_handle t, const float* noisy, const float* clean, int B, int H, int W, void* ws,
                         size_t ws_bytes, void* stream) {
    CID_REQUIRE(t && noisy && clean, "cid_train_g_backward: null argument");
    CID_REQUIRE(t->ws == ws && t->B == B && t->H == H && t->W == W,
                "cid_train_g_backward continues the cid_train_d_backward call of the same batch and workspace");
    DeviceGuard dg(t->device);
    CID_REQUIRE(dg.ok, "cannot select CUDA device %d", t->device);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    TrainRun r;
    if (train_make_run(t, B, H, W, ws, ws_bytes, st, &r)) return 1;
    CID_CHECK_CUDA(cudaMemsetAsync(t->state + train_P(t), 0, t->n_g * 4, st));
    CID_CHECK_CUDA(cudaMemsetAsync(t->dsc + DS_LOSS + 2, 0, 2 * 8, st));
    // fake_pred = D(denoised); adv = BCE(fake_pred, 1); g_loss = MSE + 0.001 * adv  (:419-423)
    const float* fake = static_cast<const float*>(r.tbuf(TB_Y));
    if (d_forward(r, fake)) return 1;
    if (d_backward(r, fake, 1.f, 0.001f, 2, false, true)) return 1;
    {
        GLossParams p;
        p.B = B; p.H = H; p.W = W;
        p.y = fake;
        p.clean = clean;
        p.gx = static_cast<const float*>(r.tbuf(TB_GX));
        p.gz = static_cast<float*>(r.tbuf(TB_GZ));
        p.gz_nchw = static_cast<float*>(r.tbuf(TB_GZC));
        p.loss_sum = t->dsc + DS_LOSS + 3;
        p.bias_grad = train_ptr(t, "generator.upconv1.2.bias", 1);
        p.inv_n = 1.f / (static_cast<float>(B) * 3.f * H * W);
        launch_k(t->g->pdl, gloss_kernel, static_cast<unsigned>((r.pixels(0) + 255) / 256), 256, 0, st, p);
        CID_CHECK_CUDA(cudaGetLastError());
        t->launches++;
    }
    return g_backward_layers(r, noisy);
}

int cid_train_adam(cid_train_handle t, int which, float lr, float beta1, float beta2, float eps, void* stream) {
    CID_REQUIRE(t && t->state, "cid_train_adam: state not bound");
    CID_REQUIRE(which == 0 || which == 1, "which = 0 (generator) or 1 (discriminator)");
    DeviceGuard dg(t->device);
    CID_REQUIRE(dg.ok, "cannot select CUDA device %d", t->device);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const size_t P = train_P(t), off = which ? t->n_g : 0, n = which ? t->n_d : t->n_g;
    const int64_t step = which ? ++t->step_d : ++t->step_g;
    const float bc1 = 1.f - powf(beta1, static_cast<float>(step));
    const float bc2 = sqrtf(1.f - powf(beta2, static_cast<float>(step)));
    if (t->peer_world > 1) {
        // gradient all-reduce fused into the optimiser step over NVLink peer memory (adam_allreduce_kernel)
        CID_REQUIRE(*reinterpret_cast<volatile int*>(t->peer_err_host) == 0, "%s", kPeerFaultMsg);
        PeerSet ps;
        memset(&ps, 0, sizeof(ps));
        ps.world = t->peer_world;
        ps.rank = t->peer_rank;
        for (int r = 0; r < t->peer_world; ++r) {
            ps.grads[r] = t->peer_state[r] + P + off;
            ps.flags[r] = t->peer_flags[r];
        }
        adam_allreduce_kernel<<<kPeerBlocks, 256, 0, st>>>(t->state + off, t->state + 2 * P + off, t->state + 3 * P + off, n, ps,
                                                           ++t->peer_epoch, t->peer_spin_ns, t->peer_err, t->peer_err_host_dev, lr, beta1, beta2, eps, bc1,
                                                           bc2);
    } else {
        launch_k(t->g->pdl, adam_kernel, static_cast<unsigned>((n + 255) / 256), 256, 0, st, t->state + off, t->state + P + off, t->state + 2 * P + off,
                                                                            t->state + 3 * P + off, n, lr, beta1, beta2, eps, bc1, bc2);
    }
    CID_CHECK_CUDA(cudaGetLastError());
    t->launches++;
    return train_repack(t, which ? 2 : 1, st);
}

int cid_train_peer_flag_ints(void) { return 2 * kPeerMax * kPeerBlocks; }

int cid_train_set_peers(cid_train_handle t, int world, int rank, const void* const* state_ptrs, void* const* flag_ptrs) {
    CID_REQUIRE(t && t->state, "cid_train_set_peers: state not bound");
    if (world <= 1) {
        t->peer_world = 0;
        return 0;
    }
    CID_REQUIRE(world <= kPeerMax, "cid_train_set_peers: at most %d ranks (one NVLink domain)", kPeerMax);
    CID_REQUIRE(rank >= 0 && rank < world && state_ptrs && flag_ptrs, "cid_train_set_peers: bad rank / null pointer table");
    CID_REQUIRE(state_ptrs[rank] == static_cast<const void*>(t->state),
                "cid_train_set_peers: state_ptrs[rank] must be the bound state buffer (allocate it as symmetric memory)");
    DeviceGuard dg(t->device);
    CID_REQUIRE(dg.ok, "cannot select CUDA device %d", t->device);
    for (int r = 0; r < world; ++r) {
        CID_REQUIRE(state_ptrs[r] && flag_ptrs[r], "cid_train_set_peers: null peer pointer for rank %d", r);
        t->peer_state[r] = static_cast<const float*>(state_ptrs[r]);
        t->peer_flags[r] = static_cast<int*>(flag_ptrs[r]);
    }
    {   // the per-block-index barriers need every block of every rank resident at once
        int per_sm = 0;
        CID_CHECK_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, adam_allreduce_kernel, 256, 0));
        CID_REQUIRE(per_sm * t->num_sms >= kPeerBlocks, "fused all-reduce: %d blocks cannot be co-resident on this device (%d x %d)",
                    kPeerBlocks, per_sm, t->num_sms);
    }
    if (!t->peer_err) {
        CID_CHECK_CUDA(cudaMalloc(reinterpret_cast<void**>(&t->peer_err), sizeof(int)));
        CID_CHECK_CUDA(cudaHostAlloc(reinterpret_cast<void**>(&t->peer_err_host), sizeof(int), cudaHostAllocMapped));
        CID_CHECK_CUDA(cudaHostGetDevicePointer(reinterpret_cast<void**>(&t->peer_err_host_dev), t->peer_err_host, 0));
    }
    CID_CHECK_CUDA(cudaMemset(t->peer_err, 0, sizeof(int)));
    *t->peer_err_host = 0;
    t->peer_world = world;
    t->peer_rank = rank;
    t->peer_epoch = 0;
    return 0;
}

int cid_train_losses(cid_train_handle t, double* out4) {
    CID_REQUIRE(t && out4, "cid_train_losses: null argument");
    DeviceGuard dg(t->device);
    CID_REQUIRE(dg.ok, "cannot select CUDA device %d", t->device);
    double h[4];
    CID_CHECK_CUDA(cudaDeviceSynchronize());
    CID_CHECK_CUDA(cudaMemcpy(h, t->dsc + DS_LOSS, sizeof(h), cudaMemcpyDeviceToHost));
    if (t->peer_err) {
        int e = 0;
        CID_CHECK_CUDA(cudaMemcpy(&e, t->peer_err, sizeof(int), cudaMemcpyDeviceToHost));
        CID_REQUIRE(e == 0 && *reinterpret_cast<volatile int*>(t->peer_err_host) == 0, "%s", kPeerFaultMsg);
    }
    out4[0] = h[0] + h[1];          // d_loss
    out4[1] = h[3] + 0.001 * h[2];  // g_loss
    out4[2] = h[3];                 // content (MSE)
    out4[3] = h[2];                 // adversarial (BCE)
    return 0;
}

int64_t cid_train_kernel_launches(cid_train_handle t) { return t ? t->launches : -1; }

int cid_train_debug_buffer(cid_train_handle t, const char* name, const void* ws, int B, int H, int W, void* dst_host,
                           size_t dst_bytes, int* channels, int* level, int* elem_size) {
    CID_REQUIRE(t && name && ws && dst_host, "cid_train_debug_buffer: null argument");
    static const char* names[TB_COUNT] = {"y", "gz", "gx", "dz10", "gcat1", "dz8", "dz7", "gcat2", "dz5", "dz4", "gp2", "dz3", "dz2",
                                          "gp1", "dz1", "dz0", "d_a0", "d_z1", "d_a1", "d_z2", "d_a2", "d_z3", "d_a3", "d_g3", "d_g2",
                                          "d_g1", "d_g0", "gz_nchw"};
    int b = -1;
    for (int i = 0; i < TB_COUNT; ++i)
        if (!strcmp(name, names[i])) b = i;
    CID_REQUIRE(b >= 0, "unknown training buffer '%s'", name);
    DeviceGuard dg(t->device);
    CID_REQUIRE(dg.ok, "cannot select CUDA device %d", t->device);
    size_t toff[TB_COUNT];
    train_layout(B, H, W, toff);
    const int lv = kTBufLevel[b];
    const size_t es = tbuf_elem(b);
    const size_t n = static_cast<size_t>(B) * (H >> lv) * (W >> lv) * kTBufC[b] * es;
    CID_REQUIRE(dst_bytes >= n, "destination too small: %zu < %zu", dst_bytes, n);
    CID_CHECK_CUDA(cudaDeviceSynchronize());
    CID_CHECK_CUDA(cudaMemcpy(dst_host, static_cast<const uint8_t*>(ws) + toff[b], n, cudaMemcpyDeviceToHost));
    if (channels) *channels = kTBufC[b];
    if (level) *level = lv;
    if (elem_size) *elem_size = static_cast<int>(es);
    return 0;
}

}  // extern "C"
