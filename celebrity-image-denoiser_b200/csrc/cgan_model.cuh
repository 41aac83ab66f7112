// Keras cGAN generator (the reference's only shipped trained weights) on the same tcgen05 conv kernels.
// Included by cid_api.cu (uses its static helpers).
//
// Reference: architecture backend/trainingcode/cgan_code/training5barrev.py:24-43, served by
// backend/app.py:338-345 (load) and :438-467 (run_keras_cgan).
//
//   G0 conv2d              Conv2D 3->64  k4 s2 'same' + LeakyReLU(.2)        cgan_first_conv_kernel      -> c0 [H/2,W/2,64]
//   G1 conv2d_1            Conv2D 64->128 k4 s2 'same' + BN + LeakyReLU      <128,16,2,ring,LRELU>       -> c1 [H/4,W/4,128]
//   G2 conv2d_transpose    Conv2DTranspose 128->128 k4 s2 'same' + BN + LReLU <128,8,1,ring,T4>          -> c2 [H/2,W/2,128]
//   G3 conv2d_transpose_1  Conv2DTranspose 128->64 k4 s2 'same' + BN + LReLU  <64,8,2,ring,T4>           -> c3 [H,W,64]
//   G4 conv2d_2            Conv2D 64->3 k3 'same' + tanh                      last_conv_umma_kernel      -> y
// BatchNormalization (inference, eps 1e-3) is folded into the preceding layer's weights and bias at
// cid_finalize: w' = w * gamma/sqrt(var+eps), b' = (b - mean) * gamma/sqrt(var+eps) + beta.
#pragma once

namespace cid {

static const char* kCganLayerNames[5] = {"conv2d", "conv2d_1", "conv2d_transpose", "conv2d_transpose_1", "conv2d_2"};
static const char* kCganBnNames[5] = {nullptr, "batch_normalization", "batch_normalization_1", "batch_normalization_2", nullptr};
static const int kCganCin[5] = {3, 64, 128, 128, 64};
static const int kCganCout[5] = {64, 128, 128, 64, 3};
static const int kCganK[5] = {4, 4, 4, 4, 3};
static const float kCganBnEps = 1e-3f;
static const float kCganSlope = 0.2f;

enum CganBuf { CB_C0, CB_C1, CB_C2, CB_C3, CB_COUNT };
static const int kCganBufC[CB_COUNT] = {64, 128, 128, 64};
static const int kCganBufLevel[CB_COUNT] = {1, 2, 1, 0};

static size_t cgan_layout_offsets(int B, int H, int W, size_t* off) {
    size_t cur = 0;
    for (int i = 0; i < CB_COUNT; ++i) {
        if (off) off[i] = cur;
        const int lv = kCganBufLevel[i];
        size_t n = static_cast<size_t>(B) * (H >> lv) * (W >> lv) * kCganBufC[i] * 2;
        cur += (n + 1023) & ~static_cast<size_t>(1023);
    }
    return cur;
}

// expected shape of a Keras variable of this model (Keras' own layouts); returns ndim or 0 if unknown
static int cgan_param_shape(const std::string& key, int64_t* shape) {
    const size_t dot = key.rfind('.');
    if (dot == std::string::npos) return 0;
    const std::string lname = key.substr(0, dot), what = key.substr(dot + 1);
    for (int l = 0; l < 5; ++l) {
        if (lname == kCganLayerNames[l]) {
            if (what == "bias") {
                shape[0] = kCganCout[l];
                return 1;
            }
            if (what != "kernel") return 0;
            const bool transposed = (l == 2 || l == 3);  // Conv2DTranspose kernel: [kh, kw, Cout, Cin]
            shape[0] = shape[1] = kCganK[l];
            shape[2] = transposed ? kCganCout[l] : kCganCin[l];
            shape[3] = transposed ? kCganCin[l] : kCganCout[l];
            return 4;
        }
        if (kCganBnNames[l] && lname == kCganBnNames[l]) {
            if (what != "gamma" && what != "beta" && what != "moving_mean" && what != "moving_variance") return 0;
            shape[0] = kCganCout[l];
            return 1;
        }
    }
    return 0;
}

// NHWC bf16 activation read through a STRIDE-2 box: {64 ch, 9 px (every 2nd), 17 rows (every 2nd), 1 image}
static int make_act_map_stride2(CUtensorMap* tm, const void* base, int B, int H, int W, int C) {
    EncodeTiledFn fn = get_encode_fn();
    CID_REQUIRE(fn, "cuTensorMapEncodeTiled entry point not available");
    cuuint64_t dims[4] = {static_cast<cuuint64_t>(C), static_cast<cuuint64_t>(W), static_cast<cuuint64_t>(H),
                          static_cast<cuuint64_t>(B)};
    cuuint64_t strides[3] = {static_cast<cuuint64_t>(C) * 2, static_cast<cuuint64_t>(W) * C * 2,
                             static_cast<cuuint64_t>(H) * W * C * 2};
    cuuint32_t box[4] = {64, 18, 34, 1};  // bounding box; elements loaded = ceil(box / stride) = 9 x 17
    cuuint32_t estr[4] = {1, 2, 2, 1};
    CUresult r = fn(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(base), dims, strides, box, estr,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    CID_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled(stride-2 activation %dx%dx%dx%d) failed: %d", B, H, W, C,
                static_cast<int>(r));
    return 0;
}

static int cgan_pack(cid_ctx* c) {
    auto get = [&](const std::string& k) -> const std::vector<float>* {
        auto it = c->params.find(k);
        return it == c->params.end() ? nullptr : &it->second;
    };
    for (int l = 0; l < 5; ++l) {
        const std::string ln = kCganLayerNames[l];
        const std::vector<float>* w = get(ln + ".kernel");
        const std::vector<float>* b = get(ln + ".bias");
        CID_REQUIRE(w && b, "missing variable for layer %s", ln.c_str());
        const int ci_n = kCganCin[l], co_n = kCganCout[l], k = kCganK[l];
        std::vector<float> scale(co_n, 1.f), shift(co_n, 0.f);
        if (kCganBnNames[l]) {
            const std::string bn = kCganBnNames[l];
            const std::vector<float>*g = get(bn + ".gamma"), *be = get(bn + ".beta"), *mu = get(bn + ".moving_mean"),
                               *var = get(bn + ".moving_variance");
            CID_REQUIRE(g && be && mu && var, "missing BatchNormalization variable for %s", bn.c_str());
            for (int co = 0; co < co_n; ++co) {
                scale[co] = (*g)[co] / sqrtf((*var)[co] + kCganBnEps);
                shift[co] = (*be)[co] - (*mu)[co] * scale[co];
            }
        }
        std::vector<float> bias((co_n + 31) / 32 * 32, 0.f);
        for (int co = 0; co < co_n; ++co) bias[co] = (*b)[co] * scale[co] + shift[co];
        const bool transposed = (l == 2 || l == 3);
        // folded weight of (ky,kx,ci,co) from Keras' layout
        auto W = [&](int ky, int kx, int ci, int co) {
            const size_t idx = transposed ? ((static_cast<size_t>(ky) * k + kx) * co_n + co) * ci_n + ci
                                          : ((static_cast<size_t>(ky) * k + kx) * ci_n + ci) * co_n + co;
            return (*w)[idx] * scale[co];
        };
        std::vector<uint16_t> packed;
        int rows = 0, box_rows = 0, row_len = ci_n;
        if (l == 0) {
            // [tap][k chunk][cout][8 bf16], K chunk 0 = [w_r w_g w_b | w_r w_g w_b | 0 0] (input split hi+lo along K),
            // K chunk 1 = [wlo_r wlo_g wlo_b | 0 ...] (weights split hi+lo too)
            packed.assign(kG0WBytes / 2, 0);
            for (int tap = 0; tap < 16; ++tap)
                for (int co = 0; co < 64; ++co)
                    for (int ci = 0; ci < 3; ++ci) {
                        const float wf = W(tap >> 2, tap & 3, ci, co);
                        const uint16_t v = f32_to_bf16_rne(wf);
                        packed[((tap * 2 + 0) * 64 + co) * 8 + ci] = v;
                        packed[((tap * 2 + 0) * 64 + co) * 8 + 3 + ci] = v;
                        // K chunk 1 multiplies the same pixel chunk again (LBO = 0): w_lo against x_hi
                        packed[((tap * 2 + 1) * 64 + co) * 8 + ci] = f32_to_bf16_rne(wf - bf16_to_f32(v));
                    }
        } else if (l == 4) {
            rows = box_rows = 128;
            packed.assign(static_cast<size_t>(rows) * ci_n, 0);
            for (int co = 0; co < co_n; ++co)
                for (int ci = 0; ci < ci_n; ++ci)
                    for (int t = 0; t < 9; ++t)
                        packed[static_cast<size_t>(last_conv_row(t, co)) * ci_n + ci] = f32_to_bf16_rne(W(t / 3, t % 3, ci, co));
        } else {
            rows = 16 * co_n;
            box_rows = co_n;
            packed.assign(static_cast<size_t>(rows) * ci_n, 0);
            for (int tap = 0; tap < 16; ++tap)
                for (int co = 0; co < co_n; ++co)
                    for (int ci = 0; ci < ci_n; ++ci)
                        packed[(static_cast<size_t>(tap) * co_n + co) * ci_n + ci] = f32_to_bf16_rne(W(tap >> 2, tap & 3, ci, co));
        }
        if (c->w_dev[l]) cudaFree(c->w_dev[l]);
        if (c->bias_dev[l]) cudaFree(c->bias_dev[l]);
        c->w_dev[l] = nullptr;
        c->bias_dev[l] = nullptr;
        CID_CHECK_CUDA(cudaMalloc(&c->w_dev[l], packed.size() * 2));
        CID_CHECK_CUDA(cudaMemcpy(c->w_dev[l], packed.data(), packed.size() * 2, cudaMemcpyHostToDevice));
        CID_CHECK_CUDA(cudaMalloc(reinterpret_cast<void**>(&c->bias_dev[l]), bias.size() * 4));
        CID_CHECK_CUDA(cudaMemcpy(c->bias_dev[l], bias.data(), bias.size() * 4, cudaMemcpyHostToDevice));
        if (l > 0 && make_weight_map(&c->tmW[l], c->w_dev[l], rows, row_len, box_rows)) return 1;
        // CTA-pair form (each CTA of a pair loads half of every weight tile)
        if (l > 0 && box_rows % 2 == 0 && make_weight_map(&c->tmWh[l], c->w_dev[l], rows, row_len, box_rows / 2)) return 1;
    }
    return 0;
}

static int cgan_get_plan(cid_ctx* c, int Bc, int H, int W, const void* ws, cid_plan** out) {
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
    cgan_layout_offsets(Bc, H, W, pl.off);
    const uint8_t* base = static_cast<const uint8_t*>(ws);
    // outputs: G0 -> c0 (tile store), G1 -> c1 (tile store), G2 -> c2 / G3 -> c3 (2x-upsampling scatter)
    if (make_act_map(&pl.tmO[0], base + pl.off[CB_C0], Bc, H >> 1, W >> 1, 64, 16, 8)) return 1;
    if (make_act_map(&pl.tmO[1], base + pl.off[CB_C1], Bc, H >> 2, W >> 2, 128, 16, 8)) return 1;
    if (make_convt_out_map(&pl.tmO[2], base + pl.off[CB_C2], Bc, H >> 2, W >> 2, 128)) return 1;
    if (make_convt_out_map(&pl.tmO[3], base + pl.off[CB_C3], Bc, H >> 1, W >> 1, 64)) return 1;
    // inputs: G1 reads c0 through the stride-2 box; G2/G3/G4 read 3x3 halo patches
    if (make_act_map_stride2(&pl.tmA[1], base + pl.off[CB_C0], Bc, H >> 1, W >> 1, 64)) return 1;
    if (make_act_map(&pl.tmA[2], base + pl.off[CB_C1], Bc, H >> 2, W >> 2, 128, 18, 10)) return 1;
    if (make_act_map(&pl.tmA[3], base + pl.off[CB_C2], Bc, H >> 1, W >> 1, 128, 18, 10)) return 1;
    if (make_act_map(&pl.tmA[4], base + pl.off[CB_C3], Bc, H, W, 64, 18, 10)) return 1;
    c->plans.push_back(pl);
    *out = &c->plans.back();
    return 0;
}

enum CganIo { CGAN_IO_U8 = 0, CGAN_IO_F32_NHWC = 1 };

static int cgan_run_layers(cid_ctx* c, cid_plan* pl, int nb, const IoSpec& io, cudaStream_t st) {
    const int H = pl->H, W = pl->W;
    uint8_t* ws = static_cast<uint8_t*>(const_cast<void*>(pl->ws));
    auto buf = [&](int b) { return reinterpret_cast<__nv_bfloat16*>(ws + pl->off[b]); };
    if (prof_mark(c, st)) return 1;
    {  // G0
        G0Params gp;
        memset(&gp, 0, sizeof(gp));
        gp.B = nb; gp.H = H; gp.W = W;
        gp.in = io.x; gp.in_h = io.in_h; gp.in_w = io.in_w; gp.pad_l = io.pad_l; gp.pad_t = io.pad_t;
        gp.w_packed = static_cast<const uint4*>(c->w_dev[0]);
        gp.bias = c->bias_dev[0];
        gp.slope = kCganSlope;
        gp.tiles_x = ((W >> 1) + 7) / 8;
        gp.tiles_y = ((H >> 1) + 15) / 16;
        gp.m_tiles = nb * gp.tiles_x * gp.tiles_y;
        if (!(c->attr_done & (1u << 28))) {
            CID_CHECK_CUDA(cudaFuncSetAttribute(cgan_first_conv_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kG0Smem));
            CID_CHECK_CUDA(cudaFuncSetAttribute(cgan_first_conv_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kG0Smem));
            c->attr_done |= (1u << 28);
        }
        const int grid = std::min(gp.m_tiles, c->num_sms * kG0CtasPerSm);
        if (io.u8)
            CID_CHECK_CUDA(launch_ex(c, cgan_first_conv_kernel<true>, grid, kG0Threads, kG0Smem, st, 1, pl->tmO[0], gp));
        else
            CID_CHECK_CUDA(launch_ex(c, cgan_first_conv_kernel<false>, grid, kG0Threads, kG0Smem, st, 1, pl->tmO[0], gp));
        CID_CHECK_CUDA(cudaGetLastError());
        c->launches++;
    }
    if (prof_mark(c, st)) return 1;
    auto base = [&](int l, int lv) {
        ConvParams p;
        memset(&p, 0, sizeof(p));
        p.B = nb;
        p.H = H >> lv;  // the tile grid: output pixels (G1) / input pixels (G2, G3)
        p.W = W >> lv;
        p.kb_count = kCganCin[l] / 64;
        p.bias = c->bias_dev[l];
        p.cout = kCganCout[l];
        p.w_rows_per_tap = kCganCout[l];
        p.slope = kCganSlope;
        return p;
    };
    {  // G1: 4x4 s2 conv 64 -> 128
        ConvParams p = base(1, 2);
        p.n_tiles = 1;
        p.out = buf(CB_C1); p.out_cs = 128; p.out_c0 = 0;
        if (launch_conv<128, 16, 2, false, EPI_LRELU, 9, 1>(c, 20, pl->tmA[1], c->tmW[1], pl->tmO[1], pl->tmO[1], p, st)) return 1;
    }
    if (prof_mark(c, st)) return 1;
    {  // G2: transposed 4x4 s2 128 -> 128 (work item = (pixel tile, output row phase))
        ConvParams p = base(2, 2);
        p.n_tiles = 2;
        p.out = buf(CB_C2); p.out_cs = 128; p.out_c0 = 0;
        if (launch_conv<128, 8, 1, false, EPI_T4, 10, 1>(c, 21, pl->tmA[2], c->tmW[2], pl->tmO[2], pl->tmO[2], p, st)) return 1;
    }
    if (prof_mark(c, st)) return 1;
    {  // G3: transposed 4x4 s2 128 -> 64
        ConvParams p = base(3, 1);
        p.n_tiles = 2;
        p.out = buf(CB_C3); p.out_cs = 64; p.out_c0 = 0;
        // CTA pair (bit-identical to single CTAs; 1.48 -> 1.28 ms at batch 256 @256^2; G2 did not gain from it)
        if (launch_conv<64, 8, 2, false, EPI_T4, 10, 1, 2>(c, 22, pl->tmA[3], c->tmWh[3], pl->tmO[3], pl->tmO[3], p, st)) return 1;
    }
    if (prof_mark(c, st)) return 1;
    {  // G4: 3x3 conv 64 -> 3 + tanh
        LastParams lp;
        memset(&lp, 0, sizeof(lp));
        lp.B = nb; lp.H = H; lp.W = W;
        lp.tiles_x = (W + 7) / 8;
        lp.tiles_y = (H + 15) / 16;
        lp.m_tiles = nb * lp.tiles_x * lp.tiles_y;
        lp.stages = std::min(kLastMaxStages, (c->max_smem - last_conv_smem(0)) / kLastStageBytes);
        CID_REQUIRE(lp.stages >= 2, "last conv: shared memory too small");
        lp.bias = c->bias_dev[4];
        lp.out = io.y;
        lp.crop_l = io.pad_l; lp.crop_t = io.pad_t; lp.out_h = io.in_h; lp.out_w = io.in_w;
        const int smem = last_conv_smem(lp.stages);
        if (!(c->attr_done & (1u << 27))) {
            CID_CHECK_CUDA(cudaFuncSetAttribute(last_conv_umma_kernel<EPI_TANH_U8>, cudaFuncAttributeMaxDynamicSharedMemorySize, c->max_smem));
            CID_CHECK_CUDA(cudaFuncSetAttribute(last_conv_umma_kernel<EPI_TANH_F32_NHWC>, cudaFuncAttributeMaxDynamicSharedMemorySize, c->max_smem));
            c->attr_done |= (1u << 27);
        }
        const int grid = std::min(lp.m_tiles, c->num_sms);
        if (io.u8)
            CID_CHECK_CUDA(launch_ex(c, last_conv_umma_kernel<EPI_TANH_U8>, grid, kLastThreads, smem, st, 1, pl->tmA[4], c->tmW[4], lp));
        else
            CID_CHECK_CUDA(launch_ex(c, last_conv_umma_kernel<EPI_TANH_F32_NHWC>, grid, kLastThreads, smem, st, 1, pl->tmA[4], c->tmW[4], lp));
        CID_CHECK_CUDA(cudaGetLastError());
        c->launches++;
    }
    if (prof_mark(c, st)) return 1;
    return 0;
}

}  // namespace cid
