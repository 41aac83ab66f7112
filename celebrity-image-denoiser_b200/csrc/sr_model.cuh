// ESRGAN and SRGAN generators (the other two nn.Modules behind POST /enhance) on the tcgen05 conv kernels.
// Included by cid_api.cu (uses its static helpers).
//
// Reference (backend/app.py):
//   ESRGANGenerator :203-217  initial Conv9(3->64)+PReLU; 8 x [x + BN(Conv3(PReLU(BN(Conv3(x)))))] (ResidualBlock
//                              :189-200); final Conv9(64->3) of (x1 + x9); no activation; [0,1] domain
//   SRGANGenerator  :146-184  initial Conv9+PReLU; 5 x [Conv3 BN PReLU Conv3 BN] (plain Sequential, NO per-block
//                              skip); mid Conv3 + x0; 2 x [Conv3(64->256) PixelShuffle(2) PReLU]; final Conv9; tanh
// BatchNorm2d (eval, eps 1e-5) is folded into the preceding conv at cid_finalize; nn.PReLU() has ONE parameter,
// i.e. a leaky ReLU whose slope is read from the state_dict.
//
// Layer program (every 64->64 conv is conv_umma_kernel<64,9,2,WRES,...,CG=2>, the down1.2 configuration):
//   first   k9_first_conv_kernel                      3 -> 64, PReLU
//   act     <64,9,2,true,EPI_LRELU,10,1,2>            Conv3 + BN + PReLU
//   add     <64,9,2,true,EPI_ADD,10,1,2>              Conv3 (+ BN) + up to two residuals (block skip, long skip)
//   shuffle <256,9,1,false,EPI_CONVT,10,1>            Conv3(64->256) + PixelShuffle(2) + PReLU: the weights are
//                                                     packed quadrant-major (n = (i*2+j)*64 + c  <-  co = c*4 + i*2+j)
//                                                     so the transposed-conv scatter epilogue IS the pixel shuffle
//   last    k9_last_conv_kernel                       64 -> 3 (+ tanh) (+ clamp / x255 truncation / crop)
#pragma once

namespace cid {

enum SrOp { SR_FIRST, SR_ACT, SR_ADD, SR_SHUFFLE, SR_LAST };

struct SrLayer {
    SrOp op;
    std::string conv;   // state_dict prefix of the Conv2d ("residuals.3.block.0")
    std::string bn;     // prefix of the BatchNorm2d folded into it ("" = none)
    std::string prelu;  // key of the PReLU weight ("" = none)
    int in_buf, out_buf;    // activation buffers (-1 = network input / output)
    int res_buf, res2_buf;  // SR_ADD residual buffers (-1 = none)
};

// buffers: 0 = X0 (output of the first layer, kept for the long skip), 1 = T, 2 = A, 3 = B (ping-pong),
// SRGAN only: 4 = U1 [2H,2W], 5 = U2 [4H,4W]
enum { SRB_X0, SRB_T, SRB_A, SRB_B, SRB_U1, SRB_U2, SRB_COUNT };

static std::vector<SrLayer> sr_program(int model) {
    std::vector<SrLayer> L;
    auto nm = [](const char* fmt, int i) {
        char b[64];
        snprintf(b, sizeof(b), fmt, i);
        return std::string(b);
    };
    L.push_back({SR_FIRST, "initial.0", "", "initial.1.weight", -1, SRB_X0, -1, -1});
    int cur = SRB_X0;
    if (model == CID_MODEL_ESRGAN) {
        for (int i = 0; i < 8; ++i) {
            const int nxt = (i & 1) ? SRB_B : SRB_A;
            L.push_back({SR_ACT, nm("residuals.%d.block.0", i), nm("residuals.%d.block.1", i),
                         nm("residuals.%d.block.2.weight", i), cur, SRB_T, -1, -1});
            // x + block(x); the last block also adds x1: final(x1 + x2) (backend/app.py:217)
            L.push_back({SR_ADD, nm("residuals.%d.block.3", i), nm("residuals.%d.block.4", i), "", SRB_T, nxt, cur,
                         i == 7 ? SRB_X0 : -1});
            cur = nxt;
        }
        L.push_back({SR_LAST, "final", "", "", cur, -1, -1, -1});
    } else {
        for (int i = 0; i < 5; ++i) {
            const int nxt = (i & 1) ? SRB_B : SRB_A;
            L.push_back({SR_ACT, nm("res_blocks.%d.0", i), nm("res_blocks.%d.1", i), nm("res_blocks.%d.2.weight", i), cur,
                         SRB_T, -1, -1});
            L.push_back({SR_ADD, nm("res_blocks.%d.3", i), nm("res_blocks.%d.4", i), "", SRB_T, nxt, -1, -1});
            cur = nxt;
        }
        const int m = cur == SRB_A ? SRB_B : SRB_A;
        L.push_back({SR_ADD, "mid", "", "", cur, m, SRB_X0, -1});  // mid(res) + x0 (backend/app.py:181)
        L.push_back({SR_SHUFFLE, "upscale.0", "", "upscale.2.weight", m, SRB_U1, -1, -1});
        L.push_back({SR_SHUFFLE, "upscale.3", "", "upscale.5.weight", SRB_U1, SRB_U2, -1, -1});
        L.push_back({SR_LAST, "final", "", "", SRB_U2, -1, -1, -1});
    }
    return L;
}

static int sr_buf_scale(int buf) { return buf == SRB_U1 ? 2 : buf == SRB_U2 ? 4 : 1; }
static int sr_num_bufs(int model) { return model == CID_MODEL_SRGAN ? SRB_COUNT : SRB_U1; }

static size_t sr_layout_offsets(int model, int B, int H, int W, size_t* off) {
    size_t cur = 0;
    for (int i = 0; i < sr_num_bufs(model); ++i) {
        if (off) off[i] = cur;
        const int sc = sr_buf_scale(i);
        const size_t n = static_cast<size_t>(B) * (H * sc) * (W * sc) * 64 * 2;
        cur += (n + 1023) & ~static_cast<size_t>(1023);
    }
    return cur;
}

// expected shape of a state_dict tensor of this model; returns ndim or 0 if the key is not a float parameter
static int sr_param_shape(int model, const std::string& key, int64_t* shape) {
    const std::vector<SrLayer> prog = sr_program(model);
    for (const SrLayer& l : prog) {
        const int k = (l.op == SR_FIRST || l.op == SR_LAST) ? 9 : 3;
        const int cin = l.op == SR_FIRST ? 3 : 64;
        const int cout = l.op == SR_LAST ? 3 : l.op == SR_SHUFFLE ? 256 : 64;
        if (key == l.conv + ".weight") {
            shape[0] = cout; shape[1] = cin; shape[2] = k; shape[3] = k;
            return 4;
        }
        if (key == l.conv + ".bias") {
            shape[0] = cout;
            return 1;
        }
        if (!l.prelu.empty() && key == l.prelu) {
            shape[0] = 1;
            return 1;
        }
        if (!l.bn.empty() && (key == l.bn + ".weight" || key == l.bn + ".bias" || key == l.bn + ".running_mean" ||
                              key == l.bn + ".running_var")) {
            shape[0] = cout;
            return 1;
        }
    }
    return 0;
}

static int sr_pack(cid_ctx* c) {
    const std::vector<SrLayer> prog = sr_program(c->model);
    CID_REQUIRE(static_cast<int>(prog.size()) <= kMaxLayers, "layer table overflow");
    auto get = [&](const std::string& k) -> const std::vector<float>* {
        auto it = c->params.find(k);
        return it == c->params.end() ? nullptr : &it->second;
    };
    for (size_t li = 0; li < prog.size(); ++li) {
        const SrLayer& l = prog[li];
        const int k = (l.op == SR_FIRST || l.op == SR_LAST) ? 9 : 3;
        const int cin = l.op == SR_FIRST ? 3 : 64;
        const int cout = l.op == SR_LAST ? 3 : l.op == SR_SHUFFLE ? 256 : 64;
        const std::vector<float>*w = get(l.conv + ".weight"), *b = get(l.conv + ".bias");
        CID_REQUIRE(w && b, "missing parameter for layer %s", l.conv.c_str());
        std::vector<float> scale(cout, 1.f), shift(cout, 0.f);
        if (!l.bn.empty()) {
            const std::vector<float>*g = get(l.bn + ".weight"), *be = get(l.bn + ".bias"),
                               *mu = get(l.bn + ".running_mean"), *var = get(l.bn + ".running_var");
            CID_REQUIRE(g && be && mu && var, "missing BatchNorm2d tensor for %s", l.bn.c_str());
            for (int co = 0; co < cout; ++co) {
                scale[co] = (*g)[co] / sqrtf((*var)[co] + 1e-5f);
                shift[co] = (*be)[co] - (*mu)[co] * scale[co];
            }
        }
        c->slopes[li] = 1.f;
        if (!l.prelu.empty()) {
            const std::vector<float>* a = get(l.prelu);
            CID_REQUIRE(a, "missing PReLU parameter %s", l.prelu.c_str());
            c->slopes[li] = (*a)[0];
        }
        // torch Conv2d weight [cout][cin][k][k], BN scale folded in
        auto Wt = [&](int co, int ci, int ky, int kx) {
            return (*w)[((static_cast<size_t>(co) * cin + ci) * k + ky) * k + kx] * scale[co];
        };
        // output column n of the GEMM -> torch output channel (pixel shuffle: quadrant-major)
        auto chan = [&](int n) { return l.op == SR_SHUFFLE ? (n & 63) * 4 + (n >> 6) : n; };
        std::vector<float> bias((std::max(cout, 3) + 31) / 32 * 32, 0.f);
        for (int n = 0; n < cout; ++n) bias[n] = (*b)[chan(n)] * scale[chan(n)] + shift[chan(n)];
        std::vector<uint16_t> packed;
        int rows = 0, box_rows = 0;
        if (l.op == SR_FIRST) {
            // [dy 9][dx pair 5][k chunk 2][cout 64][8 bf16]; chunk e = tap dx = 2j+e: [w_r w_g w_b | w_r w_g w_b | 0 0]
            packed.assign(kK9WBytes / 2, 0);
            for (int dy = 0; dy < 9; ++dy)
                for (int j = 0; j < 5; ++j)
                    for (int e = 0; e < 2; ++e) {
                        const int dx = 2 * j + e;
                        if (dx > 8) continue;
                        for (int co = 0; co < 64; ++co)
                            for (int ci = 0; ci < 3; ++ci) {
                                const uint16_t v = f32_to_bf16_rne(Wt(co, ci, dy, dx));
                                const size_t o = ((static_cast<size_t>(dy * 5 + j) * 2 + e) * 64 + co) * 8;
                                packed[o + ci] = v;
                                packed[o + 3 + ci] = v;
                            }
                    }
        } else if (l.op == SR_LAST) {
            rows = 9 * 32;  // [dy][dx*3 + c (27 of 32 rows)][64 ch]
            box_rows = 32;
            packed.assign(static_cast<size_t>(rows) * 64, 0);
            for (int dy = 0; dy < 9; ++dy)
                for (int dx = 0; dx < 9; ++dx)
                    for (int co = 0; co < 3; ++co)
                        for (int ci = 0; ci < 64; ++ci)
                            packed[(static_cast<size_t>(dy) * 32 + dx * 3 + co) * 64 + ci] = f32_to_bf16_rne(Wt(co, ci, dy, dx));
        } else {
            rows = 9 * cout;
            box_rows = cout;
            packed.assign(static_cast<size_t>(rows) * 64, 0);
            for (int t = 0; t < 9; ++t)
                for (int n = 0; n < cout; ++n)
                    for (int ci = 0; ci < 64; ++ci)
                        packed[(static_cast<size_t>(t) * cout + n) * 64 + ci] = f32_to_bf16_rne(Wt(chan(n), ci, t / 3, t % 3));
        }
        if (c->w_dev[li]) cudaFree(c->w_dev[li]);
        if (c->bias_dev[li]) cudaFree(c->bias_dev[li]);
        c->w_dev[li] = nullptr;
        c->bias_dev[li] = nullptr;
        CID_CHECK_CUDA(cudaMalloc(&c->w_dev[li], packed.size() * 2));
        CID_CHECK_CUDA(cudaMemcpy(c->w_dev[li], packed.data(), packed.size() * 2, cudaMemcpyHostToDevice));
        CID_CHECK_CUDA(cudaMalloc(reinterpret_cast<void**>(&c->bias_dev[li]), bias.size() * 4));
        CID_CHECK_CUDA(cudaMemcpy(c->bias_dev[li], bias.data(), bias.size() * 4, cudaMemcpyHostToDevice));
        if (l.op != SR_FIRST) {
            if (make_weight_map(&c->tmW[li], c->w_dev[li], rows, 64, box_rows)) return 1;
            if ((l.op == SR_ACT || l.op == SR_ADD) && make_weight_map(&c->tmWh[li], c->w_dev[li], rows, 64, box_rows / 2)) return 1;
        }
    }
    return 0;
}

static int sr_get_plan(cid_ctx* c, int Bc, int H, int W, const void* ws, cid_plan** out) {
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
    sr_layout_offsets(c->model, Bc, H, W, pl.off);
    const uint8_t* base = static_cast<const uint8_t*>(ws);
    for (int i = 0; i < sr_num_bufs(c->model); ++i) {
        const int sc = sr_buf_scale(i);
        const void* b = base + pl.off[i];
        if (make_act_map(&pl.tmA[i], b, Bc, H * sc, W * sc, 64, 18, 10)) return 1;   // 3x3 halo patch
        if (make_act_map(&pl.tmO2[i], b, Bc, H * sc, W * sc, 64, 24, 16)) return 1;  // 9x9 window of the last layer
        if (i == SRB_U1 || i == SRB_U2) {
            // written by the pixel-shuffle epilogue: (C, 2W, dy, H, B) view of the 2x larger image
            if (make_convt_out_map(&pl.tmO[i], b, Bc, H * sc / 2, W * sc / 2, 64)) return 1;
        } else {
            if (make_act_map(&pl.tmO[i], b, Bc, H, W, 64, 16, 8)) return 1;
        }
    }
    c->plans.push_back(pl);
    *out = &c->plans.back();
    return 0;
}

static int sr_run_layers(cid_ctx* c, cid_plan* pl, int nb, const IoSpec& io, cudaStream_t st) {
    const std::vector<SrLayer> prog = sr_program(c->model);
    const int H = pl->H, W = pl->W;
    uint8_t* ws = static_cast<uint8_t*>(const_cast<void*>(pl->ws));
    auto buf = [&](int b) { return reinterpret_cast<__nv_bfloat16*>(ws + pl->off[b]); };
    const bool srgan = c->model == CID_MODEL_SRGAN;
    if (prof_mark(c, st)) return 1;
    for (size_t li = 0; li < prog.size(); ++li) {
        const SrLayer& l = prog[li];
        if (l.op == SR_FIRST) {
            K9Params kp;
            memset(&kp, 0, sizeof(kp));
            kp.B = nb; kp.H = H; kp.W = W;
            kp.in = io.x; kp.in_h = io.in_h; kp.in_w = io.in_w; kp.pad_l = io.pad_l; kp.pad_t = io.pad_t;
            kp.norm = srgan ? 1 : 0;
            if (!srgan) kp.pad_l = kp.pad_t = 0;  // the esrgan branch feeds the un-padded image; the offsets only crop
            kp.w_packed = static_cast<const uint4*>(c->w_dev[li]);
            kp.bias = c->bias_dev[li];
            kp.slope = c->slopes[li];
            kp.tiles_x = (W + 7) / 8;
            kp.tiles_y = (H + 15) / 16;
            kp.m_tiles = nb * kp.tiles_x * kp.tiles_y;
            if (!(c->attr_done & (1u << 26))) {
                CID_CHECK_CUDA(cudaFuncSetAttribute(k9_first_conv_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kK9Smem));
                CID_CHECK_CUDA(cudaFuncSetAttribute(k9_first_conv_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kK9Smem));
                c->attr_done |= (1u << 26);
            }
            const int grid = std::min(kp.m_tiles, c->num_sms * kK9CtasPerSm);
            if (io.u8)
                CID_CHECK_CUDA(launch_ex(c, k9_first_conv_kernel<true>, grid, kK9Threads, kK9Smem, st, 1, pl->tmO[l.out_buf], kp));
            else
                CID_CHECK_CUDA(launch_ex(c, k9_first_conv_kernel<false>, grid, kK9Threads, kK9Smem, st, 1, pl->tmO[l.out_buf], kp));
            c->launches++;
        } else if (l.op == SR_LAST) {
            const int sc = sr_buf_scale(l.in_buf);
            L9Params lp;
            memset(&lp, 0, sizeof(lp));
            lp.B = nb; lp.H = H * sc; lp.W = W * sc;
            lp.tiles_x = (lp.W + 7) / 8;
            lp.tiles_y = (lp.H + 15) / 16;
            lp.m_tiles = nb * lp.tiles_x * lp.tiles_y;
            lp.stages = std::min(kL9MaxStages, (c->max_smem - k9_last_smem(0)) / kL9StageBytes);
            CID_REQUIRE(lp.stages >= 2, "9x9 last conv: shared memory too small");
            lp.bias = c->bias_dev[li];
            lp.out = io.y;
            // ESRGAN route: y_pil.crop((l, t, l+w, t+h)) of the UN-padded output (backend/app.py:473-479): a shift with
            // black fill (the caller zero-fills). SRGAN route: the 4x output of the padded frame, not cropped (:480-485).
            lp.crop_l = srgan ? 0 : io.pad_l;
            lp.crop_t = srgan ? 0 : io.pad_t;
            lp.out_h = srgan ? lp.H : io.in_h;
            lp.out_w = srgan ? lp.W : io.in_w;
            const int smem = k9_last_smem(lp.stages);
            if (!(c->attr_done & (1u << 25))) {
                CID_CHECK_CUDA(cudaFuncSetAttribute(k9_last_conv_kernel<L9_NONE_F32>, cudaFuncAttributeMaxDynamicSharedMemorySize, c->max_smem));
                CID_CHECK_CUDA(cudaFuncSetAttribute(k9_last_conv_kernel<L9_CLAMP_U8>, cudaFuncAttributeMaxDynamicSharedMemorySize, c->max_smem));
                CID_CHECK_CUDA(cudaFuncSetAttribute(k9_last_conv_kernel<L9_TANH_F32>, cudaFuncAttributeMaxDynamicSharedMemorySize, c->max_smem));
                CID_CHECK_CUDA(cudaFuncSetAttribute(k9_last_conv_kernel<L9_TANH_U8>, cudaFuncAttributeMaxDynamicSharedMemorySize, c->max_smem));
                c->attr_done |= (1u << 25);
            }
            const int grid = std::min(lp.m_tiles, c->num_sms);
            const CUtensorMap& tmA = pl->tmO2[l.in_buf];
            if (srgan && io.u8)
                CID_CHECK_CUDA(launch_ex(c, k9_last_conv_kernel<L9_TANH_U8>, grid, kL9Threads, smem, st, 1, tmA, c->tmW[li], lp));
            else if (srgan)
                CID_CHECK_CUDA(launch_ex(c, k9_last_conv_kernel<L9_TANH_F32>, grid, kL9Threads, smem, st, 1, tmA, c->tmW[li], lp));
            else if (io.u8)
                CID_CHECK_CUDA(launch_ex(c, k9_last_conv_kernel<L9_CLAMP_U8>, grid, kL9Threads, smem, st, 1, tmA, c->tmW[li], lp));
            else
                CID_CHECK_CUDA(launch_ex(c, k9_last_conv_kernel<L9_NONE_F32>, grid, kL9Threads, smem, st, 1, tmA, c->tmW[li], lp));
            c->launches++;
        } else {
            const int sc = sr_buf_scale(l.in_buf);
            ConvParams p;
            memset(&p, 0, sizeof(p));
            p.B = nb;
            p.H = H * sc;
            p.W = W * sc;
            p.kb_count = 1;
            p.n_tiles = 1;
            p.bias = c->bias_dev[li];
            p.cout = 64;  // SR_SHUFFLE: channels per quadrant
            p.w_rows_per_tap = l.op == SR_SHUFFLE ? 256 : 64;
            p.slope = c->slopes[li];
            p.out_cs = 64;
            p.out_c0 = 0;
            p.out = buf(l.out_buf);
            const CUtensorMap& tmA = pl->tmA[l.in_buf];
            const CUtensorMap& tmO = pl->tmO[l.out_buf];
            if (l.op == SR_ACT) {
                if (launch_conv<64, 9, 2, true, EPI_LRELU, 10, 1, 2>(c, 16, tmA, c->tmWh[li], tmO, tmO, p, st)) return 1;
            } else if (l.op == SR_ADD) {
                p.res = l.res_buf >= 0 ? buf(l.res_buf) : nullptr;
                p.res2 = l.res2_buf >= 0 ? buf(l.res2_buf) : nullptr;
                p.res_cs = p.res2_cs = 64;
                // tmO2 slot = tile map of the first residual tensor (TMA-loaded into the epilogue's staging slab)
                const CUtensorMap& tmR = l.res_buf >= 0 ? pl->tmO[l.res_buf] : tmO;
                if (launch_conv<64, 9, 2, true, EPI_ADD, 10, 1, 2>(c, 17, tmA, c->tmWh[li], tmO, tmR, p, st)) return 1;
            } else {
                p.bias_per_column = 1;
                if (launch_conv<256, 9, 1, false, EPI_CONVT, 10, 1>(c, 18, tmA, c->tmW[li], tmO, tmO, p, st)) return 1;
            }
        }
        if (prof_mark(c, st)) return 1;
    }
    return 0;
}

}  // namespace cid
