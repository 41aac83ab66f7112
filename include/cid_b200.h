/*
 * cid_b200.h — C ABI of the B200-native generator forward for Celebrity-Image-Denoiser.
 *
 * This is the drop-in boundary: plain pointers and sizes, no torch types. Everything the
 * reference's denoise path does between "decoded RGB image" and "denoised RGB image" is behind
 * these entry points. The reference itself has no native interface (it is pure Python); each
 * entry point cites the Python interface it replaces (file:line in flawyer/Celebrity-Image-Denoiser).
 * INTEGRATION.md shows the ctypes stub a maintainer adds to backend/app.py.
 *
 * Conventions
 *   - every function returns 0 on success, non-zero on failure; cid_last_error() gives the reason
 *     (thread-local string). There is NO CPU fallback: without a CUDA device every compute call
 *     fails loudly.
 *   - the library never frees or retains caller memory; it owns only its packed weights.
 *   - a handle is bound to one device and used from one thread at a time; handles on different
 *     devices are independent (one per GPU for batch sharding).
 *   - `stream` is a cudaStream_t passed as void* (0 = legacy default stream). All work is
 *     stream-ordered; nothing synchronises the host unless stated.
 */
#ifndef CID_B200_H
#define CID_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct cid_ctx* cid_handle;

#define CID_MAX_LAYERS 24

/* ABI version of this header (bumped on any signature change). */
int cid_version(void);

/* Reason for the most recent failure on the calling thread ("" if none). */
const char* cid_last_error(void);

/* Replaces `DenoiseGenerator().to(DEVICE)` (backend/app.py:319-320): create an empty generator
 * bound to CUDA device `device`. Weights are undefined until cid_set_param + cid_finalize. */
int cid_create(int device, cid_handle* out);
int cid_destroy(cid_handle h);

/* Same, for a named generator of the reference's model registry (backend/app.py:319-345):
 *   "denoise"     DenoiseGenerator (backend/app.py:39-103) — what cid_create makes.
 *   "keras_cgan"  the Keras cGAN generator the reference loads with
 *                 tf.keras.models.load_model("weights/cgan_epoch_500.keras") (backend/app.py:338-345;
 *                 architecture backend/trainingcode/cgan_code/training5barrev.py:24-43).
 * For "keras_cgan" cid_set_param takes the Keras variables under their Keras names and layouts:
 *   conv2d.kernel [4,4,3,64]  conv2d.bias [64]  conv2d_1.kernel [4,4,64,128]  conv2d_1.bias [128]
 *   batch_normalization{,_1,_2}.{gamma,beta,moving_mean,moving_variance} [128|128|64]
 *   conv2d_transpose.kernel [4,4,128,128] (kh,kw,Cout,Cin)  conv2d_transpose.bias [128]
 *   conv2d_transpose_1.kernel [4,4,64,128]  conv2d_transpose_1.bias [64]
 *   conv2d_2.kernel [3,3,64,3]  conv2d_2.bias [3]
 * and cid_finalize folds the inference BatchNormalization (eps 1e-3) into the weights.
 *   "esrgan"      ESRGANGenerator (backend/app.py:203-217) and
 *   "srgan"       SRGANGenerator (backend/app.py:146-184): cid_set_param takes the torch state_dict float tensors
 *                 under their own keys (initial.0.weight [64,3,9,9], initial.1.weight [1] (PReLU),
 *                 residuals.N.block.{0,3}.{weight,bias}, residuals.N.block.{1,4}.{weight,bias,running_mean,
 *                 running_var}, residuals.N.block.2.weight, final.{weight,bias}; srgan: res_blocks.N.{0..4}.*, mid.*,
 *                 upscale.{0,3}.{weight,bias}, upscale.{2,5}.weight); eval-mode BatchNorm2d (eps 1e-5) is folded at
 *                 cid_finalize. Any H, W >= 1. cid_forward_f32_nchw: esrgan y = [B,3,H,W] (no activation),
 *                 srgan y = [B,3,4H,4W] (tanh). cid_forward_u8_nhwc: esrgan = the esrgan branch of the route
 *                 (ToTensor only, un-padded input, clamp(0,1), truncating x255, then the route's crop of the UN-padded
 *                 result at the padding offsets, black outside: backend/app.py:387-397, 473-479), y = [B,H,W,3];
 *                 srgan = Pad/ToTensor/Normalize -> net -> *0.5+0.5, clamp, x255, NOT cropped (:480-485),
 *                 y = [B, 4(H+pad_t+pad_b), 4(W+pad_l+pad_r), 3]. */
int cid_create_model(int device, const char* model, cid_handle* out);

/* Replaces `model.load_state_dict(new_state, strict=False)` (backend/app.py:272): hand over one
 * state_dict tensor. `key` is one of the 24 reference keys (down1.0.weight ... upconv1.2.bias,
 * backend/app.py:42-78); `data` is fp32, host or device memory, C-contiguous in the reference's
 * own layout (Conv2d [Cout,Cin,3,3]; ConvTranspose2d [Cin,Cout,2,2]); shape is checked.
 * Unknown keys fail (the Python mirror filters them, matching strict=False). */
int cid_set_param(cid_handle h, const char* key, const float* data, const int64_t* shape, int ndim);

/* Pack all 24 tensors into the device layouts the kernels read (bf16, per-tap K-major rows for
 * the tensor-core layers; fp32 constant block for the first layer). Synchronous. Must be called
 * after the last cid_set_param and before any forward; may be called again after new params. */
int cid_finalize(cid_handle h);

/* Tunables. "chunk": images processed per pass through the layers (0 = whole batch); bounds the
 * activation working set (and the workspace). "profile": 1 = record per-layer CUDA events (switches the
 * graph replay off while set). "pdl": programmatic dependent launch between consecutive layers (default 1).
 * denoise generator only - "regime": 0 (default) = per layer, run the batch-256 tile configuration or, when its
 * pixel tiles would leave SMs idle (batch 1, the upload route's shape, backend/app.py:406), the N-split form
 * (64-column blocks of single tiles); 1 / 2 force one form (they are bit-identical). "graph": 1 (default) =
 * passes of at most 2^20 pixels are replayed as ONE CUDA graph launch from the second call of a plan on; the
 * caller's pointers are updated in the graph's first / last kernel node when they change. */
int cid_set_option(cid_handle h, const char* name, int64_t value);

/* Bytes of device scratch a forward of B images of H x W (network-input size; keras_cgan: multiples of 4)
 * needs. The caller allocates it (e.g. a torch uint8 tensor) and passes it to the forward. */
int cid_workspace_bytes(cid_handle h, int B, int H, int W, size_t* bytes);

/* Replaces `y = net(x_pt)` (backend/app.py:433; DenoiseGenerator.forward :80-103).
 * x: device fp32 NCHW [B,3,H,W] in [-1,1], H, W >= 4; y: device fp32 NCHW [B,3,H',W'] = tanh output with
 * H' = 4*floor(H/4), W' = 4*floor(W/4): for multiples of 4 (what get_padding guarantees, backend/app.py:276-281) the
 * input's size; otherwise the decoder is smaller and the skip tensors are cropped to it, as the reference's forward
 * does (:89-99). esrgan / srgan handles: any H, W >= 1, y = [B,3,H,W] / [B,3,4H,4W]. */
int cid_forward_f32_nchw(cid_handle h, const float* x, float* y, int B, int H, int W, void* ws, size_t ws_bytes,
                         void* stream);

/* keras_cgan only. Replaces `y_np = KERAS_CGAN.predict(x_np)` (backend/app.py:458):
 * x: device fp32 NHWC [B,H,W,3] in [-1,1]; y: device fp32 NHWC [B,H,W,3] = tanh output.
 * H, W multiples of 4. */
int cid_forward_f32_nhwc(cid_handle h, const float* x, float* y, int B, int H, int W, void* ws, size_t ws_bytes,
                         void* stream);

/* Replaces the whole denoise branch of the route between PIL decode and PIL encode
 * (backend/app.py:383-385 Pad, :401-406 ToTensor/Normalize, :433 net, :435 denorm+clamp,
 * :472 ToPILImage's truncating x255 cast, :474-480 crop).
 * x: device u8 NHWC [B,H,W,3] (the un-padded image); y: device u8 NHWC [B,H,W,3].
 * (pad_l,pad_t,pad_r,pad_b) is get_padding()'s tuple; H+pad_t+pad_b and W+pad_l+pad_r must be
 * multiples of 4. Workspace is sized for the PADDED dims.
 * On a "keras_cgan" handle this is the keras branch of the same route (run_keras_cgan,
 * backend/app.py:438-467: pad, /255, (x-0.5)/0.5, predict, *0.5+0.5, clip; then :472-480). */
int cid_forward_u8_nhwc(cid_handle h, const uint8_t* x, uint8_t* y, int B, int H, int W, int pad_l, int pad_t,
                        int pad_r, int pad_b, void* ws, size_t ws_bytes, void* stream);

/* Synthetic noiser (test-input builder).
 * Gaussian: replaces add_gaussian_noise (trainingcode/denoise_gan_code/noise_generation.py:6-10):
 *   out = u8(clip(in + noise, 0, 255)) in float64 with a truncating cast. If `noise` (device
 *   float64, n values, exactly what np.random.normal(mean, sigma, shape) returned) is non-NULL it
 *   is used verbatim (bit-exact parity mode); otherwise mean + sigma*z is drawn on the device
 *   from Philox4x32-10 (seed, offset) by Box-Muller (statistical parity).
 * n = number of u8 elements (H*W*C*B). */
int cid_noise_gaussian_u8(cid_handle h, const uint8_t* in, uint8_t* out, size_t n, float mean, float sigma,
                          const double* noise, uint64_t seed, uint64_t offset, void* stream);

/* The generator's other element-wise noise types (noise_generation.py:24-39), float64 arithmetic as numpy:
 *   kind 0 gaussian (p0 = mean, p1 = sigma)  = cid_noise_gaussian_u8
 *   kind 1 speckle  add_speckle_noise (:24-28)  u8(clip(x + x * N(p0, p1), 0, 255))
 *   kind 2 uniform  add_uniform_noise (:35-39)  u8(clip(x + U(p0, p1), 0, 255))
 *   kind 3 poisson  add_poisson_noise (:30-33)  np.random.poisson(x).astype(np.uint8)  (int64 -> u8 wraps mod 256)
 * draws (device float64, n values) = the reference's own random array (normal / uniform / the Poisson counts) for
 * bit-exact parity; NULL = Philox4x32-10 on the device (statistical parity). */
int cid_noise_elementwise_u8(cid_handle h, const uint8_t* in, uint8_t* out, size_t n, int kind, float p0, float p1,
                             const double* draws, uint64_t seed, uint64_t offset, void* stream);

/* Salt & pepper: replaces add_salt_pepper_noise (noise_generation.py:12-22). Per image
 * [H,W,C]: n_salt = int(H*W*C*salt_prob) pixel coordinates set to 255 on all channels, then
 * n_pepper set to 0. If coords (device int32 [B][n_salt+n_pepper][2] = (row,col)) is non-NULL
 * it is used verbatim (parity mode; coords_per_image must equal n_salt+n_pepper, else an error); otherwise
 * coordinates are Philox draws in [0,H-1) x [0,W-1) — the reference's randint(0, dim-1) excludes the last
 * row/col. The probabilities are float64 like the reference's Python floats: the draw counts are the
 * float64 product int(H*W*C*prob). */
int cid_noise_salt_pepper_u8(cid_handle h, const uint8_t* in, uint8_t* out, int B, int H, int W, int C,
                             double salt_prob, double pepper_prob, const int32_t* coords, int64_t coords_per_image,
                             uint64_t seed, uint64_t offset, void* stream);

/* The [0,1]-domain noiser of the ESRGAN data-prep script (backend/trainingcode/esrgan_code/esrgan_addNoise.py), u8 in -> u8
 * out including the script's conversions img = u8 / 255.0 (:66) and (noisy * 255).astype(np.uint8) (:79):
 *   kind 0 add_gaussian (:7-13)  clip(img + N(p0, p1), 0, 1)   reference p0 = 0, p1 = sqrt(0.01) = 0.1 (the sigma
 *                                                              BASELINE.json configs[0] names)
 *   kind 1 add_speckle (:27-30)  clip(img + img * N(p0, p1), 0, 1)   reference N(0, 1)
 *   kind 2 add_uniform (:38-41)  clip(img + U(p0, p1), 0, 1)         reference U(-0.05, 0.05)
 * draws (device float64, n values) = the reference's own random array -> float64 arithmetic, bit-exact; NULL = Philox. */
int cid_noise_unit_u8(cid_handle h, const uint8_t* in, uint8_t* out, size_t n, int kind, float p0, float p1,
                      const double* draws, uint64_t seed, uint64_t offset, void* stream);

/* add_salt_pepper (esrgan_addNoise.py:15-25): per image ceil(amount * H*W*C * 0.5) ELEMENT coordinates (row, col, channel)
 * drawn from the full randint(0, dim) range are set to 1 (255), then as many to 0. coords (device int32
 * [B][2*count][3], salt draws first; coords_per_image must be 2*count) = the reference's own draws; NULL = Philox. */
int cid_noise_salt_pepper_unit_u8(cid_handle h, const uint8_t* in, uint8_t* out, int B, int H, int W, int C, double amount,
                                  const int32_t* coords, int64_t coords_per_image, uint64_t seed, uint64_t offset,
                                  void* stream);

/* Resize + normalise data-prep: replaces Image.resize(size, BICUBIC) (noise_generation.py:61,
 * denoisegan_eval.py:89) = Pillow's two-pass separable fixed-point bicubic (a = -0.5, support
 * scaled by the down-sampling factor), u8 NHWC in -> u8 NHWC out. tmp: device scratch of
 * B*in_h*out_w*C bytes. */
int cid_resize_bicubic_u8(cid_handle h, const uint8_t* in, uint8_t* out, uint8_t* tmp, int B, int in_h, int in_w,
                          int out_h, int out_w, int C, void* stream);

/* No counterpart in the reference: added for callers that cannot allocate
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc). The same two-pass resize as
 * cid_resize_bicubic_u8 (bit-identical output), without a handle and without any allocation, host or device: the
 * coefficient tables are computed on the device, and they and the intermediate image of the horizontal pass live in one
 * static device array inside the library (CID_RESIZE_STATIC_TABLE_BYTES + CID_RESIZE_STATIC_TMP_BYTES, held by every
 * GPU that loads the library until the process exits). Runs on CUDA device `device`, on `stream`. Limits, checked
 * before anything is launched (non-zero return, reason in cid_last_error):
 *   B*in_h*out_w*C                                      <= CID_RESIZE_STATIC_TMP_BYTES
 *   4*(out_w*(2 + ksize_w) + out_h*(2 + ksize_h))       <= CID_RESIZE_STATIC_TABLE_BYTES,
 *                                                          ksize = 2*ceil(2*max(in/out, 1)) + 1 per axis
 * Every call shares the array and nothing in it needs clearing (each call rewrites all it reads), but two calls that
 * overlap on the GPU (different streams, or host threads, of one process on one device) race: keep
 * calls to this entry point on one stream, or order them with events. */
#define CID_RESIZE_STATIC_TABLE_BYTES (256u << 10)
#define CID_RESIZE_STATIC_TMP_BYTES (8u << 20)
int cid_resize_bicubic_u8_static(int device, const uint8_t* in, uint8_t* out, int B, int in_h, int in_w, int out_h,
                                 int out_w, int C, void* stream);

/* Replaces ToTensor + Normalize(mean, std) (backend/app.py:401-406, denoisegan_eval.py:76-79):
 * u8 NHWC [B,H,W,3] -> fp32 NCHW [B,3,H,W], (v/255 - mean)/std, each step rounded in fp32 as
 * torch does. (The u8 forward above fuses this into its first layer; this entry point exists for
 * callers that feed cid_forward_f32_nchw.) */
int cid_normalize_u8_to_f32_nchw(cid_handle h, const uint8_t* in, float* out, int B, int H, int W, float mean,
                                 float stdv, void* stream);

/* Number of kernels this handle has launched so far (bench.py's gpu_launches evidence). */
int64_t cid_kernel_launches(cid_handle h);

/* Number of passes this handle has run as ONE CUDA-graph launch (cid_set_option "graph"; replayed kernels are counted
 * in cid_kernel_launches as well). */
int64_t cid_graph_replays(cid_handle h);

/* Per-layer device timing. After cid_set_option(h, "profile", 1) every pass through the layers
 * records 13 CUDA events on the caller's stream (one before each of the 12 layers and one after
 * the last). cid_profile_read synchronises the device, returns the summed milliseconds per layer
 * (layer_ms[12], order: down1.0 down1.2 down2.0 down2.2 bottleneck.0 bottleneck.2 up2 upconv2.0
 * upconv2.2 up1 upconv1.0 upconv1.2) and the number of passes since the previous read.
 * keras_cgan: 6 events per pass, layer_ms[0..4] = conv2d conv2d_1 conv2d_transpose conv2d_transpose_1
 * conv2d_2; esrgan: 18 layers, srgan: 15, in forward order. layer_ms must hold CID_MAX_LAYERS (24) doubles;
 * unused entries are set to 0. */
int cid_profile_read(cid_handle h, double* layer_ms, int64_t* passes);

/* Test hook: copy one intermediate activation of the most recent forward (same B,H,W,ws) to host
 * as bf16 NHWC. name in {a0,cat1,p1,a2,cat2,p2,b0,b1,u2a,u2b,u1a} (keras_cgan: {c0,c1,c2,c3});
 * *channels gets its C.
 * Synchronises the device. */
int cid_debug_copy_activation(cid_handle h, const char* name, const void* ws, int B, int H, int W, void* dst_host,
                              size_t dst_bytes, int* channels, int* height, int* width);

/* ---------------------------------------------------------------------------------------------
 * Training step of the denoise GAN (BASELINE config #5). Replaces one iteration of
 * DenoiseGANTrainer.train (backend/trainingcode/denoise_gan_code/training.py:403-427): generator
 * (training.py:19-74 = backend/app.py:39-103) and DenoiseDiscriminator (:77-98) forward, BCELoss / MSELoss
 * (:234-235), both backward passes and torch.optim.Adam (:237-238). The step is split into the phases a
 * data-parallel caller needs, so gradients can be all-reduced (NCCL, by the caller) before each Adam update:
 *
 *   cid_train_d_backward -> [all-reduce D grads] -> cid_train_adam(which=1)
 *   cid_train_g_backward -> [all-reduce G grads] -> cid_train_adam(which=0)
 * (or, with cid_train_set_peers, no caller-side collective at all: the all-reduce runs inside cid_train_adam)
 *
 * State: ONE caller-owned fp32 device buffer of cid_train_state_floats() floats, laid out as four sections of
 * P = n_generator + n_discriminator floats - parameters (reference layout of every state_dict tensor), gradients,
 * Adam exp_avg, Adam exp_avg_sq - followed by the BatchNorm running statistics. cid_train_tensor() gives the offset
 * of one named tensor ("generator.down1.0.weight", "discriminator.model.3.running_mean", ...), cid_train_section()
 * the contiguous range of a whole network inside a section (what the caller all-reduces). The caller writes the
 * initial parameters (and zeros elsewhere; running_var = 1) into the buffer, then calls cid_train_sync_weights.
 * Activations and activation gradients are bf16 NHWC in the caller's workspace; accumulation is fp32.
 * ------------------------------------------------------------------------------------------- */
typedef struct cid_train* cid_train_handle;

int cid_train_create(int device, cid_train_handle* out);
int cid_train_destroy(cid_train_handle t);
/* Options of the underlying generator context (cid_set_option), plus "peer_timeout_ms": see cid_train_set_peers. */
int cid_train_set_option(cid_train_handle t, const char* name, int64_t value);
int cid_train_state_floats(cid_train_handle t, size_t* n);
int cid_train_bind_state(cid_train_handle t, float* state_dev, size_t n);
/* section: 0 parameters, 1 gradients, 2 Adam exp_avg, 3 Adam exp_avg_sq. Offsets are in floats from state_dev. */
int cid_train_tensor(cid_train_handle t, const char* key, int section, size_t* offset, size_t* numel);
int cid_train_tensor_count(cid_train_handle t);
const char* cid_train_tensor_name(cid_train_handle t, int index);
/* net: "generator" | "discriminator" */
int cid_train_section(cid_train_handle t, const char* net, int section, size_t* offset, size_t* numel);
/* Re-build the bf16 tensor-core operands from the fp32 parameters (after the caller changed them). */
int cid_train_sync_weights(cid_train_handle t, void* stream);
int cid_train_workspace_bytes(cid_train_handle t, int B, int H, int W, size_t* bytes);
/* noisy, clean: fp32 NCHW [B,3,H,W] in [-1,1] on the device (the DataLoader batch, training.py:406-408).
 * Computes denoised = G(noisy) and the discriminator gradients of d_loss (:410-415). */
int cid_train_d_backward(cid_train_handle t, const float* noisy, const float* clean, int B, int H, int W, void* ws,
                         size_t ws_bytes, void* stream);
/* Continues the same batch: generator gradients of g_loss = MSE + 0.001 * BCE(D(denoised), 1) (:418-424). */
int cid_train_g_backward(cid_train_handle t, const float* noisy, const float* clean, int B, int H, int W, void* ws,
                         size_t ws_bytes, void* stream);
/* optimizer.step() of one network (which: 0 generator, 1 discriminator) on the gradients section as it stands,
 * then re-packs that network's tensor-core operands. Reference values: lr 1e-4, betas (0.9, 0.999), eps 1e-8. */
int cid_train_adam(cid_train_handle t, int which, float lr, float beta1, float beta2, float eps, void* stream);
/* Data-parallel replicas on one NVLink domain (the reference trains on one device, training.py:394-437; this is the
 * exchange step of SURVEY §8(e) for config #5). With peers set, cid_train_adam FUSES the gradient all-reduce into the
 * optimiser step: one kernel reads the gradient section of every rank through peer memory (rank order, so every
 * replica computes the same bits), averages, and applies Adam; two cross-GPU flag barriers inside the kernel order it
 * against the peers' backward passes. No separate collective is then needed (or allowed) between *_backward and adam,
 * and the gradients section keeps the rank-LOCAL gradients. Requirements: the bound state buffer and one int buffer of
 * cid_train_peer_flag_ints() zeroed ints are symmetric allocations mapped into every rank (e.g. torch symmetric memory);
 * state_ptrs[r] / flag_ptrs[r] are rank r's buffers as THIS process addresses them; every rank calls cid_train_adam the
 * same number of times in the same order. world <= 8; world <= 1 switches the fused path off. All
 * cid_train_peer_flag_ints()/16 blocks of the kernel must be co-resident (checked here). A peer that does not arrive
 * within the timeout (cid_train_set_option "peer_timeout_ms", default 600000 = 10 min, the order of NCCL's watchdog) is a
 * fault: the optimiser update is SKIPPED from then on (sticky), the peers time out in turn, and every following
 * cid_train_adam / cid_train_losses call on every rank fails; the caller restores a checkpoint. */
int cid_train_peer_flag_ints(void);
int cid_train_set_peers(cid_train_handle t, int world, int rank, const void* const* state_ptrs, void* const* flag_ptrs);
/* Synchronises; out4 = {d_loss, g_loss, content (MSE), adversarial (BCE)} of the last step. */
int cid_train_losses(cid_train_handle t, double* out4);
int64_t cid_train_kernel_launches(cid_train_handle t);
/* Test hook: copy a named training buffer of the last step to the host. Names: y gz gx (fp32, 3 channels; y is
 * NCHW, gz/gx NHWC), dz10 gcat1 dz8 dz7 gcat2 dz5 dz4 gp2 dz3 dz2 gp1 dz1 dz0 (generator activation gradients,
 * bf16 NHWC), d_a0 d_z1 d_a1 d_z2 d_a2 d_z3 d_a3 d_g3 d_g2 d_g1 d_g0 (discriminator). */
int cid_train_debug_buffer(cid_train_handle t, const char* name, const void* ws, int B, int H, int W, void* dst_host,
                           size_t dst_bytes, int* channels, int* level, int* elem_size);

#ifdef __cplusplus
}
#endif
#endif /* CID_B200_H */
