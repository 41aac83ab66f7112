// Shared host-side declarations of the C-ABI library (context object, error plumbing).
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdarg.h>
#include <stdint.h>
#include <stdio.h>

#include <map>
#include <string>
#include <vector>

#include "../../include/cid_b200.h"

namespace cid {

void set_error(const char* fmt, ...);

#define CID_CHECK_CUDA(expr)                                                                       \
    do {                                                                                           \
        cudaError_t _e = (expr);                                                                   \
        if (_e != cudaSuccess) {                                                                   \
            cid::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
            return 1;                                                                              \
        }                                                                                          \
    } while (0)

#define CID_REQUIRE(cond, ...)        \
    do {                              \
        if (!(cond)) {                \
            cid::set_error(__VA_ARGS__); \
            return 1;                 \
        }                             \
    } while (0)

struct DeviceGuard {
    int prev = -1;
    bool ok = false;
    explicit DeviceGuard(int dev) {
        if (cudaGetDevice(&prev) != cudaSuccess) return;
        if (prev != dev && cudaSetDevice(dev) != cudaSuccess) return;
        ok = true;
    }
    ~DeviceGuard() {
        if (prev >= 0) cudaSetDevice(prev);
    }
};

}  // namespace cid

// One (chunk, H, W, workspace) execution plan: buffer offsets and the TMA descriptors of every
// tensor-core layer's activation operand.
struct cid_plan {
    int Bc = 0, H = 0, W = 0;
    const void* ws = nullptr;
    size_t off[11] = {0};   // byte offsets of a0,cat1,p1,a2,cat2,p2,b0,b1,u2a,u2b,u1a (other models: their own buffers)
    CUtensorMap tmA[12];    // indexed by layer (0 unused); ESRGAN/SRGAN: indexed by BUFFER (3x3 halo-patch map)
    CUtensorMap tmO[12];    // TMA-store maps of each layer's output buffer (box 64ch x 8px x 16 rows); SR: by buffer
    CUtensorMap tmO2[12];   // pooled output (box 64ch x 4px x 8 rows); layers 1 and 3 only; SR: 9x9 window maps by buffer
};

enum { CID_MODEL_DENOISE = 0, CID_MODEL_KERAS_CGAN = 1, CID_MODEL_ESRGAN = 2, CID_MODEL_SRGAN = 3 };
constexpr int kMaxLayers = 24;  // ESRGAN: 18 convolutions

struct cid_ctx {
    int device = 0;
    int model = CID_MODEL_DENOISE;
    int num_sms = 0;
    int max_smem = 0;
    int64_t launches = 0;
    int64_t chunk = 0;
    bool finalized = false;
    std::map<std::string, std::vector<float>> params;  // host fp32 copies, reference layout
    void* w_dev[kMaxLayers] = {nullptr};      // packed bf16 weights per layer (0 unused)
    float* bias_dev[kMaxLayers] = {nullptr};  // fp32 bias per layer
    CUtensorMap tmW[kMaxLayers];
    CUtensorMap tmWh[kMaxLayers];            // same weights, box of N/2 rows (CTA-pair kernels load half a tile each)
    void* first_w_host = nullptr;     // cid::FirstConvW (kernel parameter block)
    float slopes[kMaxLayers] = {0};   // ESRGAN / SRGAN: PReLU parameter of each layer (1 = none)
    std::vector<cid_plan> plans;
    uint32_t attr_done = 0;           // bit per kernel instantiation: max-dynamic-smem attribute set
    std::map<uint64_t, void*> resize_coeffs;  // (in_size,out_size) -> device coefficient block
    // per-layer timing ("profile" option): 13 events bracket the 12 layers of every pass
    int debug_flags = 0;
    int variant = 34 | 4096 | 8192;                // layers run as CTA pairs ("variant" option, see cid_set_option)
    bool first_tc = true;             // layer 0 on tensor cores (false: CUDA-core direct conv)
    bool patch = true;                // 3x3 layers load one halo patch per K block instead of 3 boxes
    bool pdl = true;                  // programmatic dependent launch between layers ("pdl" option)
    bool profile = false;
    std::vector<cudaEvent_t> prof_events;  // 13 per recorded pass
    size_t prof_used = 0;                  // events recorded since the last cid_profile_read
};
