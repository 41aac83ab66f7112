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

struct cid_ctx;
namespace cid {
void clear_plans(cid_ctx* c);  // releases the plans' graphs, then the plans
}

// CUDA-graph replay of one pass of a plan (denoise generator, small batches; cid_api.cu run_pass). Handles are kept as
// void* so that this header stays free of graph types for the files that do not use them.
struct cid_graph {
    void* graph = nullptr;   // cudaGraph_t
    void* exec = nullptr;    // cudaGraphExec_t
    void* first = nullptr;   // cudaGraphNode_t of the first / last kernel: they carry the caller's pointers
    void* last = nullptr;
    int launches = 0;        // kernels per replay
    int eager_runs = 0;
    bool io_u8 = false;      // what is baked into those two nodes
    const void* x = nullptr;
    void* y = nullptr;
    int in_h = 0, in_w = 0, pad_l = 0, pad_t = 0;
};

// One (chunk, H, W, workspace) execution plan: buffer offsets and the TMA descriptors of every
// tensor-core layer's activation operand.
struct cid_plan {
    int Bc = 0, H = 0, W = 0;
    const void* ws = nullptr;
    size_t off[11] = {0};   // byte offsets of a0,cat1,p1,a2,cat2,p2,b0,b1,u2a,u2b,u1a (other models: their own buffers)
    CUtensorMap tmA[12];    // indexed by layer (0 unused); ESRGAN/SRGAN: indexed by BUFFER (3x3 halo-patch map)
    CUtensorMap tmO[12];    // TMA-store maps of each layer's output buffer (box 64ch x 8px x 16 rows); SR: by buffer
    CUtensorMap tmO2[12];   // pooled output (box 64ch x 4px x 8 rows); layers 1 and 3 only; SR: 9x9 window maps by buffer
    cid_graph graph[2];     // [fp32 NCHW entry point, u8 NHWC entry point]; released by cid::clear_plans
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
    CUtensorMap tmW64[kMaxLayers];           // same weights, box of 64 rows (N-split small-batch form of the 3x3 layers)
    float slopes[kMaxLayers] = {0};   // ESRGAN / SRGAN: PReLU parameter of each layer (1 = none)
    std::vector<cid_plan> plans;
    uint32_t attr_done = 0;           // bit per kernel instantiation: max-dynamic-smem attribute set
    std::map<uint64_t, void*> resize_coeffs;  // (in_size,out_size) -> device coefficient block
    int regime = 0;                   // "regime" option: 0 per-layer choice, 1 batch-256 configuration, 2 N-split form
    int64_t graph_max_px = 1 << 20;   // "graph_max_px" option: passes of at most this many pixels replay as a CUDA graph
    int ast[16] = {0, 4};             // "ast<l>" options (tuning): cap on the A-ring depth of layer l (0 = none). down1.2 runs 4 patches deep:
                                      // started from an idle GPU (the per-layer profiled pass) it is 0.272 ms instead of 0.310 ms with 6 -
                                      // inside the back-to-back chain of the real step the depth makes no measurable difference
    int wst[16] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0};  // "wst<l>" options: weight-ring depth of the streamed-weight
                                      // layers in their batch-256 form (0 = launch_conv's default), see run_layers
    bool sp_scatter = false;          // "sp_scatter" option: salt & pepper as copy + two scatter launches (the fallback form)
    bool use_graph = true;            // "graph" option: CUDA-graph replay of launch-bound passes
    bool graph_broken = false;        // capture / node update failed once: eager launches from then on
    void* capture_stream = nullptr;   // cudaStream_t the graphs are captured on (nothing ever executes on it)
    int64_t graph_replays = 0;
    // per-layer timing ("profile" option): 13 events bracket the 12 layers of every pass
    bool pdl = true;                  // programmatic dependent launch between layers ("pdl" option)
    bool profile = false;
    std::vector<cudaEvent_t> prof_events;  // 13 per recorded pass
    size_t prof_used = 0;                  // events recorded since the last cid_profile_read
};
