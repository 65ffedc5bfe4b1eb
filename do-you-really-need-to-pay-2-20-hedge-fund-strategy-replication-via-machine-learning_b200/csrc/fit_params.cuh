// Declarations shared by the fit kernels (fit_kernel.cu: one CTA per task / several tasks per CTA; fit_cluster.cu: a
// cluster of six CTAs per task).
#pragma once
#include "common.cuh"

namespace hfrep {

constexpr int EP = 24;    // padded ETF width
constexpr int XS = 28;    // row stride of Xs and dZy (8 consecutive rows -> 8 distinct 16-byte bank groups)
constexpr int WDS = 28;   // row stride of Wd

struct FitParams {
    const float* X;
    const int* task_L;
    const int* task_x;
    const float* W0;
    const int* perm;
    const unsigned* task_seed;   // per-task shuffle stream (device-side epoch permutations) or NULL
    const float4* sched;
    float* W_out;
    float* loss_hist;
    float* val_hist;
    int* epochs_run;
    int* status;
    float* grad_out;
    long long perm_task_stride;
    int PS;              // ints reserved per team for the device-side shuffle (0 or 2*n_fit rounded up to 4)
    int n_tasks, E, T, n_fit, n_val, batch, BP, max_epochs, patience, early_stop, w_stride, shared_x, dump_step;
    int force_tile;      // testing: batch-tile mode even when the block could be resident (env HFREP_FORCE_TILE)
    int private_x;       // resident mode with a training block PER TEAM (tasks with their own blocks that fit shared memory)
    int debug_skip;      // profiling only (env HFREP_DEBUG_SKIP): bit0 rows, bit1 grads, bit2 update phases skipped
    float alpha, lr, b1, b2, eps;
};

__device__ __forceinline__ float4 ld4(const float* p) { return *reinterpret_cast<const float4*>(p); }
__device__ __forceinline__ void st4(float* p, float4 v) { *reinterpret_cast<float4*>(p) = v; }
__device__ __forceinline__ float2 ld2(const float* p) { return *reinterpret_cast<const float2*>(p); }
__device__ __forceinline__ void st2(float* p, float2 v) { *reinterpret_cast<float2*>(p) = v; }

// fit_cluster.cu: the latency mapping over thread-block clusters (six CTAs per task).  Returns HFREP_OK after launching,
// or a negative error; `fits` tells whether the shape is one the cluster kernel takes at all.
bool fit_cluster_takes(const FitParams& p, int L_max);
int fit_cluster_dispatch(const FitParams& p, int L_max, cudaStream_t stream);

}  // namespace hfrep
