// common.cuh -- shared host/device plumbing for the sm_100a compositing kernels.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include "../../include/mltcuda.h"

struct mltcuda_ctx
{
    int device;
    cudaStream_t stream;
    cudaMemPool_t pool;
    unsigned long long launches;
    int sm_count;
    uint32_t *noise_tab; // device: skip-ahead tables for the noise producer (lazy)
    // pinned double-buffered staging for uploads from pageable host memory (mltcuda_upload_staged)
    void *stage[2];
    cudaEvent_t stage_done[2];
    int stage_next;
    // opacity query (mltcuda_plane_opaque_query): device flag + its pinned host mirror
    int *opaque_dev, *opaque_host;
};
#define MLTCUDA_STAGE_BYTES (4u << 20)

struct mltcuda_event
{
    cudaEvent_t ev;
};

void mltcuda_set_error(const char *fmt, ...);

#define MC_CHECK(expr) \
    do { \
        cudaError_t _e = (expr); \
        if (_e != cudaSuccess) { \
            mltcuda_set_error("%s:%d %s -> %s", __FILE__, __LINE__, #expr, cudaGetErrorString(_e)); \
            return (int) _e; \
        } \
    } while (0)

#define MC_BADARG(cond, msg) \
    do { \
        if (cond) { \
            mltcuda_set_error("%s: %s", __func__, msg); \
            return MLTCUDA_E_BADARG; \
        } \
    } while (0)

static inline int mc_after_launch(mltcuda_ctx *ctx, const char *what)
{
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        mltcuda_set_error("launch %s -> %s", what, cudaGetErrorString(e));
        return (int) e;
    }
    ctx->launches++;
    return 0;
}

static inline int mc_use(mltcuda_ctx *ctx)
{
    if (!ctx) {
        mltcuda_set_error("null context");
        return MLTCUDA_E_BADARG;
    }
    cudaError_t e = cudaSetDevice(ctx->device);
    if (e != cudaSuccess) {
        mltcuda_set_error("cudaSetDevice(%d) -> %s", ctx->device, cudaGetErrorString(e));
        return (int) e;
    }
    return 0;
}

static inline unsigned mc_blocks(unsigned long long items, unsigned per_block)
{
    return (unsigned) ((items + per_block - 1) / per_block);
}
