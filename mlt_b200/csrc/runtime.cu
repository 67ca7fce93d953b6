// runtime.cu -- contexts, stream-ordered device memory, pinned staging, events.
// One mltcuda_ctx per (host thread, GPU): this is what MLT's real_time worker threads
// (reference src/framework/mlt_consumer.c:1079-1170) bind to; frames shard by worker, no
// collective.
#include "common.cuh"
#include <stdarg.h>
#include <atomic>

static thread_local char g_err[512] = "";
static std::atomic<int> g_next_device{0};

void mltcuda_set_error(const char *fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

extern "C" {

int mltcuda_abi_version(void)
{
    return MLTCUDA_ABI_VERSION;
}

const char *mltcuda_last_error(void)
{
    return g_err;
}

int mltcuda_device_count(void)
{
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    return n;
}

mltcuda_ctx *mltcuda_ctx_create(int device)
{
    int n = mltcuda_device_count();
    if (n <= 0) {
        mltcuda_set_error("no CUDA device visible (there is no CPU fallback)");
        return nullptr;
    }
    if (device < 0)
        device = g_next_device.fetch_add(1) % n;
    if (device >= n) {
        mltcuda_set_error("device %d out of range (%d visible)", device, n);
        return nullptr;
    }
    if (cudaSetDevice(device) != cudaSuccess) {
        mltcuda_set_error("cudaSetDevice(%d) failed", device);
        return nullptr;
    }
    mltcuda_ctx *ctx = new mltcuda_ctx();
    ctx->device = device;
    ctx->launches = 0;
    ctx->noise_tab = nullptr;
    if (cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking) != cudaSuccess) {
        mltcuda_set_error("cudaStreamCreate failed: %s", cudaGetErrorString(cudaGetLastError()));
        delete ctx;
        return nullptr;
    }
    cudaDeviceGetAttribute(&ctx->sm_count, cudaDevAttrMultiProcessorCount, device);
    // Keep freed blocks cached in the pool: frame buffers recycle every frame.
    cudaDeviceGetDefaultMemPool(&ctx->pool, device);
    unsigned long long keep = ~0ull;
    cudaMemPoolSetAttribute(ctx->pool, cudaMemPoolAttrReleaseThreshold, &keep);
    return ctx;
}

void mltcuda_ctx_destroy(mltcuda_ctx *ctx)
{
    if (!ctx)
        return;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    if (ctx->noise_tab)
        cudaFree(ctx->noise_tab);
    cudaStreamDestroy(ctx->stream);
    delete ctx;
}

int mltcuda_ctx_device(const mltcuda_ctx *ctx)
{
    return ctx ? ctx->device : -1;
}

void *mltcuda_ctx_stream(const mltcuda_ctx *ctx)
{
    return ctx ? (void *) ctx->stream : nullptr;
}

int mltcuda_ctx_sync(mltcuda_ctx *ctx)
{
    if (int e = mc_use(ctx))
        return e;
    MC_CHECK(cudaStreamSynchronize(ctx->stream));
    return 0;
}

uint64_t mltcuda_ctx_launches(const mltcuda_ctx *ctx)
{
    return ctx ? ctx->launches : 0;
}

// 256-byte aligned (cudaMallocAsync guarantees it) + 64 bytes of tail slack for the vector kernels.
void *mltcuda_malloc(mltcuda_ctx *ctx, size_t bytes)
{
    if (mc_use(ctx))
        return nullptr;
    void *p = nullptr;
    size_t padded = ((bytes + 63) & ~(size_t) 63) + 64;
    cudaError_t e = cudaMallocAsync(&p, padded, ctx->stream);
    if (e != cudaSuccess) {
        mltcuda_set_error("cudaMallocAsync(%zu) -> %s", padded, cudaGetErrorString(e));
        return nullptr;
    }
    return p;
}

void mltcuda_free(mltcuda_ctx *ctx, void *p)
{
    if (!p || mc_use(ctx))
        return;
    cudaFreeAsync(p, ctx->stream);
}

void *mltcuda_host_alloc(size_t bytes)
{
    void *p = nullptr;
    cudaError_t e = cudaHostAlloc(&p, bytes ? bytes : 1, cudaHostAllocPortable);
    if (e != cudaSuccess) {
        mltcuda_set_error("cudaHostAlloc(%zu) -> %s", bytes, cudaGetErrorString(e));
        return nullptr;
    }
    return p;
}

void mltcuda_host_free(void *p)
{
    if (p)
        cudaFreeHost(p);
}

int mltcuda_host_register(void *p, size_t bytes)
{
    MC_CHECK(cudaHostRegister(p, bytes, cudaHostRegisterPortable));
    return 0;
}

int mltcuda_host_unregister(void *p)
{
    MC_CHECK(cudaHostUnregister(p));
    return 0;
}

int mltcuda_upload(mltcuda_ctx *ctx, void *dst, const void *src, size_t bytes)
{
    if (int e = mc_use(ctx))
        return e;
    MC_CHECK(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, ctx->stream));
    return 0;
}

int mltcuda_download(mltcuda_ctx *ctx, void *dst, const void *src, size_t bytes)
{
    if (int e = mc_use(ctx))
        return e;
    MC_CHECK(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, ctx->stream));
    return 0;
}

int mltcuda_copy(mltcuda_ctx *ctx, void *dst, const void *src, size_t bytes)
{
    if (int e = mc_use(ctx))
        return e;
    MC_CHECK(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToDevice, ctx->stream));
    return 0;
}

int mltcuda_memset(mltcuda_ctx *ctx, void *dst, int value, size_t bytes)
{
    if (int e = mc_use(ctx))
        return e;
    MC_CHECK(cudaMemsetAsync(dst, value, bytes, ctx->stream));
    return 0;
}

mltcuda_event *mltcuda_event_create(void)
{
    mltcuda_event *ev = new mltcuda_event();
    if (cudaEventCreate(&ev->ev) != cudaSuccess) {
        mltcuda_set_error("cudaEventCreate failed");
        delete ev;
        return nullptr;
    }
    return ev;
}

void mltcuda_event_destroy(mltcuda_event *ev)
{
    if (!ev)
        return;
    cudaEventDestroy(ev->ev);
    delete ev;
}

int mltcuda_event_record(mltcuda_ctx *ctx, mltcuda_event *ev)
{
    if (int e = mc_use(ctx))
        return e;
    MC_CHECK(cudaEventRecord(ev->ev, ctx->stream));
    return 0;
}

int mltcuda_ctx_wait_event(mltcuda_ctx *ctx, mltcuda_event *ev)
{
    if (int e = mc_use(ctx))
        return e;
    MC_BADARG(!ev, "null event");
    MC_CHECK(cudaStreamWaitEvent(ctx->stream, ev->ev, 0));
    return 0;
}

int mltcuda_event_sync(mltcuda_event *ev)
{
    MC_CHECK(cudaEventSynchronize(ev->ev));
    return 0;
}

float mltcuda_event_elapsed_ms(mltcuda_event *a, mltcuda_event *b)
{
    float ms = -1.f;
    if (cudaEventElapsedTime(&ms, a->ev, b->ev) != cudaSuccess) {
        mltcuda_set_error("cudaEventElapsedTime failed");
        return -1.f;
    }
    return ms;
}

} // extern "C"
