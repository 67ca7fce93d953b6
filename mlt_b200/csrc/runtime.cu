// runtime.cu -- contexts, stream-ordered device memory, pinned staging, events.
// One mltcuda_ctx per (host thread, GPU): this is what MLT's real_time worker threads
// (reference src/framework/mlt_consumer.c:1079-1170) bind to; frames shard by worker, no
// collective.
#include "common.cuh"
#include <mutex>
#include <vector>
#include <string.h>
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
    ctx->stage[0] = ctx->stage[1] = nullptr, ctx->stage_next = 0;
    ctx->opaque_dev = ctx->opaque_host = nullptr;
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
    if (ctx->opaque_dev)
        cudaFree(ctx->opaque_dev);
    if (ctx->opaque_host)
        cudaFreeHost(ctx->opaque_host);
    for (int i = 0; i < 2; i++)
        if (ctx->stage[i]) {
            cudaFreeHost(ctx->stage[i]);
            cudaEventDestroy(ctx->stage_done[i]);
        }
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

// Recycling pool of pinned host buffers for images handed back to host code (a download straight
// into pinned memory needs no staging copy).  cudaHostAlloc costs milliseconds, frames repeat the
// same few sizes: freed buffers are kept (exact-size match) up to a cap.
namespace {
struct PinnedEntry
{
    void *p;
    size_t bytes;
    bool busy;
};
std::mutex g_pin_lock;
std::vector<PinnedEntry> g_pinned;
size_t g_pinned_idle_bytes = 0;
const size_t PINNED_IDLE_CAP = 4ull << 30;
} // namespace

void *mltcuda_host_pool_alloc(size_t bytes)
{
    if (!bytes)
        bytes = 1;
    {
        std::lock_guard<std::mutex> g(g_pin_lock);
        for (auto &e : g_pinned)
            if (!e.busy && e.bytes == bytes) {
                e.busy = true;
                g_pinned_idle_bytes -= bytes;
                return e.p;
            }
    }
    void *p = mltcuda_host_alloc(bytes);
    if (!p)
        return nullptr;
    std::lock_guard<std::mutex> g(g_pin_lock);
    g_pinned.push_back({p, bytes, true});
    return p;
}

void mltcuda_host_pool_release(void *p)
{
    if (!p)
        return;
    void *to_free = nullptr;
    {
        std::lock_guard<std::mutex> g(g_pin_lock);
        for (size_t i = 0; i < g_pinned.size(); i++)
            if (g_pinned[i].p == p) {
                if (g_pinned_idle_bytes + g_pinned[i].bytes > PINNED_IDLE_CAP) {
                    to_free = p;
                    g_pinned.erase(g_pinned.begin() + i);
                } else {
                    g_pinned[i].busy = false;
                    g_pinned_idle_bytes += g_pinned[i].bytes;
                }
                break;
            }
    }
    if (to_free)
        cudaFreeHost(to_free);
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

// Upload from PAGEABLE host memory (the reference's mlt_pool buffers) at PCIe speed: the source is
// copied by the calling thread into two pinned 4 MiB slots in turn, each followed by an asynchronous
// DMA on the context stream, so the CPU copy of chunk k+1 overlaps the DMA of chunk k.  A plain
// cudaMemcpyAsync from pageable memory goes through the driver's own staging buffer, under a
// process-wide lock: 16 worker threads together reached 3.7 GB/s (tools/melt_bench.py).  The
// source may be released as soon as the call returns.
int mltcuda_upload_staged(mltcuda_ctx *ctx, void *dst, const void *src, size_t bytes)
{
    if (int e = mc_use(ctx))
        return e;
    MC_BADARG(!dst || !src, "null pointer");
    for (int i = 0; i < 2; i++)
        if (!ctx->stage[i]) {
            MC_CHECK(cudaHostAlloc(&ctx->stage[i], MLTCUDA_STAGE_BYTES, cudaHostAllocPortable));
            MC_CHECK(cudaEventCreateWithFlags(&ctx->stage_done[i], cudaEventDisableTiming));
            MC_CHECK(cudaEventRecord(ctx->stage_done[i], ctx->stream));
        }
    const char *s = (const char *) src;
    char *d = (char *) dst;
    while (bytes) {
        const size_t n = bytes < MLTCUDA_STAGE_BYTES ? bytes : (size_t) MLTCUDA_STAGE_BYTES;
        const int i = ctx->stage_next;
        ctx->stage_next ^= 1;
        MC_CHECK(cudaEventSynchronize(ctx->stage_done[i])); // the slot's previous DMA has left
        memcpy(ctx->stage[i], s, n);
        MC_CHECK(cudaMemcpyAsync(d, ctx->stage[i], n, cudaMemcpyHostToDevice, ctx->stream));
        MC_CHECK(cudaEventRecord(ctx->stage_done[i], ctx->stream));
        s += n, d += n, bytes -= n;
    }
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

// is_opaque() of the reference (src/modules/core/transition_luma.c:34-41) as a stream-ordered query:
// a plane whose every byte is 0xff need not cross PCIe, the host can write it itself.
namespace {
__global__ void __launch_bounds__(256) k_plane_not_ff(const uint4 *p, size_t n16, const uint8_t *tail, int ntail, int *flag)
{
    size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x;
    const size_t step = (size_t) gridDim.x * blockDim.x;
    uint32_t acc = 0xffffffffu;
    for (; i < n16; i += step) {
        const uint4 v = __ldg(p + i);
        acc &= v.x & v.y & v.z & v.w;
    }
    if (blockIdx.x == 0 && (int) threadIdx.x < ntail)
        acc &= 0xffffff00u | tail[threadIdx.x];
    if (__any_sync(0xffffffffu, acc != 0xffffffffu) && (threadIdx.x & 31) == 0)
        *flag = 1;
}
} // namespace

int mltcuda_plane_opaque_query(mltcuda_ctx *ctx, const uint8_t *plane, size_t bytes)
{
    if (int e = mc_use(ctx))
        return e;
    MC_BADARG(!plane || ((uintptr_t) plane & 15), "plane must be a 16-byte aligned device pointer");
    if (!ctx->opaque_dev) {
        MC_CHECK(cudaMalloc(&ctx->opaque_dev, sizeof(int)));
        MC_CHECK(cudaHostAlloc(&ctx->opaque_host, sizeof(int), cudaHostAllocPortable));
    }
    MC_CHECK(cudaMemsetAsync(ctx->opaque_dev, 0, sizeof(int), ctx->stream));
    const size_t n16 = bytes / 16;
    k_plane_not_ff<<<ctx->sm_count * 8, 256, 0, ctx->stream>>>((const uint4 *) plane, n16, plane + n16 * 16, (int) (bytes - n16 * 16),
                                                             ctx->opaque_dev);
    if (int e = mc_after_launch(ctx, "k_plane_not_ff"))
        return e;
    MC_CHECK(cudaMemcpyAsync(ctx->opaque_host, ctx->opaque_dev, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
    return 0;
}

int mltcuda_plane_opaque_result(const mltcuda_ctx *ctx)
{
    return ctx && ctx->opaque_host ? *ctx->opaque_host == 0 : 0;
}

// The same question with the verdict in memory the caller owns.  The per-context slot above is shared by every
// query on the context: when two threads download images that live on one context, each can read the other's
// verdict (or have its device flag cleared under it) and an alpha plane with real transparency is then replaced
// by 0xff.  Here the kernel raises the caller's pinned word directly (pinned host memory is device-addressable
// under unified addressing), so a query owns nothing but its own slot: no device flag, no memset, no copy.
int mltcuda_plane_opaque_query_into(mltcuda_ctx *ctx, const uint8_t *plane, size_t bytes, int *translucent_pinned)
{
    if (int e = mc_use(ctx))
        return e;
    MC_BADARG(!plane || ((uintptr_t) plane & 15), "plane must be a 16-byte aligned device pointer");
    MC_BADARG(!translucent_pinned, "null verdict slot");
    int *dev_alias = nullptr;
    cudaError_t e = cudaHostGetDevicePointer((void **) &dev_alias, translucent_pinned, 0);
    if (e != cudaSuccess) {
        cudaGetLastError();
        mltcuda_set_error("mltcuda_plane_opaque_query_into: the verdict slot must be pinned host memory "
                          "(mltcuda_host_alloc / mltcuda_host_pool_alloc / mltcuda_host_register): %s", cudaGetErrorString(e));
        return (int) e;
    }
    *translucent_pinned = 0; // the caller's slot: nothing on the GPU touches it before the kernel below
    const size_t n16 = bytes / 16;
    k_plane_not_ff<<<ctx->sm_count * 8, 256, 0, ctx->stream>>>((const uint4 *) plane, n16, plane + n16 * 16, (int) (bytes - n16 * 16),
                                                             dev_alias);
    return mc_after_launch(ctx, "k_plane_not_ff");
}

// ---- CUDA graphs: a launch-bound sequence (hundreds of per-frame launches on one or several contexts) recorded
// once and replayed with one call, so the host's launch rate is out of the picture.
struct mltcuda_graph
{
    cudaGraph_t graph;
    cudaGraphExec_t exec;
    size_t kernel_nodes;
};

int mltcuda_graph_begin(mltcuda_ctx *ctx)
{
    if (int e = mc_use(ctx))
        return e;
    MC_CHECK(cudaStreamBeginCapture(ctx->stream, cudaStreamCaptureModeThreadLocal));
    return 0;
}

mltcuda_graph *mltcuda_graph_end(mltcuda_ctx *ctx)
{
    if (mc_use(ctx))
        return nullptr;
    mltcuda_graph *g = new mltcuda_graph();
    g->graph = nullptr, g->exec = nullptr, g->kernel_nodes = 0;
    cudaError_t e = cudaStreamEndCapture(ctx->stream, &g->graph);
    if (e == cudaSuccess)
        e = cudaGraphInstantiate(&g->exec, g->graph, 0);
    if (e != cudaSuccess) {
        mltcuda_set_error("graph capture / instantiate -> %s", cudaGetErrorString(e));
        cudaGetLastError();
        if (g->graph)
            cudaGraphDestroy(g->graph);
        delete g;
        return nullptr;
    }
    size_t n = 0;
    if (cudaGraphGetNodes(g->graph, nullptr, &n) == cudaSuccess && n) {
        std::vector<cudaGraphNode_t> nodes(n);
        cudaGraphGetNodes(g->graph, nodes.data(), &n);
        for (size_t i = 0; i < n; i++) {
            cudaGraphNodeType t;
            if (cudaGraphNodeGetType(nodes[i], &t) == cudaSuccess && t == cudaGraphNodeTypeKernel)
                g->kernel_nodes++;
        }
    }
    return g;
}

int mltcuda_graph_launch(mltcuda_ctx *ctx, mltcuda_graph *g)
{
    if (int e = mc_use(ctx))
        return e;
    MC_BADARG(!g || !g->exec, "null graph");
    MC_CHECK(cudaGraphLaunch(g->exec, ctx->stream));
    ctx->launches += g->kernel_nodes;
    return 0;
}

uint64_t mltcuda_graph_kernel_nodes(const mltcuda_graph *g)
{
    return g ? g->kernel_nodes : 0;
}

void mltcuda_graph_destroy(mltcuda_graph *g)
{
    if (!g)
        return;
    if (g->exec)
        cudaGraphExecDestroy(g->exec);
    if (g->graph)
        cudaGraphDestroy(g->graph);
    delete g;
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

int mltcuda_ctx_wait_ctx(mltcuda_ctx *ctx, mltcuda_ctx *other)
{
    MC_BADARG(!ctx || !other, "null context");
    if (ctx == other)
        return 0;
    if (int e = mc_use(other))
        return e;
    cudaEvent_t ev; // an event is cheap to create and may be destroyed while still pending
    MC_CHECK(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
    cudaError_t e = cudaEventRecord(ev, other->stream);
    if (e == cudaSuccess && (e = cudaSetDevice(ctx->device)) == cudaSuccess)
        e = cudaStreamWaitEvent(ctx->stream, ev, 0);
    cudaEventDestroy(ev);
    MC_CHECK(e);
    return 0;
}

int mltcuda_copy_peer(mltcuda_ctx *dst_ctx, void *dst, mltcuda_ctx *src_ctx, const void *src, size_t bytes)
{
    MC_BADARG(!dst_ctx || !src_ctx || !dst || !src, "null argument");
    if (int e = mltcuda_ctx_wait_ctx(dst_ctx, src_ctx))
        return e;
    if (int e = mc_use(dst_ctx))
        return e;
    MC_CHECK(cudaMemcpyPeerAsync(dst, dst_ctx->device, src, src_ctx->device, bytes, dst_ctx->stream));
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
