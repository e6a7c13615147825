// Runtime part of the C ABI: contexts, device memory, the host-mirror registry, shape arithmetic.
#include <math.h>

#include "common.cuh"

static thread_local char g_err[1024] = "";
unsigned long long g_dl4j_launches = 0;

void dl4j_set_error(const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

void *dl4j_workspace(dl4jb200_ctx *ctx, int slot, size_t bytes) {
    if (bytes <= ctx->ws_bytes[slot] && ctx->ws[slot]) return ctx->ws[slot];
    if (ctx->ws_frozen) {
        dl4j_set_error("workspace: slot %d would have to grow from %zu to %zu bytes, but the context's scratch is frozen "
                       "(a captured CUDA graph references it): warm up with every shape before capturing, or unfreeze and "
                       "re-capture", slot, ctx->ws_bytes[slot], bytes);
        return nullptr;
    }
    // grow: the old region may still be in use by enqueued kernels -> drain the stream first
    size_t want = bytes + bytes / 4 + (1 << 20);
    if (cudaStreamSynchronize(ctx->stream) != cudaSuccess) { dl4j_set_error("workspace: stream sync failed"); return nullptr; }
    if (ctx->ws[slot]) cudaFree(ctx->ws[slot]);
    ctx->ws[slot] = nullptr;
    ctx->ws_bytes[slot] = 0;
    cudaError_t e = cudaMalloc(&ctx->ws[slot], want);
    if (e != cudaSuccess) {
        dl4j_set_error("workspace: cudaMalloc(%zu) failed: %s", want, cudaGetErrorString(e));
        return nullptr;
    }
    ctx->ws_bytes[slot] = want;
    ctx->ws_generation++;
    return ctx->ws[slot];
}

extern "C" {

const char *dl4jb200_last_error(void) { return g_err; }
int dl4jb200_version(void) { return 100; }
unsigned long long dl4jb200_launch_count(void) { return g_dl4j_launches; }

int dl4jb200_ctx_create(int device, void *stream, dl4jb200_ctx **out) {
    DL4J_REQUIRE(out != nullptr, "ctx_create: out is NULL");
    int n = 0;
    DL4J_CUDA_TRY(cudaGetDeviceCount(&n));
    DL4J_REQUIRE(device >= 0 && device < n, "ctx_create: device %d out of range (%d devices)", device, n);
    DL4J_CUDA_TRY(cudaSetDevice(device));
    cudaDeviceProp prop;
    DL4J_CUDA_TRY(cudaGetDeviceProperties(&prop, device));
    if (prop.major != 10) {
        dl4j_set_error("ctx_create: device %d is sm_%d%d; this library is built for sm_100a (B200) only", device,
                       prop.major, prop.minor);
        return DL4JB200_NOT_SUPPORTED;
    }
    dl4jb200_ctx *c = new dl4jb200_ctx();
    c->device = device;
    c->sm_count = prop.multiProcessorCount;
    c->round_act = 0;
    const char *pt = getenv("DL4JB200_PAIR_TILES");  // default of DL4JB200_OPT_PAIR_TILES for contexts of this process (A/B runs, tests)
    c->pair_tiles = (pt && pt[0] == '1') ? 1 : 0;
    for (int i = 0; i < DL4J_WS_SLOTS; i++) { c->ws[i] = nullptr; c->ws_bytes[i] = 0; }
    if (stream) {
        c->stream = (cudaStream_t)stream;
        c->own_stream = false;
    } else {
        cudaError_t e = cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking);
        if (e != cudaSuccess) { delete c; dl4j_set_error("ctx_create: stream: %s", cudaGetErrorString(e)); return DL4JB200_ERR_CUDA; }
        c->own_stream = true;
    }
    *out = c;
    return DL4JB200_OK;
}

int dl4jb200_ctx_destroy(dl4jb200_ctx *ctx) {
    if (!ctx) return DL4JB200_OK;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    for (auto &kv : ctx->mirrors) cudaFree(kv.second.dev);
    for (auto &kv : ctx->mirror_pool) cudaFree(kv.second);
    conv_filters_free(ctx);
    for (int i = 0; i < DL4J_WS_SLOTS; i++) if (ctx->ws[i]) cudaFree(ctx->ws[i]);
    if (ctx->own_stream) cudaStreamDestroy(ctx->stream);
    delete ctx;
    return DL4JB200_OK;
}

int dl4jb200_ctx_set_option(dl4jb200_ctx *ctx, int option, int value) {
    DL4J_REQUIRE(ctx, "ctx is NULL");
    DL4J_REQUIRE(option == DL4JB200_OPT_TF32_ACTIVATIONS || option == DL4JB200_OPT_PAIR_TILES, "ctx_set_option: unknown option %d", option);
    if (option == DL4JB200_OPT_TF32_ACTIVATIONS) ctx->round_act = value ? 1 : 0;
    else ctx->pair_tiles = value ? 1 : 0;
    return DL4JB200_OK;
}

int dl4jb200_ctx_set_stream(dl4jb200_ctx *ctx, void *stream) {
    DL4J_REQUIRE(ctx, "ctx is NULL");
    if (ctx->own_stream) { cudaStreamSynchronize(ctx->stream); cudaStreamDestroy(ctx->stream); ctx->own_stream = false; }
    ctx->stream = (cudaStream_t)stream;
    return DL4JB200_OK;
}

void *dl4jb200_ctx_stream(dl4jb200_ctx *ctx) { return ctx ? (void *)ctx->stream : nullptr; }

int dl4jb200_ctx_sync(dl4jb200_ctx *ctx) {
    DL4J_REQUIRE(ctx, "ctx is NULL");
    DL4J_CUDA_TRY(cudaStreamSynchronize(ctx->stream));
    return DL4JB200_OK;
}

size_t dl4jb200_ctx_workspace_bytes(dl4jb200_ctx *ctx) {
    size_t s = 0;
    if (ctx) for (int i = 0; i < DL4J_WS_SLOTS; i++) s += ctx->ws_bytes[i];
    return s;
}

int dl4jb200_ctx_sm_count(dl4jb200_ctx *ctx) { return ctx ? ctx->sm_count : 0; }

int dl4jb200_ctx_workspace_freeze(dl4jb200_ctx *ctx, int frozen) {
    DL4J_REQUIRE(ctx, "ctx is NULL");
    ctx->ws_frozen = frozen != 0;
    return DL4JB200_OK;
}

unsigned long long dl4jb200_ctx_workspace_generation(dl4jb200_ctx *ctx) { return ctx ? ctx->ws_generation : 0; }

int dl4jb200_malloc(dl4jb200_ctx *ctx, size_t bytes, void **out) {
    DL4J_REQUIRE(ctx && out, "malloc: NULL argument");
    DL4J_CUDA_TRY(cudaSetDevice(ctx->device));
    DL4J_CUDA_TRY(cudaMalloc(out, bytes ? bytes : 16));
    return DL4JB200_OK;
}

int dl4jb200_free(dl4jb200_ctx *ctx, void *ptr) {
    DL4J_REQUIRE(ctx, "ctx is NULL");
    if (ptr) {
        DL4J_CUDA_TRY(cudaStreamSynchronize(ctx->stream));
        DL4J_CUDA_TRY(cudaFree(ptr));
    }
    return DL4JB200_OK;
}

int dl4jb200_memset(dl4jb200_ctx *ctx, void *dst, int value, size_t bytes) {
    DL4J_REQUIRE(ctx && dst, "memset: NULL argument");
    DL4J_CUDA_TRY(cudaMemsetAsync(dst, value, bytes, ctx->stream));
    return DL4JB200_OK;
}

int dl4jb200_memcpy_h2d(dl4jb200_ctx *ctx, void *dst, const void *src_host, size_t bytes) {
    DL4J_REQUIRE(ctx && dst && src_host, "memcpy_h2d: NULL argument");
    DL4J_CUDA_TRY(cudaMemcpyAsync(dst, src_host, bytes, cudaMemcpyHostToDevice, ctx->stream));
    ctx->h2d_bytes += bytes;
    return DL4JB200_OK;
}

int dl4jb200_memcpy_d2h(dl4jb200_ctx *ctx, void *dst_host, const void *src, size_t bytes) {
    DL4J_REQUIRE(ctx && dst_host && src, "memcpy_d2h: NULL argument");
    DL4J_CUDA_TRY(cudaMemcpyAsync(dst_host, src, bytes, cudaMemcpyDeviceToHost, ctx->stream));
    DL4J_CUDA_TRY(cudaStreamSynchronize(ctx->stream));  // the host may read dst as soon as we return
    ctx->d2h_bytes += bytes;
    return DL4JB200_OK;
}

int dl4jb200_memcpy_d2d(dl4jb200_ctx *ctx, void *dst, const void *src, size_t bytes) {
    DL4J_REQUIRE(ctx && dst && src, "memcpy_d2d: NULL argument");
    DL4J_CUDA_TRY(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToDevice, ctx->stream));
    return DL4JB200_OK;
}

// ---- host-mirror registry. An INDArray under nd4j-native lives in host memory; its device copy is owned here, keyed by
// the host address, with a state that says which side is current. Helper calls acquire their operands: an input is uploaded
// only if the host side is newer, an output just becomes "device newer"; the host gets the bytes when (and only when) it
// asks (sync_to_host). Chained helper calls therefore move nothing across the host link.
static int mirror_get(dl4jb200_ctx *ctx, const void *host_ptr, size_t bytes, MirrorBuf **out) {
    auto it = ctx->mirrors.find(host_ptr);
    if (it != ctx->mirrors.end() && it->second.bytes >= bytes) { *out = &it->second; return DL4JB200_OK; }
    if (it != ctx->mirrors.end()) {  // grown: the old copy goes back to the pool (same stream: ordering keeps it safe)
        ctx->mirror_pool.emplace(it->second.bytes, it->second.dev);
        ctx->mirrors.erase(it);
    }
    void *d = nullptr;
    const size_t want = bytes ? bytes : 16;
    auto p = ctx->mirror_pool.find(want);
    if (p != ctx->mirror_pool.end()) { d = p->second; ctx->mirror_pool.erase(p); }
    else DL4J_CUDA_TRY(cudaMalloc(&d, want));
    *out = &(ctx->mirrors[host_ptr] = MirrorBuf{d, want, DL4JB200_BUF_HOST_NEWER});
    return DL4JB200_OK;
}

int dl4jb200_buf_register(dl4jb200_ctx *ctx, const void *host_ptr, size_t bytes, void **device_out) {
    DL4J_REQUIRE(ctx && host_ptr && device_out, "buf_register: NULL argument");
    MirrorBuf *m;
    int rc = mirror_get(ctx, host_ptr, bytes, &m);
    if (rc) return rc;
    *device_out = m->dev;
    return DL4JB200_OK;
}

int dl4jb200_buf_acquire(dl4jb200_ctx *ctx, const void *host_ptr, size_t bytes, int access, void **device_out) {
    DL4J_REQUIRE(ctx && host_ptr && device_out, "buf_acquire: NULL argument");
    DL4J_REQUIRE(access >= 1 && access <= 3, "buf_acquire: access must be READ (1), WRITE (2) or both (3)");
    MirrorBuf *m;
    int rc = mirror_get(ctx, host_ptr, bytes, &m);
    if (rc) return rc;
    if ((access & DL4JB200_BUF_READ) && m->state == DL4JB200_BUF_HOST_NEWER) {
        DL4J_CUDA_TRY(cudaMemcpyAsync(m->dev, host_ptr, bytes, cudaMemcpyHostToDevice, ctx->stream));
        ctx->h2d_bytes += bytes;
        m->state = DL4JB200_BUF_SYNCED;
    }
    if (access & DL4JB200_BUF_WRITE) m->state = DL4JB200_BUF_DEVICE_NEWER;
    *device_out = m->dev;
    return DL4JB200_OK;
}

int dl4jb200_buf_mark_host_dirty(dl4jb200_ctx *ctx, const void *host_ptr) {
    DL4J_REQUIRE(ctx && host_ptr, "buf_mark_host_dirty: NULL argument");
    auto it = ctx->mirrors.find(host_ptr);
    if (it != ctx->mirrors.end()) it->second.state = DL4JB200_BUF_HOST_NEWER;
    return DL4JB200_OK;
}

int dl4jb200_buf_sync_to_host(dl4jb200_ctx *ctx, void *host_ptr) {
    DL4J_REQUIRE(ctx && host_ptr, "buf_sync_to_host: NULL argument");
    auto it = ctx->mirrors.find(host_ptr);
    if (it == ctx->mirrors.end()) { dl4j_set_error("buf_sync_to_host: %p is not registered", host_ptr); return DL4JB200_ERR_INVALID_ARG; }
    if (it->second.state != DL4JB200_BUF_DEVICE_NEWER) return DL4JB200_OK;  // the host copy is current: nothing moves
    DL4J_CUDA_TRY(cudaMemcpyAsync(host_ptr, it->second.dev, it->second.bytes, cudaMemcpyDeviceToHost, ctx->stream));
    DL4J_CUDA_TRY(cudaStreamSynchronize(ctx->stream));
    ctx->d2h_bytes += it->second.bytes;
    it->second.state = DL4JB200_BUF_SYNCED;
    return DL4JB200_OK;
}

int dl4jb200_buf_state(dl4jb200_ctx *ctx, const void *host_ptr, int *state_out) {
    DL4J_REQUIRE(ctx && host_ptr && state_out, "buf_state: NULL argument");
    auto it = ctx->mirrors.find(host_ptr);
    *state_out = it == ctx->mirrors.end() ? -1 : it->second.state;
    return DL4JB200_OK;
}

int dl4jb200_ctx_transfer_bytes(dl4jb200_ctx *ctx, unsigned long long *h2d_out, unsigned long long *d2h_out) {
    DL4J_REQUIRE(ctx, "ctx is NULL");
    if (h2d_out) *h2d_out = ctx->h2d_bytes;
    if (d2h_out) *d2h_out = ctx->d2h_bytes;
    return DL4JB200_OK;
}

int dl4jb200_buf_lookup(dl4jb200_ctx *ctx, const void *host_ptr, void **device_out, size_t *bytes_out) {
    DL4J_REQUIRE(ctx && host_ptr, "buf_lookup: NULL argument");
    auto it = ctx->mirrors.find(host_ptr);
    if (it == ctx->mirrors.end()) { dl4j_set_error("buf_lookup: %p is not registered", host_ptr); return DL4JB200_ERR_INVALID_ARG; }
    if (device_out) *device_out = it->second.dev;
    if (bytes_out) *bytes_out = it->second.bytes;
    return DL4JB200_OK;
}

int dl4jb200_buf_upload(dl4jb200_ctx *ctx, const void *host_ptr) {
    DL4J_REQUIRE(ctx && host_ptr, "buf_upload: NULL argument");
    auto it = ctx->mirrors.find(host_ptr);
    if (it == ctx->mirrors.end()) { dl4j_set_error("buf_upload: %p is not registered", host_ptr); return DL4JB200_ERR_INVALID_ARG; }
    DL4J_CUDA_TRY(cudaMemcpyAsync(it->second.dev, host_ptr, it->second.bytes, cudaMemcpyHostToDevice, ctx->stream));
    ctx->h2d_bytes += it->second.bytes;
    it->second.state = DL4JB200_BUF_SYNCED;
    return DL4JB200_OK;
}

int dl4jb200_buf_download(dl4jb200_ctx *ctx, void *host_ptr) {
    DL4J_REQUIRE(ctx && host_ptr, "buf_download: NULL argument");
    auto it = ctx->mirrors.find(host_ptr);
    if (it == ctx->mirrors.end()) { dl4j_set_error("buf_download: %p is not registered", host_ptr); return DL4JB200_ERR_INVALID_ARG; }
    DL4J_CUDA_TRY(cudaMemcpyAsync(host_ptr, it->second.dev, it->second.bytes, cudaMemcpyDeviceToHost, ctx->stream));
    DL4J_CUDA_TRY(cudaStreamSynchronize(ctx->stream));
    ctx->d2h_bytes += it->second.bytes;
    it->second.state = DL4JB200_BUF_SYNCED;
    return DL4JB200_OK;
}

int dl4jb200_buf_release(dl4jb200_ctx *ctx, const void *host_ptr) {
    DL4J_REQUIRE(ctx && host_ptr, "buf_release: NULL argument");
    auto it = ctx->mirrors.find(host_ptr);
    if (it == ctx->mirrors.end()) return DL4JB200_OK;
    // back to the pool, not to the driver: every use of the buffer is ordered on the context's stream, and a steady-state
    // loop (same shapes every iteration) then allocates nothing
    ctx->mirror_pool.emplace(it->second.bytes, it->second.dev);
    ctx->mirrors.erase(it);
    return DL4JB200_OK;
}

// NN/util/ConvolutionUtils.java:96-131, 133-211, 213-221, 243-266 (Java int division truncates, as C++ does)
int dl4jb200_out_size(int in, int kernel, int stride, int pad, int dilation, int mode, int *out, int *pad_lo,
                      int *pad_hi) {
    DL4J_REQUIRE(out && pad_lo && pad_hi, "out_size: NULL output");
    DL4J_REQUIRE(stride > 0 && dilation > 0 && in > 0, "out_size: stride/dilation/input must be positive");
    DL4J_REQUIRE(mode >= 0 && mode <= 2, "out_size: unknown ConvolutionMode %d", mode);
    int ke = dilation == 1 ? kernel : kernel + (kernel - 1) * (dilation - 1);
    if (mode != DL4JB200_MODE_SAME && (ke <= 0 || ke > in + 2 * pad)) {
        dl4j_set_error("Invalid input data or configuration: (effective) kernel = %d, input = %d and padding = %d "
                       "do not satisfy 0 < %d <= %d", ke, in, pad, ke, in + 2 * pad);
        return DL4JB200_ERR_INVALID_INPUT;
    }
    if (mode == DL4JB200_MODE_STRICT && (in - ke + 2 * pad) % stride != 0) {
        dl4j_set_error("Invalid input data or configuration: ConvolutionMode.Strict requires (input - kernelSize + "
                       "2*padding)/stride + 1 to be an integer. Got: (%d - %d + 2*%d)/%d + 1 = %.2f", in, ke, pad,
                       stride, (in - ke + 2 * pad) / (double)stride + 1.0);
        return DL4JB200_ERR_INVALID_CONFIG;
    }
    if (mode == DL4JB200_MODE_SAME) {
        int o = (int)ceil(in / (double)stride);
        *out = o;
        *pad_lo = ((o - 1) * stride + ke - in) / 2;
        *pad_hi = ((o - 1) * stride + ke - in + 1) / 2;
    } else {
        *out = (in - ke + 2 * pad) / stride + 1;
        *pad_lo = pad;
        *pad_hi = pad;
    }
    return DL4JB200_OK;
}

}  // extern "C"
