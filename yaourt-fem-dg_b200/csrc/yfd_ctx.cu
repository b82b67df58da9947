/*
 * yfd_ctx.cu -- context lifetime, status strings, error plumbing (include/yfd.h).
 * No reference counterpart: the reference is a single-threaded host program.
 */
#include "yfd_internal.hpp"
#include <cstring>

namespace yfd {

yfd_status fail(yfd_ctx *ctx, yfd_status s, const char *fmt, ...)
{
    char buf[1024];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    if (ctx) ctx->err = buf;
    return s;
}

void *pool_alloc(yfd_ctx *ctx, size_t bytes)
{
    /* best fit, but never hand out a block more than 25 % larger than asked (keeps big blocks for big requests) */
    auto it = ctx->pool.lower_bound(bytes);
    if (it != ctx->pool.end() && it->first <= bytes + bytes / 4) {
        void *p = it->second;
        ctx->pool_bytes -= it->first;
        ctx->pool.erase(it);
        return p;
    }
    void *p = nullptr;
    if (cudaMalloc(&p, bytes) == cudaSuccess) return p;
    cudaGetLastError();
    pool_trim(ctx);                      /* out of memory: give the cache back and retry once */
    if (cudaMalloc(&p, bytes) == cudaSuccess) return p;
    cudaGetLastError();
    return nullptr;
}

void pool_free(yfd_ctx *ctx, void *p, size_t bytes)
{
    if (!p) return;
    /* callers that lie about the size cannot exist: dbuf remembers what pool_alloc gave it */
    ctx->pool.emplace(bytes, p);
    ctx->pool_bytes += bytes;
}

void pool_trim(yfd_ctx *ctx)
{
    cudaStreamSynchronize(ctx->stream);
    for (auto &kv : ctx->pool) cudaFree(kv.second);
    ctx->pool.clear();
    ctx->pool_bytes = 0;
}

} // namespace yfd

namespace yfd {
/* FP64 FMA peak of this device ("measure a DFMA microbenchmark before quoting a compute fraction"):
 * 16 independent multiply-add chains per thread, 8 CTAs of 256 threads per SM, nothing but DFMA in the loop */
__global__ void __launch_bounds__(256) dfma_peak(int iters, double m, double *__restrict__ sink)
{
    double a[16];
#pragma unroll
    for (int k = 0; k < 16; k++) a[k] = 1.0 + 1e-9 * (threadIdx.x + k);
    const double c = 1e-12;
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int k = 0; k < 16; k++) a[k] = fma(a[k], m, c);
    }
    double s = 0.0;
#pragma unroll
    for (int k = 0; k < 16; k++) s += a[k];
    if (s == 12345.678) sink[0] = s;          /* never true: keeps the chains alive */
}

/* FP64 TENSOR-core peak (north_star: "FP64 tensor cores are used only if ncu shows a win over FP64 FMA"): 8 independent
 * mma.sync m8n8k4 f64 accumulator chains per warp (SASS: DMMA.8x8x4), 256 multiply-adds per instruction per warp */
__global__ void __launch_bounds__(256) dmma_peak(int iters, double m, double *__restrict__ sink)
{
    double c[8][2];
#pragma unroll
    for (int k = 0; k < 8; k++) { c[k][0] = 1e-9 * (threadIdx.x + k); c[k][1] = 2e-9 * (threadIdx.x + k); }
    const double a = m, b = 1.0 / 4.0;
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int k = 0; k < 8; k++)
            asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                         : "+d"(c[k][0]), "+d"(c[k][1]) : "d"(a), "d"(b));
    }
    double s = 0.0;
#pragma unroll
    for (int k = 0; k < 8; k++) s += c[k][0] + c[k][1];
    if (s == 12345.678) sink[0] = s;
}
}

extern "C" {

const char *yfd_status_string(yfd_status s)
{
    switch (s) {
    case YFD_OK:                   return "ok";
    case YFD_ERR_INVALID_ARGUMENT: return "invalid argument";
    case YFD_ERR_INVALID_STATE:    return "Assembler in invalid state";
    case YFD_ERR_CELL_NOT_FOUND:   return "Mesh cell not found";
    case YFD_ERR_FACE_NOT_FOUND:   return "Mesh face not found";
    case YFD_ERR_DEGREE_TOO_HIGH:  return "Dunavant quadrature: degree too high";
    case YFD_ERR_NO_NEIGHBOURS:    return "No neighbour information.";
    case YFD_ERR_TOO_MANY_OWNERS:  return "BUG: a face has max 2 owners";
    case YFD_ERR_SHAPE_MISMATCH:   return "shape mismatch";
    case YFD_ERR_UNSUPPORTED:      return "unsupported";
    case YFD_ERR_NO_DEVICE:        return "no CUDA device (this library has no CPU path)";
    case YFD_ERR_CUDA:             return "CUDA error";
    case YFD_ERR_COMM:             return "NCCL error";
    case YFD_ERR_OUT_OF_MEMORY:    return "out of device memory";
    }
    return "unknown status";
}

yfd_status yfd_ctx_create(int device, void *stream, yfd_ctx **out)
{
    if (!out) return YFD_ERR_INVALID_ARGUMENT;
    *out = nullptr;
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev <= 0) {
        cudaGetLastError();
        return YFD_ERR_NO_DEVICE;
    }
    if (device < 0 || device >= ndev) return YFD_ERR_INVALID_ARGUMENT;
    if (cudaSetDevice(device) != cudaSuccess) return YFD_ERR_NO_DEVICE;

    yfd_ctx *ctx = new yfd_ctx();
    ctx->device = device;
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) { delete ctx; return YFD_ERR_NO_DEVICE; }
    ctx->sm_count = prop.multiProcessorCount;
    ctx->cc_major = prop.major;
    ctx->cc_minor = prop.minor;
    ctx->smem_optin = prop.sharedMemPerBlockOptin;
    ctx->smem_per_sm = prop.sharedMemPerMultiprocessor;
    ctx->name = prop.name;
    if (stream) {
        ctx->stream = (cudaStream_t) stream;
        ctx->own_stream = false;
    } else {
        if (cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking) != cudaSuccess) { delete ctx; return YFD_ERR_CUDA; }
        ctx->own_stream = true;
    }
    *out = ctx;
    return YFD_OK;
}

void yfd_comm_destroy_internal(yfd_ctx *ctx);   /* yfd_comm.cpp */

void yfd_ctx_destroy(yfd_ctx *ctx)
{
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    yfd_comm_destroy_internal(ctx);
    yfd::pool_trim(ctx);
    if (ctx->tab_gauss_x) cudaFree(ctx->tab_gauss_x);
    if (ctx->tab_gauss_w) cudaFree(ctx->tab_gauss_w);
    if (ctx->tab_dunavant) cudaFree(ctx->tab_dunavant);
    if (ctx->own_stream) cudaStreamDestroy(ctx->stream);
    delete ctx;
}

const char *yfd_last_error(const yfd_ctx *ctx) { return ctx ? ctx->err.c_str() : "null context"; }

yfd_status yfd_ctx_synchronize(yfd_ctx *ctx)
{
    if (!ctx) return YFD_ERR_INVALID_ARGUMENT;
    YFD_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return YFD_OK;
}

yfd_status yfd_ctx_trim(yfd_ctx *ctx, size_t *released_bytes)
{
    if (!ctx) return YFD_ERR_INVALID_ARGUMENT;
    yfd::scoped_device dev(ctx->device);
    if (released_bytes) *released_bytes = ctx->pool_bytes;
    yfd::pool_trim(ctx);
    return YFD_OK;
}

int64_t yfd_ctx_launch_count(const yfd_ctx *ctx) { return ctx ? ctx->launches : 0; }

yfd_status yfd_ctx_device_info(const yfd_ctx *ctx, char *name, size_t cap, int *sm_count, int *cc_major, int *cc_minor)
{
    if (!ctx) return YFD_ERR_INVALID_ARGUMENT;
    if (name && cap) { std::strncpy(name, ctx->name.c_str(), cap - 1); name[cap - 1] = 0; }
    if (sm_count) *sm_count = ctx->sm_count;
    if (cc_major) *cc_major = ctx->cc_major;
    if (cc_minor) *cc_minor = ctx->cc_minor;
    return YFD_OK;
}

yfd_status yfd_ctx_measure_fp64(yfd_ctx *ctx, double *tflops)
{
    using namespace yfd;
    if (!ctx || !tflops) return YFD_ERR_INVALID_ARGUMENT;
    scoped_device dev(ctx->device);
    dbuf<double> sink;
    YFD_TRY(sink.alloc(ctx, 1));
    const int iters = 4096, grid = ctx->sm_count * 8;
    YFD_LAUNCH(ctx, yfd::dfma_peak, grid, 256, 0, iters, 1.000000001, sink.p);        /* warm-up */
    gpu_timer t;
    double best = 0.0;
    for (int rep = 0; rep < 5; rep++) {
        t.start(ctx->stream);
        YFD_LAUNCH(ctx, yfd::dfma_peak, grid, 256, 0, iters, 1.000000001, sink.p);
        t.stop(ctx->stream);
        const double flops = 2.0 * 16.0 * (double) iters * 256.0 * (double) grid;
        const double tf = flops / (t.ms() * 1e-3) / 1e12;
        if (tf > best) best = tf;
    }
    *tflops = best;
    return YFD_OK;
}

yfd_status yfd_ctx_measure_fp64_tensor(yfd_ctx *ctx, double *tflops)
{
    using namespace yfd;
    if (!ctx || !tflops) return YFD_ERR_INVALID_ARGUMENT;
    scoped_device dev(ctx->device);
    dbuf<double> sink;
    YFD_TRY(sink.alloc(ctx, 1));
    const int iters = 4096, grid = ctx->sm_count * 8;
    YFD_LAUNCH(ctx, yfd::dmma_peak, grid, 256, 0, iters, 1.000000001, sink.p);
    gpu_timer t;
    double best = 0.0;
    for (int rep = 0; rep < 5; rep++) {
        t.start(ctx->stream);
        YFD_LAUNCH(ctx, yfd::dmma_peak, grid, 256, 0, iters, 1.000000001, sink.p);
        t.stop(ctx->stream);
        /* per warp and iteration: 8 instructions x (8 x 8 x 4) multiply-adds */
        const double flops = 2.0 * 8.0 * 256.0 * (double) iters * 8.0 /* warps per CTA */ * (double) grid;
        const double tf = flops / (t.ms() * 1e-3) / 1e12;
        if (tf > best) best = tf;
    }
    *tflops = best;
    return YFD_OK;
}

int yfd_ctx_rank(const yfd_ctx *ctx) { return ctx ? ctx->rank : 0; }
int yfd_ctx_nranks(const yfd_ctx *ctx) { return ctx ? ctx->nranks : 1; }

} /* extern "C" */
