/*
 * yfd_sort.cu -- device scans and a stable LSD radix sort, hand-written for sm_100a.
 *
 * Used by the deterministic triplet -> CSR build (yfd_triplets.cu; replaces the
 * std::sort + fold of blaze::reduce_triplets, core/blaze_sparse_init.hpp:101-122) and
 * by the Z-curve cell renumbering (yfd_mesh.cu).  HBM-bound integer work: coalesced
 * tile loads, shared-memory digit histograms, warp match/ballot ranking.
 */
#include "yfd_internal.hpp"

namespace yfd {

/* ------------------------------------------------------------------------------- */
/* exclusive scan int32 -> int64, n+1 outputs                                        */
/* ------------------------------------------------------------------------------- */
constexpr int SCAN_THREADS = 256;
constexpr int SCAN_ITEMS   = 8;
constexpr int SCAN_TILE    = SCAN_THREADS * SCAN_ITEMS;

__device__ __forceinline__ int64_t warp_incl_scan(int64_t v)
{
    const int lane = threadIdx.x & 31;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        int64_t t = __shfl_up_sync(0xffffffffu, v, o);
        if (lane >= o) v += t;
    }
    return v;
}

/* block-wide exclusive scan of one value per thread; returns exclusive prefix, *total = block sum */
__device__ __forceinline__ int64_t block_excl_scan(int64_t v, int64_t *total)
{
    __shared__ int64_t wsum[32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int nw = (blockDim.x + 31) >> 5;
    int64_t inc = warp_incl_scan(v);
    if (lane == 31) wsum[warp] = inc;
    __syncthreads();
    if (warp == 0) {
        int64_t w = lane < nw ? wsum[lane] : 0;
        int64_t wi = warp_incl_scan(w);
        wsum[lane] = wi - w;                  /* exclusive prefix of the warp sums */
    }
    __syncthreads();
    int64_t res = inc - v + wsum[warp];
    /* total: sum over warps */
    if (total) {
        /* recompute from last warp's entries: exclusive prefix of last warp + its sum */
        __shared__ int64_t tot;
        if (threadIdx.x == blockDim.x - 1) tot = res + v;
        __syncthreads();
        *total = tot;
    }
    __syncthreads();
    return res;
}

__global__ void scan_tile_sums(const int32_t *__restrict__ in, int64_t n, int64_t *__restrict__ tile_sums)
{
    const int64_t base = (int64_t) blockIdx.x * SCAN_TILE;
    int64_t s = 0;
#pragma unroll
    for (int i = 0; i < SCAN_ITEMS; i++) {
        int64_t idx = base + (int64_t) i * SCAN_THREADS + threadIdx.x;
        if (idx < n) s += in[idx];
    }
    int64_t total;
    block_excl_scan(s, &total);
    if (threadIdx.x == 0) tile_sums[blockIdx.x] = total;
}

/* single block: in-place exclusive scan of m int64 values, total appended at [m] */
__global__ void scan_sums_inplace(int64_t *__restrict__ sums, int64_t m)
{
    int64_t carry = 0;
    for (int64_t base = 0; base < m; base += blockDim.x) {
        int64_t idx = base + threadIdx.x;
        int64_t v = idx < m ? sums[idx] : 0;
        int64_t total;
        int64_t ex = block_excl_scan(v, &total);
        if (idx < m) sums[idx] = carry + ex;
        carry += total;
    }
    if (threadIdx.x == 0) sums[m] = carry;
}

__global__ void scan_tile_write(const int32_t *__restrict__ in, int64_t n,
                                const int64_t *__restrict__ tile_offs, int64_t *__restrict__ out)
{
    /* thread t owns SCAN_ITEMS consecutive inputs so its running sum is a plain loop */
    const int64_t base = (int64_t) blockIdx.x * SCAN_TILE + (int64_t) threadIdx.x * SCAN_ITEMS;
    int32_t v[SCAN_ITEMS];
    int64_t s = 0;
#pragma unroll
    for (int i = 0; i < SCAN_ITEMS; i++) {
        int64_t idx = base + i;
        v[i] = idx < n ? in[idx] : 0;
        s += v[i];
    }
    int64_t ex = block_excl_scan(s, nullptr) + tile_offs[blockIdx.x];
#pragma unroll
    for (int i = 0; i < SCAN_ITEMS; i++) {
        int64_t idx = base + i;
        if (idx < n) out[idx] = ex;
        ex += v[i];
    }
    if (blockIdx.x == gridDim.x - 1 && threadIdx.x == 0)
        out[n] = tile_offs[gridDim.x];
}

__global__ void write_zero_i64(int64_t *p) { *p = 0; }

yfd_status exclusive_scan_i32_to_i64(yfd_ctx *ctx, const int32_t *d_in, int64_t *d_out, int64_t n)
{
    if (n <= 0) {
        YFD_LAUNCH(ctx, write_zero_i64, 1, 1, 0, d_out);
        return YFD_OK;
    }
    const int64_t tiles = ceil_div(n, SCAN_TILE);
    dbuf<int64_t> sums;
    YFD_TRY(sums.alloc(ctx, (size_t) tiles + 1));
    YFD_LAUNCH(ctx, scan_tile_sums, (unsigned) tiles, SCAN_THREADS, 0, d_in, n, sums.p);
    YFD_LAUNCH(ctx, scan_sums_inplace, 1, 512, 0, sums.p, tiles);
    YFD_LAUNCH(ctx, scan_tile_write, (unsigned) tiles, SCAN_THREADS, 0, d_in, n, sums.p, d_out);
    YFD_CUDA(ctx, cudaStreamSynchronize(ctx->stream));   /* sums freed on return */
    return YFD_OK;
}

/* ------------------------------------------------------------------------------- */
/* stable LSD radix sort, 8-bit digits                                               */
/* ------------------------------------------------------------------------------- */
constexpr int RS_THREADS = 256;
constexpr int RS_WARPS   = RS_THREADS / 32;
constexpr int RS_ITEMS   = 16;                       /* per thread */
constexpr int RS_TILE    = RS_THREADS * RS_ITEMS;    /* 4096 keys per block */
constexpr int RS_RADIX   = 256;

__global__ void __launch_bounds__(RS_THREADS)
rs_histogram(const uint64_t *__restrict__ keys, int64_t n, int shift, int32_t *__restrict__ ghist, int nblocks)
{
    __shared__ int32_t hist[RS_RADIX];
    hist[threadIdx.x] = 0;
    __syncthreads();
    const int64_t base = (int64_t) blockIdx.x * RS_TILE;
#pragma unroll
    for (int i = 0; i < RS_ITEMS; i++) {
        int64_t idx = base + (int64_t) i * RS_THREADS + threadIdx.x;
        if (idx < n) atomicAdd(&hist[(int) ((keys[idx] >> shift) & (RS_RADIX - 1))], 1);
    }
    __syncthreads();
    /* digit-major so that one exclusive scan over the table yields global offsets */
    ghist[(int64_t) threadIdx.x * nblocks + blockIdx.x] = hist[threadIdx.x];
}

__global__ void __launch_bounds__(RS_THREADS)
rs_scatter(const uint64_t *__restrict__ keys_in, const uint32_t *__restrict__ vals_in,
           uint64_t *__restrict__ keys_out, uint32_t *__restrict__ vals_out,
           int64_t n, int shift, const int64_t *__restrict__ goffs, int nblocks)
{
    __shared__ int32_t whist[RS_WARPS][RS_RADIX];
    __shared__ int64_t gbase[RS_RADIX];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int i = threadIdx.x; i < RS_WARPS * RS_RADIX; i += RS_THREADS) (&whist[0][0])[i] = 0;
    gbase[threadIdx.x] = goffs[(int64_t) threadIdx.x * nblocks + blockIdx.x];
    __syncthreads();

    /* warp w owns the contiguous slice [w*512, (w+1)*512) of the tile; round t, lane l -> element t*32+l */
    const int64_t wbase = (int64_t) blockIdx.x * RS_TILE + (int64_t) warp * (RS_ITEMS * 32);
    uint64_t key[RS_ITEMS];
    uint32_t val[RS_ITEMS];
    int32_t  loc[RS_ITEMS];
    const uint32_t lt = (1u << lane) - 1u;
#pragma unroll
    for (int t = 0; t < RS_ITEMS; t++) {
        const int64_t idx = wbase + t * 32 + lane;
        const bool valid = idx < n;
        key[t] = valid ? keys_in[idx] : 0;
        val[t] = valid ? vals_in[idx] : 0;
        const int digit = valid ? (int) ((key[t] >> shift) & (RS_RADIX - 1)) : RS_RADIX;   /* sentinel groups the tail */
        const uint32_t peers = __match_any_sync(0xffffffffu, digit);
        const int rank = __popc(peers & lt);
        int32_t b = 0;
        if (valid) b = whist[warp][digit];
        __syncwarp();
        if (valid && rank == 0) whist[warp][digit] = b + __popc(peers);
        __syncwarp();
        loc[t] = b + rank;
    }
    __syncthreads();
    {   /* exclusive prefix over warps, one digit per thread */
        const int d = threadIdx.x;
        int32_t run = 0;
#pragma unroll
        for (int w = 0; w < RS_WARPS; w++) {
            int32_t c = whist[w][d];
            whist[w][d] = run;
            run += c;
        }
    }
    __syncthreads();
#pragma unroll
    for (int t = 0; t < RS_ITEMS; t++) {
        const int64_t idx = wbase + t * 32 + lane;
        if (idx < n) {
            const int digit = (int) ((key[t] >> shift) & (RS_RADIX - 1));
            const int64_t pos = gbase[digit] + whist[warp][digit] + loc[t];
            keys_out[pos] = key[t];
            vals_out[pos] = val[t];
        }
    }
}

yfd_status radix_sort_pairs(yfd_ctx *ctx, uint64_t *d_keys, uint32_t *d_vals,
                            uint64_t *d_keys_tmp, uint32_t *d_vals_tmp, int64_t n, int key_bits)
{
    if (n <= 1 || key_bits <= 0) return YFD_OK;
    if (key_bits > 64) key_bits = 64;
    const int passes = (key_bits + 7) / 8;
    const int64_t nblocks64 = ceil_div(n, RS_TILE);
    if (nblocks64 > 0x7fffffff) return fail(ctx, YFD_ERR_UNSUPPORTED, "radix sort: too many keys");
    const int nblocks = (int) nblocks64;
    dbuf<int32_t> ghist;
    dbuf<int64_t> goffs;
    YFD_TRY(ghist.alloc(ctx, (size_t) RS_RADIX * nblocks));
    YFD_TRY(goffs.alloc(ctx, (size_t) RS_RADIX * nblocks + 1));
    uint64_t *kin = d_keys, *kout = d_keys_tmp;
    uint32_t *vin = d_vals, *vout = d_vals_tmp;
    for (int p = 0; p < passes; p++) {
        const int shift = 8 * p;
        YFD_LAUNCH(ctx, rs_histogram, nblocks, RS_THREADS, 0, kin, n, shift, ghist.p, nblocks);
        YFD_TRY(exclusive_scan_i32_to_i64(ctx, ghist.p, goffs.p, (int64_t) RS_RADIX * nblocks));
        YFD_LAUNCH(ctx, rs_scatter, nblocks, RS_THREADS, 0, kin, vin, kout, vout, n, shift, goffs.p, nblocks);
        uint64_t *tk = kin; kin = kout; kout = tk;
        uint32_t *tv = vin; vin = vout; vout = tv;
    }
    if (kin != d_keys) {
        YFD_CUDA(ctx, cudaMemcpyAsync(d_keys, kin, (size_t) n * sizeof(uint64_t), cudaMemcpyDeviceToDevice, ctx->stream));
        YFD_CUDA(ctx, cudaMemcpyAsync(d_vals, vin, (size_t) n * sizeof(uint32_t), cudaMemcpyDeviceToDevice, ctx->stream));
    }
    YFD_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return YFD_OK;
}

__global__ void iota_u32(uint32_t *p, int64_t n)
{
    for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t) gridDim.x * blockDim.x)
        p[i] = (uint32_t) i;
}

yfd_status fill_iota_u32(yfd_ctx *ctx, uint32_t *d, int64_t n)
{
    if (n <= 0) return YFD_OK;
    int grid = (int) (ceil_div(n, 256) < (int64_t) ctx->sm_count * 8 ? ceil_div(n, 256) : (int64_t) ctx->sm_count * 8);
    YFD_LAUNCH(ctx, iota_u32, grid, 256, 0, d, n);
    return YFD_OK;
}

} // namespace yfd
