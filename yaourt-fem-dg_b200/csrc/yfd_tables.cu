/*
 * yfd_tables.cu -- the reference's tabulated quadrature rules, resident on the device.
 * Data: include/yfd_quad_tables.h (generated from core/quadratures.hpp:182-274, :320-538).
 * Rule selection follows gauss_legendre_tab (core/quadratures.hpp:296-311) and
 * triangle_quadrature_dunavant (core/quadratures.hpp:570-581).
 */
#include "yfd_internal.hpp"
#include "yfd_quad_tables.h"

namespace yfd {

yfd_status ensure_tables(yfd_ctx *ctx)
{
    if (ctx->tab_dunavant) return YFD_OK;
    YFD_CUDA(ctx, cudaMalloc((void **) &ctx->tab_gauss_x, sizeof(yfd_gauss_x)));
    YFD_CUDA(ctx, cudaMalloc((void **) &ctx->tab_gauss_w, sizeof(yfd_gauss_w)));
    YFD_CUDA(ctx, cudaMalloc((void **) &ctx->tab_dunavant, sizeof(yfd_dunavant)));
    YFD_CUDA(ctx, cudaMemcpyAsync(ctx->tab_gauss_x, yfd_gauss_x, sizeof(yfd_gauss_x), cudaMemcpyHostToDevice, ctx->stream));
    YFD_CUDA(ctx, cudaMemcpyAsync(ctx->tab_gauss_w, yfd_gauss_w, sizeof(yfd_gauss_w), cudaMemcpyHostToDevice, ctx->stream));
    YFD_CUDA(ctx, cudaMemcpyAsync(ctx->tab_dunavant, yfd_dunavant, sizeof(yfd_dunavant), cudaMemcpyHostToDevice, ctx->stream));
    YFD_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return YFD_OK;
}

yfd_status gauss_rule_for_degree(yfd_ctx *ctx, int degree, quad_rule *out)
{
    YFD_TRY(ensure_tables(ctx));
    const int n = degree / 2 + 1;
    if (degree < 0 || n > YFD_GAUSS_MAX_POINTS)
        return fail(ctx, YFD_ERR_DEGREE_TOO_HIGH, "Gauss-Legendre table: degree %d needs %d points, table holds %d", degree, n, YFD_GAUSS_MAX_POINTS);
    out->x = ctx->tab_gauss_x + yfd_gauss_off[n];
    out->w = ctx->tab_gauss_w + yfd_gauss_off[n];
    out->n = n;
    return YFD_OK;
}

yfd_status dunavant_rule_for_degree(yfd_ctx *ctx, int degree, tri_rule *out)
{
    YFD_TRY(ensure_tables(ctx));
    if (degree < 0 || degree > YFD_DUNAVANT_MAX_DEGREE)
        return fail(ctx, YFD_ERR_DEGREE_TOO_HIGH, "Dunavant quadrature: degree too high");
    const int d = degree == 0 ? 1 : degree;
    out->cw = ctx->tab_dunavant + 4 * yfd_dunavant_off[d];
    out->n = yfd_dunavant_off[d + 1] - yfd_dunavant_off[d];
    return YFD_OK;
}

} // namespace yfd
