/* TEST INFRASTRUCTURE ONLY -- plain-C restatement of the aggregation arithmetic on the
 * STEM-GNN encoder hot path, written as the sequential per-edge loops that
 * `index_select -> message -> scatter_add_ -> / count.clamp(1)` amounts to on one CPU thread
 * (reference call sites: STEM-GNN/model/encoder.py:82,94-97 (MySAGEConv.propagate/message),
 * :124 (scatter_mean), :197/:199 (GCN/GIN sums); third-party arithmetic: torch_geometric 2.3.0
 * MeanAggregation + torch_scatter 2.1.0 scatter_sum, environment.yml:292,314).
 * It pins the SUMMATION ORDER (original edge order per destination) that the CUDA kernels
 * reproduce, and serves as a single-thread CPU reference.  Built by oracle/c/Makefile into
 * oracle/_build/liboracle_agg.so.  Never linked into the product library.
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

/* out[dst] (+)= act(x[src] + xe[e]) in edge order, then mean division.
 * act: 0 none, 1 relu.  reduce: 0 sum, 1 mean.  xe may be NULL.  w (per-edge weight, GCN)
 * may be NULL; when given the message is w[e] * x[src] (rounded) before the add. */
void oracle_aggregate_fwd(const int64_t* src, const int64_t* dst, int64_t E, int64_t N,
                          int64_t F, const float* x, const float* xe, const float* w,
                          int act, int reduce, float* out)
{
    float* cnt = (float*)calloc((size_t)N, sizeof(float));
    memset(out, 0, (size_t)N * (size_t)F * sizeof(float));
    for (int64_t e = 0; e < E; ++e) {
        const float* xs = x + src[e] * F;
        float* o = out + dst[e] * F;
        for (int64_t f = 0; f < F; ++f) {
            volatile float m = xs[f];
            if (xe) m = m + xe[e * F + f];
            if (act == 1) m = m > 0.f ? m : 0.f;
            if (w) m = w[e] * m;
            o[f] = o[f] + m;
        }
        cnt[dst[e]] += 1.f;
    }
    if (reduce == 1)
        for (int64_t i = 0; i < N; ++i) {
            float c = cnt[i] < 1.f ? 1.f : cnt[i];
            for (int64_t f = 0; f < F; ++f) out[i * F + f] = out[i * F + f] / c;
        }
    free(cnt);
}

/* autograd of the above w.r.t. x: g_s = g_out / clamp(count,1) (mean), gathered per edge,
 * relu mask (x[src] + xe > 0), index_add_ into g_x[src] in edge order. */
void oracle_aggregate_bwd(const int64_t* src, const int64_t* dst, int64_t E, int64_t N,
                          int64_t F, const float* x, const float* xe, const float* w,
                          int act, int reduce, const float* g_out, float* g_x)
{
    float* cnt = (float*)calloc((size_t)N, sizeof(float));
    for (int64_t e = 0; e < E; ++e) cnt[dst[e]] += 1.f;
    memset(g_x, 0, (size_t)N * (size_t)F * sizeof(float));
    for (int64_t e = 0; e < E; ++e) {
        const float* go = g_out + dst[e] * F;
        const float* xs = x + src[e] * F;
        float c = cnt[dst[e]] < 1.f ? 1.f : cnt[dst[e]];
        float* gx = g_x + src[e] * F;
        for (int64_t f = 0; f < F; ++f) {
            volatile float g = go[f];
            if (reduce == 1) g = g / c;
            if (w) g = w[e] * g;
            if (act == 1) {
                float pre = xs[f];
                if (xe) pre = pre + xe[e * F + f];
                if (!(pre > 0.f)) g = 0.f;
            }
            gx[f] = gx[f] + g;
        }
    }
    free(cnt);
}
