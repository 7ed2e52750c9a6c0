/*
 * oracle/mas_oracle.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * Plain-C restatement of the reference's monotonic alignment search, used only
 * as the checker in tests/, __graft_entry__.smoke() and bench.py's cpu_baseline
 * / --impl reference legs.  Nothing under glowtts_b200/ may link or call this.
 *
 * Follows /root/reference/monotonic_align/core.pyx:
 *   - mas_oracle_each  <- maximum_path_each  (core.pyx:9-35)
 *   - mas_oracle_batch <- maximum_path_c     (core.pyx:40-45; the shipped
 *     setup.py:5-9 passes no -fopenmp, so the prange loop is serial)
 *   - mas_oracle_wrapper_prep <- the value*mask / length extraction of
 *     monotonic_align/__init__.py:11-19
 * The max() is the select form Cython emits ((v_prev > v_cur) ? v_prev : v_cur).
 *
 * Parity pin: tests/test_oracle_mas.py checks this file bit-for-bit against the
 * reference's own compiled core.pyx (oracle/_ref, built by oracle/build.py from
 * the sources where they lie) and against tests/golden/mas_*.npz, which were
 * produced by that compiled reference (tests/golden/make_golden.py).
 */
#include <stdint.h>
#include <stddef.h>

static inline int imax(int a, int b) { return a > b ? a : b; }
static inline int imin(int a, int b) { return a < b ? a : b; }

/* value: [t_x_stride rows][row_stride cols], y fastest; mutated in place.
 * path : same geometry, int32, must be pre-zeroed by the caller. */
void mas_oracle_each(int32_t *path, float *value, int t_x, int t_y,
                     ptrdiff_t row_stride, float max_neg_val)
{
    int x, y;
    int index = t_x - 1;

    for (y = 0; y < t_y; y++) {
        int lo = imax(0, t_x + y - t_y);
        int hi = imin(t_x, y + 1);
        for (x = lo; x < hi; x++) {
            float v_cur, v_prev, best;
            if (x == y) v_cur = max_neg_val;
            else        v_cur = value[(ptrdiff_t)x * row_stride + (y - 1)];
            if (x == 0) v_prev = (y == 0) ? 0.0f : max_neg_val;
            else        v_prev = value[(ptrdiff_t)(x - 1) * row_stride + (y - 1)];
            best = (v_prev > v_cur) ? v_prev : v_cur;
            value[(ptrdiff_t)x * row_stride + y] = best + value[(ptrdiff_t)x * row_stride + y];
        }
    }

    for (y = t_y - 1; y >= 0; y--) {
        path[(ptrdiff_t)index * row_stride + y] = 1;
        if (index != 0 &&
            (index == y ||
             value[(ptrdiff_t)index * row_stride + (y - 1)] <
             value[(ptrdiff_t)(index - 1) * row_stride + (y - 1)]))
            index = index - 1;
    }
}

/* Batch driver: serial over b, exactly like the shipped (OpenMP-less) build. */
void mas_oracle_batch(int32_t *paths, float *values, const int32_t *t_xs,
                      const int32_t *t_ys, int b, int t_x_max, int t_y_max,
                      float max_neg_val)
{
    int i;
    ptrdiff_t plane = (ptrdiff_t)t_x_max * t_y_max;
    for (i = 0; i < b; i++) {
        if (t_xs[i] <= 0 || t_ys[i] <= 0) continue;   /* range(t_y) empty; nothing written */
        mas_oracle_each(paths + i * plane, values + i * plane, t_xs[i], t_ys[i],
                        (ptrdiff_t)t_y_max, max_neg_val);
    }
}

/* Same as mas_oracle_batch but parallel over utterances when built with
 * -fopenmp: "what prange intended".  Used only for the all-cores CPU baseline. */
void mas_oracle_batch_omp(int32_t *paths, float *values, const int32_t *t_xs,
                          const int32_t *t_ys, int b, int t_x_max, int t_y_max,
                          float max_neg_val)
{
    int i;
    ptrdiff_t plane = (ptrdiff_t)t_x_max * t_y_max;
#ifdef _OPENMP
#pragma omp parallel for schedule(dynamic, 1)
#endif
    for (i = 0; i < b; i++) {
        if (t_xs[i] <= 0 || t_ys[i] <= 0) continue;
        mas_oracle_each(paths + i * plane, values + i * plane, t_xs[i], t_ys[i],
                        (ptrdiff_t)t_y_max, max_neg_val);
    }
}

/* __init__.py:11,18-19: value *= mask ; t_x = mask.sum(1)[:,0] ; t_y = mask.sum(2)[:,0] */
void mas_oracle_wrapper_prep(float *values, const float *mask, int32_t *t_xs,
                             int32_t *t_ys, int b, int t_x_max, int t_y_max)
{
    ptrdiff_t plane = (ptrdiff_t)t_x_max * t_y_max;
    int i, x, y;
    for (i = 0; i < b; i++) {
        const float *m = mask + i * plane;
        float *v = values + i * plane;
        float sx = 0.0f, sy = 0.0f;
        for (x = 0; x < t_x_max; x++) sx += m[(ptrdiff_t)x * t_y_max];   /* column 0 summed over x */
        for (y = 0; y < t_y_max; y++) sy += m[y];                         /* row 0 summed over y */
        t_xs[i] = (int32_t)sx;
        t_ys[i] = (int32_t)sy;
        for (x = 0; x < t_x_max; x++)
            for (y = 0; y < t_y_max; y++)
                v[(ptrdiff_t)x * t_y_max + y] *= m[(ptrdiff_t)x * t_y_max + y];
    }
}
