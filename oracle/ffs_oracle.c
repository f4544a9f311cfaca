/*
 * ffs_oracle.c -- CPU ORACLE (test infrastructure, never shipped, never linked
 * into libffs_b200.so).  Plain-C restatement of the reference's CPU algorithm for
 * the spot-finding hot path.  All file:line citations are relative to the
 * reference tree (DiamondLightSource/fast-feedback-service @ d4b779e).
 *
 * Build: gcc -O2 -ffp-contract=off (no -march=native, no -ffast-math) so that
 * every double operation is a separately rounded IEEE-754 operation, exactly as
 * in the reference's x86-64 build.
 *
 * Parity status: every part is PINNED against the reference's own sources compiled in
 * place under oracle/_ref/ -- standard dispersion: standalone.cc (+ tests/golden/);
 * dispersion_extended and the gain variants: baseline.cpp; connected components, spots,
 * filters, 3D components: connected_components.cc -- see ffs_oracle.h and DESIGN.md 6.
 */
#include "ffs_oracle.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>

/* Summed-area-table cell: standalone.cc:34-38 / baseline.cpp:334-339. */
typedef struct {
    int m;    /* number of valid pixels   */
    double x; /* sum of valid pixels      */
    double y; /* sum of squares           */
} sat_cell;

void ffs_oracle_default_params(ffs_oracle_params *p) {
    /* standalone.cc:16-20, spotfind_test_utils.h:13-16 */
    p->kernel_half_x = 3;
    p->kernel_half_y = 3;
    p->min_count = 2;
    p->nsig_b = 6.0;
    p->nsig_s = 3.0;
    p->threshold = 0.0;
}

/* Effective min_count: standalone.cc:58-64 (<= 0 means "whole kernel"). */
static int effective_min_count(const ffs_oracle_params *p) {
    int num_kernel = (2 * p->kernel_half_x + 1) * (2 * p->kernel_half_y + 1);
    return p->min_count <= 0 ? num_kernel : p->min_count;
}

/* u16/u32 -> double conversion of the whole frame: baseline.cpp:866-868
 * (the C API converts image_t_type to double before thresholding). */
static double *frame_to_double(const void *image, int bpp, size_t n) {
    double *src = (double *)malloc(n * sizeof(double));
    if (!src) return NULL;
    if (bpp == 2) {
        const uint16_t *in = (const uint16_t *)image;
        for (size_t i = 0; i < n; ++i) src[i] = in[i];
    } else {
        const uint32_t *in = (const uint32_t *)image;
        for (size_t i = 0; i < n; ++i) src[i] = in[i];
    }
    return src;
}

/* compute_sat: standalone.cc:74-105 == baseline.cpp:375-406.
 * Row-running sums plus the cell above; pixels >= 2^24 or masked contribute 0. */
static void build_sat(sat_cell *table, const double *src, const uint8_t *valid, int xsize,
                      int ysize) {
    const double BIG = (double)(1 << 24);
    size_t k = 0;
    for (int j = 0; j < ysize; ++j) {
        int m = 0;
        double x = 0, y = 0;
        for (int i = 0; i < xsize; ++i, ++k) {
            int mm = (valid[k] && src[k] < BIG) ? 1 : 0;
            m += mm;
            x += mm * src[k];
            y += mm * src[k] * src[k];
            if (j == 0) {
                table[k].m = m;
                table[k].x = x;
                table[k].y = y;
            } else {
                table[k].m = table[k - xsize].m + m;
                table[k].x = table[k - xsize].x + x;
                table[k].y = table[k - xsize].y + y;
            }
        }
    }
}

/* Window sums from the SAT, in the reference's exact operation order:
 * standalone.cc:128-161 == baseline.cpp:428-461 (and :596-620 without y).
 * The window is [i-kx, i+kx] x [j-ky, j+ky] clipped to the image. */
static void sat_window(const sat_cell *table, int xsize, int ysize, int i, int j, int kx,
                       int ky, double *pm, double *px, double *py) {
    int i0 = i - kx - 1, i1 = i + kx;
    int j0 = j - ky - 1, j1 = j + ky;
    i1 = i1 < xsize ? i1 : xsize - 1;
    j1 = j1 < ysize ? j1 : ysize - 1;
    double m = 0, x = 0, y = 0;
    if (i0 >= 0 && j0 >= 0) {
        const sat_cell *d00 = &table[(size_t)j0 * xsize + i0];
        const sat_cell *d10 = &table[(size_t)j1 * xsize + i0];
        const sat_cell *d01 = &table[(size_t)j0 * xsize + i1];
        m += d00->m - (d10->m + d01->m);
        x += d00->x - (d10->x + d01->x);
        y += d00->y - (d10->y + d01->y);
    } else if (i0 >= 0) {
        const sat_cell *d10 = &table[(size_t)j1 * xsize + i0];
        m -= d10->m;
        x -= d10->x;
        y -= d10->y;
    } else if (j0 >= 0) {
        const sat_cell *d01 = &table[(size_t)j0 * xsize + i1];
        m -= d01->m;
        x -= d01->x;
        y -= d01->y;
    }
    const sat_cell *d11 = &table[(size_t)j1 * xsize + i1];
    m += d11->m;
    x += d11->x;
    y += d11->y;
    *pm = m;
    *px = x;
    *py = y;
}

/* The two dispersion tests, standalone.cc:164-171.  with_signal_test = 0 gives
 * the extended algorithm's first pass (baseline.cpp:465-469: a > c only, and no
 * "src > threshold" requirement). */
static int dispersion_test(double m, double x, double y, double v, int mask_k,
                           const ffs_oracle_params *p, int min_count, int with_signal_test) {
    if (!(mask_k && m >= min_count && x >= 0)) return 0;
    if (with_signal_test && !(v > p->threshold)) return 0;
    double a = m * y - x * x - x * (m - 1);
    double c = x * p->nsig_b * sqrt(2 * (m - 1));
    if (!with_signal_test) return a > c;
    double b = m * v - x;
    double d = p->nsig_s * sqrt(x * m);
    return a > c && b > d;
}

/* final-pass test, baseline.cpp:630-643 */
static int final_test(double m, double x, double v, int mask_k, int eroded_k,
                      const ffs_oracle_params *p) {
    if (!(mask_k && m >= 0 && x >= 0)) return 0;
    int dispersion_mask = !eroded_k;
    int global_mask = v > p->threshold;
    double mean = (m >= 2 ? (x / m) : 0);
    int local_mask = v >= (mean + p->nsig_s * sqrt(mean));
    return dispersion_mask && global_mask && local_mask;
}

/* the gain overloads of the two extended tests: compute_dispersion_threshold(table, src,
 * mask, gain, dst) baseline.cpp:484-541 and compute_final_threshold(table, src, mask, gain,
 * dst) :657-714 */
static int dispersion_test_gain(double m, double x, double y, int mask_k, double g,
                                const ffs_oracle_params *p, int min_count) {
    if (!(mask_k && m >= min_count && x >= 0)) return 0;
    double a = m * y - x * x;
    double c = g * x * (m - 1 + p->nsig_b * sqrt(2 * (m - 1)));
    return a > c;
}

static int final_test_gain(double m, double x, double v, int mask_k, int eroded_k, double g,
                           const ffs_oracle_params *p) {
    if (!(mask_k && m >= 0 && x >= 0)) return 0;
    int dispersion_mask = !eroded_k;
    int global_mask = v > p->threshold;
    double mean = (m >= 2 ? (x / m) : 0);
    int local_mask = v >= (mean + p->nsig_s * sqrt(g * mean));
    return dispersion_mask && global_mask && local_mask;
}

static int check_args(const void *image, int bpp, const uint8_t *mask, int w, int h,
                      const ffs_oracle_params *p, const uint8_t *dst) {
    if (!image || !mask || !p || !dst) return -1;
    if (bpp != 2 && bpp != 4) return -2;
    if (w <= 0 || h <= 0) return -3;
    if (p->kernel_half_x <= 0 || p->kernel_half_y <= 0) return -4;
    if (p->nsig_b < 0 || p->nsig_s < 0 || p->threshold < 0) return -5;
    return 0;
}

/* DispersionThreshold::threshold, standalone.cc:182-196. */
int64_t ffs_oracle_dispersion(const void *image, int bpp, const uint8_t *mask, int width,
                              int height, const ffs_oracle_params *p, uint8_t *dst) {
    int rc = check_args(image, bpp, mask, width, height, p, dst);
    if (rc) return rc;
    size_t n = (size_t)width * height;
    double *src = frame_to_double(image, bpp, n);
    sat_cell *table = (sat_cell *)malloc(n * sizeof(sat_cell));
    if (!src || !table) {
        free(src);
        free(table);
        return -10;
    }
    build_sat(table, src, mask, width, height);
    int min_count = effective_min_count(p);
    int64_t count = 0;
    size_t k = 0;
    for (int j = 0; j < height; ++j)
        for (int i = 0; i < width; ++i, ++k) {
            double m, x, y;
            sat_window(table, width, height, i, j, p->kernel_half_x, p->kernel_half_y, &m, &x,
                       &y);
            dst[k] = (uint8_t)dispersion_test(m, x, y, src[k], mask[k] != 0, p, min_count, 1);
            count += dst[k];
        }
    free(table);
    free(src);
    return count;
}

/* DispersionThreshold::threshold_w_gain, baseline.cpp:289-310 with compute_threshold(table,
 * src, mask, gain, dst) :187-250: same SAT, per-pixel gain map,
 *     a = m*y - x*x;  b = m*src - x;
 *     c = gain*x*(m - 1 + nsig_b*sqrt(2*(m - 1)));  d = nsig_s*sqrt(gain*x*m);
 * Pinned bit for bit against the reference's own class (tests/test_oracle_refext.py). */
int64_t ffs_oracle_dispersion_gain(const void *image, int bpp, const uint8_t *mask, int width,
                                   int height, const ffs_oracle_params *p, const double *gain,
                                   uint8_t *dst) {
    int rc = check_args(image, bpp, mask, width, height, p, dst);
    if (rc) return rc;
    if (!gain) return -1;
    size_t n = (size_t)width * height;
    double *src = frame_to_double(image, bpp, n);
    sat_cell *table = (sat_cell *)malloc(n * sizeof(sat_cell));
    if (!src || !table) {
        free(src);
        free(table);
        return -10;
    }
    build_sat(table, src, mask, width, height);
    int min_count = effective_min_count(p);
    int64_t count = 0;
    size_t k = 0;
    for (int j = 0; j < height; ++j)
        for (int i = 0; i < width; ++i, ++k) {
            double m, x, y;
            sat_window(table, width, height, i, j, p->kernel_half_x, p->kernel_half_y, &m, &x,
                       &y);
            dst[k] = 0;
            if (mask[k] && m >= min_count && x >= 0 && src[k] > p->threshold) {
                double a = m * y - x * x;
                double b = m * src[k] - x;
                double c = gain[k] * x * (m - 1 + p->nsig_b * sqrt(2 * (m - 1)));
                double d = p->nsig_s * sqrt(gain[k] * x * m);
                dst[k] = a > c && b > d;
            }
            count += dst[k];
        }
    free(table);
    free(src);
    return count;
}

/* chebyshev_distance(src, false, dst): DIALS 3.25 dials/algorithms/image/filter/
 * distance.h (un-vendored; call site baseline.cpp:558).  Published algorithm: a
 * two-pass 8-neighbour chamfer transform, out-of-image neighbours count as
 * "far" (height + width).  Result = exact L-infinity distance to the nearest
 * pixel whose value is 0 inside the image. */
static void chebyshev_to_zero(const uint8_t *plane, int width, int height, int *dist) {
    const int far = width + height;
#define DIST_AT(J, I) dist[(size_t)(J) * width + (I)]
    for (int j = 0; j < height; ++j)
        for (int i = 0; i < width; ++i) {
            if (!plane[(size_t)j * width + i]) {
                DIST_AT(j, i) = 0;
                continue;
            }
            int n = j > 0 ? DIST_AT(j - 1, i) : far;
            int w = i > 0 ? DIST_AT(j, i - 1) : far;
            int nw = (j > 0 && i > 0) ? DIST_AT(j - 1, i - 1) : far;
            int ne = (j > 0 && i + 1 < width) ? DIST_AT(j - 1, i + 1) : far;
            int best = n < w ? n : w;
            if (nw < best) best = nw;
            if (ne < best) best = ne;
            DIST_AT(j, i) = best + 1;
        }
    for (int j = height - 1; j >= 0; --j)
        for (int i = width - 1; i >= 0; --i) {
            int s = j + 1 < height ? DIST_AT(j + 1, i) : far;
            int e = i + 1 < width ? DIST_AT(j, i + 1) : far;
            int se = (j + 1 < height && i + 1 < width) ? DIST_AT(j + 1, i + 1) : far;
            int sw = (j + 1 < height && i > 0) ? DIST_AT(j + 1, i - 1) : far;
            int best = s < e ? s : e;
            if (se < best) best = se;
            if (sw < best) best = sw;
            if (best + 1 < DIST_AT(j, i)) DIST_AT(j, i) = best + 1;
        }
#undef DIST_AT
}

/* erode_dispersion_mask, baseline.cpp:552-571.  In: plane = 1 where the first
 * pass fired (a > c).  Out: plane = 1 for valid pixels that are NOT in the eroded
 * signal core, 0 for masked pixels and for the core. */
static int erode_plane(const uint8_t *mask, uint8_t *plane, int width, int height,
                       const ffs_oracle_params *p) {
    size_t n = (size_t)width * height;
    int *dist = (int *)malloc(n * sizeof(int));
    if (!dist) return -10;
    chebyshev_to_zero(plane, width, height, dist);
    int erosion_distance =
      p->kernel_half_x < p->kernel_half_y ? p->kernel_half_x : p->kernel_half_y;
    for (size_t k = 0; k < n; ++k) {
        if (mask[k])
            plane[k] = !(plane[k] && dist[k] >= erosion_distance);
        else
            plane[k] = 0;
    }
    free(dist);
    return 0;
}

/* Exported for the independent cross-checks in tests/test_oracle.py
 * (scipy.ndimage.distance_transform_cdt(metric="chessboard") and a direct 5x5 predicate). */
void ffs_oracle_chebyshev_to_zero(const uint8_t *plane, int width, int height, int32_t *dist) {
    chebyshev_to_zero(plane, width, height, (int *)dist);
}
int ffs_oracle_erode(const uint8_t *mask, uint8_t *plane, int width, int height,
                     const ffs_oracle_params *p) {
    return erode_plane(mask, plane, width, height, p);
}

/* DispersionExtendedThreshold::threshold, baseline.cpp:730-761; with gain != NULL
 * threshold_w_gain, :772-805 (same sequence through the gain overloads). */
static int64_t extended_impl(const void *image, int bpp, const uint8_t *mask, int width,
                             int height, const ffs_oracle_params *p, int stop_after_stage,
                             const double *gain, uint8_t *dst) {
    int rc = check_args(image, bpp, mask, width, height, p, dst);
    if (rc) return rc;
    size_t n = (size_t)width * height;
    double *src = frame_to_double(image, bpp, n);
    sat_cell *table = (sat_cell *)malloc(n * sizeof(sat_cell));
    if (!src || !table) {
        free(src);
        free(table);
        return -10;
    }
    int min_count = effective_min_count(p);
    int64_t count = 0;
    size_t k;

    /* :746 compute_sat(table, src, mask); :751 compute_dispersion_threshold */
    build_sat(table, src, mask, width, height);
    k = 0;
    for (int j = 0; j < height; ++j)
        for (int i = 0; i < width; ++i, ++k) {
            double m, x, y;
            sat_window(table, width, height, i, j, p->kernel_half_x, p->kernel_half_y, &m, &x,
                       &y);
            dst[k] = gain ? (uint8_t)dispersion_test_gain(m, x, y, mask[k] != 0, gain[k], p, min_count)
                          : (uint8_t)dispersion_test(m, x, y, src[k], mask[k] != 0, p, min_count, 0);
            count += dst[k];
        }
    if (stop_after_stage == 1) goto done;

    /* :754 erode_dispersion_mask */
    rc = erode_plane(mask, dst, width, height, p);
    if (rc) {
        count = rc;
        goto done;
    }
    if (stop_after_stage == 2) {
        count = 0;
        for (k = 0; k < n; ++k) count += dst[k];
        goto done;
    }

    /* :757 compute_sat(table, src, dst) -- eroded plane is the validity now;
     * :760 compute_final_threshold, window half-size + 2 (baseline.cpp:589-590) */
    build_sat(table, src, dst, width, height);
    count = 0;
    k = 0;
    for (int j = 0; j < height; ++j)
        for (int i = 0; i < width; ++i, ++k) {
            double m, x, y;
            sat_window(table, width, height, i, j, p->kernel_half_x + 2, p->kernel_half_y + 2,
                       &m, &x, &y);
            dst[k] = gain ? (uint8_t)final_test_gain(m, x, src[k], mask[k] != 0, dst[k], gain[k], p)
                          : (uint8_t)final_test(m, x, src[k], mask[k] != 0, dst[k], p);
            count += dst[k];
        }
done:
    free(table);
    free(src);
    return count;
}

int64_t ffs_oracle_dispersion_extended(const void *image, int bpp, const uint8_t *mask,
                                       int width, int height, const ffs_oracle_params *p,
                                       uint8_t *dst) {
    return extended_impl(image, bpp, mask, width, height, p, 3, NULL, dst);
}

/* DispersionExtendedThreshold::threshold_w_gain, baseline.cpp:772-805.  stage as in
 * ffs_oracle_extended_stage.  Pinned stage by stage (tests/test_oracle_refext.py). */
int64_t ffs_oracle_dispersion_extended_gain(const void *image, int bpp, const uint8_t *mask,
                                            int width, int height, const ffs_oracle_params *p,
                                            const double *gain, int stage, uint8_t *dst) {
    if (!gain) return -1;
    if (stage < 1 || stage > 3) return -6;
    return extended_impl(image, bpp, mask, width, height, p, stage, gain, dst);
}

int64_t ffs_oracle_extended_stage(const void *image, int bpp, const uint8_t *mask, int width,
                                  int height, const ffs_oracle_params *p, int stage,
                                  uint8_t *dst) {
    if (stage < 1 || stage > 3) return -6;
    return extended_impl(image, bpp, mask, width, height, p, stage, NULL, dst);
}

/* ------------------------------------------------------------------------- */
/* Brute-force cross-checks (exact integer window sums, no SAT).              */

static uint64_t pix_at(const void *image, int bpp, size_t k) {
    return bpp == 2 ? ((const uint16_t *)image)[k] : ((const uint32_t *)image)[k];
}

static void brute_window(const void *image, int bpp, const uint8_t *valid, int width,
                         int height, int i, int j, int kx, int ky, double *pm, double *px,
                         double *py) {
    uint64_t m = 0, x = 0;
    unsigned __int128 y = 0;
    for (int jj = j - ky; jj <= j + ky; ++jj) {
        if (jj < 0 || jj >= height) continue;
        for (int ii = i - kx; ii <= i + kx; ++ii) {
            if (ii < 0 || ii >= width) continue;
            size_t k = (size_t)jj * width + ii;
            uint64_t v = pix_at(image, bpp, k);
            if (valid[k] && v < (1u << 24)) {
                m += 1;
                x += v;
                y += (unsigned __int128)v * v;
            }
        }
    }
    *pm = (double)m;
    *px = (double)x;
    *py = (double)y;
}

int64_t ffs_oracle_dispersion_bruteforce(const void *image, int bpp, const uint8_t *mask,
                                         int width, int height, const ffs_oracle_params *p,
                                         uint8_t *dst) {
    int rc = check_args(image, bpp, mask, width, height, p, dst);
    if (rc) return rc;
    int min_count = effective_min_count(p);
    int64_t count = 0;
    size_t k = 0;
    for (int j = 0; j < height; ++j)
        for (int i = 0; i < width; ++i, ++k) {
            double m, x, y;
            brute_window(image, bpp, mask, width, height, i, j, p->kernel_half_x,
                         p->kernel_half_y, &m, &x, &y);
            dst[k] = (uint8_t)dispersion_test(m, x, y, (double)pix_at(image, bpp, k),
                                              mask[k] != 0, p, min_count, 1);
            count += dst[k];
        }
    return count;
}

int64_t ffs_oracle_dispersion_extended_bruteforce(const void *image, int bpp,
                                                  const uint8_t *mask, int width, int height,
                                                  const ffs_oracle_params *p, uint8_t *dst) {
    int rc = check_args(image, bpp, mask, width, height, p, dst);
    if (rc) return rc;
    size_t n = (size_t)width * height;
    uint8_t *pass1 = (uint8_t *)malloc(n);
    uint8_t *eroded = (uint8_t *)malloc(n);
    if (!pass1 || !eroded) {
        free(pass1);
        free(eroded);
        return -10;
    }
    int min_count = effective_min_count(p);
    int er = (p->kernel_half_x < p->kernel_half_y ? p->kernel_half_x : p->kernel_half_y) - 1;
    size_t k = 0;
    for (int j = 0; j < height; ++j)
        for (int i = 0; i < width; ++i, ++k) {
            double m, x, y;
            brute_window(image, bpp, mask, width, height, i, j, p->kernel_half_x,
                         p->kernel_half_y, &m, &x, &y);
            pass1[k] = (uint8_t)dispersion_test(m, x, y, (double)pix_at(image, bpp, k),
                                                mask[k] != 0, p, min_count, 0);
        }
    /* erosion as a neighbourhood predicate: a first-pass pixel stays in the signal
     * core iff no 0-pixel lies within L-inf distance (erosion_distance - 1) inside
     * the image; the eroded plane is then the complement on valid pixels. */
    k = 0;
    for (int j = 0; j < height; ++j)
        for (int i = 0; i < width; ++i, ++k) {
            if (!mask[k]) {
                eroded[k] = 0;
                continue;
            }
            int core = pass1[k];
            for (int jj = j - er; core && jj <= j + er; ++jj) {
                if (jj < 0 || jj >= height) continue;
                for (int ii = i - er; ii <= i + er; ++ii) {
                    if (ii < 0 || ii >= width) continue;
                    if (!pass1[(size_t)jj * width + ii]) {
                        core = 0;
                        break;
                    }
                }
            }
            eroded[k] = !core;
        }
    int64_t count = 0;
    k = 0;
    for (int j = 0; j < height; ++j)
        for (int i = 0; i < width; ++i, ++k) {
            double m, x, y;
            brute_window(image, bpp, eroded, width, height, i, j, p->kernel_half_x + 2,
                         p->kernel_half_y + 2, &m, &x, &y);
            dst[k] = (uint8_t)final_test(m, x, (double)pix_at(image, bpp, k), mask[k] != 0,
                                         eroded[k], p);
            count += dst[k];
        }
    free(pass1);
    free(eroded);
    return count;
}

/* ------------------------------------------------------------------------- */
/* Resolution mask: kernels/masking.cu:37-53 (distance), :64-73 (resolution),
 * :99-147 (kernel body).  float arithmetic; "0.5 * atanf(..)" is a double
 * product narrowed back to float, as in the reference.  NOTE: the reference
 * evaluates this with CUDA's sqrtf/atanf/sinf; libm may differ in the last ulp,
 * which can move individual pixels on the dmin/dmax ring.     */
void ffs_oracle_resolution_mask(uint8_t *mask, int width, int height, float wavelength,
                                float distance_m, float bcx, float bcy, float psx, float psy,
                                float dmin, float dmax) {
    for (int y = 0; y < height; ++y)
        for (int x = 0; x < width; ++x) {
            size_t k = (size_t)y * width + x;
            if (mask[k] == 0) continue;
            float dx = (((float)x + 0.5f) - bcx) * psx;
            float dy = (((float)y + 0.5f) - bcy) * psy;
            float r = sqrtf(dx * dx + dy * dy);
            float theta = (float)(0.5 * atanf(r / distance_m));
            float d = wavelength / (2 * sinf(theta));
            if (dmin > 0 && d < dmin) {
                mask[k] = 0;
                continue;
            }
            if (dmax > 0 && d > dmax) {
                mask[k] = 0;
                continue;
            }
            mask[k] = 1;
        }
}

/* ------------------------------------------------------------------------- */
/* Connected components.  Reference: connected_components.cc:17-139 (2D graph:
 * one vertex per strong pixel in ascending linear index, edges to index+1 and
 * index+width when those are strong -- NO row-wrap check on index+1, :62-70),
 * :270-470 (3D: same plus an edge between identical linear indices of
 * consecutive slices, :352-370).  boost::connected_components labels
 * components in order of their lowest vertex id, i.e. ascending minimum
 * (slice, linear index); a union-find with "smaller root wins" gives the same
 * partition, and sorting roots ascending gives the same order.               */

static int64_t uf_find(int64_t *parent, int64_t i) {
    while (parent[i] != i) {
        parent[i] = parent[parent[i]];
        i = parent[i];
    }
    return i;
}

static void uf_union(int64_t *parent, int64_t a, int64_t b) {
    a = uf_find(parent, a);
    b = uf_find(parent, b);
    if (a == b) return;
    if (a < b)
        parent[b] = a;
    else
        parent[a] = b;
}

/* Reflection3D::is_signal_preferred, connected_components.cc:143-157 */
static int peak_preferred(int cz, uint32_t cy, uint32_t cx, int pz, uint32_t py, uint32_t px) {
    if (cz != pz) return cz < pz;
    if (cy != py) return cy < py;
    return cx < px;
}

int64_t ffs_oracle_connected_components(const uint8_t *strong, const void *image, int bpp,
                                        int width, int height, int n_slices, int row_wrap,
                                        int64_t *labels_out, ffs_oracle_spot *spots_out,
                                        int64_t max_spots) {
    if (!strong || !image || width <= 0 || height <= 0 || n_slices <= 0) return -1;
    const size_t px = (size_t)width * height;
    const size_t total = px * (size_t)n_slices;

    /* vertex ids in (slice, linear index) order: connected_components.cc:292-309 */
    int64_t n_vertices = 0;
    for (size_t g = 0; g < total; ++g) n_vertices += strong[g] != 0;
    int64_t *vertex_of = (int64_t *)malloc(total * sizeof(int64_t)); /* global idx -> vertex */
    int64_t *gidx = (int64_t *)malloc((size_t)(n_vertices + 1) * sizeof(int64_t));
    int64_t *parent = (int64_t *)malloc((size_t)(n_vertices + 1) * sizeof(int64_t));
    if (!vertex_of || !gidx || !parent) {
        free(vertex_of);
        free(gidx);
        free(parent);
        return -10;
    }
    int64_t v = 0;
    for (size_t g = 0; g < total; ++g) {
        if (strong[g]) {
            vertex_of[g] = v;
            gidx[v] = (int64_t)g;
            parent[v] = v;
            ++v;
        } else
            vertex_of[g] = -1;
    }

    for (int64_t a = 0; a < n_vertices; ++a) {
        size_t g = (size_t)gidx[a];
        size_t z = g / px, k = g % px;
        size_t xk = k % (size_t)width;
        /* right neighbour: linear_index + 1 (cc:62, 66-70) */
        if (k + 1 < px && strong[g + 1] && (row_wrap || xk + 1 < (size_t)width))
            uf_union(parent, a, vertex_of[g + 1]);
        /* below neighbour: linear_index + width (cc:63, 73-77) */
        if (k + (size_t)width < px && strong[g + width]) uf_union(parent, a, vertex_of[g + width]);
        /* next slice, same linear index (cc:352-370) */
        if (z + 1 < (size_t)n_slices && strong[g + px]) uf_union(parent, a, vertex_of[g + px]);
    }

    /* component ids in ascending root order (== Boost label order) */
    int64_t *comp_of_root = (int64_t *)malloc((size_t)(n_vertices + 1) * sizeof(int64_t));
    if (!comp_of_root) {
        free(vertex_of);
        free(gidx);
        free(parent);
        return -10;
    }
    int64_t n_comp = 0;
    for (int64_t a = 0; a < n_vertices; ++a) {
        int64_t r = uf_find(parent, a);
        if (r == a) comp_of_root[a] = n_comp++;
    }

    if (labels_out) {
        for (size_t g = 0; g < total; ++g) labels_out[g] = -1;
        for (int64_t a = 0; a < n_vertices; ++a) labels_out[gidx[a]] = gidx[uf_find(parent, a)];
    }

    if (spots_out && max_spots > 0) {
        int64_t n_out = n_comp < max_spots ? n_comp : max_spots;
        /* accumulators (doubles as in Reflection3D::center_of_mass, hpp:81-99) */
        double *wx = (double *)calloc((size_t)n_out * 4, sizeof(double));
        if (!wx) {
            free(vertex_of);
            free(gidx);
            free(parent);
            free(comp_of_root);
            return -10;
        }
        double *wy = wx + n_out, *wz = wy + n_out, *wt = wz + n_out;
        for (int64_t c = 0; c < n_out; ++c) {
            ffs_oracle_spot *s = &spots_out[c];
            memset(s, 0, sizeof(*s));
            s->x_min = UINT32_MAX;
            s->y_min = UINT32_MAX;
            s->z_min = INT32_MAX;
            s->z_max = INT32_MIN;
        }
        /* signals are added in (z, linear index) order: cc:409-446 / :245-252 */
        for (int64_t a = 0; a < n_vertices; ++a) {
            int64_t r = uf_find(parent, a);
            int64_t c = comp_of_root[r];
            if (c >= n_out) continue;
            ffs_oracle_spot *s = &spots_out[c];
            size_t g = (size_t)gidx[a];
            int z = (int)(g / px);
            size_t k = g % px;
            uint32_t x = (uint32_t)(k % (size_t)width), y = (uint32_t)(k / (size_t)width);
            uint64_t inten = pix_at(image, bpp, g);
            if (s->num_pixels == 0) s->label = (uint64_t)gidx[r];
            /* add_signal, hpp:43-64 */
            if (x < s->x_min) s->x_min = x;
            if (x > s->x_max) s->x_max = x;
            if (y < s->y_min) s->y_min = y;
            if (y > s->y_max) s->y_max = y;
            if (z < s->z_min) s->z_min = z;
            if (z > s->z_max) s->z_max = z;
            /* peak search, hpp:122-176: strictly greater replaces; tie -> preferred */
            if (s->num_pixels == 0 || inten > s->peak_intensity ||
                (inten == s->peak_intensity &&
                 peak_preferred(z, y, x, s->peak_z, s->peak_y, s->peak_x))) {
                s->peak_intensity = inten;
                s->peak_x = x;
                s->peak_y = y;
                s->peak_z = z;
            }
            s->num_pixels += 1;
            s->sum_intensity += inten;
            wx[c] += ((double)x + 0.5) * (double)inten;
            wy[c] += ((double)y + 0.5) * (double)inten;
            wz[c] += (z + 0.5) * (double)inten;
            wt[c] += (double)inten;
        }
        for (int64_t c = 0; c < n_out; ++c) {
            ffs_oracle_spot *s = &spots_out[c];
            /* center_of_mass(): double quotient stored as float (hpp:90-99).  The
             * reference throws when total intensity is 0; strong pixels always have
             * v > threshold >= 0 so this cannot occur for threshold output. */
            if (wt[c] == 0) {
                s->com_x = s->com_y = s->com_z = NAN;
                s->peak_centroid_distance = NAN;
                continue;
            }
            s->com_x = (float)(wx[c] / wt[c]);
            s->com_y = (float)(wy[c] / wt[c]);
            s->com_z = (float)(wz[c] / wt[c]);
            /* peak_centroid_distance(), hpp:190-205: float arithmetic */
            float dx = ((float)s->peak_x + 0.5f) - s->com_x;
            float dy = ((float)s->peak_y + 0.5f) - s->com_y;
            float dz = (s->peak_z + 0.5f) - s->com_z;
            s->peak_centroid_distance = sqrtf(dx * dx + dy * dy + dz * dz);
        }
        free(wx);
    }

    free(comp_of_root);
    free(vertex_of);
    free(gidx);
    free(parent);
    return n_comp;
}

/* filter_reflections, connected_components.cc:207-236: two stable remove_if
 * passes -- size first, then peak-centroid separation. */
int64_t ffs_oracle_filter_spots(ffs_oracle_spot *spots, int64_t n, uint32_t min_spot_size,
                                float max_sep, int64_t *n_removed_size,
                                int64_t *n_removed_sep) {
    int64_t kept = n, rs = 0, rp = 0;
    if (min_spot_size > 0) {
        int64_t o = 0;
        for (int64_t i = 0; i < kept; ++i)
            if (!(spots[i].num_pixels < min_spot_size)) spots[o++] = spots[i];
        rs = kept - o;
        kept = o;
    }
    if (max_sep > 0) {
        int64_t o = 0;
        for (int64_t i = 0; i < kept; ++i)
            if (!(spots[i].peak_centroid_distance > max_sep)) spots[o++] = spots[i];
        rp = kept - o;
        kept = o;
    }
    if (n_removed_size) *n_removed_size = rs;
    if (n_removed_sep) *n_removed_sep = rp;
    return kept;
}

/* Per-image CPU hot path as in the reference frame loop (spotfinder.cc:859-928,
 * with the CPU threshold standing in for the CUDA kernel as --validate does,
 * :1012-1037). */
int ffs_oracle_process_frame(const void *image, int bpp, const uint8_t *mask, int width,
                             int height, const ffs_oracle_params *p, int algorithm,
                             uint32_t min_spot_size, float max_sep, int row_wrap,
                             int64_t counts[4]) {
    size_t n = (size_t)width * height;
    uint8_t *strong = (uint8_t *)malloc(n);
    if (!strong) return -10;
    int64_t ns = algorithm == 0
                   ? ffs_oracle_dispersion(image, bpp, mask, width, height, p, strong)
                   : ffs_oracle_dispersion_extended(image, bpp, mask, width, height, p, strong);
    if (ns < 0) {
        free(strong);
        return (int)ns;
    }
    int64_t cap = ns > 0 ? ns : 1;
    ffs_oracle_spot *spots = (ffs_oracle_spot *)malloc((size_t)cap * sizeof(ffs_oracle_spot));
    if (!spots) {
        free(strong);
        return -10;
    }
    int64_t nc = ffs_oracle_connected_components(strong, image, bpp, width, height, 1, row_wrap,
                                                 NULL, spots, cap);
    /* generate_boxes, connected_components.cc:87-139 */
    int64_t n_boxes = 0, n_strong_filtered = 0;
    for (int64_t c = 0; c < nc; ++c) {
        if (min_spot_size == 0 || spots[c].num_pixels >= min_spot_size) {
            ++n_boxes;
            n_strong_filtered += spots[c].num_pixels;
        }
    }
    int64_t kept = ffs_oracle_filter_spots(spots, nc, min_spot_size, max_sep, NULL, NULL);
    counts[0] = ns;
    counts[1] = n_boxes;
    counts[2] = n_strong_filtered;
    counts[3] = kept;
    free(spots);
    free(strong);
    return 0;
}
