/*
 * ffs_oracle.h -- CPU ORACLE for the DIALS dispersion spot-finding hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  This is a plain-C restatement of the reference's
 * CPU algorithm (file:line cited per function in ffs_oracle.c).  Only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
 * may load it.  The product library (libffs_b200.so) never links, loads or
 * calls anything in oracle/.
 *
 * Parity status (see DESIGN.md "Oracle"):
 *   - standard dispersion : PINNED  -- checked bit-for-bit against the
 *     reference's own baseline/spotfinder/standalone.cc compiled in place
 *     (oracle/_ref/libffs_ref.so) and against tests/golden/ fixtures produced
 *     by it.
 *   - dispersion_extended and the gain-map variants of both algorithms : PINNED
 *     -- checked bit for bit, stage by stage, against the reference's own
 *     baseline/spotfinder/baseline.cpp classes compiled in place
 *     (oracle/_ref/libffs_refext.so; scitbx array views and DIALS's
 *     chebyshev_distance are the stand-ins of oracle/stub_dials, the latter
 *     also checked against scipy's chessboard distance transform).
 *   - connected components, spot fields, the two filters, 3D components,
 *     Kabsch-space variances : PINNED -- same spots, same order, same floats
 *     as the reference's own connected_components.cc + dx2 Panel / Scan
 *     compiled in place (oracle/_ref/libffs_refcc*.so; Boost.Graph and Eigen
 *     are the stand-ins of oracle/stub_cc).
 *   Both pins need /root/reference at build time; tests/test_oracle_refext.py
 *   and tests/test_oracle_refcc.py skip where the built libraries are absent.
 */
#ifndef FFS_ORACLE_H
#define FFS_ORACLE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct ffs_oracle_params {
    int32_t kernel_half_x; /* 3  (standalone.cc:16)  */
    int32_t kernel_half_y; /* 3 */
    int32_t min_count;     /* 2  (standalone.cc:17)  */
    double nsig_b;         /* 6  (standalone.cc:19)  */
    double nsig_s;         /* 3  (standalone.cc:20)  */
    double threshold;      /* 0  (standalone.cc:18)  */
} ffs_oracle_params;

/* One connected component ("Reflection3D", connected_components.hpp:31-268). */
typedef struct ffs_oracle_spot {
    uint64_t label;          /* canonical label: min (z * width*height + linear index) */
    uint32_t x_min, x_max, y_min, y_max;
    int32_t z_min, z_max;
    uint32_t num_pixels;
    uint32_t peak_x, peak_y;
    int32_t peak_z;
    uint64_t peak_intensity;
    uint64_t sum_intensity;
    float com_x, com_y, com_z;      /* center_of_mass(): double math, float storage */
    float peak_centroid_distance;   /* peak_centroid_distance(): float math        */
} ffs_oracle_spot;

void ffs_oracle_default_params(ffs_oracle_params *p);

/* image: width*height pixels, bytes_per_pixel 2 (uint16) or 4 (uint32).
 * mask : width*height bytes, non-zero = valid.
 * dst  : width*height bytes, written 0/1.
 * Returns the number of strong pixels, or <0 on argument error.            */
int64_t ffs_oracle_dispersion(const void *image, int bytes_per_pixel,
                              const uint8_t *mask, int width, int height,
                              const ffs_oracle_params *p, uint8_t *dst);

/* threshold_w_gain (baseline.cpp:187-250,289-310): gain = width*height doubles. */
int64_t ffs_oracle_dispersion_gain(const void *image, int bytes_per_pixel, const uint8_t *mask,
                                   int width, int height, const ffs_oracle_params *p,
                                   const double *gain, uint8_t *dst);
/* DispersionExtendedThreshold::threshold_w_gain (baseline.cpp:484-541,657-714,772-805);
 * stage 1/2/3 as in ffs_oracle_extended_stage (3 = the final result). */
int64_t ffs_oracle_dispersion_extended_gain(const void *image, int bytes_per_pixel,
                                            const uint8_t *mask, int width, int height,
                                            const ffs_oracle_params *p, const double *gain,
                                            int stage, uint8_t *dst);
int64_t ffs_oracle_dispersion_extended(const void *image, int bytes_per_pixel,
                                       const uint8_t *mask, int width, int height,
                                       const ffs_oracle_params *p, uint8_t *dst);

/* Intermediate planes of the extended algorithm, for stage-by-stage parity.
 * stage 1: dst = dispersion mask (a > c);  stage 2: dst = eroded mask
 * (1 = background/valid, 0 = masked or eroded signal core).                */
int64_t ffs_oracle_extended_stage(const void *image, int bytes_per_pixel,
                                  const uint8_t *mask, int width, int height,
                                  const ffs_oracle_params *p, int stage, uint8_t *dst);

/* Brute-force (no summed-area table) versions: exact integer window sums.
 * Used only to cross-check the SAT restatement on small frames.            */
int64_t ffs_oracle_dispersion_bruteforce(const void *image, int bytes_per_pixel,
                                         const uint8_t *mask, int width, int height,
                                         const ffs_oracle_params *p, uint8_t *dst);
int64_t ffs_oracle_dispersion_extended_bruteforce(const void *image, int bytes_per_pixel,
                                                  const uint8_t *mask, int width,
                                                  int height, const ffs_oracle_params *p,
                                                  uint8_t *dst);

/* chebyshev_distance(src, false, dst) as restated from DIALS (call site baseline.cpp:558) and
 * erode_dispersion_mask (baseline.cpp:552-571), exported so tests can cross-check them against
 * independent implementations. */
void ffs_oracle_chebyshev_to_zero(const uint8_t *plane, int width, int height, int32_t *dist);
int ffs_oracle_erode(const uint8_t *mask, uint8_t *plane_inout, int width, int height,
                     const ffs_oracle_params *p);

/* bshuf_decompress_lz4(chunk + 12, out, n, elem_size, 0) as the frame loop calls it
 * (spotfinder/spotfinder.cc:830-834): one HDF5 bitshuffle-LZ4 filter chunk (12-byte header +
 * body) -> n dense elements.  0 on success (bslz4_oracle.c). */
int ffs_oracle_bslz4_decode(const void *chunk, size_t chunk_bytes, void *out, size_t n_elements,
                            size_t elem_size);

/* Resolution mask (kernels/masking.cu:37-147), applied in place to a u8 mask. */
void ffs_oracle_resolution_mask(uint8_t *mask, int width, int height, float wavelength,
                                float distance_m, float beam_center_x_px,
                                float beam_center_y_px, float pixel_size_x_m,
                                float pixel_size_y_m, float dmin, float dmax);

/* Connected components over a stack of n_slices strong planes (n_slices == 1
 * gives the 2D per-image case with z = 0, connected_components.cc:238-266;
 * n_slices > 1 the 3D case, :270-470).
 *   strong      : n_slices * width*height bytes (non-zero = strong)
 *   image       : n_slices * width*height pixels
 *   row_wrap    : 1 = reproduce the reference's "linear_index + 1" link without a
 *                 row-wrap check (connected_components.cc:62-70); 0 = true 4-conn.
 *   labels_out  : optional, n_slices*width*height int64; -1 for non-strong, else the
 *                 component's canonical label (min z*px + linear index).
 *   spots_out   : optional array of capacity max_spots; components in ascending
 *                 canonical-label order (== Boost label order), UNFILTERED.
 * Returns number of components (may exceed max_spots; only max_spots written). */
int64_t ffs_oracle_connected_components(const uint8_t *strong, const void *image,
                                        int bytes_per_pixel, int width, int height,
                                        int n_slices, int row_wrap, int64_t *labels_out,
                                        ffs_oracle_spot *spots_out, int64_t max_spots);

/* filter_reflections (connected_components.cc:207-236): stable in-place filter.
 * Returns the number kept; n_removed_size / n_removed_sep may be NULL.       */
int64_t ffs_oracle_filter_spots(ffs_oracle_spot *spots, int64_t n, uint32_t min_spot_size,
                                float max_peak_centroid_separation,
                                int64_t *n_removed_size, int64_t *n_removed_sep);

/* Whole per-image hot path on the CPU, as the reference frame loop does it
 * (spotfinder.cc:859-928): threshold -> CC -> boxes (min-size filter) ->
 * find_2d_components (both filters).  Used as the timed CPU baseline "port".
 * Outputs: counts[0]=num_strong_pixels, [1]=n_spots_total (boxes),
 *          [2]=num_strong_pixels_filtered, [3]=n_spots after both filters.   */
int ffs_oracle_process_frame(const void *image, int bytes_per_pixel, const uint8_t *mask,
                             int width, int height, const ffs_oracle_params *p,
                             int algorithm, uint32_t min_spot_size,
                             float max_peak_centroid_separation, int row_wrap,
                             int64_t counts[4]);

#ifdef __cplusplus
}
#endif
#endif
