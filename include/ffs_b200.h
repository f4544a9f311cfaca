/*
 * ffs_b200.h -- C ABI of libffs_b200.so, the B200-native (sm_100a) implementation of
 * fast-feedback-service's DIALS dispersion spot-finding hot path.
 *
 * The reference has no in-process plugin boundary for this path: its replaceable
 * seams are two launch wrappers, one host class and one CPU C API.  Each entry point
 * below names the reference interface it replaces (paths relative to the reference
 * tree).  All functions return 0 (FFS_OK) or a negative ffs_status; no exception
 * crosses the ABI; no torch / C++ type appears in a signature.
 *
 *   reference seam                                              replaced by
 *   ----------------------------------------------------------  -------------------------------
 *   call_do_spotfinding_dispersion   spotfinder/spotfinder.cuh:9-20   ffs_threshold / ffs_process
 *   call_do_spotfinding_extended     spotfinder/spotfinder.cuh:22-34  ffs_threshold / ffs_process
 *                                                                     (params.algorithm = 1)
 *   upload_mask                      spotfinder/spotfinder.cc:62-108  ffs_create (mask argument)
 *   apply_resolution_filtering       spotfinder/spotfinder.cc:110-137 ffs_apply_resolution_mask
 *   call_apply_resolution_mask       spotfinder/kernels/masking.cuh:92-97
 *   ConnectedComponents ctor/get_boxes/find_2d_components
 *        spotfinder/connected_components/connected_components.hpp:299-397
 *                                                                     ffs_process + ffs_fetch
 *   ConnectedComponents::find_3d_components  (same header)            ffs_sweep_*
 *   per-thread frame loop (pinned buffers, H2D, kernels, D2H, CC)
 *        spotfinder/spotfinder.cc:725-1092                            ffs_acquire / ffs_submit /
 *                                                                     ffs_wait  (streaming API)
 *   spotfinder_create / _free / _standard_dispersion
 *        baseline/spotfinder/baseline.h:9-13                          same names (shim section)
 *
 * Data conventions
 *   frame      : width*height pixels, row-major, uint16 or uint32 (bytes_per_pixel 2|4),
 *                dense on the host; on the device rows are padded to a 128-byte pitch.
 *   mask       : width*height bytes, non-zero = valid (VALID_PIXEL = 1, include/common.hpp).
 *   strong bits: 1 bit per pixel, bit (x & 7) of byte y*ceil(width/8) + (x >> 3) on the
 *                host side (numpy.unpackbits(..., bitorder="little") row by row).
 *   spots      : ascending canonical label = minimum (z, linear index) of the component,
 *                which is the order boost::connected_components gives the reference
 *                (connected_components.cc:241-252,396-399).
 */
#ifndef FFS_B200_H
#define FFS_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define FFS_ALGO_DISPERSION 0          /* DispersionThreshold,         standalone.cc:27-206  */
#define FFS_ALGO_DISPERSION_EXTENDED 1 /* DispersionExtendedThreshold, baseline.cpp:325-813  */

typedef enum ffs_status {
    FFS_OK = 0,
    FFS_ERR_ARG = -1,      /* bad argument                                     */
    FFS_ERR_CUDA = -2,     /* CUDA runtime/driver error (see ffs_last_error)   */
    FFS_ERR_NO_DEVICE = -3,/* no CUDA device / wrong architecture              */
    FFS_ERR_CAPACITY = -4, /* a per-frame capacity (strong pixels, spots) hit  */
    FFS_ERR_STATE = -5,    /* call sequence error (slot busy, nothing pending) */
    FFS_ERR_NOMEM = -6,
    FFS_ERR_FORMAT = -7,   /* malformed / unsupported compressed chunk          */
} ffs_status;

typedef struct ffs_ctx ffs_ctx;

typedef struct ffs_params {
    uint32_t width;           /* fast axis, pixels                                              */
    uint32_t height;          /* slow axis                                                      */
    uint32_t bytes_per_pixel; /* 2 (spotfinder) or 4 (spotfinder32, PIXEL_DATA_32BIT)           */
    int32_t algorithm;        /* FFS_ALGO_*                                                     */
    int32_t kernel_half;      /* 3 => 7x7 window (standalone.cc:16; KERNEL_RADIUS)              */
    int32_t min_count;        /* 2 (standalone.cc:17)                                           */
    double nsig_b;            /* 6 (standalone.cc:19)                                           */
    double nsig_s;            /* 3 (standalone.cc:20)                                           */
    double threshold;         /* 0 (standalone.cc:18)                                           */
    uint32_t min_spot_size;   /* 3   (--min-spot-size,    spotfinder.cc:324)                    */
    uint32_t min_spot_size_3d;/* 3   (--min-spot-size-3d, spotfinder.cc:330)                    */
    float max_peak_centroid_separation; /* 2.0 (spotfinder.cc:336)                              */
    int32_t cc_row_wrap;      /* 1 = link index k and k+1 with no row-wrap check, exactly as
                                 connected_components.cc:62-70 does; 0 = true 4-connectivity     */
    int32_t device;           /* CUDA device ordinal (-d/--device, cuda_arg_parser.cc:30)       */
    int32_t n_slots;          /* frames resident on the device at once (batch / ring depth)     */
    uint32_t max_strong_per_frame; /* capacity of the per-frame strong-pixel list (0 = default) */
    uint32_t max_spots_per_frame;  /* capacity of the per-frame spot list         (0 = default) */
    int64_t trusted_max;      /* upper end of the trusted range, reader.get_trusted_range()[1] as
                                 spotfinder.cc:482,868,879 passes it to the kernels
                                 (max_valid_pixel_value); <= 0 = none.  A mask-valid pixel ABOVE
                                 it is treated exactly like a masked pixel: never strong
                                 (thresholding.cu:208-209) and -- unlike the reference GPU
                                 kernel, like the CPU algorithm run on mask && v <= trusted_max --
                                 not part of any window sum either.             */
} ffs_params;

/* One connected component.  Mirrors Reflection3D (connected_components.hpp:31-268):
 * bbox, pixel count, centre of mass (double math, float storage, +0.5 pixel-centre
 * offset, hpp:71-103), peak pixel (max intensity, ties -> smallest (z, y, x),
 * connected_components.cc:143-157) and peak-centroid distance (float math, hpp:111-206). */
typedef struct ffs_spot {
    uint64_t label; /* min (z * width*height + y*width + x) over the component */
    uint32_t x_min, x_max, y_min, y_max;
    int32_t z_min, z_max;
    uint32_t num_pixels;
    uint32_t peak_x, peak_y;
    int32_t peak_z;
    uint64_t peak_intensity;
    uint64_t sum_intensity;
    float com_x, com_y, com_z;
    float peak_centroid_distance;
} ffs_spot;

/* Per-frame counters, as the reference frame loop reports them (spotfinder.cc:905-909,
 * 997-1005):  num_strong_pixels = all strong pixels; n_spots_total = boxes that survive
 * the min_spot_size filter only; num_strong_pixels_filtered = pixels in those boxes;
 * n_spots_filtered = spots that also survive the peak-centroid filter (find_2d_components). */
typedef struct ffs_frame_result {
    uint64_t frame_id;
    uint32_t num_strong_pixels;
    uint32_t n_components; /* before any filter */
    uint32_t n_spots_total;
    uint32_t num_strong_pixels_filtered;
    uint32_t n_spots_filtered;
    int32_t status; /* FFS_OK or FFS_ERR_CAPACITY for this frame */
} ffs_frame_result;

/* ---- lifetime ----------------------------------------------------------------------- */
void ffs_default_params(ffs_params *p);
/* mask: width*height bytes (non-zero = valid) or NULL for "all valid" (upload_mask,
 * spotfinder.cc:62-108). */
int ffs_create(const ffs_params *p, const uint8_t *mask, ffs_ctx **out);
void ffs_destroy(ffs_ctx *ctx);
const char *ffs_last_error(const ffs_ctx *ctx); /* NULL ctx: last ffs_create error */
const char *ffs_version(void);
/* Number of CUDA devices, and "<name> (CUDA M.m)" for --list-devices
 * (cuda_arg_parser.cc:38-53). */
int ffs_device_count(void);
int ffs_device_description(int device, char *buf, size_t buflen);

/* ---- mask ---------------------------------------------------------------------------- */
/* apply_resolution_mask (kernels/masking.cu:99-147): float math, applied in place to
 * the device mask; dmin/dmax <= 0 disable that bound.  Units as the kernel takes them:
 * distance and pixel sizes in metres, beam centre in pixels, wavelength in Angstrom. */
int ffs_apply_resolution_mask(ffs_ctx *ctx, float wavelength, float distance_m,
                              float beam_center_x_px, float beam_center_y_px,
                              float pixel_size_x_m, float pixel_size_y_m, float dmin, float dmax);
int ffs_get_mask(ffs_ctx *ctx, uint8_t *mask_out);
/* DispersionThreshold::threshold_w_gain (baseline/spotfinder/baseline.cpp:187-250,289-310): a
 * per-pixel gain map (width*height doubles, row-major) for the standard dispersion test:
 *     a = m*y - x*x;  c = gain*x*(m - 1 + nsig_b*sqrt(2(m-1)));  d = nsig_s*sqrt(gain*x*m).
 * With FFS_ALGO_DISPERSION_EXTENDED the map enters as in
 * DispersionExtendedThreshold::threshold_w_gain (baseline.cpp:484-541,657-714,772-805):
 *     first pass  a = m*y - x*x > c = gain*x*(m - 1 + nsig_b*sqrt(2(m-1)));
 *     final pass  src >= mean + nsig_s*sqrt(gain*mean).
 * The map is copied to the device; NULL removes it.  Not reachable from any reference
 * executable (there is no gain flag), provided for the library users of baseline.h. */
int ffs_set_gain(ffs_ctx *ctx, const double *gain); /* width*height doubles, row-major */
uint64_t ffs_mask_valid_count(ffs_ctx *ctx);

/* ---- device-resident (slot) API: the kernels with inputs already in HBM --------------- */
/* Copy n dense host frames into device slots first_slot.. (synchronous). */
int ffs_upload(ffs_ctx *ctx, int first_slot, int n, const void *host_frames);
/* Threshold only: slots -> strong bit-planes.  stream: a cudaStream_t passed as void*
 * (NULL = the context's own stream).  Asynchronous on that stream. */
int ffs_threshold(ffs_ctx *ctx, int first_slot, int n, void *stream);
/* Whole device path: threshold -> connected components -> spot reduction -> filters. */
int ffs_process(ffs_ctx *ctx, int first_slot, int n, void *stream);
/* Wait for everything queued on `stream` (or the context stream). */
int ffs_sync(ffs_ctx *ctx, void *stream);
/* Parity hooks (synchronous). */
int ffs_get_strong_bits(ffs_ctx *ctx, int slot, uint8_t *bits_out /* height*ceil(width/8) */);
int ffs_get_strong_u8(ffs_ctx *ctx, int slot, uint8_t *plane_out /* width*height, 0/1 */);
/* No counterpart in the reference.  The same result as
 * ffs_get_strong_u8 for callers that cannot call cudaMalloc (CUDA graph capture, tight serving
 * loops, harnesses that forbid it): the unpack kernel's scratch is a file-scope device array of
 * FFS_NOALLOC_MAX_PIXELS bytes (4.5 MiB, held by every GPU that loads the library until exit)
 * instead of a cudaMalloc'ed buffer, and the call makes no cudaMalloc / cudaFree.  width*height
 * above the maximum returns FFS_ERR_CAPACITY (ffs_last_error says so).  All calls on one GPU share
 * the array: calls that overlap on the GPU race (this entry point serialises its own callers). */
#define FFS_NOALLOC_MAX_PIXELS 4718592 /* 4.5 Mi pixels: Eiger2 4M (2070x2167), Jungfrau 4M (2068x2164) */
int ffs_get_strong_u8_static(ffs_ctx *ctx, int slot, uint8_t *plane_out /* width*height, 0/1 */);
size_t ffs_noalloc_max_pixels(void);
/* Extended algorithm intermediates: stage 1 = first-pass dispersion mask, stage 2 =
 * eroded plane (1 = valid pixel outside the signal core).  width*height bytes. */
int ffs_get_extended_stage_u8(ffs_ctx *ctx, int slot, int stage, uint8_t *plane_out);
/* Results of ffs_process for one slot.  spots may be NULL; at most max_spots written, in
 * canonical order, UNFILTERED (use the flags below or ffs_frame_result counts).
 * *n_spots receives the number of components. */
int ffs_fetch(ffs_ctx *ctx, int slot, ffs_frame_result *result, ffs_spot *spots,
              size_t max_spots, size_t *n_spots);
/* Canonical label per pixel (int64, -1 = not strong), for label parity. */
int ffs_get_labels(ffs_ctx *ctx, int slot, int64_t *labels_out /* width*height */);

/* ---- streaming API: host frames in, per-frame results out ------------------------------ */
/* The pinned-host ring of the reference's frame loop (spotfinder.cc:729-742), shared by
 * all producer threads.  ffs_acquire blocks until a ring slot is free and returns its
 * pinned buffer; the caller fills it (reader / decompressor output) and calls ffs_submit,
 * which queues H2D + kernels + the small result D2H on one of the context's streams.
 * ffs_wait blocks for the oldest submitted frame and returns its result.  Thread-safe. */
int ffs_acquire(ffs_ctx *ctx, int *slot, void **pinned_pixels);
int ffs_submit(ffs_ctx *ctx, int slot, uint64_t frame_id);
/* Hand an acquired slot back without submitting it (a producer that stops while it waits
 * for its frame, spotfinder.cc:789-791). */
int ffs_release(ffs_ctx *ctx, int slot);
/* Convenience: acquire + memcpy from a pageable buffer + submit. */
int ffs_submit_copy(ffs_ctx *ctx, uint64_t frame_id, const void *pixels);
/* spots/max_spots/n_spots as in ffs_fetch; returns FFS_ERR_STATE if nothing is pending. */
int ffs_wait(ffs_ctx *ctx, ffs_frame_result *result, ffs_spot *spots, size_t max_spots,
             size_t *n_spots);
int ffs_pending(ffs_ctx *ctx);

/* ---- compressed ingestion: bitshuffle-LZ4 decoded on the device ------------------------------
 * Replaces the host-side bshuf_decompress_lz4(buffer + 12, host_image, px, sizeof(pixel_t), 0)
 * + cudaMemcpy2DAsync of the dense frame (spotfinder/spotfinder.cc:823-846): the raw HDF5
 * filter chunk (12-byte header + body, exactly what Reader::get_raw_chunk returns,
 * h5read.h:173-204) is uploaded as it is and decoded by bslz4_decode_kernel, so 4-10x fewer
 * bytes cross the H2D link.  ffs_acquire_compressed returns the slot's pinned chunk buffer
 * and its capacity; ffs_submit_compressed walks the 4-byte block headers on the host (the
 * only serial step), queues H2D + decode + the same kernels as ffs_submit.  Results arrive
 * through ffs_wait; a malformed chunk gives result.status == FFS_ERR_FORMAT.
 * Supported: elem size = bytes_per_pixel, block size <= 8192 bytes (bitshuffle's default). */
int ffs_acquire_compressed(ffs_ctx *ctx, int *slot, void **pinned_chunk, size_t *capacity);
int ffs_submit_compressed(ffs_ctx *ctx, int slot, uint64_t frame_id, size_t chunk_bytes);
/* Both pinned buffers of an acquired slot (any pointer may be NULL): lets a caller fall back
 * to host decompression + ffs_submit when ffs_submit_compressed returns FFS_ERR_FORMAT
 * (e.g. a bitshuffle block size above 8192 bytes); the slot stays acquired in that case. */
int ffs_slot_buffers(ffs_ctx *ctx, int slot, void **pinned_pixels, void **pinned_chunk,
                     size_t *chunk_capacity);
/* Parity hook: decode a chunk into device slot `slot` (no further processing) and copy the
 * dense frame back to host memory. */
int ffs_decode_chunk_device(ffs_ctx *ctx, int slot, const void *chunk, size_t chunk_bytes,
                            void *frame_out);

/* Diagnostics (--writeout, spotfinder.cc:939-994 writes pixels_NNNNN.txt and image_NNNNN.png from
 * the strong-pixel mask it copied back): when switched on, ffs_wait also keeps the frame's
 * strong-pixel list (ascending linear indices y*width + x); *index stays valid until the next
 * ffs_wait on this context.  on = 2 also keeps the dense frame as the device saw it (after the
 * device decode, if any), for the diagnostic picture. */
int ffs_keep_pixel_lists(ffs_ctx *ctx, int on);

/* --validate (spotfinder.cc:1010-1052: the reference runs its CPU DIALS code on the frame a second
 * time and compares the strong pixels).  This library contains no CPU implementation, so the second
 * opinion is the device's own exact per-pixel routine: when switched on, ffs_wait decides every
 * pixel of the frame again from direct window sums (no streaming sums, no bounds, no pruning) into
 * a plane of its own and counts the pixels on which the two planes differ.  Standard dispersion
 * only (like the reference's --validate); costs a few ms per 16M frame.  ffs_last_self_check
 * reports the last ffs_wait frame: strong pixels found by the exact pass, differing pixels. */
int ffs_self_check(ffs_ctx *ctx, int on);
int ffs_last_self_check(ffs_ctx *ctx, uint32_t *n_strong_exact, uint32_t *n_mismatch);
/* The per-image timing block the reference prints with --threads 1 (spotfinder.cc:1055-1076: Copy /
 * Kernel / Post Copy): while on, every streamed frame is bracketed by CUDA events (and replayed
 * without its graph); ffs_last_frame_timing returns the times of the last ffs_wait frame in ms:
 * host->device copy (or upload + device decode), threshold + connected components, result copy. */
int ffs_frame_timing(ffs_ctx *ctx, int on);
int ffs_last_frame_timing(ffs_ctx *ctx, float *copy_ms, float *kernel_ms, float *post_copy_ms);
int ffs_last_pixel_list(ffs_ctx *ctx, const uint32_t **index, size_t *n);
int ffs_last_frame(ffs_ctx *ctx, const void **pixels, size_t *bytes);

/* ---- rotation sweeps: 3D connected components (find_3d_components) ---------------------- */
/* Frames processed with sweep recording on keep their strong-pixel lists on the device;
 * ffs_sweep_finish labels the whole stack (6-connectivity: index+1, index+width, same
 * index in the next slice, connected_components.cc:316-370), z = position of the frame
 * in ascending frame_id order (cc:409,441), and returns all components UNFILTERED in
 * canonical order.  The caller frees *spots with ffs_free. */
int ffs_sweep_begin(ffs_ctx *ctx, uint64_t max_frames);
/* Slot-API counterpart of the recording ffs_wait does: add the last ffs_process result of
 * `slot` to the sweep as frame `frame_id`. */
int ffs_sweep_record_slot(ffs_ctx *ctx, int slot, uint64_t frame_id);
int ffs_sweep_finish(ffs_ctx *ctx, ffs_spot **spots, size_t *n_spots);
/* Boundary exchange for slab-sharded sweeps (one slab of consecutive frames per GPU).
 * ffs_sweep_export_plane packs the strong pixels of the slab's first (which=0) or last
 * (which=1) frame as {uint32 linear index, uint32 local component id}; a neighbour rank
 * intersects it with its own boundary plane (ffs_sweep_match_plane) to get cross-slab
 * equivalence pairs.  Partial per-component sums are merged by the caller (see
 * ffs_b200.sweep in the Python host layer). */
int ffs_sweep_export_plane(ffs_ctx *ctx, int which, uint32_t **index_and_component,
                           size_t *n_pixels);
/* Exact integer sums of each 3D component of the last ffs_sweep_finish, in the same order:
 * 4 values per spot: sum I, sum (2x+1) I, sum (2y+1) I, sum (2z+1) I (z local to the slab).
 * These are what slab merging needs; centroids alone cannot be merged. */
int ffs_sweep_export_sums(ffs_ctx *ctx, uint64_t **sums, size_t *n_spots);

/* ---- slab-sharded sweeps across GPUs: native merge, NCCL transport (csrc/api_multi.inl) ------------
 * One rank (process or context) per GPU owns a contiguous slab of frames, streams it with sweep
 * recording on and labels it with ffs_sweep_finish.  Components continue across a slab boundary
 * only through identical linear indices of consecutive frames (connected_components.cc:352-370):
 * ffs_sweep_merge sends each slab's last plane to the next rank (ncclSend / ncclRecv), all-gathers
 * the boundary components' records and the equivalence pairs (ncclAllGather), and rebuilds the
 * merged components from their exact integer sums.  Each rank gets ITS share of the sweep's
 * components (those whose canonical root -- minimum (z, linear index) -- lies in its slab), z made
 * global, canonical order; the concatenation in rank order is the order find_3d_components gives
 * on one host (cc:292-309,396-399).  ffs_sweep_export_sums / _moments afterwards return the
 * merged values aligned with that list.  NCCL is loaded at run time (libnccl.so.2). */
typedef struct ffs_comm ffs_comm;
int ffs_comm_unique_id(uint8_t *id128);  /* ncclGetUniqueId: call on one rank, hand the 128 bytes to all */
int ffs_comm_create(const uint8_t *id128, int rank, int world, int device, ffs_comm **out);
void ffs_comm_destroy(ffs_comm *comm);
int ffs_comm_rank(const ffs_comm *comm);
int ffs_comm_world(const ffs_comm *comm);
const char *ffs_comm_last_error(const ffs_comm *comm); /* NULL: last ffs_comm_create error */
/* spots/n_spots: what ffs_sweep_finish returned on this rank (not modified); z0 = global index of
 * the slab's first frame, n_frames = frames in the slab.  *out_spots is freed with ffs_free.
 * exchange_seconds (optional): time spent in the NCCL exchange. */
int ffs_sweep_merge(ffs_ctx *ctx, ffs_comm *comm, uint64_t z0, uint64_t n_frames, const ffs_spot *spots,
                    size_t n_spots, ffs_spot **out_spots, size_t *n_out, double *exchange_seconds);
/* All ranks' final lists concatenated in rank order, on every rank (ncclAllGather). */
int ffs_sweep_gather(ffs_comm *comm, const ffs_spot *mine, size_t n, ffs_spot **all, size_t *n_all);
/* The same merge for slabs held in one address space (several contexts in one process, e.g.
 * `spotfinder --devices`): plain host arrays per slab -- spots, sums (4 per spot), moments (4 per
 * spot, may be NULL), first / last plane as {index, component} pairs -- no transport, no GPU.
 * Outputs (ffs_free): the whole sweep's components in canonical order with merged sums/moments. */
int ffs_sweep_merge_host(int n_slabs, uint32_t width, uint32_t height, const uint64_t *z0,
                         const uint64_t *n_frames, const ffs_spot *const *spots, const size_t *n_spots,
                         const uint64_t *const *sums, const uint64_t *const *moments,
                         const uint32_t *const *first, const size_t *n_first,
                         const uint32_t *const *last, const size_t *n_last, ffs_spot **out_spots,
                         uint64_t **out_sums, uint64_t **out_moments, size_t *n_out);

/* ---- per-spot variances in Kabsch space (Reflection3D::variances_in_kabsch_space,
 *      connected_components.cc:159-203; caller loop spotfinder.cc:1152-1216) ----------------
 * Second moments of each 3D component of the last ffs_sweep_finish, same order, 4 per spot,
 * taken relative to the spot's own bounding-box minimum (x_min, y_min, z_min) so that they
 * stay small and exact:  sum I u^2, sum I v^2, sum I u v, sum I w^2   with
 * u = x - x_min, v = y - y_min, w = z - z_min.  Together with ffs_sweep_export_sums they are
 * everything the variances need; like the sums they can be merged across slabs. */
int ffs_sweep_export_moments(ffs_ctx *ctx, uint64_t **moments, size_t *n_spots);

/* The experiment description the reference builds at spotfinder.cc:1157-1168: a dx2 Panel
 * from distance / beam centre / pixel size with fast = +x, slow = -y (dx2 detector.cxx:243-268,
 * no parallax correction: thickness = mu = 0), s0 = (0, 0, -1/wavelength), rotation axis +x,
 * oscillation in degrees. */
typedef struct ffs_kabsch_geometry {
    double distance_mm;
    double beam_center_x_px, beam_center_y_px;
    double pixel_size_x_mm, pixel_size_y_mm;
    double wavelength;
    double oscillation_width_deg;
} ffs_kabsch_geometry;

/* For every spot: sigma_b_variance = (var_e1 + var_e2) / 2, sigma_m_variance = var_e3,
 * bbox_depth = z_max - z_min + 1 (the tuple variances_in_kabsch_space returns), evaluated on
 * the device from {spots, sums, moments} (host arrays of n, 4n and 4n entries; any output
 * pointer may be NULL).  s1 and phi are taken at the spot's centre of mass as the caller
 * loop does.  Spots with zero total intensity give NaN, as the reference's 0/0 does. */
int ffs_kabsch_variances(ffs_ctx *ctx, const ffs_kabsch_geometry *geometry, const ffs_spot *spots,
                         const uint64_t *sums, const uint64_t *moments, size_t n,
                         double *sigma_b_variance, double *sigma_m_variance, int32_t *bbox_depth);
void ffs_free(void *p);

/* filter_reflections (connected_components.cc:207-236): stable, in place; returns kept. */
size_t ffs_filter_spots(ffs_spot *spots, size_t n, uint32_t min_spot_size,
                        float max_peak_centroid_separation, size_t *n_removed_size,
                        size_t *n_removed_separation);

/* ---- instrumentation --------------------------------------------------------------------- */
/* Number of kernels this library has launched since ffs_create (all streams). */
uint64_t ffs_kernel_launches(const ffs_ctx *ctx);
/* Device pitch (bytes) of a frame row / bytes of one device frame / of one bit-plane. */
/* Per-kernel timing for roofline reporting: while enabled, the threshold kernel of every launch
 * (fused_dispersion_kernel; for dispersion_extended the first pass + erosion + final pass
 * together) is bracketed by CUDA events on the stream it is launched on.
 * ffs_profile_read waits for the recorded events, returns their summed duration and the
 * number of bracketed launches, and resets both. */
int ffs_profile_enable(ffs_ctx *ctx, int on);
int ffs_profile_read(ffs_ctx *ctx, double *total_ms, uint64_t *n_launches);
size_t ffs_frame_pitch(const ffs_ctx *ctx);
size_t ffs_frame_bytes_device(const ffs_ctx *ctx);

/* ---- ingestion helper ----------------------------------------------------------------------- */
/* bshuf_decompress_lz4(chunk + 12, out, n, elem_size, 0) as the reference frame loop calls it
 * (spotfinder/spotfinder.cc:830-834): decode one bitshuffle-LZ4 HDF5-filter chunk (12-byte
 * header + body) into n dense elements (host memory, e.g. a pinned ffs_acquire slot). */
int ffs_bslz4_decode_chunk(const void *chunk, size_t chunk_bytes, void *out, size_t n_elements,
                           size_t elem_size);

/* ---- baseline.h-compatible shim (baseline/spotfinder/baseline.h:9-13) -------------------- */
/* image_t as h5read/include/h5read.h:34-39 declares it; image_t_type is uint16_t here
 * (uint32_t when the caller was built with PIXEL_DATA_32BIT: use the *32 entry points). */
#ifdef __cplusplus
typedef bool ffs_bool;
#else
typedef _Bool ffs_bool;
#endif
typedef struct ffs_image_t {
    void *data;    /* image_t_type*  */
    uint8_t *mask; /* 1 = valid      */
    size_t slow;
    size_t fast;
} ffs_image_t;
void *spotfinder_create(size_t width, size_t height);
void spotfinder_free(void *context);
/* Returns the number of strong pixels; *destination borrows context-owned storage
 * (width*height bool) valid until the next call on the context. */
uint32_t spotfinder_standard_dispersion(void *context, ffs_image_t *image, ffs_bool **destination);
void *spotfinder_create32(size_t width, size_t height);
uint32_t spotfinder_standard_dispersion32(void *context, ffs_image_t *image, ffs_bool **destination);

#ifdef __cplusplus
}
#endif
#endif /* FFS_B200_H */
