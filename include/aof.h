/*
 * aof.h -- C ABI of libaof_b200.so: the B200 (sm_100a) implementation of the
 * frame-to-frame calcFlow() hot path of aero-optical-flow's modules/OpticalFlow.
 *
 * Plain C: opaque handles, POD structs, pointers and sizes.  No exceptions, no
 * STL, no torch types cross this boundary.  Every entry point names the
 * reference interface it replaces (paths relative to
 * /root/reference/modules/OpticalFlow/).
 *
 * Conventions
 *   - "stream" = one independent camera / frame sequence with its own state
 *     (previous frame or pyramid, feature slots, rate-limiter sums).  A handle
 *     owns n_streams of them on ONE device; calls on one handle are serialised
 *     by the caller, different handles may be used concurrently.
 *   - calc functions return the reference's value (quality 0..255, or -1 = "do
 *     not publish yet", src/optical_flow.cpp:92); batch functions return AOF_OK
 *     and write one quality per stream.  Failures are <= AOF_ERR_BASE.
 *   - images are 8-bit luma, row-major, stride == width (src/mainloop.cpp:317).
 *   - there is NO CPU fallback: without a CUDA device every create fails with
 *     AOF_ERR_CUDA / AOF_ERR_NO_DEVICE.
 */
#ifndef AOF_B200_H
#define AOF_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define AOF_OK 0
#define AOF_ERR_BASE (-100)
#define AOF_ERR_INVALID (-100)      /* bad argument */
#define AOF_ERR_UNSUPPORTED (-101)  /* geometry outside what the kernels cover */
#define AOF_ERR_CUDA (-102)         /* a CUDA call failed; see aof_last_error() */
#define AOF_ERR_NO_DEVICE (-103)
#define AOF_ERR_STATE (-104)        /* call not valid for this handle */

#define AOF_PX4_MAX_BLOCKS 25

const char *aof_version(void);
/* thread-local text of the last failure (CUDA error string etc.) */
const char *aof_last_error(void);
int aof_device_count(void);

/* ------------------------------------------------------------------ SAD path */

/* ctor arguments of OpticalFlowPX4 (include/flow_px4.hpp:62-65; defaults
 * :47-49 and include/optical_flow.hpp:47-49). */
typedef struct aof_px4_params {
	float focal_length_x, focal_length_y; /* px */
	int output_rate;                      /* Hz; <= 0: publish every frame */
	int image_width, image_height;
	int search_size;                      /* +-px; 1..127 (4, 6, 8 have specialised kernels) */
	int flow_feature_threshold;           /* texture gate (30) */
	int flow_value_threshold;             /* SAD acceptance (3000) */
} aof_px4_params;

void aof_px4_default_params(aof_px4_params *p);

/* what one calcFlow call yields for one stream (16 B, the "host gather" unit) */
typedef struct aof_flow_out {
	int32_t quality; /* 0..255, -1 = rate-limited */
	int32_t dt_us;   /* valid when quality >= 0 (or always if output_rate <= 0) */
	float flow_x;    /* rad */
	float flow_y;
} aof_flow_out;

/* raw per-pair result of PX4Flow::compute_flow (src/px4flow.cpp:349) */
typedef struct aof_px4_raw {
	float px_x, px_y; /* pixel flow */
	int32_t quality;  /* 0..255 */
	int32_t count;    /* accepted blocks (meancount) */
} aof_px4_raw;

/* per-block diagnostics, AOF_PX4_MAX_BLOCKS per pair, raster order over the grid
 * (compute_diff :114, compute_sad_8x8 :300, compute_subpixel :155) */
typedef struct aof_px4_block {
	int16_t x, y;       /* block origin */
	int16_t dx, dy;     /* arg-min displacement */
	uint32_t texture;   /* compute_diff value */
	uint32_t best_sad;  /* 0xFFFFFFFF if the texture gate rejected the block */
	uint32_t acc[8];    /* half-pixel SADs (valid when accepted) */
	int32_t subdir;     /* 0..7, 8 = none */
	int32_t accepted;   /* best_sad < flow_value_threshold */
} aof_px4_block;

typedef struct aof_px4 aof_px4;

/* replaces OpticalFlowPX4::OpticalFlowPX4 (src/flow_px4.cpp:47-62) for
 * n_streams independent instances resident on CUDA device `device`. */
int aof_px4_create(const aof_px4_params *p, int n_streams, int device, aof_px4 **out);
void aof_px4_destroy(aof_px4 *h);
int aof_px4_num_streams(const aof_px4 *h);

/* replaces OpticalFlowPX4::calcFlow (src/flow_px4.cpp:69-98) for a 1-stream
 * handle.  Synchronous; `img` may be reused as soon as it returns.  First call
 * returns 0 and leaves the outputs untouched. */
int aof_px4_calc_flow(aof_px4 *h, const uint8_t *img, uint32_t img_time_us, int *dt_us, float *flow_x,
		      float *flow_y);

/* the same call for all n_streams at once: imgs = n_streams frames back to back
 * in HOST memory (pageable or pinned), t_us[n_streams]; out[n_streams].  Staging
 * to pinned memory, the async H2D copy, the kernels and the 16 B/stream D2H
 * gather all happen inside.  Streams that are on their first frame report
 * quality 0 with dt/flow = 0. */
int aof_px4_calc_flow_batch(aof_px4 *h, const uint8_t *imgs, const uint32_t *t_us, aof_flow_out *out);

/* device-resident form: d_imgs / d_t_us / d_out are DEVICE pointers on the
 * handle's device; launched on `cuda_stream` (a cudaStream_t, NULL = default
 * stream) without synchronising.  d_imgs must stay valid until the NEXT call
 * returns (it becomes the previous frame; no copy is made).  d_raw / d_blocks
 * may be NULL.  d_imgs must be 16-byte aligned (the frames are staged with bulk
 * copies of the TMA engine, which need 16-byte aligned sources; width * height is
 * a multiple of 16 by construction): a misaligned pointer is AOF_ERR_INVALID.
 * One CUDA stream per handle: successive calls must be issued on the same stream
 * (or be ordered by the caller), because they read and write handle-owned state. */
int aof_px4_calc_flow_device(aof_px4 *h, const uint8_t *d_imgs, const uint32_t *d_t_us, aof_flow_out *d_out,
			     aof_px4_raw *d_raw, aof_px4_block *d_blocks, void *cuda_stream);

/* stateless PX4Flow::compute_flow (src/px4flow.cpp:349-527) over n pairs held in
 * HOST memory; raw[n]; blocks may be NULL or n*AOF_PX4_MAX_BLOCKS entries. */
int aof_px4_compute_flow_pairs(aof_px4 *h, const uint8_t *prev, const uint8_t *cur, int n, aof_px4_raw *raw,
			       aof_px4_block *blocks);
/* same with DEVICE pointers (16-byte aligned, see above), asynchronous on cuda_stream */
int aof_px4_compute_flow_pairs_device(aof_px4 *h, const uint8_t *d_prev, const uint8_t *d_cur, int n,
				      aof_px4_raw *d_raw, aof_px4_block *d_blocks, void *cuda_stream);

/* forget previous frames and limiter state (as a freshly constructed object) */
int aof_px4_reset(aof_px4 *h);

/* The flow of n independent (previous, current) pairs WITHOUT ANY ALLOCATION: no handle, no cudaMalloc / cudaFree, no
 * allocating library routine -- for callers that cannot allocate (CUDA graph capture, tight serving loops, harnesses
 * that forbid it).  No counterpart in the reference.
 * out[i] is what a freshly constructed OpticalFlowPX4(p) returns from its SECOND calcFlow call when fed prev[i] and
 * then cur[i] at time t_us[i] (= aof_px4_calc_flow_device on a new handle, called with prev, then with cur): quality,
 * dt_us, flow_x, flow_y under p->output_rate.  The scratch between the two kernels (16 B of raw flow + 24 B of limiter
 * state per pair) lives in one fixed __device__ array of AOF_STATIC_MAX_PAIRS pairs (320 KiB, held by every GPU that
 * loads the library until exit): n beyond that is AOF_ERR_INVALID.  All calls on a GPU share the array, so calls that
 * overlap on the GPU (two streams / two threads on one device) race -- serialise them.  d_prev / d_cur: n frames each,
 * DEVICE pointers, 16-byte aligned; d_t_us, d_out: n entries, DEVICE pointers; asynchronous on cuda_stream. */
#define AOF_STATIC_MAX_PAIRS 8192
int aof_px4_static_max_pairs(void);
int aof_px4_calc_flow_pairs_static(const aof_px4_params *p, int device, const uint8_t *d_prev, const uint8_t *d_cur,
				   const uint32_t *d_t_us, int n, aof_flow_out *d_out, void *cuda_stream);

/* replaces OpticalFlow::setFocalLengthX / setFocalLengthY / setOutputRate (include/optical_flow.hpp:76-78) on a
 * live handle: plain assignments, as in the reference -- previous frames, limiter sums and the time of the last
 * publish are kept; the new values apply from the next calc call, to all streams. */
int aof_px4_set_params(aof_px4 *h, float focal_length_x, float focal_length_y, int output_rate);

/* ------------------------------------------------------------------- LK path */

/* ctor arguments of OpticalFlowOpenCV (include/flow_opencv.hpp:81-83; defaults
 * :53-54 and include/optical_flow.hpp:47-49). */
typedef struct aof_lk_params {
	float focal_length_x, focal_length_y; /* px */
	int output_rate;                      /* Hz; <= 0: publish every frame */
	int image_width, image_height;
	int num_features;                     /* 1..100 (trackFeatures.cpp:56) */
	float confidence_multiplier;          /* 1.645 */
} aof_lk_params;

void aof_lk_default_params(aof_lk_params *p);

/* per-call diagnostics of one stream (the values flow_opencv.cpp:121-202 derives) */
typedef struct aof_lk_raw {
	float px_x, px_y;  /* mean pixel flow after the confidence gate */
	int32_t inliers;   /* meancount after the gate */
	int32_t active;    /* features with status 1 before the gate */
	int32_t fast_bins; /* bins FAST ran on in initMorePoints */
	int32_t quality;   /* round(255 * inliers / N) before limitRate */
} aof_lk_raw;

typedef struct aof_lk aof_lk;

/* replaces OpticalFlowOpenCV::OpticalFlowOpenCV (src/flow_opencv.cpp:47-65) for
 * n_streams independent instances on CUDA device `device`.  The tracker state
 * that the reference keeps in file statics (trackFeatures.cpp:53-56) is per stream. */
int aof_lk_create(const aof_lk_params *p, int n_streams, int device, aof_lk **out);
void aof_lk_destroy(aof_lk *h);
int aof_lk_num_streams(const aof_lk *h);
int aof_lk_num_features(const aof_lk *h);
/* diagnostics of the lazy FAST path since create / reset (synchronises the device): out[0] = bins whose picks were found
 * without scoring every pixel, out[1] = candidate-pool rebuilds, out[2] = bins that overflowed the lazy buffers,
 * out[3] = stream-frames handed to the complete path (second-chance pass of initMorePoints).  All zero when the
 * shape or AOF_FAST_LAZY=0 selects the complete path. */
int aof_lk_lazy_stats(aof_lk *h, uint64_t out[4]);
int aof_lk_reset(aof_lk *h);

/* replaces OpticalFlowOpenCV::calcFlow (src/flow_opencv.cpp:89-210), 1-stream
 * handle, synchronous.  Always writes flow_x/flow_y; dt_us only when it
 * publishes (return >= 0) or output_rate <= 0. */
int aof_lk_calc_flow(aof_lk *h, const uint8_t *img, uint32_t img_time_us, int *dt_us, float *flow_x, float *flow_y);

/* all n_streams at once from HOST frames (back to back); out[n_streams]; raw may be NULL */
int aof_lk_calc_flow_batch(aof_lk *h, const uint8_t *imgs, const uint32_t *t_us, aof_flow_out *out, aof_lk_raw *raw);

/* device-resident form, asynchronous on cuda_stream.  d_imgs must stay valid
 * and unchanged until the NEXT call has completed (it is level 0 of the previous
 * pyramid; no copy is made).  d_raw may be NULL.  One CUDA stream per handle:
 * successive calls (and aof_lk_calc_flow_batch, which uses the handle's own stream)
 * read and write handle-owned state, so issue them on the same stream or order them. */
int aof_lk_calc_flow_device(aof_lk *h, const uint8_t *d_imgs, const uint32_t *d_t_us, aof_flow_out *d_out,
			    aof_lk_raw *d_raw, void *cuda_stream);

/* replaces setFocalLengthX / setFocalLengthY / setOutputRate (include/optical_flow.hpp:76-78) and
 * OpticalFlowOpenCV::setConfMultiplier (include/flow_opencv.hpp:73) on a live handle: plain assignments, as in the
 * reference -- tracker state, feature slots, limiter sums and the time of the last publish are kept. */
int aof_lk_set_params(aof_lk *h, float focal_length_x, float focal_length_y, int output_rate,
		      float confidence_multiplier);

/* replace OpticalFlowOpenCV::setCameraMatrix / setCameraDistortion (src/flow_opencv.cpp:72-87; note the
 * reference's argument order k1, k2, k3, p1, p2).  As in the reference, feature positions are undistorted
 * (cv::undistortPoints + re-projection, :108-119) only once BOTH have been called; applies to all streams. */
int aof_lk_set_camera_matrix(aof_lk *h, float focal_len_x, float focal_len_y, float principal_point_x,
			     float principal_point_y);
int aof_lk_set_camera_distortion(aof_lk *h, float k1, float k2, float k3, float p1, float p2);

/* state read-back for tests: per stream N entries of updateVector, features_current
 * (x, y), prev_corners (x, y) and prev_status.  Any pointer may be NULL. */
int aof_lk_get_state(aof_lk *h, int32_t *update_vec, float *feat_cur, float *prev_corners, uint8_t *prev_status);

/* kernel-level entry points (known-answer tests; all HOST pointers, synchronous) */
/* levels of cv::buildOpticalFlowPyramid for n frames: returns the number of levels L,
 * writes level sizes to lw[4]/lh[4] and, if levels != NULL, the levels of frame i
 * back to back at levels + i * (sum of level sizes). */
int aof_lk_pyramid(aof_lk *h, const uint8_t *imgs, int n, int *lw, int *lh, uint8_t *levels);
/* cv::calcOpticalFlowPyrLK(prev, cur, pts, Size(9,9), 3) for n frame pairs with n_pts points each */
int aof_lk_track_points(aof_lk *h, const uint8_t *prev, const uint8_t *cur, int n, const float *pts, int n_pts,
			float *next_pts, uint8_t *status);
/* cv::FAST(bin ROI, 10, nonmax) on all 16 bins of n frames: score[n * w * h], 0 = no keypoint */
int aof_lk_fast_scores(aof_lk *h, const uint8_t *imgs, int n, uint8_t *score);

/* ------------------------------------------------------------------ multi-GPU */

/* Both paths shard by whole streams: a stream's state is private and nothing is reduced across streams
 * (src/flow_px4.cpp:69-98, src/flow_opencv.cpp:89-210), so the streams of a batch are partitioned over the GPUs of
 * one box, each GPU owning a contiguous range of whole streams, with a host-side gather of the 16 B / stream results
 * and no collective (no NCCL, no peer access).  A multi handle owns, per device, one single-device handle and one host
 * worker thread (bound to the CPUs local to that GPU when the kernel exposes them); a batch call hands every worker
 * its slice of the host frames, the worker runs aof_*_calc_flow_batch on its device (pinned staging, chunked
 * asynchronous H2D under the kernels of the previous chunk) and its D2H lands in its slice of ONE pinned result
 * array: that array is the gather.  devices[] are CUDA ordinals, all different; device i owns streams
 * [first_i, first_i + count_i) with count_i = n_streams_total / n_devices (+ 1 for the first n_streams_total %
 * n_devices devices).  Calls on one multi handle are serialised by the caller. */
typedef struct aof_px4_multi aof_px4_multi;
typedef struct aof_lk_multi aof_lk_multi;

int aof_px4_create_multi(const aof_px4_params *p, int n_streams_total, const int *devices, int n_devices, aof_px4_multi **out);
void aof_px4_destroy_multi(aof_px4_multi *m);
/* imgs: n_streams_total frames back to back in HOST memory, t_us[n_streams_total].  out may be NULL (read the gathered
 * results in place with aof_px4_multi_results) or n_streams_total entries, which receive a copy. */
int aof_px4_multi_calc_flow_batch(aof_px4_multi *m, const uint8_t *imgs, const uint32_t *t_us, aof_flow_out *out);
const aof_flow_out *aof_px4_multi_results(const aof_px4_multi *m);
int aof_px4_multi_range(const aof_px4_multi *m, int i, int *device, int *first, int *count);
int aof_px4_multi_set_params(aof_px4_multi *m, float focal_length_x, float focal_length_y, int output_rate);

int aof_lk_create_multi(const aof_lk_params *p, int n_streams_total, const int *devices, int n_devices, aof_lk_multi **out);
void aof_lk_destroy_multi(aof_lk_multi *m);
/* raw may be NULL, aof_lk_multi_raw(m) itself (diagnostics gathered in place) or n_streams_total entries (copy) */
int aof_lk_multi_calc_flow_batch(aof_lk_multi *m, const uint8_t *imgs, const uint32_t *t_us, aof_flow_out *out, aof_lk_raw *raw);
const aof_flow_out *aof_lk_multi_results(const aof_lk_multi *m);
const aof_lk_raw *aof_lk_multi_raw(const aof_lk_multi *m);
int aof_lk_multi_range(const aof_lk_multi *m, int i, int *device, int *first, int *count);
int aof_lk_multi_set_params(aof_lk_multi *m, float focal_length_x, float focal_length_y, int output_rate,
			    float confidence_multiplier);
/* state read-back for tests, all streams in global stream order (see aof_lk_get_state) */
int aof_lk_multi_get_state(aof_lk_multi *m, int32_t *update_vec, float *feat_cur, float *prev_corners, uint8_t *prev_status);

/* ------------------------------------------------------------ dense SAD field */

/* Per-block flow field over every 8x8 block of the 8-px lattice whose search window (+ the 1-px
 * sub-pixel halo) lies inside the image: the reference's per-block primitives compute_diff
 * (src/px4flow.cpp:114), compute_sad_8x8 (:300, search order :394-405) and compute_subpixel (:155,
 * direction choice :413-422) with 32-bit offsets, for images the reference itself cannot address
 * (BASELINE config 4: 1280x720, +-8). */
typedef struct aof_dense_params {
	int image_width, image_height; /* width a multiple of 4 */
	int search_size;               /* 4, 6 or 8 */
	int flow_feature_threshold;    /* texture gate (30) */
	int flow_value_threshold;      /* SAD acceptance (3000) */
} aof_dense_params;

typedef struct aof_dense_block {
	int16_t dx, dy;    /* arg-min displacement (valid when flags & 1) */
	int16_t subdir;    /* half-pixel direction 0..7, 8 = none */
	int16_t flags;     /* 1 = texture gate passed (searched), 2 = accepted (best_sad < value threshold) */
	uint32_t best_sad; /* 0xFFFFFFFF when not searched */
	uint32_t texture;  /* compute_diff */
} aof_dense_block;

typedef struct aof_sad_dense aof_sad_dense;

int aof_sad_dense_create(const aof_dense_params *p, int device, aof_sad_dense **out);
void aof_sad_dense_destroy(aof_sad_dense *h);
/* lattice: nbx x nby blocks, block (i, j) has its origin at (x0 + 8 i, y0 + 8 j) */
int aof_sad_dense_grid(const aof_sad_dense *h, int *nbx, int *nby, int *x0, int *y0);
/* n_pairs frame pairs in HOST memory -> field[n_pairs * nby * nbx], synchronous */
int aof_sad_dense_compute(aof_sad_dense *h, const uint8_t *prev, const uint8_t *cur, int n_pairs, aof_dense_block *field);
/* DEVICE pointers (4-byte aligned), asynchronous on cuda_stream */
int aof_sad_dense_compute_device(aof_sad_dense *h, const uint8_t *d_prev, const uint8_t *d_cur, int n_pairs,
				 aof_dense_block *d_field, void *cuda_stream);

/* ------------------------------------------------- Shi-Tomasi / goodFeaturesToTrack */

/* The corner selector north_star names (the reference selects with FAST):
 * cv::cornerMinEigenVal(img, blockSize 3, Sobel 3) and cv::goodFeaturesToTrack(img, maxCorners,
 * qualityLevel, minDistance) on n frames of width x height (HOST pointers, synchronous). */
typedef struct aof_gftt aof_gftt;
int aof_gftt_create(int width, int height, int device, aof_gftt **out);
void aof_gftt_destroy(aof_gftt *g);
/* eig[n * height * width] */
int aof_gftt_min_eig(aof_gftt *g, const uint8_t *imgs, int n, float *eig);
/* xy[n * max_corners * 2] (x, y), counts[n]; corners in cv's order (response desc, address desc) */
int aof_gftt_detect(aof_gftt *g, const uint8_t *imgs, int n, int max_corners, double quality_level, double min_distance,
		    float *xy, int *counts);

/* ------------------------------------------- callers / consumers either side of the path */

/* (f1) centre crop + staging.  The reference wraps the camera buffer in a Mat, takes the centre ROI of
 * image_width x image_height and copies it to a contiguous Mat before calcFlow (src/mainloop.cpp:294-298,317-320).
 * Here the crop is the DMA itself: one strided host->device copy per camera frame, straight from the camera buffer
 * (register page-aligned V4L2 USERPTR buffers once with aof_host_register, src/camera.cpp:97-115) into the contiguous
 * device batch the calc_flow_device entry points take.  frames[i] = top-left of camera frame i (cam_w x cam_h, stride
 * cam_w); d_batch receives n crops of w x h.  Asynchronous on cuda_stream. */
int aof_host_register(void *ptr, size_t bytes);
int aof_host_unregister(void *ptr);
int aof_crop_to_device(const uint8_t *const *frames, int n, int cam_w, int cam_h, int w, int h, uint8_t *d_batch,
		       void *cuda_stream);

/* (f3) auto-exposure statistics of Mainloop::_exposure_update (src/mainloop.cpp:197-220): 10-bin histogram
 * (cv::calcHist, range [0, 255), i.e. value 255 is not counted) of the central 128 x 128 pixels of each frame and the
 * mean sample value msv = sum_i (i + 1) * hist[i] / 16384 accumulated in float in that order. */
typedef struct aof_ae_stats {
	uint32_t hist[10];
	float msv;
	uint32_t pad;
} aof_ae_stats;
/* n frames of w x h (w, h >= 128) in DEVICE memory -> d_out[n], asynchronous on cuda_stream */
int aof_ae_stats_device(const uint8_t *d_frames, int n, int w, int h, aof_ae_stats *d_out, void *cuda_stream);
/* HOST frames, synchronous */
int aof_ae_stats_host(const uint8_t *frames, int n, int w, int h, int device, aof_ae_stats *out);

/* (f4) result wire format: the OPTICAL_FLOW_RAD message (id 106, 44-byte payload, CRC_EXTRA 138;
 * modules/mavlink_c/common/mavlink_msg_optical_flow_rad.h:4-22) the reference fills at src/mainloop.cpp:359-371 and
 * frames with mavlink_msg_optical_flow_rad_encode + mavlink_msg_to_send_buffer (src/mavlink_tcp.cpp:142-149), i.e. a
 * MAVLink 2 frame: STX 0xFD, payload length after trimming trailing zero bytes, incompat 0, compat 0, seq, sysid,
 * compid, 24-bit msgid, payload, X.25 CRC over everything after STX plus CRC_EXTRA.
 * For each stream i with res[i].quality >= 0 (the reference returns early on -1, src/mainloop.cpp:328) one frame is
 * appended to buf: time_usec = time_offset_us + img_time_us[i], integration_time_us = dt_us, integrated_x / y = flow,
 * integrated_xgyro = -gyro[3i+1], integrated_ygyro = gyro[3i], integrated_zgyro = gyro[3i+2], distance = -1,
 * everything else 0, quality as is; seq = (*seq)++ per frame.  frame_off[i] = offset of stream i's frame in buf or -1.
 * Returns the number of bytes written (<= 56 per frame), or AOF_ERR_INVALID if buf_cap is too small.  Host-side. */
int aof_pack_optical_flow_rad(const aof_flow_out *res, int n, uint64_t time_offset_us, const uint32_t *img_time_us,
			      const double *gyro_xyz, uint8_t system_id, uint8_t component_id, uint8_t *seq,
			      uint8_t *buf, int buf_cap, int *frame_off);

/* optional per-kernel timing (CUDA events around every launch of this library; off by default).
 * aof_timing_read(k, ...) for k in [0, aof_timing_kernels()) waits for the recorded launches and
 * returns the kernel's name, its summed device time and its launch count since the last reset. */
void aof_timing_enable(int on);
void aof_timing_reset(void);
int aof_timing_kernels(void);
int aof_timing_read(int kernel, const char **name, double *total_ms, uint64_t *launches);

/* number of kernel launches this library has issued in this process */
uint64_t aof_launch_count(void);

#ifdef __cplusplus
}
#endif
#endif /* AOF_B200_H */
