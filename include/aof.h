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
	int search_size;                      /* +-px; 4, 6 or 8 */
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
 * may be NULL. */
int aof_px4_calc_flow_device(aof_px4 *h, const uint8_t *d_imgs, const uint32_t *d_t_us, aof_flow_out *d_out,
			     aof_px4_raw *d_raw, aof_px4_block *d_blocks, void *cuda_stream);

/* stateless PX4Flow::compute_flow (src/px4flow.cpp:349-527) over n pairs held in
 * HOST memory; raw[n]; blocks may be NULL or n*AOF_PX4_MAX_BLOCKS entries. */
int aof_px4_compute_flow_pairs(aof_px4 *h, const uint8_t *prev, const uint8_t *cur, int n, aof_px4_raw *raw,
			       aof_px4_block *blocks);
/* same with DEVICE pointers, asynchronous on cuda_stream */
int aof_px4_compute_flow_pairs_device(aof_px4 *h, const uint8_t *d_prev, const uint8_t *d_cur, int n,
				      aof_px4_raw *d_raw, aof_px4_block *d_blocks, void *cuda_stream);

/* forget previous frames and limiter state (as a freshly constructed object) */
int aof_px4_reset(aof_px4 *h);

/* number of kernel launches this library has issued in this process */
uint64_t aof_launch_count(void);

#ifdef __cplusplus
}
#endif
#endif /* AOF_B200_H */
