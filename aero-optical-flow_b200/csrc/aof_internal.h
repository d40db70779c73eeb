// Internal declarations shared by the kernels (.cu) and the C-ABI host code.
#pragma once
#include <climits>
#include <cstdint>
#include <cuda_runtime.h>
#include <utility>

#include "../../include/aof.h"

#if defined(__CUDACC__)
#define AOF_HD __host__ __device__ __forceinline__
#else
#define AOF_HD inline
#endif

namespace aof {

// geometry of the SAD grid (src/px4flow.cpp:359-361) + thresholds
struct SadGeom {
	int width;             // pixels per row == row pitch (multiple of 4)
	uint32_t stage_bytes;  // bytes of each frame brought to shared memory (multiple of 16)
	size_t frame_stride;   // bytes between consecutive frames of the batch in global memory
	int nb;                // grid positions per axis (<= 5)
	int lo, step;          // first position, spacing
	uint32_t feat_thr, val_thr;
	// A/B and grid-shaping switches, read from the environment ONCE when the handle is created (env_int)
	int tun_v1;            // AOF_SAD_V1 = 1: the CTA-per-pair kernel
	int tun_v2;            // AOF_SAD_V2 = 1: never sad_flow3_kernel
	int tun_ctas;          // AOF_SAD_CTAS: cap on resident CTAs per SM (0 = occupancy)
};

// integer environment switch, read where a handle is created (never on a launch path)
int env_int(const char *name, int dflt);

// per-stream OpticalFlow rate limiter (include/optical_flow.hpp:61-64 plus the
// function-static time_last_pub of src/optical_flow.cpp:55, made per stream)
struct LimiterState {
	float sum_x, sum_y;
	int sum_q, n_valid;
	uint32_t t_last;
	uint32_t pad;
};

// src/optical_flow.cpp:51-96
AOF_HD int limit_rate(LimiterState &l, int output_rate, int quality, uint32_t t_us, int32_t &dt_us, float &fx,
		      float &fy)
{
	if (output_rate <= 0) {
		dt_us = (int32_t)(t_us - l.t_last);
		l.t_last = t_us;
		return quality;
	}
	if (quality > 0) {
		l.sum_x += fx; l.sum_y += fy; l.sum_q += quality; l.n_valid++;
	}
	if ((float)(uint32_t)(t_us - l.t_last) > 1.0e6f / (float)output_rate) {
		int q = 0;
		if (l.n_valid > 0) q = l.sum_q / l.n_valid;
		fx = l.sum_x; fy = l.sum_y;
		l.sum_x = 0.f; l.sum_y = 0.f; l.sum_q = 0; l.n_valid = 0;
		dt_us = (int32_t)(t_us - l.t_last);
		l.t_last = t_us;
		return q;
	}
	dt_us = 0;
	return -1;
}

void count_launch(uint64_t n = 1);

// kernel ids of the optional per-kernel event timing (aof_timing_* in include/aof.h)
enum KernelId {
	KID_SAD_FLOW = 0, KID_PX4_EPILOGUE, KID_LK_PYR_DOWN, KID_LK_TRACK_UPDATE, KID_LK_FAST, KID_LK_SELECT_EPILOGUE,
	KID_LK_TRACK_POINTS, KID_SAD_DENSE, KID_GFTT_MIN_EIG, KID_GFTT_SELECT, KID_LK_PYR_FUSED, KID_LK_FAST_PICK, KID_LK_FAST_FLAGGED, KID_LK_FAST_PLAN, KID_COUNT
};

// brackets one launch with two events when timing is enabled; a no-op otherwise
bool timing_on(); // per-kernel event timing requested (aof_timing_enable)

// kernel<<<grid, block, smem, st>>>(args...), optionally with programmatic stream serialisation (the kernel must call
// pdl_enter() before it touches global memory)
template <typename... Exp, typename... Act>
inline cudaError_t launch_k(bool pdl, void (*kernel)(Exp...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Act &&... args)
{
	cudaLaunchConfig_t cfg = {};
	cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
	cudaLaunchAttribute at[1];
	at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
	at[0].val.programmaticStreamSerializationAllowed = 1;
	cfg.attrs = at; cfg.numAttrs = pdl ? 1 : 0;
	return cudaLaunchKernelEx(&cfg, kernel, std::forward<Act>(args)...);
}

class KernelTimer {
public:
	// `symbol`: the __global__ function this launch runs (what a profiler lists); aof_timing_read reports it for the slot
	KernelTimer(int kid, cudaStream_t st, const char *symbol = nullptr);
	~KernelTimer();
private:
	int kid_;
	cudaStream_t st_;
	void *a_ = nullptr, *b_ = nullptr;
	bool active_ = false;
};

cudaError_t sad_flow_launch(int search, const uint8_t *prev, const uint8_t *cur, const SadGeom &g, int n,
			    aof_px4_raw *raw, aof_px4_block *blocks, cudaStream_t st);
// second version (sad_flow2.cu): pre-shifted copies of the current frame, 8-lane half-pixel stage
bool sad_flow2_supported(const SadGeom &g);
cudaError_t sad_flow2_launch(int search, const uint8_t *prev, const uint8_t *cur, const SadGeom &g, int n,
			     aof_px4_raw *raw, aof_px4_block *blocks, cudaStream_t st);
// third version (sad_flow3.cu): 64x64, search 4 only; everything static, paired x-displacements
bool sad_flow3_supported(int search, const SadGeom &g);
cudaError_t sad_flow3_launch(const uint8_t *prev, const uint8_t *cur, const SadGeom &g, int n, aof_px4_raw *raw,
			     aof_px4_block *blocks, cudaStream_t st);
cudaError_t px4_epilogue_launch(const aof_px4_raw *raw, const uint32_t *t_us, LimiterState *lim, aof_flow_out *out,
				int n, int output_rate, float fx, float fy, cudaStream_t st);

} // namespace aof

// ---------------------------------------------------------------------------
// LK path (OpticalFlowOpenCV::calcFlow, trackFeatures, calcOpticalFlowPyrLK, FAST)
namespace aof {

// band FAST kernel (lk_fast_band.cu): output rows per band, threads of the masked kernel on 640-px frames.  Measured (r2Y,
// masked kernel per 512 LK streams): 38 rows x 320 threads (4 CTAs / SM) 274.6 us; 19 x 160 (8 / SM) 283.1; 26 x 256 (5 / SM)
// 311.1; 19 x 320 336.9 -- more, smaller CTAs do not pay for their extra halo rows.
#ifndef AOF_FB_ROWS
#define AOF_FB_ROWS 38
#endif
#ifndef AOF_FB_MTHREADS
#define AOF_FB_MTHREADS 320
#endif
constexpr int FB_ROWS = AOF_FB_ROWS;
constexpr int LK_MAX_LEVELS = 4;   // maxLevel = 3 (trackFeatures.cpp:98)
constexpr int LK_WIN = 9;          // Size(9, 9)
constexpr int LK_MAX_FEATURES = 100; // trackFeatures.cpp:56
constexpr int LK_BINS = 16;        // 4 x 4 (trackFeatures.cpp:156-157)
constexpr int LK_MAX_TPB = (LK_MAX_FEATURES - 1) / LK_BINS + 1; // targetFeaturesPerBin <= 7

struct LkGeom {
	int w[LK_MAX_LEVELS], h[LK_MAX_LEVELS];
	int top;          // index of the coarsest level (OpenCV stops when the next one would be <= the window)
	int n_feat;       // N = updateVector.size()
	int bw, bh;       // bin size (cols/4, rows/4)
	int tpb;          // targetFeaturesPerBin
	int dist;         // binWidth / targetFeaturesPerBin
	int kp_cap;       // keypoint-list capacity per bin
	int fast_thr;     // 10
	int fb_bands;     // 38-row bands per bin of the band FAST kernel (lk_fast_band.cu)
	int fb_ucap;      // live-unit capacity of one band: 40 response rows x segments of 16 pixel pairs
	// A/B and tuning switches, read from the environment ONCE in aof_lk_create (env_int)
	int tun_pyr_bands; // AOF_PYR_BANDS: forced band count of the streaming pyramid kernel (0 = auto)
	int tun_pyr_occ;   // AOF_PYR_OCC: CTAs per SM the pyramid kernel is compiled for (16 / 20 / 24)
	int tun_pyr_v1;    // AOF_PYR_V1 = 1: the generic streaming kernel
	int tun_fast_v1;   // AOF_FAST_V1 = 1: the tile FAST kernel
	int tun_track_v1;  // AOF_TRACK_V1 = 1: the CTA-per-stream tracker
	int tun_pdl;       // AOF_LK_PDL = 0: plain stream-ordered launches (default 1: programmatic dependent launch, pdl_enter())
	int tun_fast_lazy; // AOF_FAST_LAZY = 0: score every pixel of the needed bins; 1: lazy arg-max (lk_fast_lazy.cu); 2 (default): masked band FAST
};

// level base pointers of one pyramid slot; level l of stream s is at lvl[l] + s * w[l] * h[l]
struct LkPyr {
	const uint8_t *lvl[LK_MAX_LEVELS];
};

// per-stream tracker state, structure of arrays over n_streams
struct LkState {
	float2 *prev_corners;   // [n * N]   trackFeatures.cpp:54
	uint8_t *prev_status;   // [n * N]   :56
	int *n_prev_corners;    // [n]       prev_corners.size(): 0 or N
	int *update_vec;        // [n * N]   flow_opencv.hpp:66
	float2 *feat_cur;       // [n * N]   features_current
	float2 *feat_prev;      // [n * N]   features_previous
	LimiterState *lim;      // [n]
	int *bin_needed;        // [n * 16]  neededFeatures per bin (bin = x * 4 + y)
	int *target;            // [n]       targetNumPoints
	int *kp_count;          // [n * 16]
	uint32_t *kp;           // [n * 16 * kp_cap]  score << 24 | y << 12 | x (bin-local, so raster order == numeric order)
	float2 *lk_next;        // [n * N]   LK result per slot, handed from the tracking CTAs to the bookkeeping CTA
	uint8_t *lk_stat;       // [n * N]
	int *ticket;            // [n]       tracking CTAs of the stream that have finished (0 between launches)
	// lazy FAST (lk_fast_lazy.cu): the per-bin picks of initMorePoints' first stage, found without scoring every pixel
	uint32_t *lz_pick;      // [n * 16 * LK_MAX_TPB]  x | y << 16 (image coordinates), in pick order
	int *lz_npick;          // [n * 16]
	int *lz_floor;          // [n * 16]  adaptive pool threshold of the bin (persistent; 0 = not yet set)
	int *need_full;         // [n]       1 = the selection needs the complete keypoint lists of this stream (second chance)
	int *n_full;            // [n]       element 0 of a range counts its flagged streams (reset by the fallback selection)
	int *full_list;         // [n]       range-local indices of the flagged streams
	unsigned long long *lz_stats; // [4] diagnostics, whole handle: bins picked lazily, pool rebuilds, bins that overflowed
	                              //     the lazy buffers, streams sent to the complete path
	                              //     (masked FAST: [0] bands with work, [1] their live units)
	// masked band FAST (lk_fast_band.cu): lk_fast_plan_kernel lists, per (stream, bin, band), the units of 16 pixel
	// pairs x 1 response row that the first selection stage can accept from, and the bands that have any
	uint16_t *plan_units;   // [n * 16 * fb_bands * fb_ucap]  response row << 8 | segment
	uint32_t *plan_hdr;     // [n * 16 * fb_bands]  units | first live row << 16 | last live row << 24
	uint32_t *work_list;    // [n * 16 * fb_bands]  range-local (stream * 16 + bin) * fb_bands + band of the bands with work
	int *n_work;            // [n]       element 0 of a range counts its work items (reset by the selection kernel)
};

// camera model of setCameraMatrix / setCameraDistortion (flow_opencv.cpp:72-87); on only when both were set
struct LkCamera {
	int on;
	float fx, fy, cx, cy;
	float k1, k2, p1, p2, k3;
};

cudaError_t lk_pyr_down_launch(const uint8_t *src, uint8_t *dst, int sw, int sh, int n, cudaStream_t st);
cudaError_t lk_pyr_fused_launch(const LkGeom &g, const uint8_t *l0, uint8_t *l1, uint8_t *l2, uint8_t *l3, int n, cudaStream_t st);
// streaming 3-level pyramid (lk_pyr_stream.cu); needs W % 32 == 0, H % 8 == 0, an 8-byte aligned batch
bool lk_pyr_stream_supported(const LkGeom &g, const uint8_t *l0);
cudaError_t lk_pyr_stream_launch(const LkGeom &g, const uint8_t *l0, uint8_t *l1, uint8_t *l2, uint8_t *l3, int n, cudaStream_t st);
cudaError_t lk_track_update_launch(const LkGeom &g, const LkPyr &prev, const LkPyr &next, const LkState &s, int n,
				   int have_prev, cudaStream_t st);
cudaError_t lk_fast_launch(const LkGeom &g, const uint8_t *l0, const LkState &s, int n, uint8_t *score_map,
			   cudaStream_t st);
// banded FAST (lk_fast_band.cu); lk_fast_launch dispatches to it when the shape allows
bool lk_fast_band_supported(const LkGeom &g, const uint8_t *l0);
cudaError_t lk_fast_band_launch(const LkGeom &g, const uint8_t *l0, const LkState &s, int n, uint8_t *score_map, cudaStream_t st);
// the same for the streams listed in s.full_list only (small persistent grid)
cudaError_t lk_fast_band_flagged_launch(const LkGeom &g, const uint8_t *l0, const LkState &s, int n, cudaStream_t st);
// masked band FAST: the plan (live units per band, work list), then a persistent grid over the work list
bool lk_fast_masked_supported(const LkGeom &g, const uint8_t *l0);
cudaError_t lk_fast_masked_launch(const LkGeom &g, const uint8_t *l0, const LkState &s, int n, cudaStream_t st);
// lazy FAST (lk_fast_lazy.cu): the per-bin picks of initMorePoints without scoring every pixel
bool lk_fast_lazy_supported(const LkGeom &g, const uint8_t *l0);
cudaError_t lk_fast_pick_launch(const LkGeom &g, const uint8_t *l0, const LkState &s, int n, cudaStream_t st);
// mode of lk_select_epilogue_launch: the complete path; the picks of lk_fast_pick_kernel (flags the streams that need
// the complete path and leaves them untouched); the complete path for the flagged streams only
enum { LK_SEL_COMPLETE = 0, LK_SEL_LAZY = 1, LK_SEL_FLAGGED = 2, LK_SEL_MASKED = 3 };
cudaError_t lk_select_epilogue_launch(const LkGeom &g, const LkState &s, const uint32_t *t_us, aof_flow_out *out,
				      aof_lk_raw *raw, int n, int have_feat_prev, int output_rate, float conf, float fx,
				      float fy, const LkCamera &cam, cudaStream_t st, int mode = LK_SEL_COMPLETE);
// stateless pyramidal LK over given points (KAT entry): pts[n * n_pts]
cudaError_t lk_track_points_launch(const LkGeom &g, const LkPyr &prev, const LkPyr &next, const float2 *pts, int n_pts,
				   float2 *out, uint8_t *status, int n, cudaStream_t st);

} // namespace aof

// ---------------------------------------------------------------------------
// dense SAD field (BASELINE config 4)
namespace aof {
struct DenseGeom {
	int width, height;
	size_t frame_stride;
	int i_lo, i_hi, j_lo, j_hi; // lattice indices (block origin = 8 * index) whose search window + 1 px lies inside the image
	int nbx, nby;
	uint32_t feat_thr, val_thr;
};
cudaError_t sad_dense_launch(int search, const uint8_t *prev, const uint8_t *cur, const DenseGeom &g, int n,
			     aof_dense_block *field, cudaStream_t st);
} // namespace aof
