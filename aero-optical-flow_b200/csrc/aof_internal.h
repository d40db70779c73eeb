// Internal declarations shared by the kernels (.cu) and the C-ABI host code.
#pragma once
#include <climits>
#include <cstdint>
#include <cuda_runtime.h>

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
};

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

cudaError_t sad_flow_launch(int search, const uint8_t *prev, const uint8_t *cur, const SadGeom &g, int n,
			    aof_px4_raw *raw, aof_px4_block *blocks, cudaStream_t st);
cudaError_t px4_epilogue_launch(const aof_px4_raw *raw, const uint32_t *t_us, LimiterState *lim, aof_flow_out *out,
				int n, int output_rate, float fx, float fy, cudaStream_t st);

} // namespace aof
