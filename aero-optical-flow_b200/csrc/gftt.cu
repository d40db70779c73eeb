// K8  Shi-Tomasi min-eigenvalue response + goodFeaturesToTrack selection -- the corner selector that
// north_star names (the reference itself selects with FAST).  Arithmetic spec:
// cv::cornerMinEigenVal(blockSize 3, Sobel 3) and cv::goodFeaturesToTrack as restated and checked against
// cv2 4.13 in oracle/lk_oracle.c (olk_min_eig, olk_gftt); these kernels match that oracle bit for bit
// (-fmad=false).
//   gftt_min_eig_kernel : tile -> Sobel (f32, reflect-101) -> covariance -> 3x3 box sum (double) -> lambda_min,
//                         plus the per-frame maximum (ordered-int atomicMax)
//   gftt_select_kernel  : threshold quality * max, strict-or-equal 3x3 maximum, interior only -> candidate list;
//                         then picks = repeated masked arg-max (value desc, address desc) under the Euclidean
//                         minDistance gate -- the same as walking cv's sorted list, because the accepted set
//                         only grows.
#include <new>

#include "aof_common.cuh"
#include "aof_internal.h"
#include "aof_host.h"

namespace aof {

__device__ __forceinline__ int gf_reflect(int p, int n)
{
	if (n == 1) return 0;
	while (p < 0 || p >= n) p = p < 0 ? -p : 2 * n - 2 - p;
	return p;
}

// monotone map float -> uint32 (so atomicMax orders floats, negatives included)
__device__ __forceinline__ uint32_t gf_ord(float f)
{
	const uint32_t u = __float_as_uint(f);
	return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float gf_unord(uint32_t k)
{
	return __uint_as_float((k & 0x80000000u) ? (k & 0x7FFFFFFFu) : ~k);
}

constexpr int GF_TW = 32, GF_TH = 16;

__global__ void __launch_bounds__(256)
gftt_min_eig_kernel(const uint8_t *__restrict__ imgs, int w, int h, float *__restrict__ eig, uint32_t *__restrict__ frame_max)
{
	__shared__ uint8_t s_pix[GF_TH + 4][GF_TW + 4];   // [r][c] <-> image (reflect(y0 - 2 + r), reflect(x0 - 2 + c))
	__shared__ float s_cov[3][GF_TH + 2][GF_TW + 2];  // [r][c] <-> covariance at (reflect(y0 - 1 + r), reflect(x0 - 1 + c))
	__shared__ uint32_t s_max;
	const size_t frame = blockIdx.z;
	const uint8_t *img = imgs + frame * (size_t)w * h;
	const int x0 = blockIdx.x * GF_TW, y0 = blockIdx.y * GF_TH;
	const int tid = threadIdx.x;
	if (tid == 0) s_max = 0;
	for (int i = tid; i < (GF_TH + 4) * (GF_TW + 4); i += 256) {
		const int r = i / (GF_TW + 4), c = i - r * (GF_TW + 4);
		s_pix[r][c] = img[(size_t)gf_reflect(y0 - 2 + r, h) * w + gf_reflect(x0 - 2 + c, w)];
	}
	__syncthreads();
	const float s = (float)(1. / 3060.), s2 = 2.f * s;
	for (int i = tid; i < (GF_TH + 2) * (GF_TW + 2); i += 256) {
		const int r = i / (GF_TW + 2), c = i - r * (GF_TW + 2);
		// the covariance the box filter reads here is the one AT the reflected coordinate
		const int ya = gf_reflect(y0 - 1 + r, h), xa = gf_reflect(x0 - 1 + c, w);
		const int pr = ya - (y0 - 2), pc = xa - (x0 - 2);       // its centre in the pixel tile
		float dx = 0.f, dy = 0.f;
		if (pr >= 1 && pr <= GF_TH + 2 && pc >= 1 && pc <= GF_TW + 2) {
			float rx[3], ry[3];
#pragma unroll
			for (int j = 0; j < 3; j++) {
				const int a = s_pix[pr + j - 1][pc - 1], b = s_pix[pr + j - 1][pc], cc = s_pix[pr + j - 1][pc + 1];
				rx[j] = (float)(cc - a);
				ry[j] = (float)(a + cc) * s + (float)b * s2;
			}
			dx = (rx[0] + rx[2]) * s + rx[1] * s2;
			dy = ry[2] - ry[0];
		}
		s_cov[0][r][c] = dx * dx; s_cov[1][r][c] = dx * dy; s_cov[2][r][c] = dy * dy;
	}
	__syncthreads();
	uint32_t my = 0;
	for (int i = tid; i < GF_TH * GF_TW; i += 256) {
		const int r = i / GF_TW, c = i - r * GF_TW;
		const int x = x0 + c, y = y0 + r;
		if (x >= w || y >= h) continue;
		double acc[3] = {0., 0., 0.};
#pragma unroll
		for (int j = 0; j < 3; j++)
#pragma unroll
			for (int k = 0; k < 3; k++) {
				acc[0] += (double)s_cov[0][r + j][c + k]; acc[1] += (double)s_cov[1][r + j][c + k]; acc[2] += (double)s_cov[2][r + j][c + k];
			}
		const float a = (float)acc[0] * 0.5f, b = (float)acc[1], cc = (float)acc[2] * 0.5f;
		const float v = (a + cc) - sqrtf((a - cc) * (a - cc) + b * b);
		eig[frame * (size_t)w * h + (size_t)y * w + x] = v;
		my = max(my, gf_ord(v));
	}
	my = __reduce_max_sync(0xFFFFFFFFu, my);
	if ((tid & 31) == 0) atomicMax(&s_max, my);
	__syncthreads();
	if (tid == 0 && s_max) atomicMax(&frame_max[frame], s_max);
}

constexpr int GS_THREADS = 512;

// one CTA per frame.  cand: scratch list of w*h/2 (value bits, index) pairs per frame.
__global__ void __launch_bounds__(GS_THREADS)
gftt_select_kernel(const float *__restrict__ eig_all, int w, int h, const uint32_t *__restrict__ frame_max, int max_corners,
		   double quality, double min_dist, uint64_t *__restrict__ cand_all, int cand_cap, float2 *__restrict__ xy_all,
		   int *__restrict__ count_all)
{
	__shared__ int s_n, s_found, s_done;
	__shared__ unsigned long long s_red[GS_THREADS / 32];
	extern __shared__ float2 s_acc[];  // accepted corners (max_corners)
	const size_t frame = blockIdx.x;
	const float *eig = eig_all + frame * (size_t)w * h;
	uint64_t *cand = cand_all + frame * (size_t)cand_cap;
	float2 *xy = xy_all + frame * (size_t)max_corners;
	const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
	const float mx = gf_unord(frame_max[frame]);
	const float thr = (float)((double)mx * quality);
	if (tid == 0) { s_n = 0; s_found = 0; s_done = 0; }
	__syncthreads();
	// candidates: above the threshold, equal to the 3x3 maximum of the thresholded map, interior
	for (int i = tid; i < w * h; i += GS_THREADS) {
		const int y = i / w, x = i - y * w;
		if (x < 1 || x >= w - 1 || y < 1 || y >= h - 1) continue;
		const float v = eig[i];
		if (!(v > thr) || v == 0.f) continue;
		bool is_max = true;
#pragma unroll
		for (int j = -1; j <= 1; j++)
#pragma unroll
			for (int k = -1; k <= 1; k++) {
				const float u = eig[i + j * w + k];
				if (u > thr && u > v) is_max = false;
			}
		if (!is_max) continue;
		const int slot = atomicAdd(&s_n, 1);
		if (slot < cand_cap) cand[slot] = ((uint64_t)gf_ord(v) << 32) | (uint32_t)i;
	}
	__syncthreads();
	const int n = min(s_n, cand_cap);
	const float md2 = (float)(min_dist * min_dist);
	const bool gate = min_dist >= 1.0;
	unsigned long long below = ~0ull; // only candidates ranked after the previous pick can be picked next
	while (!s_done) {
		const int found = s_found;
		unsigned long long best = 0;
		for (int e = tid; e < n; e += GS_THREADS) {
			const unsigned long long key = cand[e];
			if (key >= below || key <= best) continue;
			bool ok = true;
			if (gate) {
				const int idx = (int)(key & 0xFFFFFFFFu);
				const float x = (float)(idx % w), y = (float)(idx / w);
				for (int j = 0; j < found && ok; j++) {
					const float dx = x - s_acc[j].x, dy = y - s_acc[j].y;
					ok = !(dx * dx + dy * dy < md2);
				}
			}
			if (ok) best = key;
		}
#pragma unroll
		for (int o = 16; o; o >>= 1) {
			const unsigned long long other = __shfl_xor_sync(0xFFFFFFFFu, best, o);
			best = other > best ? other : best;
		}
		if (lane == 0) s_red[warp] = best;
		__syncthreads();
		unsigned long long m = 0;
		for (int k = 0; k < GS_THREADS / 32; k++) m = s_red[k] > m ? s_red[k] : m;
		if (m) below = m;
		if (tid == 0) {
			if (!m) s_done = 1;
			else {
				const int idx = (int)(m & 0xFFFFFFFFu);
				const float2 p = make_float2((float)(idx % w), (float)(idx / w));
				s_acc[found] = p; xy[found] = p;
				s_found = found + 1;
				if (found + 1 == max_corners) s_done = 1;
			}
		}
		__syncthreads();
	}
	if (tid == 0) count_all[frame] = s_found;
}

} // namespace aof

using namespace aof;

struct aof_gftt {
	int w = 0, h = 0, device = 0, cap = 0, cap_corners = 0;
	cudaStream_t st = nullptr;
	uint8_t *d_img = nullptr;
	float *d_eig = nullptr;
	uint32_t *d_max = nullptr;
	uint64_t *d_cand = nullptr;
	float2 *d_xy = nullptr;
	int *d_cnt = nullptr;
};

static int gftt_reserve(aof_gftt *g, int n, int max_corners)
{
	const size_t px = (size_t)g->w * g->h;
	if (n > g->cap) {
		cudaFree(g->d_img); cudaFree(g->d_eig); cudaFree(g->d_max); cudaFree(g->d_cand); cudaFree(g->d_cnt);
		g->d_img = nullptr; g->d_eig = nullptr; g->d_max = nullptr; g->d_cand = nullptr; g->d_cnt = nullptr; g->cap = 0;
		AOF_CUDA(cudaMalloc(&g->d_img, px * n));
		AOF_CUDA(cudaMalloc(&g->d_eig, sizeof(float) * px * n));
		AOF_CUDA(cudaMalloc(&g->d_max, sizeof(uint32_t) * n));
		AOF_CUDA(cudaMalloc(&g->d_cand, sizeof(uint64_t) * (px / 2 + 1) * n));
		AOF_CUDA(cudaMalloc(&g->d_cnt, sizeof(int) * n));
		g->cap = n;
	}
	if ((size_t)n * max_corners > (size_t)g->cap_corners) {
		cudaFree(g->d_xy); g->d_xy = nullptr; g->cap_corners = 0;
		AOF_CUDA(cudaMalloc(&g->d_xy, sizeof(float2) * (size_t)n * max_corners));
		g->cap_corners = n * max_corners;
	}
	return AOF_OK;
}

static int gftt_run(aof_gftt *g, const uint8_t *imgs, int n, bool select, int max_corners, double quality, double min_dist)
{
	const size_t px = (size_t)g->w * g->h;
	AOF_CUDA(cudaMemcpyAsync(g->d_img, imgs, px * n, cudaMemcpyHostToDevice, g->st));
	AOF_CUDA(cudaMemsetAsync(g->d_max, 0, sizeof(uint32_t) * n, g->st));
	{
		dim3 grid((g->w + GF_TW - 1) / GF_TW, (g->h + GF_TH - 1) / GF_TH, n);
		KernelTimer kt(KID_GFTT_MIN_EIG, g->st);
		gftt_min_eig_kernel<<<grid, 256, 0, g->st>>>(g->d_img, g->w, g->h, g->d_eig, g->d_max);
		count_launch();
		AOF_CUDA(cudaGetLastError());
	}
	if (select) {
		KernelTimer kt(KID_GFTT_SELECT, g->st);
		gftt_select_kernel<<<n, GS_THREADS, sizeof(float2) * max_corners, g->st>>>(g->d_eig, g->w, g->h, g->d_max, max_corners, quality,
											     min_dist, g->d_cand, (int)(px / 2 + 1), g->d_xy, g->d_cnt);
		count_launch();
		AOF_CUDA(cudaGetLastError());
	}
	return AOF_OK;
}

extern "C" {

int aof_gftt_create(int width, int height, int device, aof_gftt **out)
{
	if (!out || width < 3 || height < 3) { set_error("aof_gftt_create: bad argument"); return AOF_ERR_INVALID; }
	*out = nullptr;
	int rc = select_device(device);
	if (rc != AOF_OK) return rc;
	aof_gftt *g = new (std::nothrow) aof_gftt();
	if (!g) return AOF_ERR_INVALID;
	g->w = width; g->h = height; g->device = device;
	cudaError_t e = cudaStreamCreateWithFlags(&g->st, cudaStreamNonBlocking);
	if (e != cudaSuccess) { set_error("aof_gftt_create: %s", cudaGetErrorString(e)); delete g; return AOF_ERR_CUDA; }
	*out = g;
	return AOF_OK;
}

void aof_gftt_destroy(aof_gftt *g)
{
	if (!g) return;
	cudaSetDevice(g->device);
	if (g->st) cudaStreamSynchronize(g->st);
	cudaFree(g->d_img); cudaFree(g->d_eig); cudaFree(g->d_max); cudaFree(g->d_cand); cudaFree(g->d_xy); cudaFree(g->d_cnt);
	if (g->st) cudaStreamDestroy(g->st);
	delete g;
}

int aof_gftt_min_eig(aof_gftt *g, const uint8_t *imgs, int n, float *eig)
{
	if (!g || !imgs || !eig || n < 1 || n > 65535) { set_error("aof_gftt_min_eig: bad argument"); return AOF_ERR_INVALID; }
	AOF_CUDA(cudaSetDevice(g->device));
	int rc = gftt_reserve(g, n, 1);
	if (rc != AOF_OK) return rc;
	rc = gftt_run(g, imgs, n, false, 0, 0., 0.);
	if (rc != AOF_OK) return rc;
	AOF_CUDA(cudaMemcpyAsync(eig, g->d_eig, sizeof(float) * (size_t)g->w * g->h * n, cudaMemcpyDeviceToHost, g->st));
	AOF_CUDA(cudaStreamSynchronize(g->st));
	return AOF_OK;
}

int aof_gftt_detect(aof_gftt *g, const uint8_t *imgs, int n, int max_corners, double quality_level, double min_distance,
		    float *xy, int *counts)
{
	if (!g || !imgs || !xy || !counts || n < 1 || n > 65535 || max_corners < 1 || max_corners > 4096) {
		set_error("aof_gftt_detect: bad argument (1 <= max_corners <= 4096)");
		return AOF_ERR_INVALID;
	}
	AOF_CUDA(cudaSetDevice(g->device));
	int rc = gftt_reserve(g, n, max_corners);
	if (rc != AOF_OK) return rc;
	rc = gftt_run(g, imgs, n, true, max_corners, quality_level, min_distance);
	if (rc != AOF_OK) return rc;
	AOF_CUDA(cudaMemcpyAsync(xy, g->d_xy, sizeof(float2) * (size_t)n * max_corners, cudaMemcpyDeviceToHost, g->st));
	AOF_CUDA(cudaMemcpyAsync(counts, g->d_cnt, sizeof(int) * n, cudaMemcpyDeviceToHost, g->st));
	AOF_CUDA(cudaStreamSynchronize(g->st));
	return AOF_OK;
}

} // extern "C"
