// LK-path kernels (sm_100a).  Replace, for batches of independent streams,
//   K3 lk_pyr_down        cv::pyrDown inside calcOpticalFlowPyrLK            (trackFeatures.cpp:98)
//   K6 lk_track_*         cv::calcOpticalFlowPyrLK(prev, cur, pts, 9x9, 3)    (trackFeatures.cpp:98)
//      + the trackFeatures bookkeeping after it                              (trackFeatures.cpp:77-125)
//      + the per-bin demand of initMorePoints                                (trackFeatures.cpp:139-183)
//   K4 lk_fast            cv::FAST(bin ROI, kps, 10) score + non-max suppression (trackFeatures.cpp:196)
//   K5 lk_select_epilogue initMorePoints selection                           (trackFeatures.cpp:185-362)
//   K7   + OpticalFlowOpenCV::calcFlow statistics, status machine, limitRate, atan2
//                                                                            (flow_opencv.cpp:121-209)
// Arithmetic spec: SURVEY.md Appendix A (verified bit-exact against cv2 4.13 by
// oracle/lk_oracle.c, which these kernels must match bit for bit).  The file is
// compiled with -fmad=false: every float operation rounds as written.
#include <climits>
#include <cstdlib>

#include "aof_common.cuh"
#include "aof_internal.h"

namespace aof {

__device__ __forceinline__ int reflect101(int p, int n)
{
	p = p < 0 ? -p : p;
	p = p >= n ? 2 * n - 2 - p : p;
	return min(max(p, 0), n - 1); // the clamp only ever acts on taps whose value is never used
}

// ---------------------------------------------------------------------------
// K3: one pyramid level.  CTA = 64 x 16 output pixels of one stream.
// tile of the source (35 x 131 px, reflect-101) -> horizontal [1 4 6 4 1] on even
// columns -> vertical, (sum + 128) >> 8.
constexpr int PD_TW = 64, PD_TH = 16;
constexpr int PD_SW = 2 * PD_TW + 4, PD_SH = 2 * PD_TH + 3; // 132 (131 used), 35

__global__ void __launch_bounds__(256)
lk_pyr_down_kernel(const uint8_t *__restrict__ src, uint8_t *__restrict__ dst, int sw, int sh, int dw, int dh)
{
	__shared__ uint8_t s_src[PD_SH][PD_SW];
	__shared__ uint16_t s_h[PD_SH][PD_TW];
	const int tid = threadIdx.x;
	const int ox0 = blockIdx.x * PD_TW, oy0 = blockIdx.y * PD_TH;
	const uint8_t *img = src + (size_t)blockIdx.z * sw * sh;
	uint8_t *out = dst + (size_t)blockIdx.z * dw * dh;
	for (int i = tid; i < PD_SH * PD_SW; i += 256) {
		const int r = i / PD_SW, c = i - r * PD_SW;
		const int y = reflect101(2 * oy0 - 2 + r, sh), x = reflect101(2 * ox0 - 2 + c, sw);
		s_src[r][c] = img[(size_t)y * sw + x];
	}
	__syncthreads();
	for (int i = tid; i < PD_SH * PD_TW; i += 256) {
		const int r = i / PD_TW, c = i - r * PD_TW;
		const uint8_t *p = &s_src[r][2 * c];
		s_h[r][c] = (uint16_t)(p[0] + 4 * p[1] + 6 * p[2] + 4 * p[3] + p[4]);
	}
	__syncthreads();
	for (int i = tid; i < PD_TH * PD_TW; i += 256) {
		const int r = i / PD_TW, c = i - r * PD_TW;
		const int ox = ox0 + c, oy = oy0 + r;
		if (ox < dw && oy < dh) {
			const int v = s_h[2 * r][c] + 4 * s_h[2 * r + 1][c] + 6 * s_h[2 * r + 2][c] + 4 * s_h[2 * r + 3][c] +
				      s_h[2 * r + 4][c];
			out[(size_t)oy * dw + ox] = (uint8_t)((v + 128) >> 8);
		}
	}
}

cudaError_t lk_pyr_down_launch(const uint8_t *src, uint8_t *dst, int sw, int sh, int n, cudaStream_t st)
{
	if (n <= 0) return cudaSuccess;
	const int dw = (sw + 1) / 2, dh = (sh + 1) / 2;
	dim3 grid((dw + PD_TW - 1) / PD_TW, (dh + PD_TH - 1) / PD_TH, n);
	KernelTimer kt(KID_LK_PYR_DOWN, st);
	lk_pyr_down_kernel<<<grid, 256, 0, st>>>(src, dst, sw, sh, dw, dh);
	count_launch();
	return cudaGetLastError();
}

// ---------------------------------------------------------------------------
// K3 fused: levels 1, 2 and 3 of one frame in ONE pass over level 0 (the HBM-bound part of the
// LK path: 307 200 B read + 100 800 B written per 640 x 480 frame).  CTA = one 16 x 8 tile of
// level 3 with the level-2 / level-1 / level-0 regions it depends on (35 x 19, 73 x 41, 149 x 85)
// staged in shared memory; the halo re-reads of level 0 (1.55x) are L2 hits from the neighbouring
// CTAs.  Each level uses cv::pyrDown's own BORDER_REFLECT_101 on ITS size, so tiles are addressed
// by unreflected coordinate and every tap is reflected explicitly (an even-sized level does not
// commute with reflecting the level below).
constexpr int PF3W = 16, PF3H = 8;
constexpr int PF2W = 2 * PF3W + 3, PF2H = 2 * PF3H + 3;   // 35 x 19
constexpr int PF1W = 2 * PF2W + 3, PF1H = 2 * PF2H + 3;   // 73 x 41
constexpr int PF0W = 2 * PF1W + 3, PF0H = 2 * PF1H + 3;   // 149 x 85
constexpr int PF0P = 152;                                  // level-0 tile pitch: column 0 <-> x = o0 - 2 (8-aligned)
constexpr int PF_THREADS = 256;

// one pyrDown from a source tile to a destination tile, both in shared memory.
// tile element [r][c] <-> level coordinate (oy + r, ox + c); only coordinates inside the level are valid.
__device__ __forceinline__ void pf_down(const uint8_t *src, int sp, int sox, int soy, int slw, int slh, int sth,
					uint8_t *dst, int dp, int dox, int doy, int dlw, int dlh, int dtw, int dth, uint16_t *tmp)
{
	const int tid = threadIdx.x;
	for (int i = tid; i < sth * dtw; i += PF_THREADS) {
		const int r = i / dtw, c = i - r * dtw;
		const int y = soy + r, X = dox + c;
		if (y < 0 || y >= slh || X < 0 || X >= dlw) continue;
		const uint8_t *row = src + r * sp - sox;
		const int x = 2 * X;
		int v;
		if (x >= 2 && x + 2 < slw) v = row[x - 2] + 4 * row[x - 1] + 6 * row[x] + 4 * row[x + 1] + row[x + 2];
		else v = row[reflect101(x - 2, slw)] + 4 * row[reflect101(x - 1, slw)] + 6 * row[x] + 4 * row[reflect101(x + 1, slw)] +
			 row[reflect101(x + 2, slw)];
		tmp[i] = (uint16_t)v;
	}
	__syncthreads();
	for (int i = tid; i < dth * dtw; i += PF_THREADS) {
		const int r = i / dtw, c = i - r * dtw;
		const int Y = doy + r, X = dox + c;
		if (Y < 0 || Y >= dlh || X < 0 || X >= dlw) continue;
		const int y = 2 * Y;
		const uint16_t *col = tmp + c - soy * dtw;
		int v;
		if (y >= 2 && y + 2 < slh) v = col[(y - 2) * dtw] + 4 * col[(y - 1) * dtw] + 6 * col[y * dtw] + 4 * col[(y + 1) * dtw] + col[(y + 2) * dtw];
		else v = col[reflect101(y - 2, slh) * dtw] + 4 * col[reflect101(y - 1, slh) * dtw] + 6 * col[y * dtw] +
			 4 * col[reflect101(y + 1, slh) * dtw] + col[reflect101(y + 2, slh) * dtw];
		dst[r * dp + c] = (uint8_t)((v + 128) >> 8);
	}
	__syncthreads();
}

// copy the core [cx, cx + cw) x [cy, cy + ch) of a tile (tile coordinates) to its level image
__device__ __forceinline__ void pf_store(const uint8_t *tile, int tp, int ox, int oy, int cx, int cy, int cw, int ch,
					 uint8_t *img, int lw, int lh)
{
	for (int i = threadIdx.x; i < cw * ch; i += PF_THREADS) {
		const int r = i / cw, c = i - r * cw;
		const int x = ox + cx + c, y = oy + cy + r;
		if (x < lw && y < lh) img[(size_t)y * lw + x] = tile[(cy + r) * tp + cx + c];
	}
}

template <bool ALIGNED>
__global__ void __launch_bounds__(PF_THREADS)
lk_pyr_fused_kernel(LkGeom g, const uint8_t *__restrict__ l0, uint8_t *__restrict__ l1, uint8_t *__restrict__ l2,
		    uint8_t *__restrict__ l3)
{
	__shared__ __align__(16) uint8_t s0[PF0H * PF0P];
	__shared__ uint16_t s_tmp[PF0H * PF1W];
	__shared__ uint8_t s1[PF1H * PF1W], s2[PF2H * PF2W], s3[PF3H * PF3W];
	const size_t stream = blockIdx.z;
	const int o3x = blockIdx.x * PF3W, o3y = blockIdx.y * PF3H;
	const int o2x = 2 * o3x - 2, o2y = 2 * o3y - 2, o1x = 2 * o2x - 2, o1y = 2 * o2y - 2, o0x = 2 * o1x - 2, o0y = 2 * o1y - 2;
	const int w0 = g.w[0], h0 = g.h[0];
	const uint8_t *img = l0 + stream * (size_t)w0 * h0;
	const int bx = o0x - 2; // tile column 0, a multiple of 8
	if (ALIGNED) {
		for (int i = threadIdx.x; i < PF0H * (PF0P / 4); i += PF_THREADS) {
			const int r = i / (PF0P / 4), q = i - r * (PF0P / 4);
			const int y = o0y + r, x = bx + 4 * q;
			if (y >= 0 && y < h0 && x >= 0 && x < w0)
				reinterpret_cast<uint32_t *>(s0)[r * (PF0P / 4) + q] = __ldg(reinterpret_cast<const uint32_t *>(img + (size_t)y * w0 + x));
		}
	} else {
		for (int i = threadIdx.x; i < PF0H * PF0P; i += PF_THREADS) {
			const int r = i / PF0P, c = i - r * PF0P;
			const int y = o0y + r, x = bx + c;
			if (y >= 0 && y < h0 && x >= 0 && x < w0) s0[i] = img[(size_t)y * w0 + x];
		}
	}
	__syncthreads();
	pf_down(s0, PF0P, bx, o0y, w0, h0, PF0H, s1, PF1W, o1x, o1y, g.w[1], g.h[1], PF1W, PF1H, s_tmp);
	pf_down(s1, PF1W, o1x, o1y, g.w[1], g.h[1], PF1H, s2, PF2W, o2x, o2y, g.w[2], g.h[2], PF2W, PF2H, s_tmp);
	pf_down(s2, PF2W, o2x, o2y, g.w[2], g.h[2], PF2H, s3, PF3W, o3x, o3y, g.w[3], g.h[3], PF3W, PF3H, s_tmp);
	pf_store(s1, PF1W, o1x, o1y, 6, 6, 4 * PF3W, 4 * PF3H, l1 + stream * (size_t)g.w[1] * g.h[1], g.w[1], g.h[1]);
	pf_store(s2, PF2W, o2x, o2y, 2, 2, 2 * PF3W, 2 * PF3H, l2 + stream * (size_t)g.w[2] * g.h[2], g.w[2], g.h[2]);
	pf_store(s3, PF3W, o3x, o3y, 0, 0, PF3W, PF3H, l3 + stream * (size_t)g.w[3] * g.h[3], g.w[3], g.h[3]);
}

// levels 1..3 of n frames (only for a 4-level pyramid)
cudaError_t lk_pyr_fused_launch(const LkGeom &g, const uint8_t *l0, uint8_t *l1, uint8_t *l2, uint8_t *l3, int n, cudaStream_t st)
{
	if (n <= 0) return cudaSuccess;
	dim3 grid((g.w[3] + PF3W - 1) / PF3W, (g.h[3] + PF3H - 1) / PF3H, n);
	KernelTimer kt(KID_LK_PYR_FUSED, st, "lk_pyr_fused_kernel");
	if (g.w[0] % 4 == 0 && (reinterpret_cast<uintptr_t>(l0) & 3) == 0 && ((size_t)g.w[0] * g.h[0]) % 4 == 0)
		lk_pyr_fused_kernel<true><<<grid, PF_THREADS, 0, st>>>(g, l0, l1, l2, l3);
	else
		lk_pyr_fused_kernel<false><<<grid, PF_THREADS, 0, st>>>(g, l0, l1, l2, l3);
	count_launch();
	return cudaGetLastError();
}

// ---------------------------------------------------------------------------
// K6: pyramidal LK of one point by one warp.
// Per-warp scratch in shared memory.
struct LkWarpScratch {
	uint8_t I[12 * 12];   // template source patch (reflect-101)
	uint8_t J[10 * 10];   // current-image taps of one iteration
	short2 d[10 * 10];    // Scharr (Ix, Iy) at the 10 x 10 taps, 0 outside the image
	uint8_t pad[4];
};

__device__ __forceinline__ int descale(int x, int n) { return (x + (1 << (n - 1))) >> n; }

__device__ __forceinline__ void q14_weights(float a, float b, int &w00, int &w01, int &w10, int &w11)
{
	const float ia = 1.f - a, ib = 1.f - b;
	w00 = __float2int_rn(ia * ib * 16384.f);
	w01 = __float2int_rn(a * ib * 16384.f);
	w10 = __float2int_rn(ia * b * 16384.f);
	w11 = 16384 - w00 - w01 - w10;
}

// exact sum over the warp of per-lane partial sums that fit i32 but whose total may not
__device__ __forceinline__ float warp_sum_exact_f32(int v)
{
	const int lo = __reduce_add_sync(0xFFFFFFFFu, v & 0xFFFF);
	const int hi = __reduce_add_sync(0xFFFFFFFFu, v >> 16);
	const long long t = ((long long)hi << 16) + lo;
	// same rounding either way; the 32-bit conversion is an ALU instruction, the 64-bit one a slow XU one on the
	// critical path of every LK iteration (the total almost always fits)
	return t == (long long)(int)t ? (float)(int)t : __ll2float_rn(t);
}

// returns status (1/0); next point in (nx, ny).  All lanes run the same scalar math.
__device__ int lk_track_point(const LkGeom &g, const LkPyr &prev, const LkPyr &next, size_t stream, float ptx, float pty,
			      LkWarpScratch &sc, float &out_x, float &out_y)
{
	const int lane = threadIdx.x & 31;
	const float half = (LK_WIN - 1) * 0.5f;
	const float FLT_SCALE = 1.f / (1 << 20);
	int status = 1;
	float nx = 0.f, ny = 0.f;
	// window pixels owned by this lane: idx = lane + 32 k < 81
	int wy[3], wx[3];
#pragma unroll
	for (int k = 0; k < 3; k++) { const int idx = lane + 32 * k; wy[k] = idx / 9; wx[k] = idx - wy[k] * 9; }
	// taps gathered by this lane: 12 x 12 template patch (i = lane + 32 k < 144), 10 x 10 window of J (i < 100)
	int pr[5], pcol[5], jr[4], jc[4];
#pragma unroll
	for (int k = 0; k < 5; k++) { const int i = lane + 32 * k; pr[k] = i / 12; pcol[k] = i - pr[k] * 12; }
#pragma unroll
	for (int k = 0; k < 4; k++) { const int i = lane + 32 * k; jr[k] = i / 10; jc[k] = i - jr[k] * 10; }
	for (int level = g.top; level >= 0; --level) {
		const int w = g.w[level], h = g.h[level];
		const uint8_t *I = prev.lvl[level] + stream * (size_t)w * h;
		const uint8_t *J = next.lvl[level] + stream * (size_t)w * h;
		const float scale = 1.f / (float)(1 << level);
		float px = ptx * scale, py = pty * scale;
		if (level == g.top) { nx = px; ny = py; }
		else { nx = nx * 2.f; ny = ny * 2.f; }
		px -= half; py -= half;
		const int ipx = (int)floorf(px), ipy = (int)floorf(py);
		if (ipx < -LK_WIN || ipx >= w || ipy < -LK_WIN || ipy >= h) {
			if (level == 0) status = 0;
			continue;
		}
		int w00, w01, w10, w11;
		q14_weights(px - (float)ipx, py - (float)ipy, w00, w01, w10, w11);
		__syncwarp();
		if (ipx >= 1 && ipx + 10 < w && ipy >= 1 && ipy + 10 < h) { // the patch is inside the image: no border arithmetic
			const uint8_t *base = I + (size_t)(ipy - 1) * w + (ipx - 1);
#pragma unroll
			for (int k = 0; k < 5; k++)
				if (lane + 32 * k < 144) sc.I[lane + 32 * k] = __ldg(base + pr[k] * w + pcol[k]);
		} else {
#pragma unroll
			for (int k = 0; k < 5; k++)
				if (lane + 32 * k < 144)
					sc.I[lane + 32 * k] = I[(size_t)reflect101(ipy - 1 + pr[k], h) * w + reflect101(ipx - 1 + pcol[k], w)];
		}
		__syncwarp();
		for (int i = lane; i < 100; i += 32) {
			const int r = i / 10, c = i - r * 10;
			const int X = ipx + c, Y = ipy + r;
			short2 v = make_short2(0, 0);
			if (X >= 0 && X < w && Y >= 0 && Y < h) {
				const uint8_t *p = &sc.I[r * 12 + c]; // top-left of the 3x3 around patch (r+1, c+1)
				const int a0 = p[0], a1 = p[1], a2 = p[2], b0 = p[12], b2 = p[14], c0 = p[24], c1 = p[25], c2 = p[26];
				const int t0l = 3 * (a0 + c0) + 10 * b0, t0r = 3 * (a2 + c2) + 10 * b2;
				const int t1l = c0 - a0, t1m = c1 - a1, t1r = c2 - a2;
				v.x = (short)(t0r - t0l);
				v.y = (short)(3 * (t1l + t1r) + 10 * t1m);
			}
			sc.d[i] = v;
		}
		__syncwarp();
		int Iw[3], Ixw[3], Iyw[3];
		int s11 = 0, s12 = 0, s22 = 0;
#pragma unroll
		for (int k = 0; k < 3; k++) {
			Iw[k] = Ixw[k] = Iyw[k] = 0;
			if (lane + 32 * k < 81) {
				const int y = wy[k], x = wx[k];
				const uint8_t *p = &sc.I[(y + 1) * 12 + x + 1];
				const short2 d00 = sc.d[y * 10 + x], d01 = sc.d[y * 10 + x + 1], d10 = sc.d[y * 10 + 10 + x],
					     d11 = sc.d[y * 10 + 11 + x];
				Iw[k] = descale(p[0] * w00 + p[1] * w01 + p[12] * w10 + p[13] * w11, 9);
				Ixw[k] = (short)descale(d00.x * w00 + d01.x * w01 + d10.x * w10 + d11.x * w11, 14);
				Iyw[k] = (short)descale(d00.y * w00 + d01.y * w01 + d10.y * w10 + d11.y * w11, 14);
				s11 += Ixw[k] * Ixw[k]; s12 += Ixw[k] * Iyw[k]; s22 += Iyw[k] * Iyw[k];
			}
		}
		// 81 terms of at most 4080^2 fit i32
		const float A11 = (float)__reduce_add_sync(0xFFFFFFFFu, s11) * FLT_SCALE;
		const float A12 = (float)__reduce_add_sync(0xFFFFFFFFu, s12) * FLT_SCALE;
		const float A22 = (float)__reduce_add_sync(0xFFFFFFFFu, s22) * FLT_SCALE;
		float D = A11 * A22 - A12 * A12;
		const float dd = A11 - A22;
		const float min_eig = __fdiv_rn(A22 + A11 - __fsqrt_rn(dd * dd + 4.f * A12 * A12), (float)(2 * LK_WIN * LK_WIN));
		if (min_eig < 1e-4f || D < 1.1920928955078125e-7f) {
			if (level == 0) status = 0;
			continue;
		}
		D = __fdiv_rn(1.f, D);
		float cx = nx - half, cy = ny - half;
		float pdx = 0.f, pdy = 0.f;
		int win_x = INT_MIN, win_y = INT_MIN;                  // integer position of the window held in sc.J
		for (int j = 0; j < 30; j++) {
			const int inx = (int)floorf(cx), iny = (int)floorf(cy);
			if (inx < -LK_WIN || inx >= w || iny < -LK_WIN || iny >= h) {
				if (level == 0) status = 0;
				break;
			}
			q14_weights(cx - (float)inx, cy - (float)iny, w00, w01, w10, w11);
			// the 10 x 10 taps depend on the integer position only: after the first step or two the update is a fraction
			// of a pixel and the window already in shared memory is the one needed (the gather is the one global
			// round trip of an iteration)
			if (inx != win_x || iny != win_y) {
				win_x = inx; win_y = iny;
				__syncwarp();
				if (inx >= 0 && inx + 9 < w && iny >= 0 && iny + 9 < h) {
					const uint8_t *base = J + (size_t)iny * w + inx;
#pragma unroll
					for (int k = 0; k < 4; k++)
						if (lane + 32 * k < 100) sc.J[lane + 32 * k] = __ldg(base + jr[k] * w + jc[k]);
				} else {
#pragma unroll
					for (int k = 0; k < 4; k++)
						if (lane + 32 * k < 100)
							sc.J[lane + 32 * k] = __ldg(&J[(size_t)reflect101(iny + jr[k], h) * w + reflect101(inx + jc[k], w)]);
				}
				__syncwarp();
			}
			int p1 = 0, p2 = 0;
#pragma unroll
			for (int k = 0; k < 3; k++)
				if (lane + 32 * k < 81) {
					const uint8_t *p = &sc.J[wy[k] * 10 + wx[k]];
					const int diff = descale(p[0] * w00 + p[1] * w01 + p[10] * w10 + p[11] * w11, 9) - Iw[k];
					p1 += diff * Ixw[k]; p2 += diff * Iyw[k];
				}
			const float b1 = warp_sum_exact_f32(p1) * FLT_SCALE, b2 = warp_sum_exact_f32(p2) * FLT_SCALE;
			const float ddx = (A12 * b2 - A22 * b1) * D, ddy = (A12 * b1 - A11 * b2) * D;
			cx += ddx; cy += ddy;
			nx = cx + half; ny = cy + half;
			// delta.ddot(delta) <= epsilon^2 is evaluated in double by the reference.  The f32 value of the same sum is
			// within 2e-7 relative of it, so only a result that close to the threshold needs the double arithmetic
			// (which is a chain of slow FP64 instructions on the critical path of every iteration).
			const float dd32 = ddx * ddx + ddy * ddy;
			bool small;
			if (dd32 < 9.9999e-5f) small = true;
			else if (dd32 > 1.00001e-4f) small = false;
			else small = (double)ddx * (double)ddx + (double)ddy * (double)ddy <= 0.01 * 0.01;
			if (small) break;
			// (double)f < 0.01 for an f32 f is f <= 0.01f: 0.01f is the largest f32 below the double 0.01
			if (j > 0 && fabsf(ddx + pdx) <= 0.01f && fabsf(ddy + pdy) <= 0.01f) {
				nx -= ddx * 0.5f; ny -= ddy * 0.5f;
				break;
			}
			pdx = ddx; pdy = ddy;
		}
	}
	// calcOpticalFlowPyrLK re-tests the window origin of the final position at level 0 when the caller asks for
	// the error vector, as trackFeatures.cpp:94,98 does: a last update that left the range clears the status
	if (status) {
		const int fx = (int)floorf(nx - half), fy = (int)floorf(ny - half);
		if (fx < -LK_WIN || fx >= g.w[0] || fy < -LK_WIN || fy >= g.h[0]) status = 0;
	}
	out_x = nx; out_y = ny;
	return status;
}

// ---------------------------------------------------------------------------
// K6 driver + trackFeatures bookkeeping + per-bin demand.  One CTA per stream.
#ifndef AOF_TU_WARPS
#define AOF_TU_WARPS 8
#endif
constexpr int TU_WARPS = AOF_TU_WARPS;

__global__ void __launch_bounds__(TU_WARPS * 32, 4) // 4 CTAs per SM: 512 streams are one wave
lk_track_update_kernel(LkGeom g, LkPyr prev, LkPyr next, LkState s, int have_prev)
{
	__shared__ LkWarpScratch s_sc[TU_WARPS];
	__shared__ float2 s_next[LK_MAX_FEATURES], s_pc[LK_MAX_FEATURES];
	__shared__ int s_uv[LK_MAX_FEATURES];
	__shared__ uint8_t s_lk_status[LK_MAX_FEATURES];
	__shared__ int s_per_bin[LK_BINS];
	const size_t stream = blockIdx.x;
	const int N = g.n_feat;
	const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	float2 *pc = s.prev_corners + stream * N;
	uint8_t *ps = s.prev_status + stream * N;
	int *uv = s.update_vec + stream * N;
	float2 *fc = s.feat_cur + stream * N;
	const int npc = s.n_prev_corners[stream];
	const bool do_lk = have_prev && npc > 0;
	// the per-stream state is staged in shared memory once: everything below that walks over the slots would
	// otherwise be a chain of dependent global loads
	for (int i = threadIdx.x; i < N; i += blockDim.x) { s_uv[i] = uv[i]; s_pc[i] = pc[i]; }
	if (threadIdx.x < LK_BINS) { s_per_bin[threadIdx.x] = 0; s.kp_count[stream * LK_BINS + threadIdx.x] = 0; }
	__syncthreads();
	if (do_lk) {
		// only slots that were active can influence anything downstream; the reference tracks
		// the stale ones too (trackFeatures.cpp:98) but never reads them again before they are reassigned
		for (int f = warp; f < npc; f += TU_WARPS) {
			if (s_uv[f] != 1) continue;
			const float2 p = s_pc[f];
			float ox, oy;
			const int st = lk_track_point(g, prev, next, stream, p.x, p.y, s_sc[warp], ox, oy);
			if (lane == 0) { s_next[f] = make_float2(ox, oy); s_lk_status[f] = (uint8_t)st; }
		}
	}
	__syncthreads();
	// trackFeatures.cpp:77-125, one thread per slot; then the demand of initMorePoints (:144-183): bin = x * 4 + y
	int wants = 0;
	for (int i = threadIdx.x; i < N; i += blockDim.x) {
		int status = s_uv[i];
		int pst = status == 1;
		float2 feat = make_float2(-100.f, -100.f);
		float2 corner = s_pc[i];
		if (do_lk && i < npc) {
			if (pst) {
				corner = s_next[i];
				pc[i] = corner;
				if (!s_lk_status[i]) pst = 0;
			}
			if (pst) {
				if (corner.x < 0.f || corner.x > (float)g.w[0] || corner.y < 0.f || corner.y > (float)g.h[0]) status = 0;
				else { feat = corner; status = 1; }
			} else if (status == 1) status = 0;
		}
		ps[i] = (uint8_t)pst;
		uv[i] = status;
		fc[i] = feat;
		wants += status == 2;
		if (i < npc && status == 1) {
			const int bx = (int)__fdiv_rn(corner.x, (float)g.bw), by = (int)__fdiv_rn(corner.y, (float)g.bh);
			if (bx < 4 && by < 4) atomicAdd(&s_per_bin[bx * 4 + by], 1);
		}
	}
	const int target = __syncthreads_count(wants); // N <= blockDim.x: at most one slot per thread
	if (threadIdx.x < LK_BINS)
		s.bin_needed[stream * LK_BINS + threadIdx.x] = target ? max(0, g.tpb - s_per_bin[threadIdx.x]) : 0;
	if (threadIdx.x == 0) s.target[stream] = target;
}

// The same step with the points of a stream spread over several small CTAs (TS_WARPS points each, one per warp):
// an LK point is a long chain of dependent gathers (up to 4 levels x 30 iterations), so the time of the CTA-per-stream
// form is the time of the warp that drew the three slowest points.  Here every point is its own warp, the CTAs of all
// streams share the SMs dynamically, and the LAST CTA of a stream to finish (ticket counter) does the bookkeeping.
// Measured (r2B, 512 streams x 20 features, B200): CTAs of 4 point-warps at 8 per SM 104.4 us, 2 at 16 103.0, ONE warp per
// CTA at 32 per SM 95.7 us (at 24 per SM, 72 registers: 99.3): a finished warp no longer holds its CTA's slots while the
// slowest point of the CTA runs on (27 % of the stall samples of the 4-warp form sat at its closing barrier).
#ifndef AOF_TS_WARPS
#define AOF_TS_WARPS 1
#endif
#ifndef AOF_TS_MINB
#define AOF_TS_MINB 32
#endif
constexpr int TS_WARPS = AOF_TS_WARPS;

__global__ void __launch_bounds__(TS_WARPS * 32, AOF_TS_MINB)
lk_track_split_kernel(LkGeom g, LkPyr prev, LkPyr next, LkState s, int have_prev, int groups)
{
	pdl_enter();
	__shared__ LkWarpScratch s_sc[TS_WARPS];
	__shared__ int s_per_bin[LK_BINS];
	__shared__ int s_last, s_target;
	const int stream_i = blockIdx.x / groups, grp = blockIdx.x - stream_i * groups;
	const size_t stream = stream_i;
	const int N = g.n_feat;
	const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	float2 *pc = s.prev_corners + stream * N;
	uint8_t *ps = s.prev_status + stream * N;
	int *uv = s.update_vec + stream * N;
	float2 *fc = s.feat_cur + stream * N;
	float2 *lk_next = s.lk_next + stream * N;
	uint8_t *lk_stat = s.lk_stat + stream * N;
	const int npc = s.n_prev_corners[stream];
	const bool do_lk = have_prev && npc > 0;
	if (do_lk) {
		// only slots that were active can influence anything downstream; the reference tracks
		// the stale ones too (trackFeatures.cpp:98) but never reads them again before they are reassigned
		const int f = grp * TS_WARPS + warp;
		if (f < npc && uv[f] == 1) {
			const float2 p = pc[f];
			float ox, oy;
			const int st = lk_track_point(g, prev, next, stream, p.x, p.y, s_sc[warp], ox, oy);
			if (lane == 0) { lk_next[f] = make_float2(ox, oy); lk_stat[f] = (uint8_t)st; __threadfence(); }
		}
	}
	__syncthreads();
	if (threadIdx.x == 0) {
		__threadfence();
		s_last = atomicAdd(&s.ticket[stream], 1) == groups - 1;
	}
	for (int b = threadIdx.x; b < LK_BINS; b += blockDim.x) s_per_bin[b] = 0;
	if (threadIdx.x == 0) s_target = 0;
	__syncthreads();
	if (!s_last) return;
	__threadfence();
	if (threadIdx.x == 0) s.ticket[stream] = 0;
	for (int b = threadIdx.x; b < LK_BINS; b += blockDim.x) s.kp_count[stream * LK_BINS + b] = 0;
	// trackFeatures.cpp:77-125, one thread per slot; then the demand of initMorePoints (:144-183): bin = x * 4 + y
	int wants = 0;
	for (int i = threadIdx.x; i < N; i += blockDim.x) {
		int status = uv[i];
		int pst = status == 1;
		float2 feat = make_float2(-100.f, -100.f);
		float2 corner = pc[i];
		if (do_lk && i < npc) {
			if (pst) {
				corner = __ldcg(lk_next + i);                 // written by other CTAs of this launch: read at L2
				pc[i] = corner;
				if (!__ldcg(lk_stat + i)) pst = 0;
			}
			if (pst) {
				if (corner.x < 0.f || corner.x > (float)g.w[0] || corner.y < 0.f || corner.y > (float)g.h[0]) status = 0;
				else { feat = corner; status = 1; }
			} else if (status == 1) status = 0;
		}
		ps[i] = (uint8_t)pst;
		uv[i] = status;
		fc[i] = feat;
		wants += status == 2;
		if (i < npc && status == 1) {
			const int bx = (int)__fdiv_rn(corner.x, (float)g.bw), by = (int)__fdiv_rn(corner.y, (float)g.bh);
			if (bx < 4 && by < 4) atomicAdd(&s_per_bin[bx * 4 + by], 1);
		}
	}
	if (wants) atomicAdd(&s_target, wants);            // (a thread may own several slots when the CTA is narrower than N)
	__syncthreads();
	const int target = s_target;
	for (int b = threadIdx.x; b < LK_BINS; b += blockDim.x)
		s.bin_needed[stream * LK_BINS + b] = target ? max(0, g.tpb - s_per_bin[b]) : 0;
	if (threadIdx.x == 0) s.target[stream] = target;
}

cudaError_t lk_track_update_launch(const LkGeom &g, const LkPyr &prev, const LkPyr &next, const LkState &s, int n,
				   int have_prev, cudaStream_t st)
{
	if (n <= 0) return cudaSuccess;
	KernelTimer kt(KID_LK_TRACK_UPDATE, st, g.tun_track_v1 == 1 ? "lk_track_update_kernel" : "lk_track_split_kernel");
	if (g.tun_track_v1 == 1) lk_track_update_kernel<<<n, TU_WARPS * 32, 0, st>>>(g, prev, next, s, have_prev);
	else {
		const int groups = (g.n_feat + TS_WARPS - 1) / TS_WARPS;
		launch_k(g.tun_pdl != 0, lk_track_split_kernel, n * groups, TS_WARPS * 32, 0, st, g, prev, next, s, have_prev, groups);
	}
	count_launch();
	return cudaGetLastError();
}

// stateless LK over given points: one warp per (pair, point)
__global__ void __launch_bounds__(256)
lk_track_points_kernel(LkGeom g, LkPyr prev, LkPyr next, const float2 *__restrict__ pts, int n_pts, float2 *__restrict__ out,
		       uint8_t *__restrict__ status, int n)
{
	__shared__ LkWarpScratch s_sc[8];
	const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	const long long gw = (long long)blockIdx.x * 8 + warp;
	if (gw >= (long long)n * n_pts) return;
	const size_t stream = (size_t)(gw / n_pts);
	const float2 p = pts[gw];
	float ox, oy;
	const int st = lk_track_point(g, prev, next, stream, p.x, p.y, s_sc[warp], ox, oy);
	if (lane == 0) { out[gw] = make_float2(ox, oy); status[gw] = (uint8_t)st; }
}

cudaError_t lk_track_points_launch(const LkGeom &g, const LkPyr &prev, const LkPyr &next, const float2 *pts, int n_pts,
				   float2 *out, uint8_t *status, int n, cudaStream_t st)
{
	const long long total = (long long)n * n_pts;
	if (total <= 0) return cudaSuccess;
	KernelTimer kt(KID_LK_TRACK_POINTS, st);
	lk_track_points_kernel<<<(unsigned)((total + 7) / 8), 256, 0, st>>>(g, prev, next, pts, n_pts, out, status, n);
	count_launch();
	return cudaGetLastError();
}

// ---------------------------------------------------------------------------
// K4: FAST-9/16 score + 3x3 non-max suppression with cv::FAST semantics on each
// bin ROI as a stand-alone image.  CTA = 32 x 24 pixels of one bin
// (a 160 x 120 bin is exactly 5 x 5 tiles).
//
// The response of EVERY pixel is computed exactly and branch-free, two horizontally
// adjacent pixels per thread in u16x2 lanes: with D_k = c - ring_k + 256 (no borrow
// between the lanes), the response is  max_arcs(min_9 D) - 257  (dark arcs) or
// 255 - min_arcs(max_9 D)  (bright arcs); the sliding 9-windows are two levels of
// VIMNMX3.U16x2 (3-input packed min / max).  On textures where a third of the pixels
// are corners (the LK benchmark texture) a test-then-score scheme diverges on almost
// every warp; this one does 1/2.7 of the instructions of the scalar version.
// Materialised max3/min3 only: ptxas 12.9 drops the negation when max(m, -x) is
// folded into VIMNMX3 (tools/ptxas_vimnmx3_neg_repro.cu).
constexpr int FT_TW = 32, FT_TH = 24;
constexpr int FT_PW = FT_TW + 8, FT_PH = FT_TH + 8;   // pixel tile 40 x 32
constexpr int FT_SW = FT_TW + 2, FT_SH = FT_TH + 2;   // score tile 34 x 26
constexpr int FT_PAIRS = FT_SW / 2;                    // 17 pixel pairs per score row

__device__ __forceinline__ uint32_t mn3(uint32_t a, uint32_t b, uint32_t c) { return __vimin3_u16x2(a, b, c); }
__device__ __forceinline__ uint32_t mx3(uint32_t a, uint32_t b, uint32_t c) { return __vimax3_u16x2(a, b, c); }

// pairs A[r][j] = (px[2j], px[2j+1]), B[r][j] = (px[2j+1], px[2j+2]) as u16x2; `row` -> A or B row r,
// `j` the pair index of the CENTRE pair in B (centre columns are odd: c = 2j + 1)
__device__ __forceinline__ uint32_t fast_score_pair(const uint32_t (*A)[FT_PW / 2], const uint32_t (*B)[FT_PW / 2], int r, int j)
{
	const uint32_t cb = B[r][j] + 0x01000100u;
	uint32_t D[16];
	// ring offsets (dx, dy); odd dx -> even start column -> A, even dx -> B.  start column = 2j + 1 + dx
	D[0] = cb - B[r + 3][j];         D[1] = cb - A[r + 3][j + 1];   D[2] = cb - B[r + 2][j + 1];  D[3] = cb - A[r + 1][j + 2];
	D[4] = cb - A[r][j + 2];         D[5] = cb - A[r - 1][j + 2];   D[6] = cb - B[r - 2][j + 1];  D[7] = cb - A[r - 3][j + 1];
	D[8] = cb - B[r - 3][j];         D[9] = cb - A[r - 3][j];       D[10] = cb - B[r - 2][j - 1]; D[11] = cb - A[r - 1][j - 1];
	D[12] = cb - A[r][j - 1];        D[13] = cb - A[r + 1][j - 1];  D[14] = cb - B[r + 2][j - 1]; D[15] = cb - A[r + 3][j];
	uint32_t lo3[16], hi3[16];
#pragma unroll
	for (int k = 0; k < 16; k++) {
		lo3[k] = mn3(D[k], D[(k + 1) & 15], D[(k + 2) & 15]);
		hi3[k] = mx3(D[k], D[(k + 1) & 15], D[(k + 2) & 15]);
	}
	uint32_t lo9[16], hi9[16];
#pragma unroll
	for (int k = 0; k < 16; k++) {
		lo9[k] = mn3(lo3[k], lo3[(k + 3) & 15], lo3[(k + 6) & 15]);
		hi9[k] = mx3(hi3[k], hi3[(k + 3) & 15], hi3[(k + 6) & 15]);
	}
	uint32_t dark = mx3(lo9[0], lo9[1], lo9[2]), bright = mn3(hi9[0], hi9[1], hi9[2]);
#pragma unroll
	for (int k = 3; k < 15; k += 2) { dark = mx3(dark, lo9[k], lo9[k + 1]); bright = mn3(bright, hi9[k], hi9[k + 1]); }
	dark = __vmaxu2(dark, lo9[15]); bright = __vminu2(bright, hi9[15]);
	return __vmaxu2(dark, 0x02000200u - bright); // per lane: m + 256
}

constexpr int FT_THREADS = 224; // 442 pixel pairs per tile = 2 rounds of 221

// ALIGNED: W % 16 == 0 and the frame base is 4-byte aligned, so every tile row starts on a word
template <bool ALIGNED>
__global__ void __launch_bounds__(FT_THREADS)
lk_fast_kernel(LkGeom g, const uint8_t *__restrict__ l0, LkState s, uint8_t *__restrict__ score_map, int tiles_x, int tiles_y)
{
	__shared__ __align__(8) uint32_t s_A[FT_PH][FT_PW / 2], s_B[FT_PH][FT_PW / 2];
	__shared__ uint8_t s_score[FT_SH][FT_SW + 2];
	const size_t stream = blockIdx.y;
	const int per_bin = tiles_x * tiles_y;
	const int bin = blockIdx.x / per_bin, t = blockIdx.x - bin * per_bin;
	if (!score_map && s.bin_needed[stream * LK_BINS + bin] == 0) return;
	const int ty = t / tiles_x, tx = t - ty * tiles_x;
	const int W = g.w[0], H = g.h[0], bw = g.bw, bh = g.bh;
	const int c0 = (bin >> 2) * bw, r0 = (bin & 3) * bh;      // bin = x * 4 + y
	const int x0 = tx * FT_TW, y0 = ty * FT_TH;                // tile origin inside the bin
	const uint8_t *img = l0 + stream * (size_t)W * H;
	const int tid = threadIdx.x;
	// pixels (x0 - 4 .. x0 + 35, y0 - 4 .. y0 + 27) of the bin as u16x2 pairs.  Pixels outside the bin (read from
	// the neighbouring bins, clamped at the image border) only ever reach positions that are not tested.
	for (int i = tid; i < FT_PH * (FT_PW / 4); i += FT_THREADS) {
		const int r = i / (FT_PW / 4), q = i - r * (FT_PW / 4);
		const int y = min(max(r0 + y0 - 4 + r, 0), H - 1);
		uint32_t p0, p1, p2, p3, p4;
		if (ALIGNED) {
			const uint32_t *row = reinterpret_cast<const uint32_t *>(img + (size_t)y * W);
			const int wq = (c0 + x0 - 4) / 4 + q, wmax = W / 4 - 1;
			const uint32_t w0 = __ldg(row + min(max(wq, 0), wmax)), w1 = __ldg(row + min(max(wq + 1, 0), wmax));
			p0 = w0 & 0xFF; p1 = (w0 >> 8) & 0xFF; p2 = (w0 >> 16) & 0xFF; p3 = w0 >> 24; p4 = w1 & 0xFF;
		} else {
			const uint8_t *row = img + (size_t)y * W;
			const int x = c0 + x0 - 4 + 4 * q;
			p0 = row[min(max(x, 0), W - 1)]; p1 = row[min(max(x + 1, 0), W - 1)]; p2 = row[min(max(x + 2, 0), W - 1)];
			p3 = row[min(max(x + 3, 0), W - 1)]; p4 = row[min(max(x + 4, 0), W - 1)];
		}
		*reinterpret_cast<uint2 *>(&s_A[r][2 * q]) = make_uint2(p0 | (p1 << 16), p2 | (p3 << 16));
		*reinterpret_cast<uint2 *>(&s_B[r][2 * q]) = make_uint2(p1 | (p2 << 16), p3 | (p4 << 16));
	}
	__syncthreads();
	// scores of columns x0 - 1 .. x0 + 32 (tile columns 3 .. 36), rows y0 - 1 .. y0 + 24 (tile rows 3 .. 28)
	const uint32_t lim = 256u + (uint32_t)g.fast_thr;
	for (int i = tid; i < FT_SH * FT_PAIRS; i += FT_THREADS) {
		const int r = i / FT_PAIRS, jp = i - r * FT_PAIRS;
		const int y = y0 - 1 + r, x = x0 - 1 + 2 * jp;
		uint32_t sc0 = 0, sc1 = 0;
		if (y >= 3 && y < bh - 3 && x + 1 >= 3 && x < bw - 3) {
			const uint32_t m = fast_score_pair(s_A, s_B, r + 3, jp + 1); // centre columns 2 (jp + 1) + 1 = 3 + 2 jp
			const uint32_t m0 = m & 0xFFFFu, m1 = m >> 16;
			if (x >= 3 && m0 > lim) sc0 = m0 - 257u;
			if (x + 1 < bw - 3 && m1 > lim) sc1 = m1 - 257u;
		}
		*reinterpret_cast<uint16_t *>(&s_score[r][2 * jp]) = (uint16_t)(sc0 | (sc1 << 8));
	}
	__syncthreads();
	for (int i = tid; i < FT_TH * FT_TW; i += FT_THREADS) {
		const int r = i >> 5, c = i & 31;
		const int y = y0 + r, x = x0 + c;
		const int v = s_score[r + 1][c + 1];
		if (!v || x >= bw || y >= bh) continue;
		const bool keep = v > s_score[r][c] && v > s_score[r][c + 1] && v > s_score[r][c + 2] && v > s_score[r + 1][c] &&
				  v > s_score[r + 1][c + 2] && v > s_score[r + 2][c] && v > s_score[r + 2][c + 1] &&
				  v > s_score[r + 2][c + 2];
		if (!keep) continue;
		if (score_map) score_map[stream * (size_t)W * H + (size_t)(r0 + y) * W + c0 + x] = (uint8_t)v;
		else {
			const int slot = atomicAdd(&s.kp_count[stream * LK_BINS + bin], 1);
			if (slot < g.kp_cap)
				s.kp[(stream * LK_BINS + bin) * (size_t)g.kp_cap + slot] = ((uint32_t)v << 24) | ((uint32_t)y << 12) | (uint32_t)x;
		}
	}
}

cudaError_t lk_fast_launch(const LkGeom &g, const uint8_t *l0, const LkState &s, int n, uint8_t *score_map, cudaStream_t st)
{
	if (n <= 0) return cudaSuccess;
	const int tiles_x = (g.bw + FT_TW - 1) / FT_TW, tiles_y = (g.bh + FT_TH - 1) / FT_TH;
	dim3 grid(tiles_x * tiles_y * LK_BINS, n);
	const bool band = lk_fast_band_supported(g, l0);
	KernelTimer kt(KID_LK_FAST, st, !band ? "lk_fast_kernel" : g.bw == 160 ? "lk_fast_band_kernel<320,84>" : "lk_fast_band_kernel");
	if (band) {
		count_launch();
		return lk_fast_band_launch(g, l0, s, n, score_map, st);
	}
	if (g.w[0] % 16 == 0 && (reinterpret_cast<uintptr_t>(l0) & 3) == 0)
		lk_fast_kernel<true><<<grid, FT_THREADS, 0, st>>>(g, l0, s, score_map, tiles_x, tiles_y);
	else
		lk_fast_kernel<false><<<grid, FT_THREADS, 0, st>>>(g, l0, s, score_map, tiles_x, tiles_y);
	count_launch();
	return cudaGetLastError();
}

// ---------------------------------------------------------------------------
// K5 + K7: initMorePoints selection (trackFeatures.cpp:185-362) and the calcFlow
// epilogue (flow_opencv.cpp:121-209, optical_flow.cpp:51-96).  One CTA per stream.
//
// The reference walks each bin's keypoints in descending response (ties: raster
// order, the documented total order) and accepts the first `needed` that keep a
// Chebyshev distance >= dist from the active points and from this bin's picks.
// "The first acceptable one in sorted order" is an arg-max under the gate, and a
// point rejected once stays rejected because the set it is tested against only
// grows -- so each pick is one masked warp/block arg-max, no sort.
#ifndef AOF_SEL_THREADS
#define AOF_SEL_THREADS 256
#endif
constexpr int SEL_THREADS = AOF_SEL_THREADS;
constexpr int SEL_MAX_TPB = LK_MAX_TPB;
constexpr int SEL_MAX_GOOD = LK_BINS * SEL_MAX_TPB + LK_MAX_FEATURES;
constexpr int SEL_STAGE = 8192 / (SEL_THREADS / 32);          // keypoints of a bin staged in shared memory (32 KB per CTA)

__device__ __forceinline__ bool cheb_close(int ax, int ay, int bx, int by, int dist)
{
	return abs(ax - bx) < dist && abs(ay - by) < dist;
}

// cv::undistortPoints(src, dst, K, D) -- 5 fixed-point iterations in double, no R / P -- followed by the
// re-projection of flow_opencv.cpp:113-118 in float
__device__ float2 lk_undistort(const LkCamera &c, float2 pt)
{
	const double fx = c.fx, fy = c.fy, cx = c.cx, cy = c.cy;
	const double k1 = c.k1, k2 = c.k2, p1 = c.p1, p2 = c.p2, k3 = c.k3;
	const double ifx = 1. / fx, ify = 1. / fy;
	const double u = pt.x, v = pt.y;
	double x = (u - cx) * ifx, y = (v - cy) * ify;
	const double x0 = x, y0 = y;
	for (int j = 0; j < 5; j++) {
		const double r2 = x * x + y * y;
		const double icdist = 1. / (1 + ((k3 * r2 + k2) * r2 + k1) * r2);
		if (icdist < 0) { x = (u - cx) * ifx; y = (v - cy) * ify; break; }
		const double dx = 2 * p1 * x * y + p2 * (r2 + 2 * x * x);
		const double dy = p1 * (r2 + 2 * y * y) + 2 * p2 * x * y;
		x = (x0 - dx) * icdist; y = (y0 - dy) * icdist;
	}
	return make_float2((float)x * c.fx + c.cx, (float)y * c.fy + c.cy);
}

__global__ void __launch_bounds__(SEL_THREADS)
lk_select_epilogue_kernel(LkGeom g, LkState s, const uint32_t *__restrict__ t_us, aof_flow_out *__restrict__ out,
			  aof_lk_raw *__restrict__ raw, int have_feat_prev, int output_rate, float conf, float focal_x,
			  float focal_y, LkCamera cam, int mode)
{
	pdl_enter();
	__shared__ int s_ex[LK_MAX_FEATURES], s_ey[LK_MAX_FEATURES];
	__shared__ int s_nex, s_ngood, s_done;
	__shared__ int s_needed[LK_BINS], s_cnt[LK_BINS], s_npick[LK_BINS];
	__shared__ int s_px[LK_BINS][SEL_MAX_TPB], s_py[LK_BINS][SEL_MAX_TPB];
	__shared__ int s_gx[SEL_MAX_GOOD], s_gy[SEL_MAX_GOOD];
	__shared__ unsigned long long s_red[SEL_THREADS / 32];
	__shared__ int s_bx[SEL_THREADS / 32][LK_MAX_FEATURES], s_by[SEL_THREADS / 32][LK_MAX_FEATURES]; // per-warp: active points near the bin
	__shared__ uint32_t s_key[SEL_THREADS / 32][SEL_STAGE];     // per-warp: keys of the bin's staged keypoints
	const size_t stream = blockIdx.x;
	const int N = g.n_feat, W = g.w[0];
	const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
	// LK_SEL_FLAGGED: the complete path for the streams the lazy pass flagged, nothing for the others
	if (mode == LK_SEL_MASKED && blockIdx.x == 0 && tid == 0) s.n_work[0] = 0;   // the masked FAST launch before this one has consumed it
	if (mode == LK_SEL_FLAGGED) {
		if (blockIdx.x == 0 && tid == 0) s.n_full[0] = 0;      // the flagged FAST launch before this one has consumed it
		if (!s.need_full[stream]) return;
	}
	float2 *pc = s.prev_corners + stream * N;
	uint8_t *ps = s.prev_status + stream * N;
	int *uv = s.update_vec + stream * N;
	float2 *fc = s.feat_cur + stream * N;
	float2 *fp = s.feat_prev + stream * N;
	const int target = s.target[stream];
	int fast_bins = 0;
	// the per-stream state is staged in shared memory: the order-dependent parts below run on one thread and
	// would otherwise be chains of dependent global loads
	__shared__ float2 s_pc[LK_MAX_FEATURES], s_fc[LK_MAX_FEATURES], s_fp[LK_MAX_FEATURES];
	__shared__ int s_uv[LK_MAX_FEATURES];
	__shared__ uint8_t s_ps[LK_MAX_FEATURES];
	__shared__ int s_fast_bins, s_lazy_fail;
	if (tid == 0) s_lazy_fail = 0;
	for (int i = tid; i < N; i += SEL_THREADS) { s_pc[i] = pc[i]; s_fc[i] = fc[i]; s_fp[i] = fp[i]; s_uv[i] = uv[i]; s_ps[i] = ps[i]; }
	__syncthreads();

	if (target > 0) { // uniform over the CTA
		const int npc = s.n_prev_corners[stream];
		if (tid == 0) {
			int n = 0;
			for (int j = 0; j < npc; j++)
				if (s_ps[j]) { s_ex[n] = (int)s_pc[j].x; s_ey[n] = (int)s_pc[j].y; n++; }
			s_nex = n;
		}
		if (tid < LK_BINS) {
			s_needed[tid] = s.bin_needed[stream * LK_BINS + tid];
			s_cnt[tid] = min(s.kp_count[stream * LK_BINS + tid], g.kp_cap);
			s_npick[tid] = 0;
		}
		__syncthreads();
		const int nex = s_nex;
		// ---- stage 1 with the picks of lk_fast_pick_kernel (lazy FAST): they ARE what the loop below would find
		if (mode == LK_SEL_LAZY && tid < LK_BINS) {
			const int b = tid;
			int np = s_needed[b] > 0 ? s.lz_npick[stream * LK_BINS + b] : 0;
			if (np < 0) { s_lazy_fail = 1; np = 0; }           // the bin overflowed the lazy kernel's buffers
			for (int j = 0; j < np; j++) {
				const uint32_t e = s.lz_pick[(stream * LK_BINS + b) * LK_MAX_TPB + j];
				s_px[b][j] = (int)(e & 0xFFFFu); s_py[b][j] = (int)(e >> 16);
			}
			s_npick[b] = np;
		}
		// ---- stage 1: one warp per bin
		for (int b = warp; b < LK_BINS && mode != LK_SEL_LAZY; b += SEL_THREADS / 32) {
			if (s_needed[b] <= 0) continue;
			const int needed = s_needed[b], cnt = s_cnt[b];
			const int c0 = (b >> 2) * g.bw, r0 = (b & 3) * g.bh;
			const uint32_t *list = s.kp + (stream * LK_BINS + b) * (size_t)g.kp_cap;
			// only the active points whose exclusion square reaches into this bin can reject one of its keypoints
			// (about a third of them for 4 x 4 bins and dist = half a bin): compact them once per bin
			int nexb = 0;
			for (int j0 = 0; j0 < nex; j0 += 32) {
				const int j = j0 + lane;
				const int ex = j < nex ? s_ex[j] : 0, ey = j < nex ? s_ey[j] : 0;
				const bool near = j < nex && ex > c0 - g.dist && ex < c0 + g.bw - 1 + g.dist && ey > r0 - g.dist &&
						  ey < r0 + g.bh - 1 + g.dist;
				const uint32_t m = __ballot_sync(0xFFFFFFFFu, near);
				if (near) {
					const int k = nexb + __popc(m & ((1u << lane) - 1u));
					s_bx[warp][k] = ex; s_by[warp][k] = ey;
				}
				nexb += __popc(m);
			}
			__syncwarp();
			// The bin's keypoints are staged in shared memory once (the first SEL_STAGE of them; a bin of a 640 x 480 frame
			// holds a few hundred), turned into keys (score desc, raster asc) and filtered against the active points: a
			// keypoint rejected once stays rejected, so every pick is then one max over the keys and one pass that
			// clears the keys inside the pick's exclusion square.  Entries beyond SEL_STAGE are tested from global memory.
			uint32_t *key = s_key[warp];
			const int cs = min(cnt, SEL_STAGE);
#pragma unroll 4
			for (int e = lane; e < cs; e += 32) key[e] = list[e];
			__syncwarp();
			for (int e = lane; e < cs; e += 32) {
				const uint32_t v = key[e];
				const int x = c0 + (int)(v & 0xFFFu), y = r0 + (int)((v >> 12) & 0xFFFu);
				bool ok = true;
				for (int j = 0; j < nexb && ok; j++) ok = !cheb_close(s_bx[warp][j], s_by[warp][j], x, y, g.dist);
				key[e] = ok ? (v & 0xFF000000u) | (0xFFFFFFu - (v & 0xFFFFFFu)) : 0u;
			}
			__syncwarp();
			int npick = 0;
			while (npick < needed) {
				uint32_t best = 0;
				for (int e = lane; e < cs; e += 32) best = max(best, key[e]);
				for (int e = SEL_STAGE + lane; e < cnt; e += 32) {
					const uint32_t v = list[e];
					const uint32_t k = (v & 0xFF000000u) | (0xFFFFFFu - (v & 0xFFFFFFu));
					if (k <= best) continue;
					const int x = c0 + (int)(v & 0xFFFu), y = r0 + (int)((v >> 12) & 0xFFFu);
					bool ok = true;
					for (int j = 0; j < nexb && ok; j++) ok = !cheb_close(s_bx[warp][j], s_by[warp][j], x, y, g.dist);
					for (int j = 0; j < npick && ok; j++) ok = !cheb_close(s_px[b][j], s_py[b][j], x, y, g.dist);
					if (ok) best = k;
				}
				best = __reduce_max_sync(0xFFFFFFFFu, best);
				if (!best) break;
				const uint32_t li = 0xFFFFFFu - (best & 0xFFFFFFu);
				const int px = c0 + (int)(li & 0xFFFu), py = r0 + (int)(li >> 12);
				if (lane == 0) { s_px[b][npick] = px; s_py[b][npick] = py; }
				npick++;
				if (npick < needed)
					for (int e = lane; e < cs; e += 32) {
						const uint32_t k = key[e];
						if (!k) continue;
						const uint32_t l2 = 0xFFFFFFu - (k & 0xFFFFFFu);
						if (cheb_close(px, py, c0 + (int)(l2 & 0xFFFu), r0 + (int)(l2 >> 12), g.dist)) key[e] = 0u;
					}
				__syncwarp();
			}
			if (lane == 0) s_npick[b] = npick;
		}
		__syncthreads();
		// ---- stage 2: concatenate in bin order (x outer, y inner == bin index), thin if too many (every index i with
		// i % step == 0 is dropped, trackFeatures.cpp:253-264).  Lane b of warp 0 places the picks of bin b: its offset is
		// a prefix sum of the pick counts, and a kept index i lands at i - (i / step + 1).
		if (warp == 0) {
			const int np = lane < LK_BINS ? s_npick[lane] : 0;
			int incl = np;
#pragma unroll
			for (int o = 1; o < 32; o <<= 1) {
				const int v = __shfl_up_sync(0xFFFFFFFFu, incl, o);
				if (lane >= o) incl += v;
			}
			const int n = __shfl_sync(0xFFFFFFFFu, incl, 31), off = incl - np;
			const bool thin = n > target;
			const int step = thin ? target / (n - target) + 2 : 1;
			for (int j = 0; j < np; j++) {
				const int i = off + j;
				if (thin && i % step == 0) continue;
				const int m = thin ? i - (i / step + 1) : i;
				s_gx[m] = s_px[lane][j]; s_gy[m] = s_py[lane][j];
			}
			if (lane == 0) {
				const int kept = thin ? n - (n + step - 1) / step : n;
				s_ngood = kept;
				s_done = kept >= target;
			}
		}
		__syncthreads();
		if ((mode == LK_SEL_LAZY || mode == LK_SEL_MASKED) && (!s_done || s_lazy_fail)) {
			// the second-chance pass needs every keypoint of every processed bin: hand the stream to the complete path
			// (lk_fast_band_flagged_kernel, then this kernel in LK_SEL_FLAGGED mode); no state has been written yet
			// (MASKED: the lists hold the free region's keypoints only; the complete pass appends from zero)
			if (mode == LK_SEL_MASKED && tid < LK_BINS) s.kp_count[stream * LK_BINS + tid] = 0;
			if (tid == 0) {
				s.need_full[stream] = 1;
				atomicAdd(&s.lz_stats[3], 1ull);
				s.full_list[atomicAdd(&s.n_full[0], 1)] = (int)blockIdx.x;
			}
			return;
		}
		// ---- stage 3: second chance over every keypoint of the processed bins at dist / 2
		const int dist2 = g.dist / 2;
		while (!s_done) {
			const int ngood = s_ngood;
			unsigned long long best = 0;
			for (int b = 0; b < LK_BINS; b++) {
				if (s_needed[b] <= 0) continue;
				const int c0 = (b >> 2) * g.bw, r0 = (b & 3) * g.bh, cnt = s_cnt[b];
				const uint32_t *list = s.kp + (stream * LK_BINS + b) * (size_t)g.kp_cap;
				for (int e = tid; e < cnt; e += SEL_THREADS) {
					const uint32_t v = list[e];
					const int y = r0 + (int)((v >> 12) & 0xFFFu), x = c0 + (int)(v & 0xFFFu);
					const unsigned long long key = ((unsigned long long)(v >> 24) << 32) | (0xFFFFFFFFu - (uint32_t)(y * W + x));
					if (key <= best) continue;
					bool ok = true;
					for (int j = 0; j < nex && ok; j++) ok = !cheb_close(s_ex[j], s_ey[j], x, y, dist2);
					for (int j = 0; j < ngood && ok; j++) ok = !cheb_close(s_gx[j], s_gy[j], x, y, dist2);
					if (ok) best = key;
				}
			}
#pragma unroll
			for (int o = 16; o; o >>= 1) {
				const unsigned long long other = __shfl_xor_sync(0xFFFFFFFFu, best, o);
				best = other > best ? other : best;
			}
			if (lane == 0) s_red[warp] = best;
			__syncthreads();
			if (tid == 0) {
				unsigned long long m = 0;
				for (int k = 0; k < SEL_THREADS / 32; k++) m = s_red[k] > m ? s_red[k] : m;
				if (!m) s_done = 1;
				else {
					const uint32_t gi = 0xFFFFFFFFu - (uint32_t)(m & 0xFFFFFFFFu);
					s_gx[ngood] = (int)(gi % (uint32_t)W); s_gy[ngood] = (int)(gi / (uint32_t)W);
					s_ngood = ngood + 1;
					if (ngood + 1 == target) s_done = 1;
				}
			}
			__syncthreads();
		}
		if (tid < LK_BINS) fast_bins = s_needed[tid] > 0;
	}
	if (tid < 32) {
		fast_bins = __reduce_add_sync(0xFFFFFFFFu, fast_bins);
		if (tid == 0) s_fast_bins = fast_bins;
	}
	if (tid == 0) {
		// ---- stage 4: slot assignment (trackFeatures.cpp:306-362)
		if (target > 0) {
			const int ngood = s_ngood;
			if (!ngood) {
				for (int i = 0; i < N; i++)
					if (s_uv[i] == 2) s_uv[i] = 0;
			} else {
				if (s.n_prev_corners[stream] < N) {
					for (int i = s.n_prev_corners[stream]; i < N; i++) s_pc[i] = make_float2(0.f, 0.f);
					s.n_prev_corners[stream] = N;
				}
				int m = 0;
				for (int i = 0; i < N; i++) {
					if (s_uv[i] != 2) continue;
					if (m < ngood) {
						const float2 p = make_float2((float)s_gx[m], (float)s_gy[m]);
						s_pc[i] = p; s_fc[i] = p; s_ps[i] = 1;
						m++;
					} else s_uv[i] = 0;
				}
			}
		}

		// ---- optional undistortion of every slot, placeholders included (flow_opencv.cpp:108-119)
		if (cam.on)
			for (int i = 0; i < N; i++) s_fc[i] = lk_undistort(cam, s_fc[i]);

		// ---- calcFlow statistics (flow_opencv.cpp:121-180)
		int meancount = 0, active = 0;
		float mx = 0.f, my = 0.f;
		if (have_feat_prev) {
			for (int i = 0; i < N; i++)
				if (s_uv[i] == 1) { mx += s_fc[i].x - s_fp[i].x; my += s_fc[i].y - s_fp[i].y; meancount++; }
			active = meancount;
			if (meancount) {
				mx /= (float)meancount; my /= (float)meancount;
				float sx = 0.f, sy = 0.f;
				for (int i = 0; i < N; i++)
					if (s_uv[i] == 1) {
						const float ex = s_fc[i].x - s_fp[i].x - mx, ey = s_fc[i].y - s_fp[i].y - my;
						sx = (float)((double)sx + (double)ex * (double)ex); // pow(float, 2) runs in double
						sy = (float)((double)sy + (double)ey * (double)ey);
					}
				sx = sqrtf(sx / (float)meancount); sy = sqrtf(sy / (float)meancount);
				float tx = 0.f, ty = 0.f;
				meancount = 0;
				for (int i = 0; i < N; i++)
					if (s_uv[i] == 1) {
						const float fx = s_fc[i].x - s_fp[i].x, fy = s_fc[i].y - s_fp[i].y;
						if (fabsf(fx - mx) < sx * conf && fabsf(fy - my) < sy * conf) { tx += fx; ty += fy; meancount++; }
						else s_uv[i] = 0;
					}
				if (meancount) { mx = tx / (float)meancount; my = ty / (float)meancount; }
			}
		}
		const int q0 = (int)round(255.0 * (double)meancount / (double)N);
		LimiterState l = s.lim[stream];
		aof_flow_out o;
		float fx = mx, fy = my;
		o.quality = limit_rate(l, output_rate, q0, t_us[stream], o.dt_us, fx, fy);
		o.flow_x = (float)atan2((double)fx, (double)focal_x);
		o.flow_y = (float)atan2((double)fy, (double)focal_y);
		s.lim[stream] = l;
		out[stream] = o;
		if (raw) {
			aof_lk_raw r;
			r.px_x = mx; r.px_y = my; r.inliers = meancount; r.active = active; r.fast_bins = s_fast_bins; r.quality = q0;
			raw[stream] = r;
		}
	}
	__syncthreads();
	if (mode == LK_SEL_FLAGGED && tid == 0) s.need_full[stream] = 0;
	// features_previous = features_current; status machine (flow_opencv.cpp:183-196); write the state back
	for (int i = tid; i < N; i += SEL_THREADS) {
		int u = s_uv[i];
		if (u == 2) u = 1;
		else if (u == 0) u = 2;
		uv[i] = u;
		fp[i] = s_fc[i];
		fc[i] = s_fc[i];
		pc[i] = s_pc[i];
		ps[i] = s_ps[i];
	}
}

cudaError_t lk_select_epilogue_launch(const LkGeom &g, const LkState &s, const uint32_t *t_us, aof_flow_out *out,
				      aof_lk_raw *raw, int n, int have_feat_prev, int output_rate, float conf, float fx,
				      float fy, const LkCamera &cam, cudaStream_t st, int mode)
{
	if (n <= 0) return cudaSuccess;
	KernelTimer kt(KID_LK_SELECT_EPILOGUE, st);
	launch_k(g.tun_pdl != 0, lk_select_epilogue_kernel, n, SEL_THREADS, 0, st, g, s, t_us, out, raw, have_feat_prev, output_rate, conf, fx, fy, cam, mode);
	count_launch();
	return cudaGetLastError();
}

} // namespace aof
