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
	return __ll2float_rn(((long long)hi << 16) + lo);
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
		for (int i = lane; i < 144; i += 32) {
			const int r = i / 12, c = i - r * 12;
			sc.I[i] = I[(size_t)reflect101(ipy - 1 + r, h) * w + reflect101(ipx - 1 + c, w)];
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
		for (int j = 0; j < 30; j++) {
			const int inx = (int)floorf(cx), iny = (int)floorf(cy);
			if (inx < -LK_WIN || inx >= w || iny < -LK_WIN || iny >= h) {
				if (level == 0) status = 0;
				break;
			}
			q14_weights(cx - (float)inx, cy - (float)iny, w00, w01, w10, w11);
			__syncwarp();
			for (int i = lane; i < 100; i += 32) {
				const int r = i / 10, c = i - r * 10;
				sc.J[i] = __ldg(&J[(size_t)reflect101(iny + r, h) * w + reflect101(inx + c, w)]);
			}
			__syncwarp();
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
			if ((double)ddx * (double)ddx + (double)ddy * (double)ddy <= 0.01 * 0.01) break;
			if (j > 0 && (double)fabsf(ddx + pdx) < 0.01 && (double)fabsf(ddy + pdy) < 0.01) {
				nx -= ddx * 0.5f; ny -= ddy * 0.5f;
				break;
			}
			pdx = ddx; pdy = ddy;
		}
	}
	out_x = nx; out_y = ny;
	return status;
}

// ---------------------------------------------------------------------------
// K6 driver + trackFeatures bookkeeping + per-bin demand.  One CTA per stream.
constexpr int TU_WARPS = 16;

__global__ void __launch_bounds__(TU_WARPS * 32)
lk_track_update_kernel(LkGeom g, LkPyr prev, LkPyr next, LkState s, int have_prev)
{
	__shared__ LkWarpScratch s_sc[TU_WARPS];
	__shared__ float2 s_next[LK_MAX_FEATURES];
	__shared__ uint8_t s_lk_status[LK_MAX_FEATURES];
	const size_t stream = blockIdx.x;
	const int N = g.n_feat;
	const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	float2 *pc = s.prev_corners + stream * N;
	uint8_t *ps = s.prev_status + stream * N;
	int *uv = s.update_vec + stream * N;
	float2 *fc = s.feat_cur + stream * N;
	const int npc = s.n_prev_corners[stream];
	const bool do_lk = have_prev && npc > 0;
	if (do_lk) {
		// only slots that were active can influence anything downstream; the reference tracks
		// the stale ones too (trackFeatures.cpp:98) but never reads them again before they are reassigned
		for (int f = warp; f < npc; f += TU_WARPS) {
			if (uv[f] != 1) continue;
			const float2 p = pc[f];
			float ox, oy;
			const int st = lk_track_point(g, prev, next, stream, p.x, p.y, s_sc[warp], ox, oy);
			if (lane == 0) { s_next[f] = make_float2(ox, oy); s_lk_status[f] = (uint8_t)st; }
		}
	}
	__syncthreads();
	// trackFeatures.cpp:77-125, one thread per slot
	for (int i = threadIdx.x; i < N; i += blockDim.x) {
		int status = uv[i];
		int pst = status == 1;
		float2 feat = make_float2(-100.f, -100.f);
		if (do_lk && i < npc) {
			if (pst) {
				pc[i] = s_next[i];
				if (!s_lk_status[i]) pst = 0;
			}
			if (pst) {
				const float2 c = s_next[i];
				if (c.x < 0.f || c.x > (float)g.w[0] || c.y < 0.f || c.y > (float)g.h[0]) status = 0;
				else { feat = c; status = 1; }
			} else if (status == 1) status = 0;
		}
		ps[i] = (uint8_t)pst;
		uv[i] = status;
		fc[i] = feat;
	}
	__syncthreads();
	// initMorePoints demand (trackFeatures.cpp:144-183): bin = x * 4 + y
	if (threadIdx.x < LK_BINS) s.kp_count[stream * LK_BINS + threadIdx.x] = 0;
	if (threadIdx.x == 0) {
		int target = 0;
		int per_bin[LK_BINS];
		for (int b = 0; b < LK_BINS; b++) per_bin[b] = 0;
		for (int i = 0; i < N; i++) target += uv[i] == 2;
		for (int i = 0; i < npc; i++)
			if (uv[i] == 1) {
				const int bx = (int)__fdiv_rn(pc[i].x, (float)g.bw), by = (int)__fdiv_rn(pc[i].y, (float)g.bh);
				if (bx >= 4 || by >= 4) continue;
				per_bin[bx * 4 + by]++;
			}
		for (int b = 0; b < LK_BINS; b++)
			s.bin_needed[stream * LK_BINS + b] = target ? max(0, g.tpb - per_bin[b]) : 0;
		s.target[stream] = target;
	}
}

cudaError_t lk_track_update_launch(const LkGeom &g, const LkPyr &prev, const LkPyr &next, const LkState &s, int n,
				   int have_prev, cudaStream_t st)
{
	if (n <= 0) return cudaSuccess;
	KernelTimer kt(KID_LK_TRACK_UPDATE, st);
	lk_track_update_kernel<<<n, TU_WARPS * 32, 0, st>>>(g, prev, next, s, have_prev);
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
// bin ROI as a stand-alone image.  CTA = 32 x 16 pixels of one bin.
constexpr int FT_TW = 32, FT_TH = 16;

__device__ __forceinline__ bool has_arc9(unsigned m)
{
	m |= m << 16;
	const unsigned a = m & (m >> 1), b = a & (a >> 2), c = b & (b >> 4);
	return (c & (m >> 8) & 0xFFFFu) != 0;
}

// p -> centre pixel in a shared tile of row pitch `pitch`; returns the response (0 = not a corner)
__device__ __forceinline__ int fast_score(const uint8_t *p, int pitch, int thr)
{
	const int c = p[0];
	int d[16];
	d[0] = c - p[3 * pitch];      d[1] = c - p[3 * pitch + 1];  d[2] = c - p[2 * pitch + 2];  d[3] = c - p[pitch + 3];
	d[4] = c - p[3];              d[5] = c - p[-pitch + 3];     d[6] = c - p[-2 * pitch + 2]; d[7] = c - p[-3 * pitch + 1];
	d[8] = c - p[-3 * pitch];     d[9] = c - p[-3 * pitch - 1]; d[10] = c - p[-2 * pitch - 2]; d[11] = c - p[-pitch - 3];
	d[12] = c - p[-3];            d[13] = c - p[pitch - 3];     d[14] = c - p[2 * pitch - 2]; d[15] = c - p[3 * pitch - 1];
	unsigned mp = 0, mn = 0;
#pragma unroll
	for (int k = 0; k < 16; k++) { mp |= (unsigned)(d[k] > thr) << k; mn |= (unsigned)(d[k] < -thr) << k; }
	if (!has_arc9(mp) && !has_arc9(mn)) return 0;
	// exact response: max over the 16 arcs of max(min d, min e), e = -d, minus 1.
	// (e is materialised: ptxas 12.9 drops the negation when it folds max(m, -hi9) into VIMNMX3 on sm_100a.)
	int e[16];
#pragma unroll
	for (int k = 0; k < 16; k++) e[k] = -d[k];
	int d2[16], e2[16], d4[16], e4[16];
#pragma unroll
	for (int k = 0; k < 16; k++) { d2[k] = min(d[k], d[(k + 1) & 15]); e2[k] = min(e[k], e[(k + 1) & 15]); }
#pragma unroll
	for (int k = 0; k < 16; k++) { d4[k] = min(d2[k], d2[(k + 2) & 15]); e4[k] = min(e2[k], e2[(k + 2) & 15]); }
	int m = -1000;
#pragma unroll
	for (int k = 0; k < 16; k++) {
		const int d9 = min(min(d4[k], d4[(k + 4) & 15]), d[(k + 8) & 15]);
		const int e9 = min(min(e4[k], e4[(k + 4) & 15]), e[(k + 8) & 15]);
		m = max(m, max(d9, e9));
	}
	return m - 1;
}

__global__ void __launch_bounds__(256)
lk_fast_kernel(LkGeom g, const uint8_t *__restrict__ l0, LkState s, uint8_t *__restrict__ score_map, int tiles_x, int tiles_y)
{
	__shared__ uint8_t s_pix[FT_TH + 8][FT_TW + 8];
	__shared__ uint8_t s_score[FT_TH + 2][FT_TW + 2];
	const size_t stream = blockIdx.y;
	const int per_bin = tiles_x * tiles_y;
	const int bin = blockIdx.x / per_bin, t = blockIdx.x - bin * per_bin;
	if (!score_map && s.bin_needed[stream * LK_BINS + bin] == 0) return;
	const int ty = t / tiles_x, tx = t - ty * tiles_x;
	const int W = g.w[0], H = g.h[0], bw = g.bw, bh = g.bh;
	const int c0 = (bin >> 2) * bw, r0 = (bin & 3) * bh;      // bin = x * 4 + y
	const int x0 = tx * FT_TW, y0 = ty * FT_TH;                // tile origin inside the bin
	const uint8_t *img = l0 + stream * (size_t)W * H + (size_t)r0 * W + c0;
	const int tid = threadIdx.x;
	for (int i = tid; i < (FT_TH + 8) * (FT_TW + 8); i += 256) {
		const int r = i / (FT_TW + 8), c = i - r * (FT_TW + 8);
		const int y = min(max(y0 - 4 + r, 0), bh - 1), x = min(max(x0 - 4 + c, 0), bw - 1);
		s_pix[r][c] = img[(size_t)y * W + x];
	}
	__syncthreads();
	for (int i = tid; i < (FT_TH + 2) * (FT_TW + 2); i += 256) {
		const int r = i / (FT_TW + 2), c = i - r * (FT_TW + 2);
		const int y = y0 - 1 + r, x = x0 - 1 + c;
		int sc = 0;
		if (x >= 3 && x < bw - 3 && y >= 3 && y < bh - 3) sc = fast_score(&s_pix[r + 3][c + 3], FT_TW + 8, g.fast_thr);
		s_score[r][c] = (uint8_t)sc;
	}
	__syncthreads();
	for (int i = tid; i < FT_TH * FT_TW; i += 256) {
		const int r = i / FT_TW, c = i - r * FT_TW;
		const int y = y0 + r, x = x0 + c;
		if (x >= bw || y >= bh) continue;
		const int v = s_score[r + 1][c + 1];
		if (!v) continue;
		const bool keep = v > s_score[r][c] && v > s_score[r][c + 1] && v > s_score[r][c + 2] && v > s_score[r + 1][c] &&
				  v > s_score[r + 1][c + 2] && v > s_score[r + 2][c] && v > s_score[r + 2][c + 1] &&
				  v > s_score[r + 2][c + 2];
		if (!keep) continue;
		if (score_map) score_map[stream * (size_t)W * H + (size_t)(r0 + y) * W + c0 + x] = (uint8_t)v;
		else {
			const int slot = atomicAdd(&s.kp_count[stream * LK_BINS + bin], 1);
			if (slot < g.kp_cap)
				s.kp[(stream * LK_BINS + bin) * (size_t)g.kp_cap + slot] = ((uint32_t)v << 24) | (uint32_t)(y * bw + x);
		}
	}
}

cudaError_t lk_fast_launch(const LkGeom &g, const uint8_t *l0, const LkState &s, int n, uint8_t *score_map, cudaStream_t st)
{
	if (n <= 0) return cudaSuccess;
	const int tiles_x = (g.bw + FT_TW - 1) / FT_TW, tiles_y = (g.bh + FT_TH - 1) / FT_TH;
	dim3 grid(tiles_x * tiles_y * LK_BINS, n);
	KernelTimer kt(KID_LK_FAST, st);
	lk_fast_kernel<<<grid, 256, 0, st>>>(g, l0, s, score_map, tiles_x, tiles_y);
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
constexpr int SEL_THREADS = 512;
constexpr int SEL_MAX_TPB = (LK_MAX_FEATURES - 1) / LK_BINS + 1; // 7
constexpr int SEL_MAX_GOOD = LK_BINS * SEL_MAX_TPB + LK_MAX_FEATURES;

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
			  float focal_y, LkCamera cam)
{
	__shared__ int s_ex[LK_MAX_FEATURES], s_ey[LK_MAX_FEATURES];
	__shared__ int s_nex, s_ngood, s_done;
	__shared__ int s_needed[LK_BINS], s_cnt[LK_BINS], s_npick[LK_BINS];
	__shared__ int s_px[LK_BINS][SEL_MAX_TPB], s_py[LK_BINS][SEL_MAX_TPB];
	__shared__ int s_gx[SEL_MAX_GOOD], s_gy[SEL_MAX_GOOD];
	__shared__ unsigned long long s_red[SEL_THREADS / 32];
	const size_t stream = blockIdx.x;
	const int N = g.n_feat, W = g.w[0];
	const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
	float2 *pc = s.prev_corners + stream * N;
	uint8_t *ps = s.prev_status + stream * N;
	int *uv = s.update_vec + stream * N;
	float2 *fc = s.feat_cur + stream * N;
	float2 *fp = s.feat_prev + stream * N;
	const int target = s.target[stream];
	int fast_bins = 0;

	if (target > 0) { // uniform over the CTA
		const int npc = s.n_prev_corners[stream];
		if (tid == 0) {
			int n = 0;
			for (int j = 0; j < npc; j++)
				if (ps[j]) { s_ex[n] = (int)pc[j].x; s_ey[n] = (int)pc[j].y; n++; }
			s_nex = n;
		}
		if (tid < LK_BINS) {
			s_needed[tid] = s.bin_needed[stream * LK_BINS + tid];
			s_cnt[tid] = min(s.kp_count[stream * LK_BINS + tid], g.kp_cap);
			s_npick[tid] = 0;
		}
		__syncthreads();
		const int nex = s_nex;
		// ---- stage 1: one warp per bin
		if (warp < LK_BINS && s_needed[warp] > 0) {
			const int b = warp, needed = s_needed[b], cnt = s_cnt[b];
			const int c0 = (b >> 2) * g.bw, r0 = (b & 3) * g.bh;
			const uint32_t *list = s.kp + (stream * LK_BINS + b) * (size_t)g.kp_cap;
			int npick = 0;
			while (npick < needed) {
				uint32_t best = 0;
				for (int e = lane; e < cnt; e += 32) {
					const uint32_t v = list[e];
					const int li = v & 0xFFFFFF;
					const uint32_t key = (v & 0xFF000000u) | (0xFFFFFFu - li);
					if (key <= best) continue;
					const int y = r0 + li / g.bw, x = c0 + li % g.bw;
					bool ok = true;
					for (int j = 0; j < nex && ok; j++) ok = !cheb_close(s_ex[j], s_ey[j], x, y, g.dist);
					for (int j = 0; j < npick && ok; j++) ok = !cheb_close(s_px[b][j], s_py[b][j], x, y, g.dist);
					if (ok) best = key;
				}
				best = __reduce_max_sync(0xFFFFFFFFu, best);
				if (!best) break;
				const int li = 0xFFFFFF - (best & 0xFFFFFF);
				if (lane == 0) { s_px[b][npick] = c0 + li % g.bw; s_py[b][npick] = r0 + li / g.bw; }
				npick++;
				__syncwarp();
			}
			if (lane == 0) s_npick[b] = npick;
		}
		__syncthreads();
		// ---- stage 2: concatenate in bin order (x outer, y inner == bin index), thin if too many
		if (tid == 0) {
			int n = 0;
			for (int b = 0; b < LK_BINS; b++)
				for (int j = 0; j < s_npick[b]; j++) { s_gx[n] = s_px[b][j]; s_gy[n] = s_py[b][j]; n++; }
			if (n > target) {
				const int n_remove = n - target, step = target / n_remove + 2;
				int m = 0;
				for (int i = 0; i < n; i++)
					if (i % step) { s_gx[m] = s_gx[i]; s_gy[m] = s_gy[i]; m++; }
				n = m;
			}
			s_ngood = n;
			s_done = n >= target;
		}
		__syncthreads();
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
					const int li = v & 0xFFFFFF;
					const int y = r0 + li / g.bw, x = c0 + li % g.bw;
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
	if (tid >= 32) return;
	fast_bins = __reduce_add_sync(0xFFFFFFFFu, fast_bins);
	if (tid != 0) return;

	// ---- stage 4: slot assignment (trackFeatures.cpp:306-362)
	if (target > 0) {
		const int ngood = s_ngood;
		if (!ngood) {
			for (int i = 0; i < N; i++)
				if (uv[i] == 2) uv[i] = 0;
		} else {
			if (s.n_prev_corners[stream] < N) {
				for (int i = s.n_prev_corners[stream]; i < N; i++) pc[i] = make_float2(0.f, 0.f);
				s.n_prev_corners[stream] = N;
			}
			int m = 0;
			for (int i = 0; i < N; i++) {
				if (uv[i] != 2) continue;
				if (m < ngood) {
					const float2 p = make_float2((float)s_gx[m], (float)s_gy[m]);
					pc[i] = p; fc[i] = p; ps[i] = 1;
					m++;
				} else uv[i] = 0;
			}
		}
	}

	// ---- optional undistortion of every slot, placeholders included (flow_opencv.cpp:108-119)
	if (cam.on)
		for (int i = 0; i < N; i++) fc[i] = lk_undistort(cam, fc[i]);

	// ---- calcFlow statistics (flow_opencv.cpp:121-180)
	int meancount = 0, active = 0;
	float mx = 0.f, my = 0.f;
	if (have_feat_prev) {
		for (int i = 0; i < N; i++)
			if (uv[i] == 1) { mx += fc[i].x - fp[i].x; my += fc[i].y - fp[i].y; meancount++; }
		active = meancount;
		if (meancount) {
			mx /= (float)meancount; my /= (float)meancount;
			float sx = 0.f, sy = 0.f;
			for (int i = 0; i < N; i++)
				if (uv[i] == 1) {
					const float ex = fc[i].x - fp[i].x - mx, ey = fc[i].y - fp[i].y - my;
					sx = (float)((double)sx + (double)ex * (double)ex); // pow(float, 2) runs in double
					sy = (float)((double)sy + (double)ey * (double)ey);
				}
			sx = sqrtf(sx / (float)meancount); sy = sqrtf(sy / (float)meancount);
			float tx = 0.f, ty = 0.f;
			meancount = 0;
			for (int i = 0; i < N; i++)
				if (uv[i] == 1) {
					const float fx = fc[i].x - fp[i].x, fy = fc[i].y - fp[i].y;
					if (fabsf(fx - mx) < sx * conf && fabsf(fy - my) < sy * conf) { tx += fx; ty += fy; meancount++; }
					else uv[i] = 0;
				}
			if (meancount) { mx = tx / (float)meancount; my = ty / (float)meancount; }
		}
	}
	// features_previous = features_current; status machine (flow_opencv.cpp:183-196)
	for (int i = 0; i < N; i++) {
		fp[i] = fc[i];
		int u = uv[i];
		if (u == 2) u = 1;
		if (u == 0) u = 2;
		uv[i] = u;
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
		r.px_x = mx; r.px_y = my; r.inliers = meancount; r.active = active; r.fast_bins = fast_bins; r.quality = q0;
		raw[stream] = r;
	}
}

cudaError_t lk_select_epilogue_launch(const LkGeom &g, const LkState &s, const uint32_t *t_us, aof_flow_out *out,
				      aof_lk_raw *raw, int n, int have_feat_prev, int output_rate, float conf, float fx,
				      float fy, const LkCamera &cam, cudaStream_t st)
{
	if (n <= 0) return cudaSuccess;
	KernelTimer kt(KID_LK_SELECT_EPILOGUE, st);
	lk_select_epilogue_kernel<<<n, SEL_THREADS, 0, st>>>(g, s, t_us, out, raw, have_feat_prev, output_rate, conf, fx, fy, cam);
	count_launch();
	return cudaGetLastError();
}

} // namespace aof
