// K4c  lk_fast_pick: the first stage of initMorePoints (external/klt_feature_tracker/src/trackFeatures.cpp:185-248)
// WITHOUT scoring every pixel of the bin.
//
// What the reference does per bin that needs features: cv::FAST on the whole bin (:196), sort by response (:198),
// then walk the list and accept the first `needed` keypoints that keep a Chebyshev distance >= dist from every
// active point and from this bin's earlier picks (:207-241).  Everything else it computed -- on the LK benchmark
// texture ~540 keypoints per 160 x 120 bin, of which it accepts 1 or 2 -- only matters if the second-chance pass
// (:266-304) runs, which needs fewer picks than requests over the WHOLE image.  lk_fast_band_kernel reproduces all
// of it (72 packed min / max per pixel pair) and was 65 % of the LK step.
//
// Pick k of a bin is "the keypoint of highest response (ties: raster order) inside the region R_k" where R_k = the
// bin's scoring area minus the exclusion squares of the active points and of picks 0..k-1.  That is an arg-max, and
// an arg-max can be found lazily with an upper bound:
//   * every 9-arc of the 16-ring contains two ADJACENT compass points (ring 0, 4, 8, 12), so with v_k the ring
//     values  max_arcs min9 <= max over the 4 adjacent compass pairs of min(v_a, v_b)  (and dually for the dark
//     arcs): a bound U >= m on the corner measure m from 5 pixels, 17 packed u16x2 operations per pixel pair
//     instead of 72, evaluated only where the free mask of R_0 has a bit (the exclusion squares are 2 dist - 1 =
//     159 px wide on a 640-px frame: in steady state most of a bin is excluded);
//   * pixels with U >= floor go to a pool in shared memory; a round takes the pool entries inside R_k with the
//     largest U (>= T), scores exactly those (scalar 16-tap network) and checks the best ones against their 8
//     neighbours' exact scores (non-maximum suppression, :FAST nms).  If the best survivor has m >= T it IS the
//     arg-max over R_k: any better keypoint q has U(q) >= m(q) >= T and was scored.  Otherwise T is lowered.
//     T never needs to go below the pool floor unless the region has no strong corner; then the pool is rebuilt
//     from the stored bounds with a lower floor (down to the FAST threshold, where "no survivor" is exact).
// Output: the picks per bin (lz_pick / lz_npick); lk_select_epilogue_kernel takes them in place of its own first
// stage and flags the stream for the complete path (lk_fast_band_kernel + its own stage 1..3) when the second-chance
// pass is needed.  Results are identical to the complete path by construction; tests/test_lk_gpu.py runs both.
#include "aof_common.cuh"
#include "aof_internal.h"

namespace aof {

constexpr int LZ_THREADS = 256;
constexpr int LZ_POOL = 2048;   // pool entries: m bound << 24 | y << 12 | x (bin-local)
constexpr int LZ_CAND = 512;    // candidates scored exactly per pick

// exact corner measure m = max over the 16 arcs of 9 of max(min_arc (c - v), -max_arc (c - v)) at the pixel t points to
// (the pixel is a corner iff m > threshold, its response is m - 1).  Same arc-pair network as
// fb_score_pair (lk_fast_band.cu), scalar.
__device__ __forceinline__ int lz_exact_m(const uint8_t *t, int P)
{
	const int c = t[0];
	int v[16];
	v[0] = t[3 * P];       v[1] = t[3 * P + 1];   v[2] = t[2 * P + 2];   v[3] = t[P + 3];
	v[4] = t[3];           v[5] = t[-P + 3];      v[6] = t[-2 * P + 2];  v[7] = t[-3 * P + 1];
	v[8] = t[-3 * P];      v[9] = t[-3 * P - 1];  v[10] = t[-2 * P - 2]; v[11] = t[-P - 3];
	v[12] = t[-3];         v[13] = t[P - 3];      v[14] = t[2 * P - 2];  v[15] = t[3 * P - 1];
	int p[8], q[8], Pm[8], Qm[8];
#pragma unroll
	for (int j = 0; j < 8; j++) {
		p[j] = min(v[2 * j + 1], v[(2 * j + 2) & 15]);
		Pm[j] = max(v[2 * j + 1], v[(2 * j + 2) & 15]);
	}
#pragma unroll
	for (int j = 0; j < 8; j++) {
		q[j] = min(p[j], p[(j + 1) & 7]);
		Qm[j] = max(Pm[j], Pm[(j + 1) & 7]);
	}
	int X = 0, Y = 255;
#pragma unroll
	for (int i = 0; i < 8; i++) {
		const int hi = max(v[2 * i], v[(2 * i + 9) & 15]), lo = min(v[2 * i], v[(2 * i + 9) & 15]);
		X = max(X, min(min(q[i], q[(i + 2) & 7]), hi));
		Y = min(Y, max(max(Qm[i], Qm[(i + 2) & 7]), lo));
	}
	return max(X - c, c - Y);
}

// bound on m for the two pixels packed in each u16x2 argument: c centre, n / e / s / w the compass ring pixels
__device__ __forceinline__ uint32_t lz_bound2(uint32_t c, uint32_t n, uint32_t e, uint32_t s, uint32_t w)
{
	const uint32_t x = __vmaxu2(__vimax3_u16x2(__vminu2(n, e), __vminu2(e, s), __vminu2(s, w)), __vminu2(w, n));
	const uint32_t y = __vminu2(__vimin3_u16x2(__vmaxu2(n, e), __vmaxu2(e, s), __vmaxu2(s, w)), __vmaxu2(w, n));
	// max(x - c, c - y) per lane without a borrow: (max(x, c) - c) and (c - min(y, c)) are both >= 0
	return __vmaxu2(__vmaxu2(x, c) - c, c - __vminu2(y, c));
}

// bits [lo, hi] (inclusive, pixel columns) that fall into the 32-bit word starting at column `base`
__device__ __forceinline__ uint32_t lz_bits(int lo, int hi, int base)
{
	lo = max(lo - base, 0);
	hi = min(hi - base, 31);
	if (lo > hi) return 0u;
	return (0xFFFFFFFFu >> (31 - hi + lo)) << lo;
}

struct LzCtl {
	int nex;        // active points whose exclusion square reaches the bin
	int npool, over;
	int ncand;
	int umax;       // largest bound among the pool entries inside the current region
	int T;          // current candidate threshold (on m)
	int floor;      // the pool holds every pixel of the current region with bound >= floor
	int found;      // 1 = pick found, 0 = none at this T
	int pick;       // the pick (pool format: m << 24 | y << 12 | x)
	int any;        // the bin has free scoring pixels at all
	int fail;       // the bin needs the complete path
	int hist[256];
};

__global__ void __launch_bounds__(LZ_THREADS)
lk_fast_pick_kernel(LkGeom g, const uint8_t *__restrict__ l0, LkState s)
{
	extern __shared__ __align__(16) uint32_t lz_smem[];
	__shared__ uint32_t s_pool[LZ_POOL];
	__shared__ uint32_t s_cand[LZ_CAND];
	__shared__ int s_cm[LZ_CAND];                       // exact m of the candidate, -1 = suppressed by a neighbour
	__shared__ int s_ex[LK_MAX_FEATURES], s_ey[LK_MAX_FEATURES];
	__shared__ int s_px[LK_MAX_TPB], s_py[LK_MAX_TPB];  // picks so far, image coordinates
	__shared__ LzCtl c;

	const size_t stream = blockIdx.y;
	const int bin = blockIdx.x;
	const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
	const int needed = s.bin_needed[stream * LK_BINS + bin];
	if (needed <= 0) {
		if (tid == 0) s.lz_npick[stream * LK_BINS + bin] = 0;
		return;
	}
	const int bw = g.bw, bh = g.bh, W = g.w[0], H = g.h[0], N = g.n_feat, dist = g.dist, thr = g.fast_thr;
	const int tw = bw >> 2, fmw = (bw + 31) >> 5;
	uint32_t *tile = lz_smem;                           // bh x tw words: the bin's pixels
	uint32_t *ub = tile + bh * tw;                      // bh x tw words: the bound on m per pixel (0 = not free)
	uint32_t *fm = ub + bh * tw;                        // bh x fmw words: free mask of R_0 (scoring area minus the active points)
	const uint8_t *tile8 = reinterpret_cast<const uint8_t *>(tile);
	const uint8_t *ub8 = reinterpret_cast<const uint8_t *>(ub);
	const int c0 = (bin >> 2) * bw, r0 = (bin & 3) * bh;  // bin = x * 4 + y

	if (tid == 0) {
		c.nex = 0; c.npool = 0; c.over = 0; c.any = 0; c.fail = 0;
		c.floor = s.lz_floor[stream * LK_BINS + bin];
	}
	__syncthreads();
	// ---- active points whose exclusion square reaches into the bin (trackFeatures.cpp:212-224: int-truncated
	// coordinates, prev_status != 0)
	{
		const int npc = s.n_prev_corners[stream];
		for (int j = tid; j < npc; j += LZ_THREADS) {
			if (!s.prev_status[stream * N + j]) continue;
			const float2 p = s.prev_corners[stream * N + j];
			const int ex = (int)p.x, ey = (int)p.y;
			if (ex > c0 - dist && ex < c0 + bw - 1 + dist && ey > r0 - dist && ey < r0 + bh - 1 + dist) {
				const int k = atomicAdd(&c.nex, 1);
				s_ex[k] = ex - c0; s_ey[k] = ey - r0;    // bin-local
			}
		}
	}
	__syncthreads();
	// ---- free mask of R_0
	{
		const int nex = c.nex;
		uint32_t any = 0;
		for (int i = tid; i < bh * fmw; i += LZ_THREADS) {
			const int y = i / fmw, w = i - y * fmw;
			uint32_t m = (y >= 3 && y < bh - 3) ? lz_bits(3, bw - 4, 32 * w) : 0u;
			for (int j = 0; j < nex && m; j++)
				if (abs(s_ey[j] - y) < dist) m &= ~lz_bits(s_ex[j] - dist + 1, s_ex[j] + dist - 1, 32 * w);
			fm[i] = m;
			any |= m;
		}
		if (any) c.any = 1;
	}
	__syncthreads();
	if (!c.any) { // every scoring pixel is excluded: no keypoint of this bin can be accepted, whatever FAST finds
		if (tid == 0) s.lz_npick[stream * LK_BINS + bin] = 0;
		return;
	}
	// ---- the bin's pixels
	{
		const uint32_t *img32 = reinterpret_cast<const uint32_t *>(l0 + stream * (size_t)W * H + (size_t)r0 * W + c0);
		const int W4 = W >> 2;
		for (int i = tid; i < bh * tw; i += LZ_THREADS) {
			const int y = i / tw, q = i - y * tw;
			tile[i] = __ldg(img32 + y * W4 + q);
		}
	}
	__syncthreads();
	// ---- bound pass: 4 pixels per item, only where the free mask has a bit
	{
		int floor = c.floor;
		if (floor <= 0) floor = 256;                    // first call: no appends, the pool is built from the histogram below
		for (int i = tid; i < (bh - 6) * tw; i += LZ_THREADS) {
			const int y = 3 + i / tw, q = i - (y - 3) * tw;
			const uint32_t bits = (fm[y * fmw + (q >> 3)] >> ((q & 7) * 4)) & 15u;
			uint32_t packed = 0u;
			if (bits) {
				const uint32_t *row = tile + y * tw + q;
				const uint32_t C = row[0], Sv = row[3 * tw], Nv = row[-3 * tw];
				const uint32_t prev = q > 0 ? row[-1] : 0u, next = q + 1 < tw ? row[1] : 0u;
				const uint32_t Wv = __funnelshift_r(prev, C, 8), Ev = __funnelshift_r(C, next, 24);
				const uint32_t lo = lz_bound2(__byte_perm(C, 0u, 0x4140), __byte_perm(Sv, 0u, 0x4140), __byte_perm(Ev, 0u, 0x4140),
							      __byte_perm(Nv, 0u, 0x4140), __byte_perm(Wv, 0u, 0x4140));
				const uint32_t hi = lz_bound2(__byte_perm(C, 0u, 0x4342), __byte_perm(Sv, 0u, 0x4342), __byte_perm(Ev, 0u, 0x4342),
							      __byte_perm(Nv, 0u, 0x4342), __byte_perm(Wv, 0u, 0x4342));
				packed = __byte_perm(lo, hi, 0x6420) & (((bits * 0x00204081u) & 0x01010101u) * 0xFFu);
#pragma unroll
				for (int e = 0; e < 4; e++) {
					const int mb = (packed >> (8 * e)) & 255;
					if (mb >= floor && mb > thr) {
						const int k = atomicAdd(&c.npool, 1);
						if (k < LZ_POOL) s_pool[k] = ((uint32_t)mb << 24) | ((uint32_t)y << 12) | (uint32_t)(4 * q + e);
						else c.over = 1;
					}
				}
			}
			ub[y * tw + q] = packed;
		}
	}
	__syncthreads();
	// inside R_k: not within dist of picks 0..k-1 (the active points are already in the free mask)
	auto in_region = [&](int x, int y, int k) -> bool {
		bool ok = true;
		for (int j = 0; j < k && ok; j++) ok = !(abs(s_px[j] - (c0 + x)) < dist && abs(s_py[j] - (r0 + y)) < dist);
		return ok;
	};
	// pool := every pixel of R_k with bound >= f, f the smallest value in (thr, below) that keeps the pool within
	// 3/4 of its capacity (histogram of the stored bounds, then one collecting pass)
	auto rebuild_pool = [&](int k, int below) {
		for (int i = tid; i < 256; i += LZ_THREADS) c.hist[i] = 0;
		if (tid == 0) atomicAdd(&s.lz_stats[1], 1ull);
		const int kept = below < 256 ? min(c.npool, LZ_POOL) : 0;   // entries >= the old floor stay (upper bound on them)
		__syncthreads();
		if (tid == 0) { c.npool = 0; c.over = 0; }
		__syncthreads();
		for (int i = tid; i < (bh - 6) * tw; i += LZ_THREADS) {
			const int y = 3 + i / tw, q = i - (y - 3) * tw;
			const uint32_t wv = ub[y * tw + q];
			if (!wv) continue;
#pragma unroll
			for (int e = 0; e < 4; e++) {
				const int mb = (wv >> (8 * e)) & 255;
				if (mb > thr && mb < below && in_region(4 * q + e, y, k)) atomicAdd(&c.hist[mb], 1);
			}
		}
		__syncthreads();
		if (tid == 0) {
			int f = below, cum = kept;
			while (f - 1 > thr && cum + c.hist[f - 1] <= LZ_POOL * 3 / 4) { f--; cum += c.hist[f]; }
			if (f == below) { // not even one more level fits
				if (below - 1 > thr) c.fail = 1;
				f = thr + 1;    // (below - 1 == thr: nothing left to add, the pool is exact down to the corner threshold)
			}
			c.floor = f;
		}
		__syncthreads();
		if (c.fail) return;
		const int f = c.floor;
		for (int i = tid; i < (bh - 6) * tw; i += LZ_THREADS) {
			const int y = 3 + i / tw, q = i - (y - 3) * tw;
			const uint32_t wv = ub[y * tw + q];
			if (!wv) continue;
#pragma unroll
			for (int e = 0; e < 4; e++) {
				const int mb = (wv >> (8 * e)) & 255;
				if (mb >= f && mb > thr && in_region(4 * q + e, y, k)) {
					const int j = atomicAdd(&c.npool, 1);
					if (j < LZ_POOL) s_pool[j] = ((uint32_t)mb << 24) | ((uint32_t)y << 12) | (uint32_t)(4 * q + e);
					else c.over = 1;
				}
			}
		}
		__syncthreads();
		if (tid == 0 && c.over) c.fail = 1;   // bounds >= the old floor alone overflow the pool: complete path
		__syncthreads();
	};

	if (c.over || c.floor <= 0) rebuild_pool(0, 256);
	else if (tid == 0 && c.floor <= thr) c.floor = thr + 1;
	__syncthreads();

	// exact score of a pixel as NMS sees it: m - 1 for corners of the scoring area, 0 elsewhere
	auto nms_m = [&](int x, int y) -> int {
		if (x < 3 || x >= bw - 3 || y < 3 || y >= bh - 3) return 0;
		const int m = lz_exact_m(tile8 + y * bw + x, bw);
		return m > thr ? m : 0;
	};

	int npick = 0;
	for (int k = 0; k < needed && !c.fail; k++) {
		bool got = false, exhausted = false;
		while (!got && !exhausted && !c.fail) {
			// ---- largest bound inside R_k
			if (tid == 0) { c.umax = 0; c.ncand = 0; }
			for (int i = tid; i < 64; i += LZ_THREADS) c.hist[i] = 0;
			__syncthreads();
			const int npool = min(c.npool, LZ_POOL);
			int um = 0;
			for (int i = tid; i < npool; i += LZ_THREADS) {
				const uint32_t e = s_pool[i];
				if (in_region(e & 0xFFF, (e >> 12) & 0xFFF, k)) um = max(um, (int)(e >> 24));
			}
			um = __reduce_max_sync(0xFFFFFFFFu, um);
			if (lane == 0 && um) atomicMax(&c.umax, um);
			__syncthreads();
			const int umax = c.umax, floor = c.floor;
			if (!umax) { // nothing in the pool lies inside R_k
				if (floor <= thr + 1) { exhausted = true; break; }   // the pool was exact: R_k holds no corner at all
				rebuild_pool(k, floor);
				continue;
			}
			// ---- threshold: the ~32 entries of R_k with the largest bounds
			for (int i = tid; i < npool; i += LZ_THREADS) {
				const uint32_t e = s_pool[i];
				const int d = umax - (int)(e >> 24);
				if (d < 64 && in_region(e & 0xFFF, (e >> 12) & 0xFFF, k)) atomicAdd(&c.hist[d], 1);
			}
			__syncthreads();
			if (tid == 0) {
				int d = 0, cum = c.hist[0];
				if (cum > LZ_CAND) c.fail = 1;          // more ties at the top than the candidate buffer holds
				while (d < 63 && cum < 32 && umax - (d + 1) >= floor && cum + c.hist[d + 1] <= LZ_CAND) { d++; cum += c.hist[d]; }
				c.T = umax - d;
				c.found = 0;
			}
			__syncthreads();
			if (c.fail) break;
			int T_hi = 256;                                 // entries with bound in [T, T_hi) are collected in this step
			for (;;) {
				const int T = c.T;
				// ---- collect and score
				for (int i = tid; i < npool; i += LZ_THREADS) {
					const uint32_t e = s_pool[i];
					const int mb = (int)(e >> 24), x = e & 0xFFF, y = (e >> 12) & 0xFFF;
					if (mb >= T && mb < T_hi && in_region(x, y, k)) {
						const int j = atomicAdd(&c.ncand, 1);
						if (j < LZ_CAND) { s_cand[j] = e; s_cm[j] = lz_exact_m(tile8 + y * bw + x, bw); }
					}
				}
				__syncthreads();
				if (c.ncand > LZ_CAND) { if (tid == 0) c.fail = 1; __syncthreads(); break; }
				// ---- warp 0: best candidate with m >= T that survives non-maximum suppression
				if (warp == 0) {
					const int ncand = c.ncand;
					for (;;) {
						uint32_t best = 0;
						int bi = -1;
						for (int i = lane; i < ncand; i += 32) {
							const int m = s_cm[i];
							if (m < T || m <= thr) continue;
							const uint32_t key = ((uint32_t)m << 24) | (0xFFFFFFu - (s_cand[i] & 0xFFFFFFu));
							if (key > best) { best = key; bi = i; }
						}
						const uint32_t top = __reduce_max_sync(0xFFFFFFFFu, best);
						if (!top) break;                        // no survivor at this T
						const int src = __ffs(__ballot_sync(0xFFFFFFFFu, best == top)) - 1;
						const int wi = __shfl_sync(0xFFFFFFFFu, bi, src);
						const uint32_t e = s_cand[wi];
						const int x = e & 0xFFF, y = (e >> 12) & 0xFFF, m = (int)(top >> 24);
						// cv::FAST keeps a corner iff its score is greater than all 8 neighbours' (strict)
						bool beaten = false;
						if (lane < 8) {
							const int o = lane < 4 ? lane : lane + 1;   // 0..8 without the centre
							beaten = nms_m(x + o % 3 - 1, y + o / 3 - 1) >= m;
						}
						if (__any_sync(0xFFFFFFFFu, beaten)) {
							if (lane == 0) s_cm[wi] = -1;
							__syncwarp();
							continue;
						}
						if (lane == 0) { c.found = 1; c.pick = (int)(((uint32_t)m << 24) | (e & 0xFFFFFFu)); }
						break;
					}
				}
				__syncthreads();
				if (c.found || T <= floor) break;
				T_hi = T;
				if (tid == 0) c.T = max(floor, T - 8);
				__syncthreads();
			}
			if (c.fail) break;
			if (c.found) { got = true; break; }
			// no keypoint of R_k has m >= floor
			if (floor <= thr + 1) { exhausted = true; break; }
			rebuild_pool(k, floor);
		}
		if (!got) break;
		if (tid == 0) {
			const int e = c.pick;
			s_px[k] = c0 + (e & 0xFFF); s_py[k] = r0 + ((e >> 12) & 0xFFF);
		}
		npick = k + 1;
		__syncthreads();
	}
	if (tid == 0) {
		int *np = s.lz_npick + stream * LK_BINS + bin;
		atomicAdd(&s.lz_stats[c.fail ? 2 : 0], 1ull);
		if (c.fail) *np = -1;                           // this bin needs the complete path
		else {
			*np = npick;
			for (int j = 0; j < npick; j++)
				s.lz_pick[(stream * LK_BINS + bin) * LK_MAX_TPB + j] = (uint32_t)s_px[j] | ((uint32_t)s_py[j] << 16);
		}
		// next frame's pool threshold: a little below this frame's last pick (it follows the scene's contrast both
		// ways; a floor that turns out too high only costs a pool rebuild), raised while the pool runs large
		int f = c.floor;
		if (npick > 0 && !c.fail) f = max(thr + 1, (int)((uint32_t)c.pick >> 24) - 8);
		if (c.npool > LZ_POOL / 2) f = min(255, max(f, c.floor + 2));
		s.lz_floor[stream * LK_BINS + bin] = f;
	}
}

// ---- host side -------------------------------------------------------------
static size_t lz_smem_bytes(const LkGeom &g)
{
	const int tw = g.bw >> 2, fmw = (g.bw + 31) >> 5;
	return ((size_t)2 * g.bh * tw + (size_t)g.bh * fmw) * sizeof(uint32_t);
}

// needs whole words per bin row, a 4-byte aligned batch and 12-bit bin-local coordinates; the complete path covers the rest
bool lk_fast_lazy_supported(const LkGeom &g, const uint8_t *l0)
{
	if (g.tun_fast_lazy == 0) return false;
	if (g.bw % 4 || g.w[0] % 4 || (reinterpret_cast<uintptr_t>(l0) & 3) || g.bh < 7 || g.bw < 8 || g.bw > 4095 || g.bh > 4095) return false;
	if (!lk_fast_band_supported(g, l0)) return false;    // the fallback of a flagged stream is the band kernel
	return lz_smem_bytes(g) <= 160 * 1024;
}

cudaError_t lk_fast_pick_launch(const LkGeom &g, const uint8_t *l0, const LkState &s, int n, cudaStream_t st)
{
	if (n <= 0) return cudaSuccess;
	static bool attr_set[64] = {};
	int dev = 0;
	cudaGetDevice(&dev);
	if (dev < 64 && !attr_set[dev]) {
		cudaError_t e = cudaFuncSetAttribute(lk_fast_pick_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024);
		if (e != cudaSuccess) return e;
		attr_set[dev] = true;
	}
	KernelTimer kt(KID_LK_FAST_PICK, st);
	lk_fast_pick_kernel<<<dim3(LK_BINS, n), LZ_THREADS, lz_smem_bytes(g), st>>>(g, l0, s);
	count_launch();
	return cudaGetLastError();
}

} // namespace aof
