// K4b  lk_fast_band: cv::FAST(bin ROI, keypoints, 10, nonmaxSuppression = true) for the bins that
// need features (external/klt_feature_tracker/src/trackFeatures.cpp:196), one CTA
// per (stream, bin, band of 38 rows).  Integer-ALU bound: the exact corner response of EVERY pixel
// is a fixed min / max network (no test-then-score branch; on the LK benchmark texture more than a
// third of the pixels are corners), two horizontally adjacent pixels per thread in u16x2 lanes.
//
//   * FAST runs on the bin as a stand-alone image, so a band needs only rows of its own bin: the
//     rows [y0 - 4, y0 + 42) are loaded once, each 4-pixel word expanded into the two pair
//     layouts A = (p[2j], p[2j+1]) and B = (p[2j+1], p[2j+2]) so that every ring tap of a pixel
//     pair is ONE aligned LDS.32 (ring offsets with odd dx come from A, even dx from B);
//   * response: with D_k = c + 256 - ring_k per lane, a dark corner of strength s has an arc of
//     9 with min D - 256 > s, a bright one max D < 256 - s.  The two arcs starting at k and k + 1
//     (k even) share ring positions k+1 .. k+8, so
//         max(min9(k), min9(k+1)) = min(min8(k+1), max(D_k, D_k+9))
//     and the 8 windows min8(odd start) are built from 8 pair-minima: 36 packed min / max per
//     polarity instead of 40 for the plain sliding window (all VIMNMX / VIMNMX3 .U16x2);
//   * responses are kept biased, b = max(m, 256 + thr) - (256 + thr) (0 = not a corner, order
//     preserved), as u16x2 in shared memory; the 3 x 3 non-maximum suppression walks down the
//     band with the horizontal maxima of the previous / next row carried in registers;
//   * survivors are collected in shared memory and appended to the bin's list with one global
//     atomic per CTA.  Entry = score << 24 | y << 12 | x, bin-local (order in the list is free:
//     the selection kernel uses the total order score desc, raster asc).
//   * MASKED mode (the default of the calcFlow path): the first stage of initMorePoints only ever accepts keypoints
//     that keep the Chebyshev distance `dist` from every active point (trackFeatures.cpp:207-241), and on a 640-px
//     frame the exclusion squares (2 dist - 1 = 159 px) cover ~70 % of a bin that needs features.  A pixel pair is
//     scored only where the mask of the band has a bit: mask = scoring area minus the exclusion squares SHRUNK by one
//     pixel, so that every free pixel still sees the exact responses of its 8 neighbours in the suppression.
//     lk_fast_plan_kernel (one small CTA per (stream, bin)) builds the mask and lists, per band, the live units of 16
//     pixel pairs x 1 response row and the bands that have any; lk_fast_band_masked_kernel is a persistent grid over
//     that work list (no CTA is launched for an empty band, the units are dealt to the half-warps, rows outside the
//     live range are neither loaded nor searched for maxima).  The lists then hold exactly the keypoints of the free
//     region (plus possibly false maxima in the one-pixel ring, all of which the selection rejects as too close);
//     a stream whose selection needs the second-chance pass (:266-304: every keypoint, dist / 2) is flagged by the
//     selection kernel and redone unmasked by lk_fast_band_flagged_kernel.
// Needs bw % 4 == 0, W % 4 == 0, a 4-byte aligned batch and bw <= 512; other shapes use lk_fast_kernel.
#include <cstdlib>

#include "aof_common.cuh"
#include "aof_internal.h"

namespace aof {

// FB_ROWS (aof_internal.h): output rows per band (38: a 120-row bin has 114 = 3 x 38)
constexpr int FB_SROWS = FB_ROWS + 2;   // response rows (one above, one below for the suppression)
constexpr int FB_PROWS = FB_ROWS + 8;   // pixel rows
constexpr int FB_MAX_GROUPS = 20;       // row groups: thread = (group, pair column); THREADS / (bw / 2) of them
constexpr int FB_MWORDS = 17;           // mask words per response row (bw <= 512, + 1 for the pair that straddles)

// bits [lo, hi] (inclusive, pixel columns) that fall into the 32-bit word starting at column `base`
__device__ __forceinline__ uint32_t fb_bits(int lo, int hi, int base)
{
	lo = max(lo - base, 0);
	hi = min(hi - base, 31);
	if (lo > hi) return 0u;
	return (0xFFFFFFFFu >> (31 - hi + lo)) << lo;
}

__device__ __forceinline__ uint32_t fb_mn(uint32_t a, uint32_t b) { return __vminu2(a, b); }
__device__ __forceinline__ uint32_t fb_mx(uint32_t a, uint32_t b) { return __vmaxu2(a, b); }
__device__ __forceinline__ uint32_t fb_mn3(uint32_t a, uint32_t b, uint32_t c) { return __vimin3_u16x2(a, b, c); }
__device__ __forceinline__ uint32_t fb_mx3(uint32_t a, uint32_t b, uint32_t c) { return __vimax3_u16x2(a, b, c); }

// response + 257 per lane of the pixel pair whose centre is B[0][0] (A, B point at the centre row / pair column,
// P = row pitch in words).  The network runs on the ring values v_k themselves: min / max commute with the order-
// reversing map D = c + 256 - v, so with X = max over arcs of min9(v) and Y = min over arcs of max9(v) the dark response
// (+257) is c + 256 - Y and the bright one X + 256 - c: no per-tap subtraction.
// Materialised min3 / max3 only (ptxas 12.9 mis-folds max(m, -x), tools/ptxas_vimnmx3_neg_repro.cu).
__device__ __forceinline__ uint32_t fb_score_pair(const uint32_t *A, const uint32_t *B, int P)
{
	const uint32_t cb = B[0] + 0x01000100u;
	uint32_t D[16];
	D[0] = B[3 * P];       D[1] = A[3 * P + 1];   D[2] = B[2 * P + 1];   D[3] = A[P + 2];
	D[4] = A[2];           D[5] = A[-P + 2];      D[6] = B[-2 * P + 1];  D[7] = A[-3 * P + 1];
	D[8] = B[-3 * P];      D[9] = A[-3 * P];      D[10] = B[-2 * P - 1]; D[11] = A[-P - 1];
	D[12] = A[-1];         D[13] = A[P - 1];      D[14] = B[2 * P - 1];  D[15] = A[3 * P];
	uint32_t p[8], q[8], Pm[8], Qm[8];
#pragma unroll
	for (int j = 0; j < 8; j++) {
		p[j] = fb_mn(D[2 * j + 1], D[(2 * j + 2) & 15]);
		Pm[j] = fb_mx(D[2 * j + 1], D[(2 * j + 2) & 15]);
	}
#pragma unroll
	for (int j = 0; j < 8; j++) {
		q[j] = fb_mn(p[j], p[(j + 1) & 7]);
		Qm[j] = fb_mx(Pm[j], Pm[(j + 1) & 7]);
	}
	uint32_t r[8], R[8];
#pragma unroll
	for (int i = 0; i < 8; i++) {
		const uint32_t hi = fb_mx(D[2 * i], D[(2 * i + 9) & 15]);
		const uint32_t lo = fb_mn(D[2 * i], D[(2 * i + 9) & 15]);
		r[i] = fb_mn3(q[i], q[(i + 2) & 7], hi);    // min9 of the better of the arcs starting at 2i and 2i + 1
		R[i] = fb_mx3(Qm[i], Qm[(i + 2) & 7], lo);  // max9 likewise
	}
	uint32_t X = fb_mx3(r[0], r[1], r[2]), Y = fb_mn3(R[0], R[1], R[2]);
	X = fb_mx3(X, r[3], r[4]); Y = fb_mn3(Y, R[3], R[4]);
	X = fb_mx3(X, r[5], r[6]); Y = fb_mn3(Y, R[5], R[6]);
	X = fb_mx(X, r[7]); Y = fb_mn(Y, R[7]);
	return fb_mx(cb - Y, X + 0x02000200u - cb);
}

// dynamic shared memory: A[FB_PROWS][P], B[FB_PROWS][P], responses[FB_SROWS][P], survivor list
// PC = row pitch of the shared-memory arrays in words when it is known at compile time (84 for the 160-px bins of a
// 640-px frame: every ring tap is then an immediate offset of one base address), 0 = run time (P_rt)
template <int THREADS, int PC>
__device__ __forceinline__ void fb_body(const LkGeom &g, const uint8_t *__restrict__ l0, const LkState &s, uint8_t *__restrict__ score_map,
					 int bands, int P_rt, int list_cap, size_t stream, int bin, int band,
					 const uint16_t *__restrict__ units, uint32_t hdr)
{
	const int P = PC ? PC : P_rt;
	extern __shared__ __align__(16) uint32_t fb_smem[];
	uint32_t *sA = fb_smem;
	uint32_t *sB = sA + FB_PROWS * P;
	uint32_t *sS = sB + FB_PROWS * P;
	uint32_t *sL = sS + FB_SROWS * P;
	__shared__ int s_n;
	__shared__ uint16_t s_units[FB_SROWS * FB_MWORDS];            // masked: live (response row << 8 | 16-pair segment) units

	const bool masked = units != nullptr;
	if (!score_map && !masked && s.bin_needed[stream * LK_BINS + bin] == 0) return;
	const int W = g.w[0], H = g.h[0], bh = g.bh;
	const int bw = PC ? (PC - 4) * 2 : g.bw;                      // compile-time bin width: the index arithmetic below folds
	const int npc = bw >> 1, nq = bw >> 2;
	const int c0 = (bin >> 2) * bw, r0 = (bin & 3) * bh;        // bin = x * 4 + y
	const int y0 = 3 + band * FB_ROWS;                            // first output row of the band (bin coordinates)
	const int y1 = min(y0 + FB_ROWS, bh - 3);                     // one past the last
	const int tid = threadIdx.x;
	// masked: the plan's live units of this band; response rows outside [rlo, rhi] hold no live unit.  The responses of
	// the pairs that are not scored are 0.
	const int nunits = masked ? (int)(hdr & 0xFFFFu) : 0;
	const int rlo = masked ? (int)((hdr >> 16) & 0xFFu) : 0, rhi = masked ? (int)(hdr >> 24) : FB_SROWS;
	if (masked) {
		for (int i = tid; i < nunits; i += THREADS) s_units[i] = units[i];
		for (int i = tid; i < (y1 - y0 + 2) * P; i += THREADS) sS[i] = 0u;
	}
	if (tid == 0) s_n = 0;
	const uint8_t *img = l0 + stream * (size_t)W * H + (size_t)r0 * W + c0;

	// ---- pixel rows y0 - 4 .. y1 + 3 of the bin -> pair layouts (tile row t <-> bin row y0 - 4 + t)
	// THREADS >= 8 nq, so 6 rounds cover the FB_PROWS x nq words; all loads of a thread are issued before the first use
	const int n_prow = y1 - y0 + 8;
	{
		constexpr int LD_ROUNDS = PC ? (FB_PROWS * ((PC - 4) / 2) + THREADS - 1) / THREADS : 6;
		// with a compile-time bin width that divides THREADS a thread keeps its word column and steps whole rows
		constexpr bool FIXED_Q = PC != 0 && THREADS % (PC ? (PC - 4) / 2 : 1) == 0;
		const int dt = THREADS / nq, dq = THREADS - dt * nq;      // (row, word) step of one round
		int t = tid / nq, q = tid - t * nq;
		uint32_t w0[LD_ROUNDS], w1[LD_ROUNDS];
		int tt[LD_ROUNDS], qq[LD_ROUNDS];
		const uint32_t *img32 = reinterpret_cast<const uint32_t *>(img);
		const int W4 = W >> 2;
#pragma unroll
		for (int k = 0; k < LD_ROUNDS; k++) {
			const int y = y0 - 4 + t;
			// rows -1 and bh only feed responses that are not computed; masked: response row sr reads tile rows sr .. sr + 6
			const bool on = t < n_prow && y >= 0 && y < bh && t >= rlo && t <= rhi + 6;
			tt[k] = on ? t : -1; qq[k] = q;
			const uint32_t *row = img32 + (on ? y * W4 : 0);      // 32-bit offsets: a bin is far below 2^31 bytes
			w0[k] = __ldg(row + q);
			w1[k] = __ldg(row + (q + 1 < nq ? q + 1 : q));        // p[bw] is never used
			t += dt;
			if (!FIXED_Q) {
				q += dq;
				if (q >= nq) { q -= nq; t++; }
			}
		}
#pragma unroll
		for (int k = 0; k < LD_ROUNDS; k++) {
			if (tt[k] < 0) continue;
			const uint32_t a0 = __byte_perm(w0[k], 0u, 0x4140), a1 = __byte_perm(w0[k], 0u, 0x4342);
			const uint32_t b0 = __byte_perm(w0[k], 0u, 0x4241), b1 = __byte_perm(__funnelshift_r(w0[k], w1[k], 24), 0u, 0x4140);
			*reinterpret_cast<uint2 *>(sA + tt[k] * P + 2 * qq[k]) = make_uint2(a0, a1);
			*reinterpret_cast<uint2 *>(sB + tt[k] * P + 2 * qq[k]) = make_uint2(b0, b1);
		}
	}
	__syncthreads();

	// ---- biased responses of rows y0 - 1 .. y1 (response row sr <-> bin row y0 - 1 + sr, tile row sr + 3)
	const int grp = tid / npc, j = tid - grp * npc;               // pair column j <-> pixels 2j + 1, 2j + 2
	const int ngrp = min(THREADS / npc, FB_MAX_GROUPS);
	const bool col_ok = grp < ngrp && j >= 1 && j <= npc - 3;
	const int n_srow = y1 - y0 + 2;
	const uint32_t lim = (256u + (uint32_t)g.fast_thr) * 0x00010001u;
	if (masked) {
		const int half = (tid >> 4) & 1, jj0 = tid & 15;
		for (int u = 2 * (tid >> 5) + half; u < nunits; u += 2 * (THREADS / 32)) {
			const int sr = s_units[u] >> 8, jj = 16 * (s_units[u] & 255) + jj0;
			if (jj < npc) {
				const uint32_t m = fb_score_pair(sA + (sr + 3) * P + jj, sB + (sr + 3) * P + jj, P);
				sS[sr * P + jj] = (jj >= 1 && jj <= npc - 3) ? fb_mx(m, lim) - lim : 0u;
			}
		}
	} else if (grp < ngrp) {
		// rows of the band whose response exists (3 <= y < bh - 3); the others (first / last band only) are zero.
		// Columns outside [1, npc - 3] run the network on the pitch padding and drop the result.
		const int sr_lo = max(0, 4 - y0), sr_hi = min(n_srow, bh - 2 - y0);
		for (int sr = grp; sr < n_srow; sr += ngrp) {
			uint32_t b = 0u;
			if (sr >= sr_lo && sr < sr_hi) {
				const uint32_t m = fb_score_pair(sA + (sr + 3) * P + j, sB + (sr + 3) * P + j, P);
				b = col_ok ? fb_mx(m, lim) - lim : 0u;
			}
			sS[sr * P + j] = b;
		}
	}
	__syncthreads();

	// ---- 3 x 3 non-maximum suppression, each thread walking down its share of the band's rows
	if (col_ok) {
		const int n_out = y1 - y0;
		const int oa = max(1, rlo), ob = min(n_out + 1, rhi + 1);          // (masked: the rows with a live unit only)
		const int chunk = (max(ob - oa, 0) + ngrp - 1) / ngrp;
		const int ra = oa + grp * chunk, rb = min(ra + chunk, ob);         // response rows of this thread's outputs
		const uint32_t *col = sS + j;
		auto hmax = [&](int sr, uint32_t &c, uint32_t &side) -> uint32_t {
			const uint32_t l = col[sr * P - 1], r = col[sr * P + 1];
			c = col[sr * P];
			const uint32_t lm = __byte_perm(l, c, 0x5432), rm = __byte_perm(c, r, 0x5432); // (x-1, x), (x+1, x+2)
			side = fb_mx(lm, rm);
			return fb_mx3(lm, c, rm);
		};
		// Rows are walked 16 at a time.  The strict-maximum bits of a row (bit 15 / bit 31 of the packed test) are shifted
		// into a per-thread mask instead of being acted on: a warp almost always holds SOME survivor, so a branch per
		// row would run its body for every row with one or two lanes active.  The mask is drained after the walk.
		for (int r16 = ra; r16 < rb; r16 += 16) {
			const int re = min(r16 + 16, rb);
			uint32_t c_cur, c_next, side_cur, side_next, c_skip, side_skip;
			uint32_t h_prev = hmax(r16 - 1, c_skip, side_skip);
			uint32_t h_cur = hmax(r16, c_cur, side_cur);
			uint32_t acc = 0u;
			for (int sr = r16; sr < re; sr++) {
				const uint32_t h_next = hmax(sr + 1, c_next, side_next);
				// strict maximum per lane: c > m.  All values are < 2^15, so bit 15 of (c | 0x8000) - m - 1 is set iff c > m
				// (no borrow crosses the lanes)
				const uint32_t m = fb_mx3(h_prev, h_next, side_cur);
				acc = (acc >> 1) | (((c_cur | 0x80008000u) - m - 0x00010001u) & 0x80008000u);
				h_prev = h_cur; h_cur = h_next; c_cur = c_next; side_cur = side_next;
			}
			// row r16 + i sits at bit 16 e + 15 - (n - 1 - i)
			const int n = re - r16;
			while (acc) {
				const int p = __ffs((int)acc) - 1;
				acc &= acc - 1u;
				const int e = p >> 4, sr = r16 + n - 16 + (p & 15);
				const uint32_t v = (col[sr * P] >> (16 * e)) & 0xFFFFu;
				const int y = y0 + sr - 1, x = 2 * j + 1 + e;
				const uint32_t score = v + (uint32_t)g.fast_thr - 1u;
				if (score_map) score_map[stream * (size_t)W * H + (size_t)(r0 + y) * W + c0 + x] = (uint8_t)score;
				else {
					const int slot = atomicAdd(&s_n, 1);
					if (slot < list_cap) sL[slot] = (score << 24) | ((uint32_t)y << 12) | (uint32_t)x;
				}
			}
		}
	}
	if (score_map) return;
	__syncthreads();
	// Warp 0 alone appends the band's survivors (a few dozen) to the bin's list: the round trip of the reserving atomic is
	// not a block-wide wait.  In the persistent kernels the other warps go on to the next item meanwhile: its first two
	// barriers include warp 0, so the survivor buffer and s_n (reset by thread 0 at the top) are not touched before the
	// copy below is done.
	if (tid < 32) {
		const int n = min(s_n, list_cap);
		if (n > 0) {
			int base = 0;
			if (tid == 0) base = atomicAdd(&s.kp_count[stream * LK_BINS + bin], n);
			base = __shfl_sync(0xFFFFFFFFu, base, 0);
			uint32_t *list = s.kp + (stream * LK_BINS + bin) * (size_t)g.kp_cap;
			for (int i = tid; i < n; i += 32)
				if (base + i < g.kp_cap) list[base + i] = sL[i];
		}
	}
}

// one CTA per (stream, bin, band)
template <int THREADS, int PC>
__global__ void __launch_bounds__(THREADS, 1280 / THREADS)
lk_fast_band_kernel(LkGeom g, const uint8_t *__restrict__ l0, LkState s, uint8_t *__restrict__ score_map, int bands, int P_rt,
		    int list_cap)
{
	pdl_enter();
	const int bin = blockIdx.x / bands;
	fb_body<THREADS, PC>(g, l0, s, score_map, bands, P_rt, list_cap, blockIdx.y, bin, blockIdx.x - bin * bands, nullptr, 0u);
}

// MASKED: a persistent grid over the plan's work list (bands with at least one live unit)
template <int THREADS, int PC>
__global__ void __launch_bounds__(THREADS, 1280 / THREADS)
lk_fast_band_masked_kernel(LkGeom g, const uint8_t *__restrict__ l0, LkState s, int bands, int P_rt, int list_cap)
{
	pdl_enter();
	const int n_work = s.n_work[0];
	int w = blockIdx.x;
	// the item and its header are two dependent global round trips: those of the NEXT item are issued before this item's
	// work and land while it runs
	uint32_t item = w < n_work ? s.work_list[w] : 0u;          // (stream * 16 + bin) * bands + band
	uint32_t hdr = w < n_work ? s.plan_hdr[item] : 0u;
	for (; w < n_work; w += gridDim.x) {
		const int wn = w + (int)gridDim.x;
		uint32_t item_n = 0u, hdr_n = 0u;
		if (wn < n_work) { item_n = s.work_list[wn]; hdr_n = s.plan_hdr[item_n]; }
		const uint32_t sb = item / (uint32_t)bands;
		fb_body<THREADS, PC>(g, l0, s, nullptr, bands, P_rt, list_cap, (size_t)(sb >> 4), (int)(sb & 15u), (int)(item - sb * bands),
				     s.plan_units + (size_t)item * g.fb_ucap, hdr);
		item = item_n; hdr = hdr_n;                         // (no barrier: see the end of fb_body)
	}
}

// The plan of the masked detector, one CTA per (stream, bin that needs features): the mask of the pixels the first
// selection stage can accept from (plus one ring pixel for the suppression), then per band the live units and the
// band's entry in the work list.
constexpr int FP_THREADS = 128;
__global__ void __launch_bounds__(FP_THREADS)
lk_fast_plan_kernel(LkGeom g, LkState s)
{
	pdl_enter();
	extern __shared__ __align__(16) uint32_t fp_mask[];           // bh x mw words (mw = mask words per row + 1)
	__shared__ short2 s_ex[LK_MAX_FEATURES];
	__shared__ int s_nex;
	__shared__ int s_cnt[64], s_lo[64], s_hi[64];                 // per band (fb_bands <= 64: lk_fast_band_supported)
	const size_t stream = blockIdx.y;
	const int bin = blockIdx.x, tid = threadIdx.x;
	const int bw = g.bw, bh = g.bh, dist = g.dist, N = g.n_feat, bands = g.fb_bands;
	// everything the CTA reads from global memory is requested at once (one round trip instead of a chain of four)
	const int needed = s.bin_needed[stream * LK_BINS + bin];
	const int npc = s.n_prev_corners[stream];
	uint8_t st0 = 0;
	float2 p0 = make_float2(0.f, 0.f);
	if (tid < N) { st0 = s.prev_status[stream * N + tid]; p0 = s.prev_corners[stream * N + tid]; }
	if (needed <= 0) return;
	const int c0 = (bin >> 2) * bw, r0 = (bin & 3) * bh;
	const int fmw = (bw + 31) >> 5, mw = fmw + 1, nseg = ((bw >> 1) + 15) >> 4;
	if (tid == 0) s_nex = 0;
	if (tid < bands) { s_cnt[tid] = 0; s_lo[tid] = FB_SROWS; s_hi[tid] = -1; }
	__syncthreads();
	// active points whose (shrunk) exclusion square reaches into the bin; int-truncated as in trackFeatures.cpp:212-224
	for (int k = tid; k < npc; k += FP_THREADS) {                 // (N <= LK_MAX_FEATURES <= FP_THREADS: one slot per thread, held above)
		const uint8_t st = k == tid ? st0 : s.prev_status[stream * N + k];
		if (!st) continue;
		const float2 p = k == tid ? p0 : s.prev_corners[stream * N + k];
		const int ex = (int)p.x - c0, ey = (int)p.y - r0;         // bin-local
		if (ex > -dist && ex < bw - 1 + dist && ey > -dist && ey < bh - 1 + dist)
			s_ex[atomicAdd(&s_nex, 1)] = make_short2((short)ex, (short)ey);
	}
	__syncthreads();
	const int nex = s_nex;
	const size_t item0 = (stream * LK_BINS + bin) * (size_t)bands;
	if (fmw <= 8) {
		// bins up to 256 px wide: the column interval of every point as 8 mask words (once per point), then one thread per
		// row ANDs the words of the points whose square covers the row in registers and emits the row's live units for
		// the one or two bands the row belongs to -- no mask array, no second pass
		uint32_t *xm = fp_mask;                                   // [nex][8]
		for (int i = tid; i < nex * 8; i += FP_THREADS) {
			const short2 e = s_ex[i >> 3];
			xm[i] = ~fb_bits(max(e.x - dist + 2, 0), min(e.x + dist - 2, bw - 1), 32 * (i & 7));
		}
		__syncthreads();
		for (int y = 3 + tid; y < bh - 3; y += FP_THREADS) {
			uint32_t m[9];
#pragma unroll
			for (int w = 0; w < 8; w++) m[w] = w < fmw ? fb_bits(3, bw - 4, 32 * w) : 0u;
			m[8] = 0u;
			for (int k = 0; k < nex; k++) {
				if (abs(s_ex[k].y - y) >= dist - 1) continue;
				const uint4 lo4 = *reinterpret_cast<const uint4 *>(xm + 8 * k), hi4 = *reinterpret_cast<const uint4 *>(xm + 8 * k + 4);
				m[0] &= lo4.x; m[1] &= lo4.y; m[2] &= lo4.z; m[3] &= lo4.w;
				m[4] &= hi4.x; m[5] &= hi4.y; m[6] &= hi4.z; m[7] &= hi4.w;
			}
			uint32_t live = 0;
#pragma unroll
			for (int seg = 0; seg < 8; seg++)
				if (seg < nseg && ((m[seg] >> 1) | (m[seg + 1] << 31))) live |= 1u << seg;    // pixels 32 seg + 1 .. 32 seg + 32
			if (!live) continue;
			// row y is response row y - (3 + 38 b - 1) of band b; rows at a band seam belong to both neighbours
			const int b_hi = min((y - 2) / FB_ROWS, bands - 1);
			for (int bnd = max(b_hi - 1, 0); bnd <= b_hi; bnd++) {
				const int y0 = 3 + bnd * FB_ROWS, y1 = min(y0 + FB_ROWS, bh - 3), sr = y - (y0 - 1);
				if (sr < 0 || sr >= y1 - y0 + 2) continue;
				uint16_t *out = s.plan_units + (item0 + bnd) * g.fb_ucap + atomicAdd(&s_cnt[bnd], __popc(live));
				for (uint32_t l = live; l; l &= l - 1u) *out++ = (uint16_t)((sr << 8) | (__ffs((int)l) - 1));
				atomicMin(&s_lo[bnd], sr);
				atomicMax(&s_hi[bnd], sr);
			}
		}
	} else {
	// one thread per row: the points whose square covers the row are found once, their column intervals cut out of all
	// the words of the row
	for (int y = tid; y < bh; y += FP_THREADS) {
		uint32_t *row = fp_mask + y * mw;
		const bool scored = y >= 3 && y < bh - 3;
		for (int w = 0; w < mw; w++) row[w] = (scored && w < fmw) ? fb_bits(3, bw - 4, 32 * w) : 0u;
		if (!scored) continue;
		for (int k = 0; k < nex; k++) {
			const short2 e = s_ex[k];
			if (abs(e.y - y) >= dist - 1) continue;
			const int lo = max(e.x - dist + 2, 0), hi = min(e.x + dist - 2, bw - 1);
			for (int w = lo >> 5; w <= (hi >> 5); w++) row[w] &= ~fb_bits(lo, hi, 32 * w);
		}
	}
	__syncthreads();
	// units: band b, response row sr <-> bin row 3 + 38 b - 1 + sr; only rows that carry a response (3 <= y < bh - 3).
	// One thread per (band, response row).
	for (int i = tid; i < bands * FB_SROWS; i += FP_THREADS) {
		const int b = i / FB_SROWS, sr = i - b * FB_SROWS;
		const int y0 = 3 + b * FB_ROWS, y1 = min(y0 + FB_ROWS, bh - 3), y = y0 - 1 + sr;
		if (sr >= y1 - y0 + 2 || y < 3 || y >= bh - 3) continue;
		const uint32_t *row = fp_mask + y * mw;
		int live = 0;
		for (int seg = 0; seg < nseg; seg++)
			if ((row[seg] >> 1) | (row[seg + 1] << 31)) {                                 // pixels 32 seg + 1 .. 32 seg + 32
				s.plan_units[(item0 + b) * g.fb_ucap + atomicAdd(&s_cnt[b], 1)] = (uint16_t)((sr << 8) | seg);
				live = 1;
			}
		if (live) { atomicMin(&s_lo[b], sr); atomicMax(&s_hi[b], sr); }
	}
	}
	__syncthreads();
	if (tid < bands && s_cnt[tid] > 0) {
		s.plan_hdr[item0 + tid] = (uint32_t)s_cnt[tid] | ((uint32_t)s_lo[tid] << 16) | ((uint32_t)s_hi[tid] << 24);
		s.work_list[atomicAdd(&s.n_work[0], 1)] = (uint32_t)(item0 + tid);
		atomicAdd(&s.lz_stats[0], 1ull);                          // diagnostics: bands with work, their live units
		atomicAdd(&s.lz_stats[1], (unsigned long long)s_cnt[tid]);
	}
}

// The complete keypoint lists of the FLAGGED streams only (lazy FAST, lk_fast_lazy.cu: the selection kernel lists the
// streams whose second-chance pass needs every keypoint in s.full_list / s.n_full[0]).  A small persistent grid walks
// over (flagged stream, bin, band): with nothing flagged -- the usual case -- the launch costs a few microseconds.
template <int THREADS, int PC>
__global__ void __launch_bounds__(THREADS, 1280 / THREADS)
lk_fast_band_flagged_kernel(LkGeom g, const uint8_t *__restrict__ l0, LkState s, int bands, int P_rt, int list_cap)
{
	pdl_enter();
	const int n_flag = s.n_full[0];
	const int per_stream = LK_BINS * bands;
	for (int w = blockIdx.x; w < n_flag * per_stream; w += gridDim.x) {
		const int f = w / per_stream, r = w - f * per_stream, bin = r / bands;
		fb_body<THREADS, PC>(g, l0, s, nullptr, bands, P_rt, list_cap, (size_t)s.full_list[f], bin, r - bin * bands, nullptr, 0u);
		__syncthreads();                                    // the next item reuses the shared-memory tiles and counters
	}
}

// ---- host side -------------------------------------------------------------
static size_t fb_smem_bytes(const LkGeom &g, int &P, int &list_cap)
{
	const int npc = g.bw / 2;
	P = npc + 4;
	list_cap = ((npc + 1) / 2 + 1) * (FB_ROWS / 2 + 1); // strict 3 x 3 maxima cannot be denser
	return ((size_t)(2 * FB_PROWS + FB_SROWS) * P + list_cap) * sizeof(uint32_t);
}

bool lk_fast_band_supported(const LkGeom &g, const uint8_t *l0)
{
	if (g.tun_fast_v1 == 1) return false;
	if (g.bw % 4 || g.w[0] % 4 || (reinterpret_cast<uintptr_t>(l0) & 3) || g.bh < 7 || g.bw < 8 || g.bw > 512) return false;
	int P, cap;
	return fb_smem_bytes(g, P, cap) <= 200 * 1024;
}

template <int THREADS, int PC>
static cudaError_t fb_launch(const LkGeom &g, const uint8_t *l0, const LkState &s, int n, uint8_t *score_map, cudaStream_t st)
{
	int P, cap;
	const size_t smem = fb_smem_bytes(g, P, cap);
	static bool attr_set[64] = {};
	int dev = 0;
	cudaGetDevice(&dev);
	if (dev < 64 && !attr_set[dev]) {
		cudaError_t e = cudaFuncSetAttribute(lk_fast_band_kernel<THREADS, PC>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
		if (e != cudaSuccess) return e;
		attr_set[dev] = true;
	}
	const int bands = (g.bh - 6 + FB_ROWS - 1) / FB_ROWS;
	dim3 grid(LK_BINS * bands, n);
	launch_k(g.tun_pdl != 0, lk_fast_band_kernel<THREADS, PC>, grid, THREADS, smem, st, g, l0, s, score_map, bands, P, cap);
	return cudaGetLastError();
}

template <int THREADS, int PC>
static cudaError_t fb_launch_flagged(const LkGeom &g, const uint8_t *l0, const LkState &s, int n, cudaStream_t st)
{
	int P, cap;
	const size_t smem = fb_smem_bytes(g, P, cap);
	static int grid_dev[64] = {};
	int dev = 0;
	cudaGetDevice(&dev);
	if (dev >= 64) return cudaErrorInvalidDevice;
	if (!grid_dev[dev]) {
		cudaError_t e = cudaFuncSetAttribute(lk_fast_band_flagged_kernel<THREADS, PC>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
		if (e != cudaSuccess) return e;
		int sms = 0;
		if ((e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev)) != cudaSuccess) return e;
		grid_dev[dev] = sms * (1280 / THREADS);
	}
	const int bands = (g.bh - 6 + FB_ROWS - 1) / FB_ROWS;
	long long items = (long long)n * LK_BINS * bands;
	const int grid = (int)(items < grid_dev[dev] ? items : grid_dev[dev]);
	launch_k(g.tun_pdl != 0, lk_fast_band_flagged_kernel<THREADS, PC>, grid, THREADS, smem, st, g, l0, s, bands, P, cap);
	return cudaGetLastError();
}

template <int THREADS, int PC>
static cudaError_t fb_launch_masked(const LkGeom &g, const uint8_t *l0, const LkState &s, int n, cudaStream_t st)
{
	int P, cap;
	const size_t smem = fb_smem_bytes(g, P, cap);
	static int grid_dev[64] = {};
	int dev = 0;
	cudaGetDevice(&dev);
	if (dev >= 64) return cudaErrorInvalidDevice;
	if (!grid_dev[dev]) {
		cudaError_t e = cudaFuncSetAttribute(lk_fast_band_masked_kernel<THREADS, PC>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
		if (e != cudaSuccess) return e;
		int sms = 0, per_sm = 0;
		if ((e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev)) != cudaSuccess) return e;
		if ((e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, lk_fast_band_masked_kernel<THREADS, PC>, THREADS, smem)) != cudaSuccess) return e;
		grid_dev[dev] = sms * (per_sm > 0 ? per_sm : 1);
	}
	const long long items = (long long)n * LK_BINS * g.fb_bands;
	const int grid = (int)(items < grid_dev[dev] ? items : grid_dev[dev]);
	launch_k(g.tun_pdl != 0, lk_fast_band_masked_kernel<THREADS, PC>, grid, THREADS, smem, st, g, l0, s, g.fb_bands, P, cap);
	return cudaGetLastError();
}

static size_t fp_smem_bytes(const LkGeom &g)
{
	const size_t rows = (size_t)g.bh * (((g.bw + 31) >> 5) + 1), pts = (size_t)LK_MAX_FEATURES * 8;   // mask rows / per-point words
	return (rows > pts ? rows : pts) * sizeof(uint32_t);
}

bool lk_fast_masked_supported(const LkGeom &g, const uint8_t *l0)
{
	return lk_fast_band_supported(g, l0) && g.fb_bands <= 64 && fp_smem_bytes(g) <= 48 * 1024 && g.dist >= 2;
}

// plan (live units per band + work list), then the persistent grid over the bands with work
cudaError_t lk_fast_masked_launch(const LkGeom &g, const uint8_t *l0, const LkState &s, int n, cudaStream_t st)
{
	if (n <= 0) return cudaSuccess;
	{
		KernelTimer kt(KID_LK_FAST_PLAN, st, "lk_fast_plan_kernel");
		launch_k(g.tun_pdl != 0, lk_fast_plan_kernel, dim3(LK_BINS, n), FP_THREADS, fp_smem_bytes(g), st, g, s);
		count_launch();
		cudaError_t e = cudaGetLastError();
		if (e != cudaSuccess) return e;
	}
	const int npc = g.bw / 2;
	KernelTimer kt(KID_LK_FAST, st, npc == 80 ? "lk_fast_band_masked_kernel<320,84>" : "lk_fast_band_masked_kernel");
	count_launch();
	if (npc == 80) return fb_launch_masked<AOF_FB_MTHREADS, 84>(g, l0, s, n, st);
	if (4 * npc <= 320) return fb_launch_masked<320, 0>(g, l0, s, n, st);
	if (4 * npc <= 640) return fb_launch_masked<640, 0>(g, l0, s, n, st);
	return fb_launch_masked<1024, 0>(g, l0, s, n, st);
}

cudaError_t lk_fast_band_flagged_launch(const LkGeom &g, const uint8_t *l0, const LkState &s, int n, cudaStream_t st)
{
	if (n <= 0) return cudaSuccess;
	const int npc = g.bw / 2;
	KernelTimer kt(KID_LK_FAST_FLAGGED, st, npc == 80 ? "lk_fast_band_flagged_kernel<320,84>" : "lk_fast_band_flagged_kernel");
	count_launch();
	if (npc == 80) return fb_launch_flagged<320, 84>(g, l0, s, n, st);
	if (4 * npc <= 320) return fb_launch_flagged<320, 0>(g, l0, s, n, st);
	if (4 * npc <= 640) return fb_launch_flagged<640, 0>(g, l0, s, n, st);
	return fb_launch_flagged<1024, 0>(g, l0, s, n, st);
}

cudaError_t lk_fast_band_launch(const LkGeom &g, const uint8_t *l0, const LkState &s, int n, uint8_t *score_map, cudaStream_t st)
{
	const int npc = g.bw / 2;
	// 640-px frames: pitch known at compile time.  (640 threads per CTA was measured 24 % slower: 5 rows per thread do
	// not amortise the per-thread set-up.)
	if (npc == 80) return fb_launch<320, 84>(g, l0, s, n, score_map, st);
	if (4 * npc <= 320) return fb_launch<320, 0>(g, l0, s, n, score_map, st);
	if (4 * npc <= 640) return fb_launch<640, 0>(g, l0, s, n, score_map, st);
	return fb_launch<1024, 0>(g, l0, s, n, score_map, st);
}

} // namespace aof
