// K1  sad_flow_sparse: PX4Flow::compute_flow for a batch of independent frame
// pairs, one CTA per pair.   Replaces
//   /root/reference/modules/OpticalFlow/src/px4flow.cpp:349-527 (compute_flow)
//   with its primitives :114 (compute_diff), :300 (compute_sad_8x8), :155
//   (compute_subpixel), bit-exactly.
// K7a px4_epilogue: OpticalFlow::limitRate (src/optical_flow.cpp:51-96) and the
//   atan2 conversion of OpticalFlowPX4::calcFlow (src/flow_px4.cpp:90-93), one
//   thread per stream.
//
// Design (B200, integer-ALU bound):
//   * both frames of the pair are brought into shared memory by two 1-D bulk
//     (TMA-engine) copies signalled on an mbarrier;
//   * phase A: the <=25 reference blocks of the previous frame are byte-aligned
//     once into 16 words each (PRMT) and texture-gated;
//   * phase B: one thread per (block, x-displacement): it slides down the
//     8+2S rows of its search column, byte-aligns each row ONCE (2 funnel
//     shifts) and feeds it to the <=8 y-displacements it belongs to
//     (VABSDIFF4.U8.ACC), 2S+1 running sums in registers; arg-min with the
//     reference's first-minimum order through the key (sad<<9 | candidate);
//   * phase C: one thread per (block, 4-px half): half-pixel refinement with
//     shared horizontal/vertical/diagonal averages (47 packed averages per
//     half instead of 96);
//   * phase D: one warp votes (integer sums of half-pixels, one IEEE divide).
#include "aof_common.cuh"
#include "aof_internal.h"

namespace aof {

template <int S> struct SadCfg {
	static constexpr int NI = 2 * S + 1;                       // displacements per axis
	static constexpr int ROWS = 8 + 2 * S;                     // rows of a search column
	static constexpr int TASKS = AOF_PX4_MAX_BLOCKS * NI;      // phase-B threads
	static constexpr int THREADS = (TASKS + 31) / 32 * 32;
};

template <int S>
__global__ void __launch_bounds__(SadCfg<S>::THREADS)
sad_flow_kernel(const uint8_t *__restrict__ prev, const uint8_t *__restrict__ cur, SadGeom g,
		aof_px4_raw *__restrict__ raw, aof_px4_block *__restrict__ blocks)
{
	using Cfg = SadCfg<S>;
	constexpr int NI = Cfg::NI;
	extern __shared__ __align__(128) uint8_t smem[];
	__shared__ __align__(16) uint32_t s_ablk[AOF_PX4_MAX_BLOCKS][16]; // aligned 8x8 blocks of prev
	__shared__ uint32_t s_key[AOF_PX4_MAX_BLOCKS * NI];
	__shared__ uint32_t s_tex[AOF_PX4_MAX_BLOCKS];
	__shared__ int s_vote[AOF_PX4_MAX_BLOCKS][2];                     // 2*dx + half, 2*dy + half (or INT_MIN)
	__shared__ __align__(8) uint64_t s_bar;

	const int tid = threadIdx.x;
	const int W = g.width;
	const uint32_t fb = g.stage_bytes;
	uint8_t *sp = smem;
	uint8_t *sc = smem + fb;
	const size_t pair = blockIdx.x;
	const int nblk = g.nb * g.nb;

	__shared__ uint32_t s_xy[AOF_PX4_MAX_BLOCKS];                     // block origin x | y << 16
	if (tid == 0) mbar_init(&s_bar, 1);
	if (tid < nblk) s_xy[tid] = (uint32_t)(g.lo + (tid % g.nb) * g.step) | ((uint32_t)(g.lo + (tid / g.nb) * g.step) << 16);
	__syncthreads();
	if (tid == 0) {
		mbar_expect_tx(&s_bar, 2 * fb);
		bulk_g2s(sp, prev + pair * g.frame_stride, fb, &s_bar);
		bulk_g2s(sc, cur + pair * g.frame_stride, fb, &s_bar);
	}
	mbar_wait(&s_bar, 0);

	// ---- phase A1: byte-align the reference blocks (thread = block, row) ----
	if (tid < nblk * 8) {
		const int b = tid >> 3, r = tid & 7;
		const int x = s_xy[b] & 0xFFFF, y = s_xy[b] >> 16;
		const uint32_t off = (uint32_t)(y + r) * W + x;
		const uint32_t *w = reinterpret_cast<const uint32_t *>(sp + (off & ~3u));
		const uint32_t sh = off & 3u;
		const uint32_t w0 = w[0], w1 = w[1], w2 = w[2];
		s_ablk[b][2 * r] = bytes_at(w0, w1, sh);
		s_ablk[b][2 * r + 1] = bytes_at(w1, w2, sh);
	}
	__syncthreads();
	// ---- phase A2: texture gate (compute_diff) on the central 4x4 ----
	if (tid < nblk) {
		uint32_t p[4];
#pragma unroll
		for (int r = 0; r < 4; r++) p[r] = __byte_perm(s_ablk[tid][2 * (r + 2)], s_ablk[tid][2 * (r + 2) + 1], 0x5432);
		uint32_t t = sad4(p[0], p[1], 0);
		t = sad4(p[1], p[2], t);
		t = sad4(p[2], p[3], t);
#pragma unroll
		for (int r = 0; r < 4; r++) t = sad4(p[r], __byte_perm(p[r], p[r], 0x3321), t); // |c0-c1|+|c1-c2|+|c2-c3|
		s_tex[tid] = t;
	}
	__syncthreads();

	// ---- phase B: integer search, thread = (block, ii) ----
	if (tid < nblk * NI) {
		const int b = tid / NI, ii = tid - b * NI;
		if (s_tex[b] >= g.feat_thr) {
			const int x = s_xy[b] & 0xFFFF, y = s_xy[b] >> 16;
			uint32_t a[16];
#pragma unroll
			for (int k = 0; k < 4; k++) {
				const uint4 v = reinterpret_cast<const uint4 *>(&s_ablk[b][0])[k];
				a[4 * k] = v.x; a[4 * k + 1] = v.y; a[4 * k + 2] = v.z; a[4 * k + 3] = v.w;
			}
			uint32_t acc[NI];
#pragma unroll
			for (int q = 0; q < NI; q++) acc[q] = 0;
			const uint32_t off = (uint32_t)(y - S) * W + (x - S + ii);
			const uint32_t sh8 = (off & 3u) * 8;
			const uint32_t *w = reinterpret_cast<const uint32_t *>(sc + (off & ~3u));
			const int wpitch = W >> 2;
#pragma unroll
			for (int r = 0; r < Cfg::ROWS; r++) {
				const uint32_t w0 = w[r * wpitch], w1 = w[r * wpitch + 1], w2 = w[r * wpitch + 2];
				const uint32_t c0 = __funnelshift_r(w0, w1, sh8);
				const uint32_t c1 = __funnelshift_r(w1, w2, sh8);
#pragma unroll
				for (int q = 0; q < NI; q++) {
					const int br = r - q; // row of the reference block this window row meets at dy = q - S
					if (br >= 0 && br < 8) {
						acc[q] = sad4(a[2 * br], c0, acc[q]);
						acc[q] = sad4(a[2 * br + 1], c1, acc[q]);
					}
				}
			}
			uint32_t best = 0xFFFFFFFFu;
#pragma unroll
			for (int q = 0; q < NI; q++) best = min(best, (acc[q] << 9) | (uint32_t)(q * NI + ii));
			s_key[tid] = best;
		}
	}
	__syncthreads();

	// ---- phase C: accept + half-pixel refinement, thread = (block, half) ----
	if (tid < 64) { // warps 0,1 whole (the shuffle below needs every lane)
		const bool live = tid < nblk * 2;
		const int b = live ? (tid >> 1) : 0, half = tid & 1;
		const int x = s_xy[b] & 0xFFFF, y = s_xy[b] >> 16;
		const uint32_t tex = s_tex[b];
		uint32_t best_sad = 0xFFFFFFFFu;
		int dx = 0, dy = 0, subdir = 8, accepted = 0;
		uint32_t acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
		if (live && tex >= g.feat_thr) {
			uint32_t key = 0xFFFFFFFFu;
#pragma unroll
			for (int k = 0; k < NI; k++) key = min(key, s_key[b * NI + k]);
			best_sad = key >> 9;
			const int cand = key & 511;
			dy = cand / NI - S;
			dx = cand - (dy + S) * NI - S;
			accepted = best_sad < g.val_thr;
		}
		if (accepted) {
			// bytes q[-1..4] of rows -1..8 of this half of the matched block
			const uint32_t off = (uint32_t)(y + dy - 1) * W + (x + dx + 4 * half - 1);
			const uint32_t sh = off & 3u;
			const uint32_t *w = reinterpret_cast<const uint32_t *>(sc + (off & ~3u));
			const int wpitch = W >> 2;
			uint32_t pQ = 0, pH0 = 0, pH4 = 0;     // previous row: centre, havg(c, c+1x), havg(c, c-1x)
			uint32_t pV = 0, pT0 = 0, pT4 = 0;     // previous row pair: vertical / diagonal averages
#pragma unroll
			for (int rr = 0; rr < 10; rr++) {      // window row rr-1
				const uint32_t u0 = w[rr * wpitch], u1 = w[rr * wpitch + 1], u2 = w[rr * wpitch + 2];
				const uint32_t X = bytes_at(u0, u1, sh);       // q[-1..2]
				const uint32_t Y = bytes_at(u1, u2, sh);       // q[3..6]
				const uint32_t Q = __byte_perm(X, Y, 0x4321);  // q[0..3]
				const uint32_t Qp = __byte_perm(X, Y, 0x5432); // q[1..4]
				const uint32_t H4 = havg4(Q, X);
				const uint32_t H0 = havg4(Q, Qp);
				if (rr >= 1) {
					const uint32_t V = havg4(pQ, Q), T0 = havg4(pH0, H0), T4 = havg4(pH4, H4);
					if (rr >= 2) { // block row i = rr-2 is the previous window row; pV.. belong to i-1
						const uint32_t ai = s_ablk[b][2 * (rr - 2) + half];
						acc[0] = sad4(ai, pH0, acc[0]);
						acc[1] = sad4(ai, T0, acc[1]);
						acc[2] = sad4(ai, V, acc[2]);
						acc[3] = sad4(ai, T4, acc[3]);
						acc[4] = sad4(ai, pH4, acc[4]);
						acc[5] = sad4(ai, pT4, acc[5]);
						acc[6] = sad4(ai, pV, acc[6]);
						acc[7] = sad4(ai, pT0, acc[7]);
					}
					pV = V; pT0 = T0; pT4 = T4;
				}
				pQ = Q; pH0 = H0; pH4 = H4;
			}
		}
		// both halves of a block sit in adjacent lanes of one warp
#pragma unroll
		for (int k = 0; k < 8; k++) acc[k] += __shfl_xor_sync(0xFFFFFFFFu, acc[k], 1);
		if (accepted) {
			uint32_t m = best_sad;
#pragma unroll
			for (int k = 0; k < 8; k++)
				if (acc[k] < m) { m = acc[k]; subdir = k; }
		}
		if (live && half == 0) {
			int vx = INT_MIN, vy = 0;
			if (accepted) {
				const int hx = (subdir == 0 || subdir == 1 || subdir == 7) - (subdir == 3 || subdir == 4 || subdir == 5);
				const int hy = (subdir == 1 || subdir == 2 || subdir == 3) - (subdir == 5 || subdir == 6 || subdir == 7);
				vx = 2 * dx + hx;
				vy = 2 * dy + hy;
			}
			s_vote[b][0] = vx;
			s_vote[b][1] = vy;
			if (blocks) {
				aof_px4_block o;
				o.x = (int16_t)x; o.y = (int16_t)y; o.dx = (int16_t)dx; o.dy = (int16_t)dy;
				o.texture = tex; o.best_sad = best_sad;
#pragma unroll
				for (int k = 0; k < 8; k++) o.acc[k] = accepted ? acc[k] : 0u;
				o.subdir = subdir; o.accepted = accepted;
				blocks[pair * AOF_PX4_MAX_BLOCKS + b] = o;
			}
		} else if (blocks && half == 0 && tid < 2 * AOF_PX4_MAX_BLOCKS) { // grid positions that do not exist
			aof_px4_block o = {};
			o.best_sad = 0xFFFFFFFFu; o.subdir = 8; o.x = -1; o.y = -1;
			blocks[pair * AOF_PX4_MAX_BLOCKS + (tid >> 1)] = o;
		}
	}
	__syncthreads();

	// ---- phase D: mean vote (px4flow.cpp:450-524) ----
	if (tid < 32) {
		const bool on = tid < nblk && s_vote[tid][0] != INT_MIN;
		const int vx = on ? s_vote[tid][0] : 0, vy = on ? s_vote[tid][1] : 0;
		const int count = __popc(__ballot_sync(0xFFFFFFFFu, on));
		const int sx = __reduce_add_sync(0xFFFFFFFFu, vx);
		const int sy = __reduce_add_sync(0xFFFFFFFFu, vy);
		if (tid == 0) {
			aof_px4_raw o;
			o.count = count;
			if (count > 10) {
				// sums of multiples of 0.5 are exact in f32; one correctly rounded divide
				o.px_x = __fdiv_rn((float)sx * 0.5f, (float)count);
				o.px_y = __fdiv_rn((float)sy * 0.5f, (float)count);
				o.quality = (count * 255 / (AOF_PX4_MAX_BLOCKS)) & 0xFF;
			} else {
				o.px_x = 0.f; o.px_y = 0.f; o.quality = 0;
			}
			raw[pair] = o;
		}
	}
}

// ---------------------------------------------------------------------------
// Run-time search size: any S >= 1 the frame can hold (include/flow_px4.hpp:62-65 takes any int; the templates above
// are the fast paths for 4, 6, 8).  Same phases as sad_flow_kernel, but the integer search is a loop over
// (block, candidate) tasks, 8 rows of two byte-aligned words each, and the first-minimum key is
// sad << 18 | candidate (sad < 2^14, candidate < (2S+1)^2 < 2^18) reduced by shared-memory atomicMin.
__global__ void __launch_bounds__(256)
sad_flow_any_kernel(const uint8_t *__restrict__ prev, const uint8_t *__restrict__ cur, SadGeom g, int S,
		    aof_px4_raw *__restrict__ raw, aof_px4_block *__restrict__ blocks)
{
	extern __shared__ __align__(128) uint8_t smem[];
	__shared__ __align__(16) uint32_t s_ablk[AOF_PX4_MAX_BLOCKS][16];
	__shared__ uint32_t s_best[AOF_PX4_MAX_BLOCKS];
	__shared__ uint32_t s_tex[AOF_PX4_MAX_BLOCKS];
	__shared__ int s_vote[AOF_PX4_MAX_BLOCKS][2];
	__shared__ uint32_t s_xy[AOF_PX4_MAX_BLOCKS];
	__shared__ __align__(8) uint64_t s_bar;

	const int tid = threadIdx.x;
	const int W = g.width, wpitch = W >> 2, NI = 2 * S + 1;
	const uint32_t fb = g.stage_bytes;
	uint8_t *sp = smem;
	uint8_t *sc = smem + fb;
	const size_t pair = blockIdx.x;
	const int nblk = g.nb * g.nb;

	if (tid == 0) mbar_init(&s_bar, 1);
	if (tid < AOF_PX4_MAX_BLOCKS) {
		s_best[tid] = 0xFFFFFFFFu;
		if (tid < nblk) s_xy[tid] = (uint32_t)(g.lo + (tid % g.nb) * g.step) | ((uint32_t)(g.lo + (tid / g.nb) * g.step) << 16);
	}
	__syncthreads();
	if (tid == 0) {
		mbar_expect_tx(&s_bar, 2 * fb);
		bulk_g2s(sp, prev + pair * g.frame_stride, fb, &s_bar);
		bulk_g2s(sc, cur + pair * g.frame_stride, fb, &s_bar);
	}
	mbar_wait(&s_bar, 0);

	if (tid < nblk * 8) { // phase A1
		const int b = tid >> 3, r = tid & 7;
		const int x = s_xy[b] & 0xFFFF, y = s_xy[b] >> 16;
		const uint32_t off = (uint32_t)(y + r) * W + x;
		const uint32_t *w = reinterpret_cast<const uint32_t *>(sp + (off & ~3u));
		const uint32_t sh = off & 3u;
		const uint32_t w0 = w[0], w1 = w[1], w2 = w[2];
		s_ablk[b][2 * r] = bytes_at(w0, w1, sh);
		s_ablk[b][2 * r + 1] = bytes_at(w1, w2, sh);
	}
	__syncthreads();
	if (tid < nblk) { // phase A2
		uint32_t p[4];
#pragma unroll
		for (int r = 0; r < 4; r++) p[r] = __byte_perm(s_ablk[tid][2 * (r + 2)], s_ablk[tid][2 * (r + 2) + 1], 0x5432);
		uint32_t t = sad4(p[0], p[1], 0);
		t = sad4(p[1], p[2], t);
		t = sad4(p[2], p[3], t);
#pragma unroll
		for (int r = 0; r < 4; r++) t = sad4(p[r], __byte_perm(p[r], p[r], 0x3321), t);
		s_tex[tid] = t;
	}
	__syncthreads();

	// ---- phase B: task = (block, candidate), candidate = (dy + S) * NI + (dx + S) in the reference's loop order ----
	for (int task = tid; task < nblk * NI * NI; task += blockDim.x) {
		const int b = task / (NI * NI), cand = task - b * NI * NI;
		if (s_tex[b] < g.feat_thr) continue;
		const int jj = cand / NI, ii = cand - jj * NI;
		const int x = s_xy[b] & 0xFFFF, y = s_xy[b] >> 16;
		const uint32_t off = (uint32_t)(y - S + jj) * W + (x - S + ii);
		const uint32_t sh8 = (off & 3u) * 8;
		const uint32_t *w = reinterpret_cast<const uint32_t *>(sc + (off & ~3u));
		uint32_t acc = 0;
#pragma unroll
		for (int br = 0; br < 8; br++) {
			const uint32_t w0 = w[br * wpitch], w1 = w[br * wpitch + 1], w2 = w[br * wpitch + 2];
			acc = sad4(s_ablk[b][2 * br], __funnelshift_r(w0, w1, sh8), acc);
			acc = sad4(s_ablk[b][2 * br + 1], __funnelshift_r(w1, w2, sh8), acc);
		}
		atomicMin(&s_best[b], (acc << 18) | (uint32_t)cand);
	}
	__syncthreads();

	// ---- phase C: accept + half-pixel refinement, thread = (block, half) ----
	if (tid < 64) {
		const bool live = tid < nblk * 2;
		const int b = live ? (tid >> 1) : 0, half = tid & 1;
		const int x = s_xy[b] & 0xFFFF, y = s_xy[b] >> 16;
		const uint32_t tex = live ? s_tex[b] : 0u;
		uint32_t best_sad = 0xFFFFFFFFu;
		int dx = 0, dy = 0, subdir = 8, accepted = 0;
		uint32_t acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
		if (live && tex >= g.feat_thr) {
			const uint32_t key = s_best[b];
			best_sad = key >> 18;
			const int cand = key & 0x3FFFF;
			dy = cand / NI - S;
			dx = cand - (dy + S) * NI - S;
			accepted = best_sad < g.val_thr;
		}
		if (accepted) {
			const uint32_t off = (uint32_t)(y + dy - 1) * W + (x + dx + 4 * half - 1);
			const uint32_t sh = off & 3u;
			const uint32_t *w = reinterpret_cast<const uint32_t *>(sc + (off & ~3u));
			uint32_t pQ = 0, pH0 = 0, pH4 = 0, pV = 0, pT0 = 0, pT4 = 0;
#pragma unroll
			for (int rr = 0; rr < 10; rr++) {
				const uint32_t u0 = w[rr * wpitch], u1 = w[rr * wpitch + 1], u2 = w[rr * wpitch + 2];
				const uint32_t X = bytes_at(u0, u1, sh), Y = bytes_at(u1, u2, sh);
				const uint32_t Q = __byte_perm(X, Y, 0x4321), Qp = __byte_perm(X, Y, 0x5432);
				const uint32_t H4 = havg4(Q, X), H0 = havg4(Q, Qp);
				if (rr >= 1) {
					const uint32_t V = havg4(pQ, Q), T0 = havg4(pH0, H0), T4 = havg4(pH4, H4);
					if (rr >= 2) {
						const uint32_t ai = s_ablk[b][2 * (rr - 2) + half];
						acc[0] = sad4(ai, pH0, acc[0]); acc[1] = sad4(ai, T0, acc[1]);
						acc[2] = sad4(ai, V, acc[2]);   acc[3] = sad4(ai, T4, acc[3]);
						acc[4] = sad4(ai, pH4, acc[4]); acc[5] = sad4(ai, pT4, acc[5]);
						acc[6] = sad4(ai, pV, acc[6]);  acc[7] = sad4(ai, pT0, acc[7]);
					}
					pV = V; pT0 = T0; pT4 = T4;
				}
				pQ = Q; pH0 = H0; pH4 = H4;
			}
		}
#pragma unroll
		for (int k = 0; k < 8; k++) acc[k] += __shfl_xor_sync(0xFFFFFFFFu, acc[k], 1);
		if (accepted) {
			uint32_t m = best_sad;
#pragma unroll
			for (int k = 0; k < 8; k++)
				if (acc[k] < m) { m = acc[k]; subdir = k; }
		}
		if (live && half == 0) {
			int vx = INT_MIN, vy = 0;
			if (accepted) {
				const int hx = (subdir == 0 || subdir == 1 || subdir == 7) - (subdir == 3 || subdir == 4 || subdir == 5);
				const int hy = (subdir == 1 || subdir == 2 || subdir == 3) - (subdir == 5 || subdir == 6 || subdir == 7);
				vx = 2 * dx + hx;
				vy = 2 * dy + hy;
			}
			s_vote[b][0] = vx;
			s_vote[b][1] = vy;
			if (blocks) {
				aof_px4_block o;
				o.x = (int16_t)x; o.y = (int16_t)y; o.dx = (int16_t)dx; o.dy = (int16_t)dy;
				o.texture = tex; o.best_sad = best_sad;
#pragma unroll
				for (int k = 0; k < 8; k++) o.acc[k] = accepted ? acc[k] : 0u;
				o.subdir = subdir; o.accepted = accepted;
				blocks[pair * AOF_PX4_MAX_BLOCKS + b] = o;
			}
		} else if (blocks && half == 0 && tid < 2 * AOF_PX4_MAX_BLOCKS) {
			aof_px4_block o = {};
			o.best_sad = 0xFFFFFFFFu; o.subdir = 8; o.x = -1; o.y = -1;
			blocks[pair * AOF_PX4_MAX_BLOCKS + (tid >> 1)] = o;
		}
	}
	__syncthreads();

	if (tid < 32) { // phase D
		const bool on = tid < nblk && s_vote[tid][0] != INT_MIN;
		const int vx = on ? s_vote[tid][0] : 0, vy = on ? s_vote[tid][1] : 0;
		const int count = __popc(__ballot_sync(0xFFFFFFFFu, on));
		const int sx = __reduce_add_sync(0xFFFFFFFFu, vx);
		const int sy = __reduce_add_sync(0xFFFFFFFFu, vy);
		if (tid == 0) {
			aof_px4_raw o;
			o.count = count;
			if (count > 10) {
				o.px_x = __fdiv_rn((float)sx * 0.5f, (float)count);
				o.px_y = __fdiv_rn((float)sy * 0.5f, (float)count);
				o.quality = (count * 255 / (AOF_PX4_MAX_BLOCKS)) & 0xFF;
			} else {
				o.px_x = 0.f; o.px_y = 0.f; o.quality = 0;
			}
			raw[pair] = o;
		}
	}
}

// limitRate + atan2, one thread per stream
__global__ void px4_epilogue_kernel(const aof_px4_raw *__restrict__ raw, const uint32_t *__restrict__ t_us,
				    LimiterState *__restrict__ lim, aof_flow_out *__restrict__ out, int n,
				    int output_rate, float focal_x, float focal_y)
{
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	const aof_px4_raw r = raw[i];
	LimiterState l = lim[i];
	aof_flow_out o;
	float fx = r.px_x, fy = r.px_y;
	o.quality = limit_rate(l, output_rate, r.quality, t_us[i], o.dt_us, fx, fy);
	o.flow_x = (float)atan2((double)fx, (double)focal_x);
	o.flow_y = (float)atan2((double)fy, (double)focal_y);
	lim[i] = l;
	out[i] = o;
}

// ---- host-side launchers ---------------------------------------------------
template <int S> static cudaError_t launch_sad(const uint8_t *prev, const uint8_t *cur, const SadGeom &g, int n,
					       aof_px4_raw *raw, aof_px4_block *blocks, cudaStream_t st)
{
	static bool attr_set[64] = {};
	const size_t dyn = 2 * (size_t)g.stage_bytes + 16;
	int dev = 0;
	cudaGetDevice(&dev);
	if (dev < 64 && !attr_set[dev]) {
		cudaError_t e = cudaFuncSetAttribute(sad_flow_kernel<S>, cudaFuncAttributeMaxDynamicSharedMemorySize,
						     200 * 1024);
		if (e != cudaSuccess) return e;
		attr_set[dev] = true;
	}
	KernelTimer kt(KID_SAD_FLOW, st, S == 4 ? "sad_flow_kernel<4>" : S == 6 ? "sad_flow_kernel<6>" : "sad_flow_kernel<8>");
	sad_flow_kernel<S><<<n, SadCfg<S>::THREADS, dyn, st>>>(prev, cur, g, raw, blocks);
	count_launch();
	return cudaGetLastError();
}

static cudaError_t launch_sad_any(int search, const uint8_t *prev, const uint8_t *cur, const SadGeom &g, int n,
				  aof_px4_raw *raw, aof_px4_block *blocks, cudaStream_t st)
{
	static bool attr_set[64] = {};
	const size_t dyn = 2 * (size_t)g.stage_bytes + 16;
	int dev = 0;
	cudaGetDevice(&dev);
	if (dev < 64 && !attr_set[dev]) {
		cudaError_t e = cudaFuncSetAttribute(sad_flow_any_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
		if (e != cudaSuccess) return e;
		attr_set[dev] = true;
	}
	KernelTimer kt(KID_SAD_FLOW, st, "sad_flow_any_kernel");
	sad_flow_any_kernel<<<n, 256, dyn, st>>>(prev, cur, g, search, raw, blocks);
	count_launch();
	return cudaGetLastError();
}

cudaError_t sad_flow_launch(int search, const uint8_t *prev, const uint8_t *cur, const SadGeom &g, int n,
			    aof_px4_raw *raw, aof_px4_block *blocks, cudaStream_t st)
{
	if (n <= 0) return cudaSuccess;
	if (sad_flow3_supported(search, g)) {
		KernelTimer kt(KID_SAD_FLOW, st, "sad_flow3_kernel");
		count_launch();
		return sad_flow3_launch(prev, cur, g, n, raw, blocks, st);
	}
	if (search != 4 && search != 6 && search != 8) return launch_sad_any(search, prev, cur, g, n, raw, blocks, st);
	if (sad_flow2_supported(g)) {
		KernelTimer kt(KID_SAD_FLOW, st, search == 4 ? "sad_flow2_kernel<4>" : search == 6 ? "sad_flow2_kernel<6>" : "sad_flow2_kernel<8>");
		count_launch();
		return sad_flow2_launch(search, prev, cur, g, n, raw, blocks, st);
	}
	switch (search) {
	case 4: return launch_sad<4>(prev, cur, g, n, raw, blocks, st);
	case 6: return launch_sad<6>(prev, cur, g, n, raw, blocks, st);
	case 8: return launch_sad<8>(prev, cur, g, n, raw, blocks, st);
	default: return launch_sad_any(search, prev, cur, g, n, raw, blocks, st);
	}
}

cudaError_t px4_epilogue_launch(const aof_px4_raw *raw, const uint32_t *t_us, LimiterState *lim, aof_flow_out *out,
				int n, int output_rate, float fx, float fy, cudaStream_t st)
{
	if (n <= 0) return cudaSuccess;
	KernelTimer kt(KID_PX4_EPILOGUE, st);
	px4_epilogue_kernel<<<(n + 127) / 128, 128, 0, st>>>(raw, t_us, lim, out, n, output_rate, fx, fy);
	count_launch();
	return cudaGetLastError();
}

} // namespace aof
