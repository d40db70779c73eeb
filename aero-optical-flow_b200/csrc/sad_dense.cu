// K2  sad_dense: per-block SAD flow field over every 8x8 block of the 8-px lattice
// (BASELINE config 4: 1280x720, +-8 search).  The per-block arithmetic is the reference's
//   compute_diff      /root/reference/modules/OpticalFlow/src/px4flow.cpp:114-142
//   compute_sad_8x8   :300-333   (+ the first-minimum search order of :394-405)
//   compute_subpixel  :155-288   (+ the direction choice of :413-422)
// with 32-bit offsets (the reference's uint16_t offsets overflow above 65 535 px and its grid
// assumes a square image); a block is computed when its search window plus
// the 1-px sub-pixel halo lies inside the image.
//
// CTA = 8 x 2 blocks.  Their 64 x 16 reference pixels and the (64 + 24) x (16 + 2S + 2) search
// region of the current frame are staged in shared memory with coalesced 32-bit loads; then the
// phases of sad_flow_kernel: texture gate, thread = (block, x-displacement) sliding search with
// one byte-alignment per row shared by all y-displacements, packed-key arg-min, half-pixel
// refinement by (block, 4-px half).  Integer-ALU bound: 289 x 16 VABSDIFF4 per block at S = 8.
#include "aof_common.cuh"
#include "aof_internal.h"

namespace aof {

#ifndef DENSE_MINB
#define DENSE_MINB 5   // CTAs per SM (40 registers, no spill).  Measured on B200, 1280x720 +-8: 8x4-block CTAs at 2 / 3 per SM
                       // 105 k / 127 k pairs/s, 8x2-block CTAs at 5 per SM 133 k: more, smaller CTAs overlap their phases better
#endif
#ifndef DENSE_FMA
#define DENSE_FMA 0   // 1 = byte alignment of the window rows on the FMA pipe.  Measured (r2D, 1280x720 +-8): 112.7 k pairs/s against
                      // 133.1 k with funnel shifts, although the ALU pipe is 84 % busy: IMAD.HI issues at half rate and the four
                      // extra instructions per row cost more issue slots than the two ALU slots they free
#endif
constexpr int DN_BX = 8, DN_BY = 2, DN_BLOCKS = DN_BX * DN_BY;
constexpr int DN_MARGIN = 12;                       // columns left of the tile in smem (>= S + 1, multiple of 4)
constexpr int DN_CP = DN_BX * 8 + 2 * DN_MARGIN;    // pitch of the current-frame tile: 88 bytes
constexpr int DN_PP = DN_BX * 8;                    // pitch of the reference tile: 64 bytes

template <int S> struct DenseCfg {
	static constexpr int NI = 2 * S + 1;
	static constexpr int ROWS = 8 + 2 * S;                  // rows of a search column
	static constexpr int CH = DN_BY * 8 + 2 * (S + 1);      // rows of the current-frame tile
	static constexpr int TASKS = DN_BLOCKS * NI;
	static constexpr int THREADS = (TASKS + 31) / 32 * 32;
};

template <int S>
__global__ void __launch_bounds__(DenseCfg<S>::THREADS, DENSE_MINB)
sad_dense_kernel(const uint8_t *__restrict__ prev, const uint8_t *__restrict__ cur, DenseGeom g, aof_dense_block *__restrict__ field)
{
	using Cfg = DenseCfg<S>;
	constexpr int NI = Cfg::NI;
	__shared__ __align__(16) uint8_t s_prev[DN_BY * 8 * DN_PP];
	__shared__ __align__(16) uint8_t s_cur[Cfg::CH * DN_CP];
	__shared__ uint32_t s_key[Cfg::TASKS];
	__shared__ uint32_t s_tex[DN_BLOCKS];
	const int tid = threadIdx.x;
	const int W = g.width, H = g.height;
	const size_t pair = blockIdx.z;
	const uint8_t *P = prev + pair * g.frame_stride, *C = cur + pair * g.frame_stride;
	const int X0 = 8 * (g.i_lo + DN_BX * blockIdx.x), Y0 = 8 * (g.j_lo + DN_BY * blockIdx.y);
	// ---- stage both tiles (32-bit loads; addresses clamped into the image, clamped values are never used).  One index
	// space over the words of both tiles, every load of a thread issued before its first store.
	{
		constexpr int NP = DN_BY * 8 * (DN_PP / 4), NC = Cfg::CH * (DN_CP / 4);
		constexpr int ROUNDS = (NP + NC + Cfg::THREADS - 1) / Cfg::THREADS;
		uint32_t v[ROUNDS];
#pragma unroll
		for (int k = 0; k < ROUNDS; k++) {
			const int i = tid + k * Cfg::THREADS;
			const uint8_t *src;
			if (i < NP) {
				const int r = i / (DN_PP / 4), q = i - r * (DN_PP / 4);
				const int y = min(Y0 + r, H - 1), x = min(X0 + 4 * q, W - 4);
				src = P + (size_t)y * W + x;
			} else {
				const int j = min(i - NP, NC - 1);
				const int r = j / (DN_CP / 4), q = j - r * (DN_CP / 4);
				const int y = min(max(Y0 - (S + 1) + r, 0), H - 1), x = min(max(X0 - DN_MARGIN + 4 * q, 0), W - 4);
				src = C + (size_t)y * W + x;
			}
			v[k] = __ldg(reinterpret_cast<const uint32_t *>(src));
		}
#pragma unroll
		for (int k = 0; k < ROUNDS; k++) {
			const int i = tid + k * Cfg::THREADS;
			if (i < NP) reinterpret_cast<uint32_t *>(s_prev)[i] = v[k];
			else if (i < NP + NC) reinterpret_cast<uint32_t *>(s_cur)[i - NP] = v[k];
		}
	}
	__syncthreads();
	// block b of the tile: lattice index (bi, bj), valid when inside the lattice
	auto block_valid = [&](int b) {
		const int bi = g.i_lo + DN_BX * blockIdx.x + (b & (DN_BX - 1)), bj = g.j_lo + DN_BY * blockIdx.y + (b >> 3);
		return bi <= g.i_hi && bj <= g.j_hi;
	};
	// ---- texture gate (compute_diff) on the central 4x4 of each block
	if (tid < DN_BLOCKS) {
		const uint32_t *w = reinterpret_cast<const uint32_t *>(s_prev) + (tid >> 3) * 8 * (DN_PP / 4) + (tid & 7) * 2;
		uint32_t p[4];
#pragma unroll
		for (int r = 0; r < 4; r++) p[r] = __byte_perm(w[(r + 2) * (DN_PP / 4)], w[(r + 2) * (DN_PP / 4) + 1], 0x5432);
		uint32_t t = sad4(p[0], p[1], 0);
		t = sad4(p[1], p[2], t);
		t = sad4(p[2], p[3], t);
#pragma unroll
		for (int r = 0; r < 4; r++) t = sad4(p[r], __byte_perm(p[r], p[r], 0x3321), t);
		s_tex[tid] = t;
	}
	__syncthreads();
	// ---- integer search, thread = (block, ii)
	if (tid < Cfg::TASKS) {
		const int b = tid / NI, ii = tid - b * NI;
		uint32_t best = 0xFFFFFFFFu;
		if (block_valid(b) && s_tex[b] >= g.feat_thr) {
			const uint32_t *pw = reinterpret_cast<const uint32_t *>(s_prev) + (b >> 3) * 8 * (DN_PP / 4) + (b & 7) * 2;
			uint32_t a[16];
#pragma unroll
			for (int r = 0; r < 8; r++) { a[2 * r] = pw[r * (DN_PP / 4)]; a[2 * r + 1] = pw[r * (DN_PP / 4) + 1]; }
			uint32_t acc[NI];
#pragma unroll
			for (int q = 0; q < NI; q++) acc[q] = 0;
			// window column of displacement ii: tile column (b & 7) * 8 + MARGIN - S + ii; first row: block row * 8 + 1
			const uint32_t off = (uint32_t)((b >> 3) * 8 + 1) * DN_CP + (b & 7) * 8 + DN_MARGIN - S + ii;
			const uint32_t sh8 = (off & 3u) * 8;
#if DENSE_FMA
			// byte alignment on the FMA pipe: bytes [k, k + 4) of w0:w1 = hi32(w0 * 2^(32 - 8k)) + lo32(w1 * 2^(32 - 8k)); k = 0 is
			// the same formula one word earlier with the multiplier 2^0 (A/B only, see DENSE_FMA)
			const uint32_t m = sh8 ? 1u << (32u - sh8) : 1u;
			const uint32_t *w = reinterpret_cast<const uint32_t *>(s_cur + (off & ~3u)) - (sh8 ? 0 : 1);
#else
			const uint32_t *w = reinterpret_cast<const uint32_t *>(s_cur + (off & ~3u));
#endif
#pragma unroll
			for (int r = 0; r < Cfg::ROWS; r++) {
				const uint32_t w0 = w[r * (DN_CP / 4)], w1 = w[r * (DN_CP / 4) + 1], w2 = w[r * (DN_CP / 4) + 2];
#if DENSE_FMA
				uint32_t c0, c1, l0, l1;
				asm("mul.lo.u32 %0, %1, %2;" : "=r"(l0) : "r"(w1), "r"(m));
				asm("mul.lo.u32 %0, %1, %2;" : "=r"(l1) : "r"(w2), "r"(m));
				asm("mad.hi.u32 %0, %1, %2, %3;" : "=r"(c0) : "r"(w0), "r"(m), "r"(l0));
				asm("mad.hi.u32 %0, %1, %2, %3;" : "=r"(c1) : "r"(w1), "r"(m), "r"(l1));
#else
				const uint32_t c0 = __funnelshift_r(w0, w1, sh8), c1 = __funnelshift_r(w1, w2, sh8);
#endif
#pragma unroll
				for (int q = 0; q < NI; q++) {
					const int br = r - q;
					if (br >= 0 && br < 8) {
						acc[q] = sad4(a[2 * br], c0, acc[q]);
						acc[q] = sad4(a[2 * br + 1], c1, acc[q]);
					}
				}
			}
#pragma unroll
			for (int q = 0; q < NI; q++) best = min(best, (acc[q] << 10) | (uint32_t)(q * NI + ii));
		}
		s_key[tid] = best;
	}
	__syncthreads();
	// ---- accept + half-pixel refinement, thread = (block, half)
	if (tid < 2 * DN_BLOCKS) {
		const int b = tid >> 1, half = tid & 1;
		const bool valid = block_valid(b);
		const uint32_t tex = s_tex[b];
		uint32_t best_sad = 0xFFFFFFFFu;
		int dx = 0, dy = 0, subdir = 8, accepted = 0, searched = 0;
		uint32_t acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
		if (valid && tex >= g.feat_thr) {
			searched = 1;
			uint32_t key = 0xFFFFFFFFu;
			for (int k = 0; k < NI; k++) key = min(key, s_key[b * NI + k]);
			best_sad = key >> 10;
			const int cand = key & 1023;
			dy = cand / NI - S;
			dx = cand - (dy + S) * NI - S;
			accepted = best_sad < g.val_thr;
		}
		if (accepted) {
			const uint32_t *pw = reinterpret_cast<const uint32_t *>(s_prev) + (b >> 3) * 8 * (DN_PP / 4) + (b & 7) * 2 + half;
			const uint32_t off = (uint32_t)((b >> 3) * 8 + S + 1 + dy - 1) * DN_CP + (b & 7) * 8 + DN_MARGIN + dx + 4 * half - 1;
			const uint32_t sh = off & 3u;
			const uint32_t *w = reinterpret_cast<const uint32_t *>(s_cur + (off & ~3u));
			uint32_t pQ = 0, pH0 = 0, pH4 = 0, pV = 0, pT0 = 0, pT4 = 0;
#pragma unroll
			for (int rr = 0; rr < 10; rr++) {
				const uint32_t u0 = w[rr * (DN_CP / 4)], u1 = w[rr * (DN_CP / 4) + 1], u2 = w[rr * (DN_CP / 4) + 2];
				const uint32_t X = bytes_at(u0, u1, sh), Y = bytes_at(u1, u2, sh);
				const uint32_t Q = __byte_perm(X, Y, 0x4321), Qp = __byte_perm(X, Y, 0x5432);
				const uint32_t H4 = havg4(Q, X), H0 = havg4(Q, Qp);
				if (rr >= 1) {
					const uint32_t V = havg4(pQ, Q), T0 = havg4(pH0, H0), T4 = havg4(pH4, H4);
					if (rr >= 2) {
						const uint32_t ai = pw[(rr - 2) * (DN_PP / 4)];
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
		if (valid && half == 0) {
			const int bi = DN_BX * blockIdx.x + (b & 7), bj = DN_BY * blockIdx.y + (b >> 3);
			aof_dense_block o;
			o.dx = (int16_t)dx; o.dy = (int16_t)dy; o.subdir = (int16_t)subdir; o.flags = (int16_t)(searched | (accepted << 1));
			o.best_sad = best_sad; o.texture = tex;
			field[(pair * g.nby + bj) * g.nbx + bi] = o;
		}
	}
}

template <int S> static cudaError_t launch_dense(const uint8_t *prev, const uint8_t *cur, const DenseGeom &g, int n,
						 aof_dense_block *field, cudaStream_t st)
{
	dim3 grid((g.nbx + DN_BX - 1) / DN_BX, (g.nby + DN_BY - 1) / DN_BY, n);
	KernelTimer kt(KID_SAD_DENSE, st);
	sad_dense_kernel<S><<<grid, DenseCfg<S>::THREADS, 0, st>>>(prev, cur, g, field);
	count_launch();
	return cudaGetLastError();
}

cudaError_t sad_dense_launch(int search, const uint8_t *prev, const uint8_t *cur, const DenseGeom &g, int n,
			     aof_dense_block *field, cudaStream_t st)
{
	if (n <= 0) return cudaSuccess;
	switch (search) {
	case 4: return launch_dense<4>(prev, cur, g, n, field, st);
	case 6: return launch_dense<6>(prev, cur, g, n, field, st);
	case 8: return launch_dense<8>(prev, cur, g, n, field, st);
	default: return cudaErrorInvalidValue;
	}
}

} // namespace aof
