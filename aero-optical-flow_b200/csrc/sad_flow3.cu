// K1c  sad_flow3: PX4Flow::compute_flow (/root/reference/modules/OpticalFlow/src/px4flow.cpp:349-527 with
// compute_diff :114, compute_sad_8x8 :300, compute_subpixel :155) for a batch of independent 64x64 frame pairs at
// search_size 4 (BASELINE configs 1, 3, 5), bit-exact.  Third version, written against the ncu profile of
// sad_flow2_kernel (profiles/r1i_sad_ncu_full_summary.txt): there only 31 % of the warp instructions were
// VABSDIFF4, the ALU pipe carried the byte alignment, the key arithmetic and the address arithmetic as well, 38 %
// of the shared-memory wavefronts were bank conflicts and every round ended in a shared-memory atomicMin (32-64
// LSU cycles).  Everything that was a run-time quantity there is a compile-time constant here (frame 64x64,
// grid {5,15,25,35,45}^2, 9 x 9 displacements), and the warp-per-pair structure is kept:
//
//   * search task = (block, PAIR of x-displacements): the three words of a window row are loaded once for two
//     displacements (LSU wavefronts per pair: 346 -> 216); the even displacement is byte-aligned by funnel shifts
//     on the ALU pipe, the odd one by IMAD.HI + IMAD on the FMA pipe (hi32(w0 * m) + w1 * m, m = 2^(24 - 8s)), so
//     the ALU pipe carries VABSDIFF4 and half of the alignment only;
//   * task -> lane is static, 4 lanes per block: 3 rounds of pair tasks (blocks 0..23), one round of single
//     displacements (dx = +4 of every block, and block 24), one leftover displacement split by dy over 9 lanes;
//     all per-task addresses, shifts and multipliers are per-lane constants of the launch;
//   * the first-minimum key (sad << 9 | candidate) is reduced over the 4 lanes of a block by two shuffles and
//     stored: no shared-memory atomics;
//   * the frame sits in shared memory with a 64-byte gap after every 10th row: the blocks of one round come from
//     two grid rows 10 image rows apart, which would otherwise hit the same banks (2-way conflicts on most loads);
//     row offsets stay immediates because every window starts at y = 1 (mod 10); the frame arrives as 7 bulk
//     (TMA-engine) copies on the warp's mbarrier.
// Other geometries keep using sad_flow2_kernel / sad_flow_kernel.
#include <cstdlib>

#include "aof_common.cuh"
#include "aof_internal.h"

namespace aof {

namespace {
constexpr int S = 4, NI = 9, W = 64, H = 64, NB = 5, LO = 5, STEP = 10, NBLK = 25;
constexpr int ROWS = 8 + 2 * S;                 // 16 window rows per search column
constexpr int GAP = 64;                         // bytes inserted after rows 9, 19, ...
constexpr int SLOT_BYTES = H * W + (H / 10) * GAP + 64;
// words per aligned reference block in shared memory: 16 used, pitch 20 (80 B, still 16-byte aligned) so that the
// 128-bit loads of the 8 blocks of a round fall into 8 different groups of 4 banks (pitch 16: two groups, 4-way conflicts)
constexpr int APITCH = 20;

__host__ __device__ constexpr uint32_t row_off(int y) { return (uint32_t)(y * W + (y / 10) * GAP); }
// window row r of any block: y = 1 + 10 by + r, so the gap count is by + (r >= 9): static in r up to a per-lane base
__host__ __device__ constexpr uint32_t wrow(int r) { return (uint32_t)(r * W + (r >= 9 ? GAP : 0)); }

// u / 5 and u / 10 for u < 64 as one multiply and one shift.  The per-lane set-up below runs ONCE per warp, and at 4096 pairs
// per launch a warp owns one pair: with plain / and % the compiler spent ~600 of the 3 300 warp instructions per pair on it
// (16-bit division sequences for seven rounds of block indices, profiles/r2R_*).
__device__ __forceinline__ uint32_t div5(uint32_t u) { return (u * 205u) >> 10; }
__device__ __forceinline__ uint32_t div10(uint32_t u) { return (u * 205u) >> 11; }
// row_off(y) for a run-time y < 64
__device__ __forceinline__ uint32_t row_off_rt(uint32_t y) { return (y + div10(y)) * (uint32_t)W; }

__device__ __forceinline__ uint32_t havg4_3(uint32_t a, uint32_t b, uint32_t half)
{
	uint32_t r;
	asm("mad.hi.u32 %0, %1, %2, %3;" : "=r"(r) : "r"((a ^ b) & 0xFEFEFEFEu), "r"(half), "r"(a & b));
	return r;
}

// bytes [s+1, s+5) of the 8-byte string w0:w1 on the FMA pipe; m = 2^(24 - 8 s), s = 0..3
__device__ __forceinline__ uint32_t align_fma(uint32_t w0, uint32_t w1, uint32_t m)
{
	uint32_t lo, r;
	asm("mul.lo.u32 %0, %1, %2;" : "=r"(lo) : "r"(w1), "r"(m));
	asm("mad.hi.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(w0), "r"(m), "r"(lo));
	return r;
}

__device__ __forceinline__ uint32_t lds32(uint32_t addr)
{
	uint32_t v;
	asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(addr));
	return v;
}

__device__ __forceinline__ uint32_t min3(uint32_t a, uint32_t b, uint32_t c) { return min(min(a, b), c); }

// min over the 9 y-displacements of sad << 9 | (q * 9): the x part of the candidate index is added by the caller
__device__ __forceinline__ uint32_t key9(const uint32_t (&acc)[NI])
{
	uint32_t k[NI];
#pragma unroll
	for (int q = 0; q < NI; q++) k[q] = acc[q] * 512u + (uint32_t)(q * NI);
	return min3(min3(k[0], k[1], k[2]), min3(k[3], k[4], k[5]), min3(k[6], k[7], k[8]));
}
} // namespace

// measured on B200 (gpurun r2e, 4096 / 32768 pairs per launch): rounds unrolled + FMA-pipe alignment 27.1 / 175.0 us,
// one copy of the round 27.2 / 168.6, funnel shifts only 27.1 / 164.7, both 26.8 / 162.5 -- the kernel is bound by issue
// slots and instruction fetch (ncu: no_instruction 14 % of the stall samples with the rounds unrolled), not by the ALU
// pipe alone, so the two extra FMA-pipe instructions per aligned word cost more than the ALU slot they free
#ifndef AOF_SAD3_UNROLLK
#define AOF_SAD3_UNROLLK 0
#endif
#ifndef AOF_SAD3_FMA
#define AOF_SAD3_FMA 0
#endif
#ifndef AOF_SAD3_MINB
#define AOF_SAD3_MINB 28 // resident warps per SM the register allocation is capped for
#endif
// warps per CTA: every warp is still autonomous (own frame slot, own mbarrier, no block barrier anywhere); the CTA only
// groups them for the block scheduler.  Measured (r2a): 1, 2 or 4 warps per CTA give the same 27.0 / 160.2 us at 4096 / 32768 pairs.
#ifndef AOF_SAD3_WPC
#define AOF_SAD3_WPC 1
#endif
struct Sad3Slot {
	alignas(128) uint8_t frame[SLOT_BYTES];
	alignas(16) uint32_t ablk[NBLK][APITCH];            // aligned 8x8 blocks of prev
	uint32_t key[NBLK];                                 // min over candidates of sad << 9 | candidate
	uint32_t tex[NBLK];
	int vote[NBLK][2];                                  // 2*dx + half, 2*dy + half (or INT_MIN)
	uint16_t rowoff[H];                                 // row_off(y), for the data-dependent rows of the half-pixel stage
	alignas(8) uint64_t bar;
};
__global__ void __launch_bounds__(32 * AOF_SAD3_WPC, AOF_SAD3_MINB / AOF_SAD3_WPC)
sad_flow3_kernel(const uint8_t *__restrict__ prev, const uint8_t *__restrict__ cur, size_t frame_stride, int n,
		 uint32_t feat_thr, uint32_t val_thr, aof_px4_raw *__restrict__ raw, aof_px4_block *__restrict__ blocks,
		 uint32_t half_mul /* 0x80000000, see havg4_3 */)
{
	__shared__ Sad3Slot s_slots[AOF_SAD3_WPC];
	Sad3Slot &SL = s_slots[threadIdx.x >> 5];
	uint8_t *const s_frame = SL.frame;
	uint32_t (*const s_ablk)[APITCH] = SL.ablk;
	uint32_t *const s_key = SL.key, *const s_tex = SL.tex;
	int (*const s_vote)[2] = SL.vote;
	uint16_t *const s_rowoff = SL.rowoff;
	uint64_t &s_bar = SL.bar;

	const uint32_t lane = threadIdx.x & 31u;
	const uint32_t sc = smem_u32(s_frame);
	const uint32_t sa = smem_u32(&s_ablk[0][0]);

	if (lane == 0) mbar_init(&s_bar, 1);
	s_rowoff[lane] = (uint16_t)row_off_rt(lane);
	s_rowoff[lane + 32] = (uint16_t)row_off_rt(lane + 32);
	__syncwarp();

	// ---- per-lane task constants (the same for every pair) ----
	// pair rounds k = 0..2: task t = lane + 32 k, block t >> 2, x-displacements 2p and 2p + 1 (p = t & 3)
	uint32_t pw[3], psh[3], pm[3], pa[3];
#pragma unroll
	for (int k = 0; k < 3; k++) {
		const uint32_t b = (lane >> 2) + 8u * k, p = lane & 3u;      // task t = lane + 32 k
		const uint32_t by = div5(b), bx = b - NB * by;
		const uint32_t o = 1u + STEP * bx + 2u * p;
		pw[k] = sc + (uint32_t)W + by * (uint32_t)(STEP * W + GAP) + (o & ~3u);   // row_off(1 + 10 by) = 64 + 704 by
		psh[k] = (o & 3u) * 8;
		pm[k] = 1u << (24 - 8 * (o & 3u));
		pa[k] = sa + b * (APITCH * 4);
	}
	// single round: lanes 0..24 = (block lane, dx index 8), lanes 25..31 = (block 24, dx index lane - 25)
	const uint32_t sb = lane < NBLK ? lane : NBLK - 1, sdx = lane < NBLK ? 8u : lane - NBLK;
	const uint32_t sby = div5(sb), sbx = sb - NB * sby;
	const uint32_t so = 1u + STEP * sbx + sdx;
	const uint32_t sw = sc + (uint32_t)W + sby * (uint32_t)(STEP * W + GAP) + (so & ~3u), ssh = (so & 3u) * 8, saddr = sa + sb * (APITCH * 4);
	// leftover: (block 24, dx index 7) split by dy over lanes 0..8
	const uint32_t lo_o = (uint32_t)(1 + STEP * 4 + 7);
	const uint32_t lw = sc + row_off_rt(1u + STEP * 4 + min(lane, 8u)) + (lo_o & ~3u), lsh = (lo_o & 3u) * 8;

	uint32_t parity = 0;
	for (int pair = blockIdx.x * AOF_SAD3_WPC + (threadIdx.x >> 5); pair < n; pair += gridDim.x * AOF_SAD3_WPC, parity ^= 1u) {
		// every lane finished reading the previous pair's frame before the __syncwarp that ended its phase C
		if (lane == 0) mbar_expect_tx(&s_bar, H * W);
		__syncwarp();
		if (lane < 7) {
			const int y0 = lane * 10, rows = lane < 6 ? 10 : H - 60;
			bulk_g2s(s_frame + row_off(y0), cur + (size_t)pair * frame_stride + y0 * W, rows * W, &s_bar);
		}

		// ---- phase A1: the 25 reference blocks straight from global memory, byte-aligned into shared memory;
		// task = (block, row).  All loads are issued before the first use and overlap the bulk copies.
		{
			const uint8_t *pf = prev + (size_t)pair * frame_stride;
			constexpr int A1_ROUNDS = (NBLK * 8 + 31) / 32;
			uint32_t w0[A1_ROUNDS], w1[A1_ROUNDS], w2[A1_ROUNDS];
#pragma unroll
			for (int k = 0; k < A1_ROUNDS; k++) {
				const uint32_t task = min(k * 32u + lane, (uint32_t)(NBLK * 8 - 1)), b = task >> 3;
				const uint32_t by = div5(b), bx = b - NB * by;
				const uint32_t off = (LO + STEP * by + (task & 7u)) * W + LO + STEP * bx;
				const uint32_t *w = reinterpret_cast<const uint32_t *>(pf + (off & ~3u));
				w0[k] = __ldg(w); w1[k] = __ldg(w + 1); w2[k] = __ldg(w + 2);
			}
#pragma unroll
			for (int k = 0; k < A1_ROUNDS; k++) {
				const uint32_t task = k * 32u + lane, b = task >> 3;
				if (task < NBLK * 8) {
					const uint32_t bx = b - NB * div5(b);
					const uint32_t sh = 1u + 2u * (bx & 1u);                   // (5 + 10 bx) & 3: 1 or 3
					s_ablk[b][2 * (task & 7)] = bytes_at(w0[k], w1[k], sh);
					s_ablk[b][2 * (task & 7) + 1] = bytes_at(w1[k], w2[k], sh);
				}
			}
		}
		__syncwarp();
		// ---- phase A2: texture gate (compute_diff) on the central 4x4 ----
		if (lane < NBLK) {
			uint32_t p[4];
#pragma unroll
			for (int r = 0; r < 4; r++) p[r] = __byte_perm(s_ablk[lane][2 * (r + 2)], s_ablk[lane][2 * (r + 2) + 1], 0x5432);
			uint32_t t = sad4(p[0], p[1], 0);
			t = sad4(p[1], p[2], t);
			t = sad4(p[2], p[3], t);
#pragma unroll
			for (int r = 0; r < 4; r++) t = sad4(p[r], __byte_perm(p[r], p[r], 0x3321), t); // |c0-c1|+|c1-c2|+|c2-c3|
			s_tex[lane] = t;
		}

		mbar_wait(&s_bar, parity);

		// ---- phase B: integer search ----
		// (the texture gate is applied in phase C only: skipping here would need every lane of a round to agree)
#if AOF_SAD3_UNROLLK
#pragma unroll
#else
#pragma unroll 1
#endif
		for (int k = 0; k < 3; k++) {
#if AOF_SAD3_UNROLLK
			const uint32_t wb = pw[k], sh8 = psh[k], m = pm[k], ab = pa[k];
#else
			// one copy of the round's code (instruction-cache footprint): the constants are selected, not indexed
			const uint32_t wb = k == 0 ? pw[0] : k == 1 ? pw[1] : pw[2];
			const uint32_t sh8 = k == 0 ? psh[0] : k == 1 ? psh[1] : psh[2];
			const uint32_t m = k == 0 ? pm[0] : k == 1 ? pm[1] : pm[2];
			const uint32_t ab = k == 0 ? pa[0] : k == 1 ? pa[1] : pa[2];
#endif
			uint32_t a[16];
#pragma unroll
			for (int j = 0; j < 4; j++)
				asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];"
					     : "=r"(a[4 * j]), "=r"(a[4 * j + 1]), "=r"(a[4 * j + 2]), "=r"(a[4 * j + 3])
					     : "r"(ab + 16 * j));
			uint32_t accE[NI], accO[NI];
#pragma unroll
			for (int q = 0; q < NI; q++) accE[q] = accO[q] = 0;
#pragma unroll
			for (int r = 0; r < ROWS; r++) {
				const uint32_t w0 = lds32(wb + wrow(r)), w1 = lds32(wb + wrow(r) + 4), w2 = lds32(wb + wrow(r) + 8);
				const uint32_t e0 = __funnelshift_r(w0, w1, sh8), e1 = __funnelshift_r(w1, w2, sh8);
#if AOF_SAD3_FMA
				const uint32_t o0 = align_fma(w0, w1, m), o1 = align_fma(w1, w2, m);
#else
				const uint32_t o0 = __funnelshift_rc(w0, w1, sh8 + 8), o1 = __funnelshift_rc(w1, w2, sh8 + 8);
				(void)m;
#endif
#pragma unroll
				for (int q = 0; q < NI; q++) {
					const int br = r - q; // row of the reference block this window row meets at dy = q - S
					if (br >= 0 && br < 8) {
						accE[q] = sad4(a[2 * br], e0, accE[q]);
						accE[q] = sad4(a[2 * br + 1], e1, accE[q]);
						accO[q] = sad4(a[2 * br], o0, accO[q]);
						accO[q] = sad4(a[2 * br + 1], o1, accO[q]);
					}
				}
			}
			// candidate index = q * 9 + dx index; dx index = 2p (even) / 2p + 1 (odd)
			uint32_t key = min(key9(accE), key9(accO) + 1u) + 2u * (uint32_t)(lane & 3);
			key = min(key, __shfl_xor_sync(0xFFFFFFFFu, key, 1));
			key = min(key, __shfl_xor_sync(0xFFFFFFFFu, key, 2));
			if ((lane & 3) == 0) s_key[(lane >> 2) + 8 * k] = key;
		}
		__syncwarp();
		{ // single round + leftover
			uint32_t a[16];
#pragma unroll
			for (int j = 0; j < 4; j++)
				asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];"
					     : "=r"(a[4 * j]), "=r"(a[4 * j + 1]), "=r"(a[4 * j + 2]), "=r"(a[4 * j + 3])
					     : "r"(saddr + 16 * j));
			uint32_t acc[NI];
#pragma unroll
			for (int q = 0; q < NI; q++) acc[q] = 0;
#pragma unroll
			for (int r = 0; r < ROWS; r++) {
				const uint32_t w0 = lds32(sw + wrow(r)), w1 = lds32(sw + wrow(r) + 4), w2 = lds32(sw + wrow(r) + 8);
				const uint32_t c0 = __funnelshift_r(w0, w1, ssh), c1 = __funnelshift_r(w1, w2, ssh);
#pragma unroll
				for (int q = 0; q < NI; q++) {
					const int br = r - q;
					if (br >= 0 && br < 8) {
						acc[q] = sad4(a[2 * br], c0, acc[q]);
						acc[q] = sad4(a[2 * br + 1], c1, acc[q]);
					}
				}
			}
			uint32_t key = key9(acc) + (uint32_t)sdx;
			if (lane < NBLK - 1) key = min(key, s_key[lane]);
			// leftover displacement (block 24, dx index 7): lane q < 9 takes dy index q, 8 rows of 2 words.
			// Its rows start at y = 41 + q, so the gap after row 49 is crossed at block row 9 - q.
			uint32_t lkey = 0xFFFFFFFFu;
			if (lane < NI) {
				uint32_t acc1 = 0;
				const uint32_t ab = sa + (NBLK - 1) * (APITCH * 4);
#pragma unroll
				for (int br = 0; br < 8; br++) {
					const uint32_t ad = lw + br * W + (br + lane >= 9 ? GAP : 0);
					const uint32_t w0 = lds32(ad), w1 = lds32(ad + 4), w2 = lds32(ad + 8);
					acc1 = sad4(lds32(ab + 8 * br), __funnelshift_r(w0, w1, lsh), acc1);
					acc1 = sad4(lds32(ab + 8 * br + 4), __funnelshift_r(w1, w2, lsh), acc1);
				}
				lkey = acc1 * 512u + (uint32_t)(lane * NI + 7);
			}
			lkey = __reduce_min_sync(0xFFFFFFFFu, lkey);
			// lanes 24..31 all belong to block 24
			uint32_t g = lane >= NBLK - 1 ? key : 0xFFFFFFFFu;
			g = min(g, __shfl_xor_sync(0xFFFFFFFFu, g, 1));
			g = min(g, __shfl_xor_sync(0xFFFFFFFFu, g, 2));
			g = min(g, __shfl_xor_sync(0xFFFFFFFFu, g, 4));
			if (lane < NBLK - 1) s_key[lane] = key;
			if (lane == NBLK - 1) s_key[NBLK - 1] = min(g, lkey);
		}
		__syncwarp();

		// ---- phase C: accept + half-pixel refinement, task = (block, 4-pixel half) walking down the 10 rows ----
#pragma unroll 1
		for (int base = 0; base < NBLK * 2; base += 32) { // whole rounds: the shuffle needs every lane
			const int task = base + lane;
			const bool live = task < NBLK * 2;
			const int b = live ? task >> 1 : 0, half = task & 1;
			const int by_ = (int)div5((uint32_t)b);
			const int x = LO + STEP * (b - NB * by_), y = LO + STEP * by_;
			const uint32_t tex = s_tex[b];
			uint32_t best_sad = 0xFFFFFFFFu;
			int dx = 0, dy = 0, subdir = 8, accepted = 0;
			uint32_t acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
			if (live && tex >= feat_thr) {
				const uint32_t key = s_key[b];
				best_sad = key >> 9;
				const int cand = key & 511;
				dy = cand / NI - S;
				dx = cand - (dy + S) * NI - S;
				accepted = best_sad < val_thr;
			}
			if (accepted) {
				// bytes q[-1..6] of rows -1..8 of this half of the matched block
				const uint32_t xo = (uint32_t)(x + dx + 4 * half - 1);
				const uint32_t sh = xo & 3u;
				const uint32_t wcol = sc + (xo & ~3u);
				const uint16_t *ro = &s_rowoff[y + dy - 1];
				uint32_t pQ = 0, pH0 = 0, pH4 = 0;     // previous row: centre, havg(c, c+1x), havg(c, c-1x)
				uint32_t pV = 0, pT0 = 0, pT4 = 0;     // previous row pair: vertical / diagonal averages
#pragma unroll
				for (int rr = 0; rr < 10; rr++) {      // window row rr - 1
					const uint32_t ad = wcol + ro[rr];
					const uint32_t u0 = lds32(ad), u1 = lds32(ad + 4), u2 = lds32(ad + 8);
					const uint32_t X = bytes_at(u0, u1, sh);       // q[-1..2]
					const uint32_t Y = bytes_at(u1, u2, sh);       // q[3..6]
					const uint32_t Q = __byte_perm(X, Y, 0x4321);  // q[0..3]
					const uint32_t Qp = __byte_perm(X, Y, 0x5432); // q[1..4]
					const uint32_t H4 = havg4_3(Q, X, half_mul);
					const uint32_t H0 = havg4_3(Q, Qp, half_mul);
					if (rr >= 1) {
						const uint32_t V = havg4_3(pQ, Q, half_mul), T0 = havg4_3(pH0, H0, half_mul), T4 = havg4_3(pH4, H4, half_mul);
						if (rr >= 2) { // block row i = rr - 2 is the previous window row; pV.. belong to rows (i - 1, i)
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
			// both halves of a block sit in adjacent lanes
#pragma unroll
			for (int k = 0; k < 8; k++) acc[k] += __shfl_xor_sync(0xFFFFFFFFu, acc[k], 1);
			if (live && half == 0) {
				if (accepted) {
					uint32_t m = best_sad;
#pragma unroll
					for (int k = 0; k < 8; k++)
						if (acc[k] < m) { m = acc[k]; subdir = k; }
				}
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
					blocks[(size_t)pair * AOF_PX4_MAX_BLOCKS + b] = o;
				}
			}
		}
		__syncwarp();

		// ---- phase D: mean vote (px4flow.cpp:450-524) ----
		{
			const bool on = lane < NBLK && s_vote[lane][0] != INT_MIN;
			const int vx = on ? s_vote[lane][0] : 0, vy = on ? s_vote[lane][1] : 0;
			const int count = __popc(__ballot_sync(0xFFFFFFFFu, on));
			const int sx = __reduce_add_sync(0xFFFFFFFFu, vx);
			const int sy = __reduce_add_sync(0xFFFFFFFFu, vy);
			if (lane == 0) {
				aof_px4_raw o;
				o.count = count;
				if (count > 10) {
					// sums of multiples of 0.5 are exact in f32; one correctly rounded divide
					o.px_x = __fdiv_rn((float)sx * 0.5f, (float)count);
					o.px_y = __fdiv_rn((float)sy * 0.5f, (float)count);
					o.quality = (count * 255 / NBLK) & 0xFF;
				} else {
					o.px_x = 0.f; o.px_y = 0.f; o.quality = 0;
				}
				raw[pair] = o;
			}
		}
		__syncwarp();
	}
}

// ---- host side -------------------------------------------------------------
bool sad_flow3_supported(int search, const SadGeom &g)
{
	return g.tun_v2 != 1 && g.tun_v1 != 1 && search == S && g.width == W && g.stage_bytes == (uint32_t)(W * H) && g.nb == NB && g.lo == LO &&
	       g.step == STEP && AOF_PX4_MAX_BLOCKS == NBLK;
}

cudaError_t sad_flow3_launch(const uint8_t *prev, const uint8_t *cur, const SadGeom &g, int n, aof_px4_raw *raw,
			     aof_px4_block *blocks, cudaStream_t st)
{
	static int grid_max[64] = {}, sms_dev[64] = {};
	int dev = 0;
	cudaGetDevice(&dev);
	if (dev >= 64) return cudaErrorInvalidDevice;
	cudaError_t e;
	if (!grid_max[dev]) {
		int sms = 0, occ = 0;
		if ((e = cudaFuncSetAttribute(sad_flow3_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, 100)) != cudaSuccess) return e;
		if ((e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev)) != cudaSuccess) return e;
		if ((e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, sad_flow3_kernel, 32 * AOF_SAD3_WPC, 0)) != cudaSuccess) return e;
		grid_max[dev] = sms * (occ < 1 ? 1 : occ);
		sms_dev[dev] = sms;
	}
	int gmax = grid_max[dev];
	if (g.tun_ctas > 0 && g.tun_ctas * sms_dev[dev] < gmax) gmax = g.tun_ctas * sms_dev[dev];
	const int ctas = (n + AOF_SAD3_WPC - 1) / AOF_SAD3_WPC;
	const int grid = ctas < gmax ? ctas : gmax;
	sad_flow3_kernel<<<grid, 32 * AOF_SAD3_WPC, 0, st>>>(prev, cur, g.frame_stride, n, g.feat_thr, g.val_thr, raw, blocks, 0x80000000u);
	return cudaGetLastError();
}

} // namespace aof
