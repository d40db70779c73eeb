// K1b  sad_flow2: PX4Flow::compute_flow (/root/reference/modules/OpticalFlow/src/px4flow.cpp:349-527 with
// compute_diff :114, compute_sad_8x8 :300, compute_subpixel :155) for a batch of independent frame pairs,
// bit-exact.  Second version of sad_flow_kernel (sad_kernels.cu), restructured after the ncu profiles of the
// CTA-per-pair kernel (profiles/r1_sad_v1_*, r1c_sad2_*): a third of its warp stalls were CTA barriers between
// phases of very different width (25 / 225 / 50 / 1 threads), and neither shifted frame copies nor a persistent
// CTA moved it.  Here ONE WARP owns a pair from the copy to the vote:
//
//   * one warp per CTA, a private 10 KB shared-memory slot (two frames + the aligned reference blocks), about 22
//     warps per SM, each walking over pairs blockIdx.x, blockIdx.x + gridDim.x, ...; lane 0 starts the two 1-D bulk
//     (TMA-engine) copies of the next pair on the warp's own mbarrier; while a warp waits the other 21 compute;
//   * no CTA barrier anywhere: the phases of a pair are loops over its tasks in rounds of 32 lanes, separated by
//     __syncwarp only, so a warp is never idle because another phase is narrow;
//   * search: task = (block, x-displacement) slides down the 8 + 2S rows, byte-aligns each row once (two funnel
//     shifts) and feeds it to every y-displacement it belongs to (VABSDIFF4.U8.ACC, 2S+1 running sums); the
//     per-block first-minimum (sad << 9 | candidate) is one shared-memory atomicMin per task;
//   * half-pixel stage: task = (block, row pair, half): 8 lanes per block, 17 packed truncating averages each
//     (3 ALU instructions per average instead of 5), direction sums added across the 8 lanes by shuffles;
//   * vote as before (integer half-pixel sums, one IEEE divide).
// Needs one pair per slot in shared memory (W * H <= 100 KB); larger frames use sad_flow_kernel.
#include <cstdlib>

#include "aof_common.cuh"
#include "aof_internal.h"

namespace aof {

// per-byte (a + b) >> 1 = (a & b) + (((a ^ b) & 0xFE..FE) >> 1): two LOP3 on the ALU pipe and one IMAD.HI on the
// otherwise idle FMA pipe (x >> 1 == hi32(x * 2^31)).  `half` must hold 0x80000000 in a register the compiler cannot
// see through, or ptxas turns the multiply back into a shift on the ALU pipe.
__device__ __forceinline__ uint32_t havg4_3(uint32_t a, uint32_t b, uint32_t half)
{
	uint32_t r;
	asm("mad.hi.u32 %0, %1, %2, %3;" : "=r"(r) : "r"((a ^ b) & 0xFEFEFEFEu), "r"(half), "r"(a & b));
	return r;
}

template <int S>
__global__ void __launch_bounds__(32)
sad_flow2_kernel(const uint8_t *__restrict__ prev, const uint8_t *__restrict__ cur, SadGeom g, int n,
		 aof_px4_raw *__restrict__ raw, aof_px4_block *__restrict__ blocks, uint32_t half_mul /* 0x80000000, see havg4_3 */)
{
	constexpr int NI = 2 * S + 1;
	constexpr int ROWS = 8 + 2 * S;
	extern __shared__ __align__(128) uint8_t smem[];
	__shared__ __align__(16) uint32_t s_ablk[AOF_PX4_MAX_BLOCKS][16]; // aligned 8x8 blocks of prev
	__shared__ uint32_t s_best[AOF_PX4_MAX_BLOCKS];                   // min over candidates of sad << 9 | candidate
	__shared__ uint32_t s_tex[AOF_PX4_MAX_BLOCKS];
	__shared__ uint32_t s_xy[AOF_PX4_MAX_BLOCKS];                     // block origin x | y << 16
	__shared__ int s_vote[AOF_PX4_MAX_BLOCKS][2];                     // 2*dx + half, 2*dy + half (or INT_MIN)
	__shared__ __align__(8) uint64_t s_bar;

	const int lane = threadIdx.x;
	const int W = g.width, wpitch = W >> 2;
	const uint32_t fb = g.stage_bytes;
	const uint8_t *sc = smem;                                         // the current frame (+16 B of slack)
	const int nblk = g.nb * g.nb;

	if (lane == 0) mbar_init(&s_bar, 1);
	if (lane < AOF_PX4_MAX_BLOCKS)
		s_xy[lane] = (uint32_t)(g.lo + (lane % g.nb) * g.step) | ((uint32_t)(g.lo + (lane / g.nb) * g.step) << 16);
	__syncwarp();

	uint32_t parity = 0;
	for (int pair = blockIdx.x; pair < n; pair += gridDim.x, parity ^= 1u) {
		// every lane finished reading the previous pair's frames before the __syncwarp that ended its phase C
		if (lane == 0) {
			mbar_expect_tx(&s_bar, fb);
			bulk_g2s(smem, cur + (size_t)pair * g.frame_stride, fb, &s_bar);
		}
		if (lane < AOF_PX4_MAX_BLOCKS) s_best[lane] = 0xFFFFFFFFu;

		// ---- phase A1: the reference blocks straight from global memory (only 25 x 8 x 8 of the previous frame's
		// bytes are ever used), byte-aligned into shared memory; task = (block, row).  The loads of all rounds are
		// issued before the first use and overlap the bulk copy of the current frame.
		{
			const uint8_t *pf = prev + (size_t)pair * g.frame_stride;
			constexpr int A1_ROUNDS = (AOF_PX4_MAX_BLOCKS * 8 + 31) / 32;
			uint32_t w0[A1_ROUNDS], w1[A1_ROUNDS], w2[A1_ROUNDS], offs[A1_ROUNDS];
#pragma unroll
			for (int k = 0; k < A1_ROUNDS; k++) {
				const int task = min(k * 32 + lane, nblk * 8 - 1);
				const uint32_t xy = s_xy[task >> 3];
				offs[k] = ((xy >> 16) + (task & 7)) * W + (xy & 0xFFFF);
				const uint32_t *w = reinterpret_cast<const uint32_t *>(pf + (offs[k] & ~3u));
				w0[k] = __ldg(w); w1[k] = __ldg(w + 1); w2[k] = __ldg(w + 2);
			}
#pragma unroll
			for (int k = 0; k < A1_ROUNDS; k++) {
				const int task = k * 32 + lane;
				if (task < nblk * 8) {
					const uint32_t sh = offs[k] & 3u;
					s_ablk[task >> 3][2 * (task & 7)] = bytes_at(w0[k], w1[k], sh);
					s_ablk[task >> 3][2 * (task & 7) + 1] = bytes_at(w1[k], w2[k], sh);
				}
			}
		}
		__syncwarp();
		// ---- phase A2: texture gate (compute_diff) on the central 4x4 ----
		if (lane < nblk) {
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
		__syncwarp();

		mbar_wait(&s_bar, parity);

		// ---- phase B: integer search, task = (block, x-displacement); whole rounds of 32 tasks ----
		const int n_tasks = nblk * NI, n_full = n_tasks & ~31;
#pragma unroll 1
		for (int task = lane; task < n_full; task += 32) {
			const int b = task / NI, ii = task - b * NI;
			if (s_tex[b] < g.feat_thr) continue;
			const uint32_t xy = s_xy[b];
			uint32_t a[16];
#pragma unroll
			for (int k = 0; k < 4; k++) {
				const uint4 v = reinterpret_cast<const uint4 *>(&s_ablk[b][0])[k];
				a[4 * k] = v.x; a[4 * k + 1] = v.y; a[4 * k + 2] = v.z; a[4 * k + 3] = v.w;
			}
			uint32_t acc[NI];
#pragma unroll
			for (int q = 0; q < NI; q++) acc[q] = 0;
			const uint32_t off = ((xy >> 16) - S) * W + ((xy & 0xFFFF) - S + ii);
			const uint32_t sh8 = (off & 3u) * 8;
			const uint32_t *w = reinterpret_cast<const uint32_t *>(sc + (off & ~3u));
#pragma unroll
			for (int r = 0; r < ROWS; r++) {
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
			for (int q = 0; q < NI; q++) best = min(best, acc[q] * 512u + (uint32_t)(q * NI + ii));
			atomicMin(&s_best[b], best);
		}
		// the last n_tasks % 32 tasks would leave most lanes idle for a whole round: split them by y-displacement
		// too, (task, dy) per lane, 16 SADs each
#pragma unroll 1
		for (int sub = lane; sub < (n_tasks - n_full) * NI; sub += 32) {
			const int task = n_full + sub / NI, q = sub - (sub / NI) * NI;
			const int b = task / NI, ii = task - b * NI;
			if (s_tex[b] < g.feat_thr) continue;
			const uint32_t xy = s_xy[b];
			const uint32_t off = ((xy >> 16) - S + q) * W + ((xy & 0xFFFF) - S + ii);
			const uint32_t sh8 = (off & 3u) * 8;
			const uint32_t *w = reinterpret_cast<const uint32_t *>(sc + (off & ~3u));
			uint32_t acc = 0;
#pragma unroll
			for (int br = 0; br < 8; br++) {
				const uint32_t w0 = w[br * wpitch], w1 = w[br * wpitch + 1], w2 = w[br * wpitch + 2];
				acc = sad4(s_ablk[b][2 * br], __funnelshift_r(w0, w1, sh8), acc);
				acc = sad4(s_ablk[b][2 * br + 1], __funnelshift_r(w1, w2, sh8), acc);
			}
			atomicMin(&s_best[b], acc * 512u + (uint32_t)(q * NI + ii));
		}
		__syncwarp();

		// ---- phase C: accept + half-pixel refinement, task = (block, 4-pixel half) walking down the 10 rows ----
#pragma unroll 1
		for (int base = 0; base < nblk * 2; base += 32) { // whole rounds: the shuffle needs every lane
			const int task = base + lane;
			const bool live = task < nblk * 2;
			const int b = live ? task >> 1 : 0, half = task & 1;
			const uint32_t xy = s_xy[b];
			const int x = xy & 0xFFFF, y = xy >> 16;
			const uint32_t tex = s_tex[b];
			uint32_t best_sad = 0xFFFFFFFFu;
			int dx = 0, dy = 0, subdir = 8, accepted = 0;
			uint32_t acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
			if (live && tex >= g.feat_thr) {
				const uint32_t key = s_best[b];
				best_sad = key >> 9;
				const int cand = key & 511;
				dy = cand / NI - S;
				dx = cand - (dy + S) * NI - S;
				accepted = best_sad < g.val_thr;
			}
			if (accepted) {
				// bytes q[-1..6] of rows -1..8 of this half of the matched block
				const uint32_t off = (uint32_t)(y + dy - 1) * W + (x + dx + 4 * half - 1);
				const uint32_t sh = off & 3u;
				const uint32_t *w = reinterpret_cast<const uint32_t *>(sc + (off & ~3u));
				uint32_t pQ = 0, pH0 = 0, pH4 = 0;     // previous row: centre, havg(c, c+1x), havg(c, c-1x)
				uint32_t pV = 0, pT0 = 0, pT4 = 0;     // previous row pair: vertical / diagonal averages
#pragma unroll
				for (int rr = 0; rr < 10; rr++) {      // window row rr - 1
					const uint32_t u0 = w[rr * wpitch], u1 = w[rr * wpitch + 1], u2 = w[rr * wpitch + 2];
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
		if (blocks && lane >= nblk && lane < AOF_PX4_MAX_BLOCKS) { // grid positions that do not exist
			aof_px4_block o = {};
			o.best_sad = 0xFFFFFFFFu; o.subdir = 8; o.x = -1; o.y = -1;
			blocks[(size_t)pair * AOF_PX4_MAX_BLOCKS + lane] = o;
		}
		__syncwarp();

		// ---- phase D: mean vote (px4flow.cpp:450-524) ----
		{
			const bool on = lane < nblk && s_vote[lane][0] != INT_MIN;
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
					o.quality = (count * 255 / (AOF_PX4_MAX_BLOCKS)) & 0xFF;
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
static size_t sad2_dyn_smem(const SadGeom &g) { return (size_t)g.stage_bytes + 16; }

bool sad_flow2_supported(const SadGeom &g)
{
	if (g.tun_v1 == 1) return false;
	return sad2_dyn_smem(g) <= 200 * 1024;
}

template <int S> static cudaError_t launch_sad2(const uint8_t *prev, const uint8_t *cur, const SadGeom &g, int n,
						aof_px4_raw *raw, aof_px4_block *blocks, cudaStream_t st)
{
	static int sms[64] = {}, occ_dev[64] = {};
	const size_t dyn = sad2_dyn_smem(g);
	int dev = 0;
	cudaGetDevice(&dev);
	if (dev >= 64) return cudaErrorInvalidDevice;
	cudaError_t e;
	if (!sms[dev]) {
		if ((e = cudaFuncSetAttribute(sad_flow2_kernel<S>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024)) != cudaSuccess) return e;
		if ((e = cudaFuncSetAttribute(sad_flow2_kernel<S>, cudaFuncAttributePreferredSharedMemoryCarveout, 100)) != cudaSuccess) return e;
		if ((e = cudaDeviceGetAttribute(&sms[dev], cudaDevAttrMultiProcessorCount, dev)) != cudaSuccess) return e;
	}
	// the occupancy query costs microseconds: once per device and frame size, not once per 30 us launch
	static size_t occ_dyn[64] = {};
	if (!occ_dev[dev] || occ_dyn[dev] != dyn) {
		int q = 0;
		if ((e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&q, sad_flow2_kernel<S>, 32, dyn)) != cudaSuccess) return e;
		occ_dev[dev] = q < 1 ? 1 : q;
		occ_dyn[dev] = dyn;
	}
	int occ = occ_dev[dev];
	if (g.tun_ctas > 0 && g.tun_ctas < occ) occ = g.tun_ctas;
	long long grid = (long long)sms[dev] * occ;
	if (grid > n) grid = n;
	sad_flow2_kernel<S><<<(unsigned)grid, 32, dyn, st>>>(prev, cur, g, n, raw, blocks, 0x80000000u);
	return cudaGetLastError();
}

cudaError_t sad_flow2_launch(int search, const uint8_t *prev, const uint8_t *cur, const SadGeom &g, int n,
			     aof_px4_raw *raw, aof_px4_block *blocks, cudaStream_t st)
{
	switch (search) {
	case 4: return launch_sad2<4>(prev, cur, g, n, raw, blocks, st);
	case 6: return launch_sad2<6>(prev, cur, g, n, raw, blocks, st);
	case 8: return launch_sad2<8>(prev, cur, g, n, raw, blocks, st);
	default: return cudaErrorInvalidValue;
	}
}

} // namespace aof
