// K3s  lk_pyr_stream: levels 1, 2 and 3 of cv::buildOpticalFlowPyramid (three cv::pyrDown, called
// inside calcOpticalFlowPyrLK at external/klt_feature_tracker/src/trackFeatures.cpp:98) in ONE
// streaming pass over level 0.  This is the HBM-bound kernel of the LK path: 307 200 B read +
// 100 800 B written per 640 x 480 frame.
//
// Design (no shared memory, no block barrier; every warp is autonomous):
//   * a warp owns a column strip of 256 level-0 pixels (8 per lane, one LDG.64 per row) and walks
//     DOWN the image.  A strip produces at most 28 level-3 columns; the 14-pixel dependency cone
//     on each side is recomputed by the neighbouring strip (640 px = 3 strips);
//   * pyrDown is separable and linear, so the vertical [1 4 6 4 1] pass runs FIRST on the raw
//     bytes, split once into even / odd columns as u16x2 lanes (no alignment work: a row below is
//     the same lane).  The horizontal pass then needs only three funnel shifts per output pair,
//     with the halo words coming from the neighbouring lanes by SHFL.  (sum + 128) >> 8 and the
//     re-packing into bytes are one IADD3 operand and one PRMT;
//   * the level-1 row a warp has just produced (4 px per lane) is re-packed to 8 px per lane in
//     the lower half-warp and pushed through the same code for level 2, and again for level 3:
//     each level keeps a four-row ring of its input in registers.  cv::pyrDown's
//     BORDER_REFLECT_101 is applied on EACH level's own size (top: the ring is mirrored when row
//     0 is emitted; bottom: mirrored when the last row arrives; left / right: the border lane
//     substitutes its own columns for the neighbour's);
//   * for small batches the image is also cut into bands of level-3 rows, each band re-running the
//     14 (level-0) rows of dependency cone above and below it.
// Needs W % 32 == 0 and H % 8 == 0 (every level then has even size and lane-aligned borders);
// other shapes use lk_pyr_fused_kernel / lk_pyr_down_kernel.
#include <cstdlib>

#include "aof_common.cuh"
#include "aof_internal.h"

namespace aof {

constexpr int PS_MAX_STRIPS = 8;
constexpr int PS_WARPS = 1; // one warp per CTA: blockIdx-derived row bounds are provably warp-uniform, so SHFL needs no WARPSYNC

struct PsStrips {
	int n;
	int s0[PS_MAX_STRIPS];   // level-0 column of lane 0 (multiple of 32)
	int oa[PS_MAX_STRIPS];   // owned level-3 columns [oa, ob), multiples of 4
	int ob[PS_MAX_STRIPS];
};

// one image row of a lane: 8 columns as u16x2 lanes, even columns (c0,c2) (c4,c6), odd (c1,c3) (c5,c7)
struct Row4 {
	uint32_t e01, e23, o01, o23;
};

__device__ __forceinline__ Row4 ps_split(uint32_t w0, uint32_t w1)
{
	Row4 r;
	r.e01 = __byte_perm(w0, 0u, 0x4240);
	r.o01 = __byte_perm(w0, 0u, 0x4341);
	r.e23 = __byte_perm(w1, 0u, 0x4240);
	r.o23 = __byte_perm(w1, 0u, 0x4341);
	return r;
}

__device__ __forceinline__ uint32_t ps_v1(uint32_t a, uint32_t b, uint32_t c, uint32_t d, uint32_t e)
{
	return a + e + 4u * (b + d) + 6u * c; // <= 16 * 255 per u16 lane
}

__device__ __forceinline__ Row4 ps_vert(const Row4 &a, const Row4 &b, const Row4 &c, const Row4 &d, const Row4 &e)
{
	Row4 v;
	v.e01 = ps_v1(a.e01, b.e01, c.e01, d.e01, e.e01);
	v.e23 = ps_v1(a.e23, b.e23, c.e23, d.e23, e.e23);
	v.o01 = ps_v1(a.o01, b.o01, c.o01, d.o01, e.o01);
	v.o23 = ps_v1(a.o23, b.o23, c.o23, d.o23, e.o23);
	return v;
}

// horizontal [1 4 6 4 1] at the even columns of the lane's 8 vertical sums, + 128, >> 8:
// 4 output pixels packed in one word.  left / right: this lane sits on the image border.
__device__ __forceinline__ uint32_t ps_horz(const Row4 &v, bool left, bool right)
{
	uint32_t le = __shfl_up_sync(0xFFFFFFFFu, v.e23, 1);    // (V-4, V-2)
	uint32_t lo = __shfl_up_sync(0xFFFFFFFFu, v.o23, 1);    // (V-3, V-1)
	uint32_t re = __shfl_down_sync(0xFFFFFFFFu, v.e01, 1);  // (V8, V10)
	if (left) { le = v.e01; lo = v.o01 << 16; }             // V-2 = V2, V-1 = V1
	if (right) re = v.e23 >> 16;                            // V8 = V6
	const uint32_t A = __byte_perm(le, v.e01, 0x5432);      // (V-2, V0)
	const uint32_t B = __byte_perm(lo, v.o01, 0x5432);      // (V-1, V1)
	const uint32_t E = __byte_perm(v.e01, v.e23, 0x5432);   // (V2, V4)
	const uint32_t Bp = __byte_perm(v.o01, v.o23, 0x5432);  // (V3, V5)
	const uint32_t Ep = __byte_perm(v.e23, re, 0x5432);     // (V6, V8)
	const uint32_t r01 = A + E + 0x00800080u + 4u * (B + v.o01) + 6u * v.e01;   // <= 65 408 per lane
	const uint32_t r23 = E + Ep + 0x00800080u + 4u * (Bp + v.o23) + 6u * v.e23;
	return __byte_perm(r01, r23, 0x7531);
}

// 4-px words of all lanes -> 8 px per lane in the lower half of the lanes
__device__ __forceinline__ Row4 ps_repack(uint32_t w, int lane)
{
	const uint32_t u0 = __shfl_sync(0xFFFFFFFFu, w, (2 * lane) & 31);
	const uint32_t u1 = __shfl_sync(0xFFFFFFFFu, w, (2 * lane + 1) & 31);
	return ps_split(u0, u1);
}

__device__ __forceinline__ int ps_reflect(int p, int n)
{
	p = p < 0 ? -p : p;
	p = p >= n ? 2 * n - 2 - p : p;
	return min(max(p, 0), n - 1);
}

// the vertical tap set of one emission: (a, b, c, d, e) normally; mirrored at the top
// (rows -2, -1 -> 2, 1) and at the bottom of an even-sized level (row H -> H - 2)
#define PS_PICK(ra, rb, rc, rd, re, top, bot)                       \
	Row4 pa = ra, pb = rb, pc = rc, pd = rd, pe = re;           \
	if (top) { pa = re; pb = rd; }                               \
	if (bot) { pa = rb; pb = rc; pc = rd; pd = re; pe = rd; }

__global__ void __launch_bounds__(PS_WARPS * 32)
lk_pyr_stream_kernel(LkGeom g, const uint8_t *__restrict__ l0, uint8_t *__restrict__ l1, uint8_t *__restrict__ l2,
		     uint8_t *__restrict__ l3, PsStrips ss, int bands, int band_rows, int n_tasks)
{
	pdl_enter();
	const int lane = threadIdx.x & 31;
	const int task = blockIdx.x * PS_WARPS + (threadIdx.x >> 5);
	if (task >= n_tasks) return;
	const int strip = task % ss.n;
	const int band = (task / ss.n) % bands;
	const size_t frame = (size_t)(task / (ss.n * bands));
	const int W0 = g.w[0], H0 = g.h[0], W1 = g.w[1], H1 = g.h[1], W2 = g.w[2], H2 = g.h[2], W3 = g.w[3], H3 = g.h[3];
	const int R0 = band * band_rows, R1 = min(H3, R0 + band_rows); // owned level-3 rows
	if (R0 >= H3) return;
	const int s0 = ss.s0[strip], oa = ss.oa[strip], ob = ss.ob[strip];
	// level-0 stage: lane owns columns c0 .. c0 + 7, produces level-1 columns x1 .. x1 + 3
	const int c0 = s0 + 8 * lane;
	const bool in0 = c0 < W0, left0 = c0 == 0, right0 = c0 + 8 == W0;
	const int x1 = (s0 >> 1) + 4 * lane;
	const bool own1 = in0 && x1 >= 4 * oa && x1 < 4 * ob;
	// level-1 stage (lanes 0..15): level-1 columns c1 .. c1 + 7 -> level-2 columns x2 .. x2 + 3
	const int c1 = (s0 >> 1) + 8 * lane;
	const bool left1 = lane < 16 && c1 == 0, right1 = lane < 16 && c1 + 8 == W1;
	const int x2 = (s0 >> 2) + 4 * lane;
	const bool own2 = lane < 16 && x2 >= 2 * oa && x2 < 2 * ob;
	// level-2 stage (lanes 0..7)
	const int c2 = (s0 >> 2) + 8 * lane;
	const bool left2 = lane < 8 && c2 == 0, right2 = lane < 8 && c2 + 8 == W2;
	const int x3 = (s0 >> 3) + 4 * lane;
	const bool own3 = lane < 8 && x3 >= oa && x3 < ob;

	const uint8_t *img = l0 + frame * (size_t)W0 * H0 + (in0 ? c0 : 0);
	uint8_t *o1 = l1 + frame * (size_t)W1 * H1 + x1;
	uint8_t *o2 = l2 + frame * (size_t)W2 * H2 + x2;
	uint8_t *o3 = l3 + frame * (size_t)W3 * H3 + x3;

	// level-1 rows this band has to produce (its own rows plus the cone of the levels above)
	const int Y0 = max(0, 4 * R0 - 6), Y1 = min(H1 - 1, 4 * R1 + 2);
	const int Z0 = Y0 == 0 ? 0 : (Y0 + 2) / 2; // first level-2 row the level-1 stage will emit

	auto load_row = [&](int y) -> uint2 {
		uint2 v = make_uint2(0u, 0u);
		if (in0) v = __ldg(reinterpret_cast<const uint2 *>(img + (size_t)ps_reflect(y, H0) * W0));
		return v;
	};
	Row4 a, b, c;
	{
		const uint2 ra = load_row(2 * Y0 - 2), rb = load_row(2 * Y0 - 1), rc = load_row(2 * Y0);
		a = ps_split(ra.x, ra.y); b = ps_split(rb.x, rb.y); c = ps_split(rc.x, rc.y);
	}
	// two iterations of rows in flight
	uint2 qa0 = load_row(2 * Y0 + 1), qb0 = load_row(2 * Y0 + 2);
	uint2 qa1 = load_row(2 * min(Y0 + 1, H1 - 1) + 1), qb1 = load_row(2 * min(Y0 + 1, H1 - 1) + 2);
	Row4 f1a = {}, f1b = {}, f1c = {}, f1d = {}, f1e = {}; // ring of level-1 rows
	Row4 f2a = {}, f2b = {}, f2c = {}, f2d = {}, f2e = {}; // ring of level-2 rows
	int cnt1 = 0, cnt2 = 0;

	for (int Y = Y0; Y <= Y1; ++Y) {
		const uint2 ca = qa0, cb = qb0;
		qa0 = qa1; qb0 = qb1;
		{
			const int yn = min(Y + 2, H1 - 1);
			qa1 = load_row(2 * yn + 1); qb1 = load_row(2 * yn + 2);
		}
		const Row4 d = ps_split(ca.x, ca.y), e = ps_split(cb.x, cb.y);
		const Row4 v0 = ps_vert(a, b, c, d, e);
		const uint32_t w1 = ps_horz(v0, left0, right0);
		if (own1 && Y >= 4 * R0 && Y < 4 * R1) *reinterpret_cast<uint32_t *>(o1 + (size_t)Y * W1) = w1;
		a = c; b = d; c = e;

		// ---- level-1 row Y enters the level-1 stage
		f1a = f1b; f1b = f1c; f1c = f1d; f1d = f1e;
		f1e = ps_repack(w1, lane);
		cnt1++;
		const bool top1 = Y == 2 && Y0 == 0, bot1 = Y == H1 - 1;
		const bool nor1 = (Y & 1) == 0 && Y >= 2 && cnt1 >= 5;
		if (!(top1 || bot1 || nor1)) continue;
		const int Z = bot1 ? H2 - 1 : (Y - 2) >> 1;
		uint32_t w2;
		{
			PS_PICK(f1a, f1b, f1c, f1d, f1e, top1, bot1)
			const Row4 v1 = ps_vert(pa, pb, pc, pd, pe);
			w2 = ps_horz(v1, left1, right1);
		}
		if (own2 && Z >= 2 * R0 && Z < 2 * R1) *reinterpret_cast<uint32_t *>(o2 + (size_t)Z * W2) = w2;

		// ---- level-2 row Z enters the level-2 stage
		f2a = f2b; f2b = f2c; f2c = f2d; f2d = f2e;
		f2e = ps_repack(w2, lane);
		cnt2++;
		const bool top2 = Z == 2 && Z0 == 0, bot2 = Z == H2 - 1;
		const bool nor2 = (Z & 1) == 0 && Z >= 2 && cnt2 >= 5;
		if (!(top2 || bot2 || nor2)) continue;
		const int R = bot2 ? H3 - 1 : (Z - 2) >> 1;
		uint32_t w3;
		{
			PS_PICK(f2a, f2b, f2c, f2d, f2e, top2, bot2)
			const Row4 v2 = ps_vert(pa, pb, pc, pd, pe);
			w3 = ps_horz(v2, left2, right2);
		}
		if (own3 && R >= R0 && R < R1) *reinterpret_cast<uint32_t *>(o3 + (size_t)R * W3) = w3;
	}
}

// ---------------------------------------------------------------------------
// Same algorithm for H % 16 == 0, with the row loop unrolled by 8 level-1 rows (= 4 level-2 rows
// = 2 level-3 rows) starting on a multiple of 8: which rows emit into which level is then static
// and nothing in the loop is conditional except the stores (outputs of a filter that has not yet
// seen its five rows are computed and dropped).  Instead of a ring of rows each level keeps two
// partial sums of the vertical filter:
//     after an even row c:  P = a + 4 b + 6 c   (the output centred on c, still missing 4 d + e),  Q = c
//     after an odd row d:   P += 4 d,  Q = c + 4 d   (the head of the NEXT output)
//     next even row e:      out = P + e;  P = Q + 6 e;  Q = e
// Top border (rows -2, -1 mirror rows 2, 1): the first output is 6 r0 + 8 r1 + 2 r2, i.e. the same
// updates with the multipliers of d and e doubled (P starts from Q = 0).  Bottom border of an
// even-sized level (row H mirrors H - 2 = c): out = P + c = P + Q - 4 d when the last (odd) row arrives.
struct PsAcc {
	Row4 P, Q;
};

// 4 instructions per register and row pair (three IMAD and one IADD)
#define PS_ODD1(P, Q, d, md) { P = (d) * (md) + P; Q = (d) * 4u + Q; }
#define PS_EVEN1(P, Q, e, me, out) { out = (e) * (me) + P; P = (e) * 6u + Q; Q = (e); }

__device__ __forceinline__ void ps_odd(PsAcc &s, const Row4 &d, uint32_t md)
{
	PS_ODD1(s.P.e01, s.Q.e01, d.e01, md) PS_ODD1(s.P.e23, s.Q.e23, d.e23, md)
	PS_ODD1(s.P.o01, s.Q.o01, d.o01, md) PS_ODD1(s.P.o23, s.Q.o23, d.o23, md)
}

__device__ __forceinline__ Row4 ps_even(PsAcc &s, const Row4 &e, uint32_t me)
{
	Row4 v;
	PS_EVEN1(s.P.e01, s.Q.e01, e.e01, me, v.e01) PS_EVEN1(s.P.e23, s.Q.e23, e.e23, me, v.e23)
	PS_EVEN1(s.P.o01, s.Q.o01, e.o01, me, v.o01) PS_EVEN1(s.P.o23, s.Q.o23, e.o23, me, v.o23)
	return v;
}

// bottom border: P + Q - 4 d
__device__ __forceinline__ Row4 ps_bottom(const PsAcc &s, const Row4 &d)
{
	Row4 v;
	v.e01 = s.P.e01 + s.Q.e01 - 4u * d.e01; v.e23 = s.P.e23 + s.Q.e23 - 4u * d.e23;
	v.o01 = s.P.o01 + s.Q.o01 - 4u * d.o01; v.o23 = s.P.o23 + s.Q.o23 - 4u * d.o23;
	return v;
}

// ps_horz with the border substitutions folded into the PRMT selectors (per-lane constants):
//   A  = (V-2, V0): 0x5432 from (le, e01);  left border  (V2, V0) = 0x5476 (e01 only)
//   B  = (V-1, V1): 0x5432 from (lo, o01);  left border  (V1, V1) = 0x5454
//   Ep = (V6, V8):  0x5432 from (e23, re);  right border (V6, V6) = 0x3232
struct PsSel {
	uint32_t a, b, ep;
};

__device__ __forceinline__ PsSel ps_make_sel(bool left, bool right)
{
	PsSel s;
	s.a = left ? 0x5476u : 0x5432u; s.b = left ? 0x5454u : 0x5432u; s.ep = right ? 0x3232u : 0x5432u;
	return s;
}

__device__ __forceinline__ uint32_t ps_horz_sel(const Row4 &v, const PsSel &s)
{
	const uint32_t le = __shfl_up_sync(0xFFFFFFFFu, v.e23, 1);
	const uint32_t lo = __shfl_up_sync(0xFFFFFFFFu, v.o23, 1);
	const uint32_t re = __shfl_down_sync(0xFFFFFFFFu, v.e01, 1);
	const uint32_t A = __byte_perm(le, v.e01, s.a);
	const uint32_t B = __byte_perm(lo, v.o01, s.b);
	const uint32_t E = __byte_perm(v.e01, v.e23, 0x5432);
	const uint32_t Bp = __byte_perm(v.o01, v.o23, 0x5432);
	const uint32_t Ep = __byte_perm(v.e23, re, s.ep);
	const uint32_t r01 = A + E + 0x00800080u + 4u * (B + v.o01) + 6u * v.e01;
	const uint32_t r23 = E + Ep + 0x00800080u + 4u * (Bp + v.o23) + 6u * v.e23;
	return __byte_perm(r01, r23, 0x7531);
}

template <int MIN_CTAS>
__global__ void __launch_bounds__(PS_WARPS * 32, MIN_CTAS)
lk_pyr_stream16_kernel(LkGeom g, const uint8_t *__restrict__ l0, uint8_t *__restrict__ l1, uint8_t *__restrict__ l2,
		       uint8_t *__restrict__ l3, PsStrips ss, int bands, int band_rows, int n_tasks)
{
	pdl_enter();
	const int lane = threadIdx.x & 31;
	const int task = blockIdx.x * PS_WARPS + (threadIdx.x >> 5);
	if (task >= n_tasks) return;
	const int strip = task % ss.n;
	const int band = (task / ss.n) % bands;
	const size_t frame = (size_t)(task / (ss.n * bands));
	const int W0 = g.w[0], H0 = g.h[0], W1 = g.w[1], H1 = g.h[1], W2 = g.w[2], H2 = g.h[2], W3 = g.w[3], H3 = g.h[3];
	const int R0 = band * band_rows, R1 = min(H3, R0 + band_rows);
	if (R0 >= H3) return;
	const int s0 = ss.s0[strip], oa = ss.oa[strip], ob = ss.ob[strip];
	const int c0 = s0 + 8 * lane;
	const bool in0 = c0 < W0, left0 = c0 == 0, right0 = c0 + 8 == W0;
	const int x1 = (s0 >> 1) + 4 * lane;
	const bool own1 = in0 && x1 >= 4 * oa && x1 < 4 * ob;
	const int c1 = (s0 >> 1) + 8 * lane;
	const bool left1 = lane < 16 && c1 == 0, right1 = lane < 16 && c1 + 8 == W1;
	const int x2 = (s0 >> 2) + 4 * lane;
	const bool own2 = lane < 16 && x2 >= 2 * oa && x2 < 2 * ob;
	const int c2 = (s0 >> 2) + 8 * lane;
	const bool left2 = lane < 8 && c2 == 0, right2 = lane < 8 && c2 + 8 == W2;
	const int x3 = (s0 >> 3) + 4 * lane;
	const bool own3 = lane < 8 && x3 >= oa && x3 < ob;
	const PsSel sel0 = ps_make_sel(left0, right0), sel1 = ps_make_sel(left1, right1), sel2 = ps_make_sel(left2, right2);

	// level-1 rows [Ya, Yb): the band's own rows plus the cone, widened to whole groups of 8
	const int Ya = max(0, 4 * R0 - 6) & ~7;
	const int Yb = min(H1, (4 * R1 + 3 + 7) & ~7);

	// lanes right of the image read column 0 (their results are never used): no predicate on the loads
	const uint8_t *img = l0 + frame * (size_t)W0 * H0 + (in0 ? c0 : 0);
	// running store pointers: row Ya of level 1, the level-2 / level-3 rows the first group emits
	uint8_t *o1 = l1 + frame * (size_t)W1 * H1 + x1 + (size_t)Ya * W1;
	uint8_t *o2 = l2 + frame * (size_t)W2 * H2 + x2 + ((ptrdiff_t)(Ya >> 1) - 1) * W2;
	uint8_t *o3 = l3 + frame * (size_t)W3 * H3 + x3 + ((ptrdiff_t)(Ya >> 2) - 1) * W3;

	// byte offset of a row inside the frame.  Odd rows are clamped to the last row (only ever hit by the
	// prefetch running past the image); even rows to row H0 - 2, which is exactly the mirror of row H0.
	const uint32_t off_odd_max = (uint32_t)(H0 - 1) * W0, off_even_max = (uint32_t)(H0 - 2) * W0;
	auto load_odd = [&](uint32_t off) -> uint2 { return __ldg(reinterpret_cast<const uint2 *>(img + min(off, off_odd_max))); };
	auto load_even = [&](uint32_t off) -> uint2 { return __ldg(reinterpret_cast<const uint2 *>(img + min(off, off_even_max))); };
	PsAcc f0, f1 = {}, f2 = {};
	uint32_t ro = (uint32_t)(2 * Ya) * W0; // row 2 Ya
	{
		Row4 a = {}, b = {};
		if (Ya > 0) {
			const uint2 ra = load_even(ro - 2 * W0), rb = load_odd(ro - W0);
			a = ps_split(ra.x, ra.y); b = ps_split(rb.x, rb.y);
		}
		const uint2 rc = load_even(ro);
		const Row4 c = ps_split(rc.x, rc.y);
		f0.P.e01 = a.e01 + 4u * b.e01 + 6u * c.e01; f0.P.e23 = a.e23 + 4u * b.e23 + 6u * c.e23;
		f0.P.o01 = a.o01 + 4u * b.o01 + 6u * c.o01; f0.P.o23 = a.o23 + 4u * b.o23 + 6u * c.o23;
		f0.Q = c;
	}
	// rows 2 Y + 1, 2 Y + 2 of two iterations in flight
	ro += W0;
	uint2 qa0 = load_odd(ro), qb0 = load_even(ro + W0);
	ro += 2 * W0;
	uint2 qa1 = load_odd(ro), qb1 = load_even(ro + W0);
	ro += 2 * W0;
	const int src0 = (2 * lane) & 31, src1 = (2 * lane + 1) & 31;

	for (int Yg = Ya; Yg < Yb; Yg += 8) {
		const bool first = Yg == 0;
		const uint32_t m8 = first ? 8u : 4u, m2 = first ? 2u : 1u;
#pragma unroll
		for (int k = 0; k < 8; k++) {
			const int Y = Yg + k;
			const uint2 ca = qa0, cb = qb0;
			qa0 = qa1; qb0 = qb1;
			qa1 = load_odd(ro); qb1 = load_even(ro + W0);
			ro += 2 * W0;
			// ---- level 0 -> level-1 row Y
			ps_odd(f0, ps_split(ca.x, ca.y), k == 0 ? m8 : 4u);
			const uint32_t w1 = ps_horz_sel(ps_even(f0, ps_split(cb.x, cb.y), k == 0 ? m2 : 1u), sel0);
			if (own1 && Y >= 4 * R0 && Y < 4 * R1) *reinterpret_cast<uint32_t *>(o1) = w1;
			o1 += W1;
			const Row4 n1 = ps_split(__shfl_sync(0xFFFFFFFFu, w1, src0), __shfl_sync(0xFFFFFFFFu, w1, src1));
			if (k & 1) { // static: odd level-1 row
				ps_odd(f1, n1, k == 1 ? m8 : 4u);
				if (k == 7 && Y == H1 - 1) { // bottom of the image: level-2 row H2 - 1, then level-3 row H3 - 1
					const uint32_t wb2 = ps_horz_sel(ps_bottom(f1, n1), sel1);
					if (own2 && H2 - 1 < 2 * R1) *reinterpret_cast<uint32_t *>(o2) = wb2;
					const Row4 nb = ps_split(__shfl_sync(0xFFFFFFFFu, wb2, src0), __shfl_sync(0xFFFFFFFFu, wb2, src1));
					ps_odd(f2, nb, 4u);
					const uint32_t wb3 = ps_horz_sel(ps_bottom(f2, nb), sel2);
					if (own3 && H3 - 1 < R1) *reinterpret_cast<uint32_t *>(o3) = wb3;
				}
				continue;
			}
			// ---- level-1 row Y (even) completes level-2 row Z = (Y - 2) / 2
			const int Z = (Y - 2) >> 1;
			const uint32_t w2 = ps_horz_sel(ps_even(f1, n1, k == 2 ? m2 : 1u), sel1);
			if (own2 && Z >= 2 * R0 && Z < 2 * R1) *reinterpret_cast<uint32_t *>(o2) = w2;
			o2 += W2;
			const Row4 n2 = ps_split(__shfl_sync(0xFFFFFFFFu, w2, src0), __shfl_sync(0xFFFFFFFFu, w2, src1));
			if ((k & 3) != 2) { // static: Z odd.  (Z = -1 of the first group does not exist: the filter must start from 0)
				if (k == 4 || !first) ps_odd(f2, n2, k == 4 ? m8 : 4u);
				continue;
			}
			const int R = (Z - 2) >> 1;
			const uint32_t w3 = ps_horz_sel(ps_even(f2, n2, k == 6 ? m2 : 1u), sel2);
			if (own3 && R >= R0 && R < R1) *reinterpret_cast<uint32_t *>(o3) = w3;
			o3 += W3;
		}
	}
}

// ---- host side -------------------------------------------------------------
// Column strips: the first starts at column 0 and owns level-3 columns [0, 28); every further strip
// owning from `oa` starts 32 level-0 columns (= the 14-px cone rounded up to the lane grid of level 2)
// to its left and owns 24 columns; the strip that contains the right image border owns up to it.
static bool ps_plan(const LkGeom &g, PsStrips &ss)
{
	const int W0 = g.w[0], W3 = g.w[3];
	ss.n = 0;
	int oa = 0;
	while (oa < W3) {
		if (ss.n == PS_MAX_STRIPS) return false;
		const int s0 = oa == 0 ? 0 : 8 * oa - 32;
		const int ob = s0 + 256 >= W0 ? W3 : (oa == 0 ? 28 : oa + 24);
		ss.s0[ss.n] = s0; ss.oa[ss.n] = oa; ss.ob[ss.n] = ob;
		ss.n++;
		oa = ob;
	}
	return true;
}

bool lk_pyr_stream_supported(const LkGeom &g, const uint8_t *l0)
{
	if (g.top != 3) return false;
	if (g.w[0] % 32 || g.h[0] % 8 || g.h[0] < 64 || g.w[0] < 64) return false;
	if (reinterpret_cast<uintptr_t>(l0) & 7) return false;
	PsStrips ss;
	return ps_plan(g, ss);
}

// bands of level-3 rows: enough warps to keep every SM busy when the batch is small, whole
// frames (no vertical cone recomputed) when it is large.  AOF_PYR_BANDS overrides (tuning).
static int ps_bands(const LkGeom &g, int strips, int n)
{
	const int forced = g.tun_pyr_bands;
	const int H3 = g.h[3];
	int bands;
	if (forced > 0) bands = forced;
	else {
		const long long want = 148LL * 24;  // warps that fill the GPU at this kernel's register budget
		const long long have = (long long)strips * n;
		bands = (int)((want + have - 1) / have);
	}
	const int max_bands = H3 / 4 > 0 ? H3 / 4 : 1;  // a band of fewer than 4 level-3 rows is mostly cone
	if (bands > max_bands) bands = max_bands;
	if (bands < 1) bands = 1;
	return bands;
}

cudaError_t lk_pyr_stream_launch(const LkGeom &g, const uint8_t *l0, uint8_t *l1, uint8_t *l2, uint8_t *l3, int n, cudaStream_t st)
{
	if (n <= 0) return cudaSuccess;
	PsStrips ss;
	if (!ps_plan(g, ss)) return cudaErrorInvalidValue;
	const int bands = ps_bands(g, ss.n, n);
	const int band_rows = (g.h[3] + bands - 1) / bands;
	const long long tasks = (long long)ss.n * bands * n;
	if (tasks > 0x7FFFFFFFLL) return cudaErrorInvalidValue;
	const unsigned blocks = (unsigned)((tasks + PS_WARPS - 1) / PS_WARPS);
	const bool rows16 = g.h[0] % 16 == 0 && g.tun_pyr_v1 != 1;
	const int o = g.tun_pyr_occ > 0 ? g.tun_pyr_occ : 16; // measured on B200: 16 CTAs/SM (<= 128 registers) is the fastest
	KernelTimer kt(KID_LK_PYR_FUSED, st, !rows16 ? "lk_pyr_stream_kernel" : o <= 16 ? "lk_pyr_stream16_kernel<16>" : o <= 20 ? "lk_pyr_stream16_kernel<20>" : "lk_pyr_stream16_kernel<24>");
	if (rows16) {
		const bool pdl = g.tun_pdl != 0;
		if (o <= 16) launch_k(pdl, lk_pyr_stream16_kernel<16>, blocks, PS_WARPS * 32, 0, st, g, l0, l1, l2, l3, ss, bands, band_rows, (int)tasks);
		else if (o <= 20) launch_k(pdl, lk_pyr_stream16_kernel<20>, blocks, PS_WARPS * 32, 0, st, g, l0, l1, l2, l3, ss, bands, band_rows, (int)tasks);
		else launch_k(pdl, lk_pyr_stream16_kernel<24>, blocks, PS_WARPS * 32, 0, st, g, l0, l1, l2, l3, ss, bands, band_rows, (int)tasks);
	} else
		launch_k(g.tun_pdl != 0, lk_pyr_stream_kernel, blocks, PS_WARPS * 32, 0, st, g, l0, l1, l2, l3, ss, bands, band_rows, (int)tasks);
	count_launch();
	return cudaGetLastError();
}

} // namespace aof
