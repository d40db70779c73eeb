/*
 * oracle/sad_oracle.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * CPU restatement of the reference's PX4Flow SAD block matcher and of the
 * OpticalFlowPX4 / rate-limiter epilogue around it.  Only tests/, smoke() and
 * bench.py's cpu_baseline / --impl reference legs may load this library; the
 * product path (libaof_b200.so) never links or calls it.
 *
 * Parity status: PINNED by execution -- tests/test_oracle_sad.py compares every
 * function here against the unmodified reference objects compiled into
 * oracle/_ref/libaof_ref.so (see oracle/Makefile) and against the known answers
 * of SURVEY.md Appendix B.2 (committed as tests/golden/sad_golden.json).
 *
 * Written from the algorithm description, byte-wise (the reference works on
 * packed 32-bit words emulating ARM USAD8/USADA8/UHADD8; the arithmetic is the
 * same per byte).  Citations are relative to
 * /root/reference/modules/OpticalFlow/.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define ORC_TILE 8
#define ORC_GRID 5

/* texture measure on the 4x4 patch at (x+2, y+2): 12 vertical-neighbour and 12
 * horizontal-neighbour absolute differences.  src/px4flow.cpp:114-142 */
uint32_t orc_block_texture(const uint8_t *img, int x, int y, int pitch)
{
	const uint8_t *p = img + (size_t)(y + 2) * pitch + (x + 2);
	uint32_t sum = 0;
	for (int r = 0; r < 3; r++)
		for (int c = 0; c < 4; c++)
			sum += (uint32_t)abs((int)p[r * pitch + c] - (int)p[(r + 1) * pitch + c]);
	for (int c = 0; c < 3; c++)
		for (int r = 0; r < 4; r++)
			sum += (uint32_t)abs((int)p[r * pitch + c] - (int)p[r * pitch + c + 1]);
	return sum;
}

/* 8x8 sum of absolute differences.  src/px4flow.cpp:300-333 */
uint32_t orc_sad8x8(const uint8_t *a, const uint8_t *b, int ax, int ay, int bx, int by, int pitch)
{
	uint32_t sum = 0;
	for (int r = 0; r < ORC_TILE; r++) {
		const uint8_t *pa = a + (size_t)(ay + r) * pitch + ax;
		const uint8_t *pb = b + (size_t)(by + r) * pitch + bx;
		for (int c = 0; c < ORC_TILE; c++)
			sum += (uint32_t)abs((int)pa[c] - (int)pb[c]);
	}
	return sum;
}

static inline int havg(int u, int v) { return (u + v) >> 1; } /* UHADD8: truncating, :91-103 */

/* SAD of the img-a block against the eight half-pixel neighbours of the matched
 * block in img-b.  Direction k: 0=+x 1=+x+y 2=+y 3=-x+y 4=-x 5=-x-y 6=-y 7=+x-y.
 * The diagonal samples are averages of two already-truncated averages
 * (two-stage), exactly as src/px4flow.cpp:202-223; accumulate as :237-244. */
void orc_subpixel(const uint8_t *a, const uint8_t *b, int ax, int ay, int bx, int by, int pitch,
		  uint32_t acc[8])
{
	for (int k = 0; k < 8; k++) acc[k] = 0;
	for (int r = 0; r < ORC_TILE; r++) {
		for (int c = 0; c < ORC_TILE; c++) {
			const uint8_t *q = b + (size_t)(by + r) * pitch + (bx + c);
			int ctr = q[0];
			int e = havg(ctr, q[1]);                          /* s0 */
			int se_src = havg(q[pitch], q[pitch + 1]);        /* s1 */
			int s = havg(ctr, q[pitch]);                      /* s2 */
			int sw_src = havg(q[pitch], q[pitch - 1]);        /* s3 */
			int w = havg(ctr, q[-1]);                         /* s4 */
			int nw_src = havg(q[-pitch], q[-pitch - 1]);      /* s5 */
			int n = havg(ctr, q[-pitch]);                     /* s6 */
			int ne_src = havg(q[-pitch], q[-pitch + 1]);      /* s7 */
			int cand[8];
			cand[0] = e;
			cand[1] = havg(e, se_src);
			cand[2] = s;
			cand[3] = havg(sw_src, w);
			cand[4] = w;
			cand[5] = havg(w, nw_src);
			cand[6] = n;
			cand[7] = havg(ne_src, e);
			int ref = a[(size_t)(ay + r) * pitch + (ax + c)];
			for (int k = 0; k < 8; k++) acc[k] += (uint32_t)abs(ref - cand[k]);
		}
	}
}

/* per-block diagnostics; one entry per grid position in raster (row-major j,i) order */
typedef struct {
	int32_t x, y;          /* block origin in img-a */
	uint32_t texture;      /* orc_block_texture */
	int32_t searched;      /* 1 if texture >= feature threshold */
	uint32_t best_sad;     /* min SAD (0xFFFFFFFF when not searched) */
	int32_t dx, dy;        /* arg-min displacement */
	int32_t accepted;      /* best_sad < value threshold */
	uint32_t acc[8];       /* half-pixel SADs (when accepted) */
	int32_t subdir;        /* 0..7 or 8 (none) */
} orc_block_info;

/* number of grid positions per axis (src/px4flow.cpp:359-361,380-381) */
int orc_grid(int width, int search, int *lo, int *step)
{
	int pix_lo = search + 1;
	int pix_hi = width - (search + 1) - ORC_TILE;
	int pix_step = (pix_hi - pix_lo) / ORC_GRID + 1;
	int n = 0;
	for (int p = pix_lo; p < pix_hi; p += pix_step) n++;
	*lo = pix_lo;
	*step = pix_step;
	return n;
}

/* PX4Flow::compute_flow, src/px4flow.cpp:349-527 (mean vote only; the
 * histogram at :429-443 is never read).  info may be NULL or hold 25 entries. */
int orc_compute_flow(const uint8_t *prev, const uint8_t *cur, int width, int search, uint32_t feat_thr,
		     uint32_t val_thr, float *flow_x, float *flow_y, orc_block_info *info)
{
	int lo, step;
	int n = orc_grid(width, search, &lo, &step);
	int dirx[ORC_GRID * ORC_GRID], diry[ORC_GRID * ORC_GRID], sub[ORC_GRID * ORC_GRID];
	int count = 0, slot = 0;

	for (int gj = 0; gj < n; gj++) {
		for (int gi = 0; gi < n; gi++, slot++) {
			int x = lo + gi * step, y = lo + gj * step;
			orc_block_info bi;
			memset(&bi, 0, sizeof bi);
			bi.x = x; bi.y = y; bi.best_sad = 0xFFFFFFFFu; bi.subdir = 8;
			bi.texture = orc_block_texture(prev, x, y, width);
			if (bi.texture >= feat_thr) {
				bi.searched = 1;
				for (int jj = -search; jj <= search; jj++)          /* :394-405 */
					for (int ii = -search; ii <= search; ii++) {
						uint32_t d = orc_sad8x8(prev, cur, x, y, x + ii, y + jj, width);
						if (d < bi.best_sad) { bi.best_sad = d; bi.dx = ii; bi.dy = jj; }
					}
				if (bi.best_sad < val_thr) {                          /* :408-427 */
					bi.accepted = 1;
					orc_subpixel(prev, cur, x, y, x + bi.dx, y + bi.dy, width, bi.acc);
					uint32_t best = bi.best_sad;
					for (int k = 0; k < 8; k++)
						if (bi.acc[k] < best) { best = bi.acc[k]; bi.subdir = k; }
					dirx[count] = bi.dx; diry[count] = bi.dy; sub[count] = bi.subdir;
					count++;
				}
			}
			if (info) info[slot] = bi;
		}
	}

	if (count <= 10) { *flow_x = 0.0f; *flow_y = 0.0f; return 0; }           /* :450,517-521 */

	float sx = 0.0f, sy = 0.0f;                                               /* :479-500 */
	for (int h = 0; h < count; h++) {
		float hx = 0.0f, hy = 0.0f;
		int k = sub[h];
		if (k == 0 || k == 1 || k == 7) hx = 0.5f;
		if (k == 3 || k == 4 || k == 5) hx = -0.5f;
		if (k == 5 || k == 6 || k == 7) hy = -0.5f;
		if (k == 1 || k == 2 || k == 3) hy = 0.5f;
		sx += (float)dirx[h] + hx;
		sy += (float)diry[h] + hy;
	}
	*flow_x = sx / (float)(uint32_t)count;
	*flow_y = sy / (float)(uint32_t)count;
	return (int)(uint8_t)(count * 255 / (ORC_GRID * ORC_GRID));                /* :524 */
}

/* ---- OpticalFlow::limitRate + OpticalFlowPX4::calcFlow, per-stream state ---- */

typedef struct {
	int output_rate;
	float sum_x, sum_y;
	int sum_q, n_valid;
	uint32_t t_last;   /* per stream here; a process-wide static in the reference
			      (src/optical_flow.cpp:55) -- identical for one stream */
} orc_limiter;

void orc_limiter_init(orc_limiter *l, int output_rate)
{
	memset(l, 0, sizeof *l);
	l->output_rate = output_rate;
}

/* src/optical_flow.cpp:51-96 */
int orc_limit_rate(orc_limiter *l, int quality, uint32_t t_us, int *dt_us, float *fx, float *fy)
{
	if (l->output_rate <= 0) {
		*dt_us = (int)(t_us - l->t_last);
		l->t_last = t_us;
		return quality;
	}
	if (quality > 0) {
		l->sum_x += *fx; l->sum_y += *fy; l->sum_q += quality; l->n_valid++;
	}
	if ((float)(uint32_t)(t_us - l->t_last) > 1.0e6f / (float)l->output_rate) {
		int q = 0;
		if (l->n_valid > 0) q = l->sum_q / l->n_valid;   /* int division, then floor() is a no-op */
		*fx = l->sum_x; *fy = l->sum_y;
		l->sum_x = l->sum_y = 0.0f; l->sum_q = 0; l->n_valid = 0;
		*dt_us = (int)(t_us - l->t_last);
		l->t_last = t_us;
		return q;
	}
	return -1;
}

typedef struct {
	int width, height, search;
	uint32_t feat_thr, val_thr;
	float focal_x, focal_y;
	int primed;
	uint8_t *prev;
	orc_limiter lim;
} orc_px4;

orc_px4 *orc_px4_new(float fx, float fy, int rate, int w, int h, int search, int feat_thr, int val_thr)
{
	orc_px4 *s = (orc_px4 *)calloc(1, sizeof *s);
	s->width = w; s->height = h; s->search = search;
	s->feat_thr = (uint32_t)feat_thr; s->val_thr = (uint32_t)val_thr;
	s->focal_x = fx; s->focal_y = fy;
	s->prev = (uint8_t *)malloc((size_t)w * h);
	orc_limiter_init(&s->lim, rate);
	return s;
}

/* setFocalLengthX/Y, setOutputRate (include/optical_flow.hpp:76-78): plain assignments, all state is kept */
void orc_px4_set_params(orc_px4 *s, float fx, float fy, int rate)
{
	s->focal_x = fx; s->focal_y = fy; s->lim.output_rate = rate;
}

void orc_px4_free(orc_px4 *s) { if (s) { free(s->prev); free(s); } }

/* src/flow_px4.cpp:69-98.  Also returns the raw pixel flow / raw quality of this
 * frame pair through px_x/px_y/raw_q (may be NULL) for kernel-level parity. */
int orc_px4_calc_flow(orc_px4 *s, const uint8_t *img, uint32_t t_us, int *dt_us, float *flow_x,
		      float *flow_y, float *px_x, float *px_y, int *raw_q)
{
	size_t n = (size_t)s->width * s->height;
	if (!s->primed) {
		memcpy(s->prev, img, n);
		s->primed = 1;
		return 0;
	}
	float fx, fy;
	int q = orc_compute_flow(s->prev, img, s->width, s->search, s->feat_thr, s->val_thr, &fx, &fy, NULL);
	if (px_x) *px_x = fx;
	if (px_y) *px_y = fy;
	if (raw_q) *raw_q = q;
	memcpy(s->prev, img, n);
	q = orc_limit_rate(&s->lim, q, t_us, dt_us, &fx, &fy);
	*flow_x = (float)atan2((double)fx, (double)s->focal_x);   /* double atan2, :92-93 */
	*flow_y = (float)atan2((double)fy, (double)s->focal_y);
	return q;
}

/* bulk helper for tests / cpu_baseline: n independent pairs, stateless */
void orc_compute_flow_batch(const uint8_t *prev, const uint8_t *cur, int n, int width, int height, int search,
			    uint32_t feat_thr, uint32_t val_thr, float *flow_x, float *flow_y, int *quality)
{
	size_t fs = (size_t)width * height;
	for (int i = 0; i < n; i++)
		quality[i] = orc_compute_flow(prev + i * fs, cur + i * fs, width, search, feat_thr, val_thr,
					      &flow_x[i], &flow_y[i], NULL);
}

/* ---- dense field (BASELINE config 4): the same three primitives on every 8x8 block of the 8-px lattice
 * whose search window + 1 px lies inside the image; 32-bit offsets.  out[nby * nbx] records of 16 bytes:
 * int16 dx, dy, subdir, flags (1 searched, 2 accepted), uint32 best_sad, uint32 texture. */
typedef struct { int16_t dx, dy, subdir, flags; uint32_t best_sad, texture; } orc_dense_block;

void orc_dense_grid(int w, int h, int search, int *i_lo, int *j_lo, int *nbx, int *nby)
{
	*i_lo = (search + 1 + 7) / 8; *j_lo = *i_lo;
	*nbx = (w - search - 9) / 8 - *i_lo + 1;
	*nby = (h - search - 9) / 8 - *j_lo + 1;
}

void orc_dense_field(const uint8_t *prev, const uint8_t *cur, int w, int h, int search, uint32_t feat_thr, uint32_t val_thr,
		     orc_dense_block *out)
{
	int i_lo, j_lo, nbx, nby;
	orc_dense_grid(w, h, search, &i_lo, &j_lo, &nbx, &nby);
	for (int bj = 0; bj < nby; bj++)
		for (int bi = 0; bi < nbx; bi++) {
			const int x = 8 * (i_lo + bi), y = 8 * (j_lo + bj);
			orc_dense_block o = {0, 0, 8, 0, 0xFFFFFFFFu, 0};
			o.texture = orc_block_texture(prev, x, y, w);
			if (o.texture >= feat_thr) {
				o.flags = 1;
				for (int jj = -search; jj <= search; jj++)
					for (int ii = -search; ii <= search; ii++) {
						const uint32_t d = orc_sad8x8(prev, cur, x, y, x + ii, y + jj, w);
						if (d < o.best_sad) { o.best_sad = d; o.dx = (int16_t)ii; o.dy = (int16_t)jj; }
					}
				if (o.best_sad < val_thr) {
					uint32_t acc[8], best = o.best_sad;
					o.flags |= 2;
					orc_subpixel(prev, cur, x, y, x + o.dx, y + o.dy, w, acc);
					for (int k = 0; k < 8; k++)
						if (acc[k] < best) { best = acc[k]; o.subdir = (int16_t)k; }
				}
			}
			out[bj * nbx + bi] = o;
		}
}
