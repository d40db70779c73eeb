/*
 * oracle/lk_oracle.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * CPU restatement of the OpenCV-backed path of aero-optical-flow:
 *   OpticalFlowOpenCV::calcFlow   modules/OpticalFlow/src/flow_opencv.cpp:89-210
 *   trackFeatures/initMorePoints  modules/OpticalFlow/external/klt_feature_tracker/src/trackFeatures.cpp:70-363
 *   OpticalFlow::limitRate        modules/OpticalFlow/src/optical_flow.cpp:51-96
 * and of the third-party arithmetic those call (OpenCV, un-vendored and
 * unpinned in the reference; oracle of record = cv2 4.13.0):
 *   cv::calcOpticalFlowPyrLK(prev,cur,pts,...,Size(9,9),3)   trackFeatures.cpp:98
 *   cv::FAST(roi, kps, 10)                                    trackFeatures.cpp:196
 *   cv::goodFeaturesToTrack (named by north_star, absent from the reference)
 *
 * PARITY: the reference has no tests or golden vectors for this path and its
 * flow_opencv.cpp cannot be compiled here (needs the OpenCV <=3 C++ SDK), so
 * parity is "unpinned" by the reference itself.  This restatement is pinned
 * instead against cv2 4.13.0 (tests/test_oracle_lk.py: pyrDown, Scharr/LK,
 * FAST scores+NMS, min-eig/GFTT) and by committed fixtures in tests/golden/.
 *
 * Documented divergences from the reference:
 *   - tracker state is per object, not file-static (trackFeatures.cpp:53-56);
 *   - ties in the response sort (unstable std::sort, trackFeatures.cpp:198,268)
 *     are broken by raster index ascending;
 *   - LK sums are accumulated exactly in integers and rounded once to f32
 *     (OpenCV adds f32 terms in a SIMD-lane dependent order).
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may load this file's library.
 */
#define _GNU_SOURCE
#include <float.h>
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

typedef uint8_t u8;

static inline int reflect101(int p, int n)
{
	if (n == 1) return 0;
	while (p < 0 || p >= n) p = p < 0 ? -p : 2 * n - 2 - p;
	return p;
}

/* ---------------------------------------------------------------- pyramid (App. A.1) */

/* number of the top level: OpenCV stops when the NEXT level would be <= the window */
int olk_pyr_top_level(int w, int h, int win, int max_level)
{
	int level = 0;
	for (; level < max_level; level++) {
		int nw = (w + 1) / 2, nh = (h + 1) / 2;
		if (nw <= win || nh <= win) break;
		w = nw; h = nh;
	}
	return level;
}

/* cv::pyrDown, u8, BORDER_REFLECT_101: [1 4 6 4 1] x [1 4 6 4 1], (sum + 128) >> 8 */
void olk_pyr_down(const u8 *src, int w, int h, u8 *dst)
{
	const int dw = (w + 1) / 2, dh = (h + 1) / 2;
	int *row = (int *)malloc(sizeof(int) * (size_t)dw * 5);
	int have[5] = {-1, -1, -1, -1, -1}; /* source row held by each of the 5 ring slots */
	for (int y = 0; y < dh; y++) {
		const int *r[5];
		for (int j = 0; j < 5; j++) {
			const int sy = reflect101(2 * y + j - 2, h), slot = sy % 5;
			int *o = row + (size_t)slot * dw;
			if (have[slot] != sy) {
				const u8 *s = src + (size_t)sy * w;
				for (int x = 0; x < dw; x++) {
					const int c = 2 * x;
					if (c >= 2 && c + 2 < w) o[x] = s[c - 2] + 4 * s[c - 1] + 6 * s[c] + 4 * s[c + 1] + s[c + 2];
					else o[x] = s[reflect101(c - 2, w)] + 4 * s[reflect101(c - 1, w)] + 6 * s[c] + 4 * s[reflect101(c + 1, w)] +
						    s[reflect101(c + 2, w)];
				}
				have[slot] = sy;
			}
			r[j] = o;
		}
		u8 *d = dst + (size_t)y * dw;
		for (int x = 0; x < dw; x++) d[x] = (u8)((r[0][x] + 4 * r[1][x] + 6 * r[2][x] + 4 * r[3][x] + r[4][x] + 128) >> 8);
	}
	free(row);
}

/* ---------------------------------------------------------------- Scharr (App. A.2) */

/* pixel with the reflect-101 pyramid border */
static inline int pix(const u8 *img, int w, int h, int x, int y)
{
	return img[(size_t)reflect101(y, h) * w + reflect101(x, w)];
}

/* interleaved (Ix, Iy) of the whole level, as calcScharrDeriv; for KATs */
void olk_scharr(const u8 *img, int w, int h, int16_t *dst)
{
	for (int y = 0; y < h; y++)
		for (int x = 0; x < w; x++) {
			int t0[3], t1[3];
			for (int i = 0; i < 3; i++) {
				const int a = pix(img, w, h, x + i - 1, y - 1), b = pix(img, w, h, x + i - 1, y),
					  c = pix(img, w, h, x + i - 1, y + 1);
				t0[i] = 3 * (a + c) + 10 * b;
				t1[i] = c - a;
			}
			dst[((size_t)y * w + x) * 2] = (int16_t)(t0[2] - t0[0]);
			dst[((size_t)y * w + x) * 2 + 1] = (int16_t)(3 * (t1[0] + t1[2]) + 10 * t1[1]);
		}
}

/* derivative tap: zero outside the image (BORDER_CONSTANT), reflect-101 inside */
static inline void deriv_at(const u8 *img, int w, int h, int x, int y, int *ix, int *iy)
{
	if (x < 0 || x >= w || y < 0 || y >= h) { *ix = 0; *iy = 0; return; }
	int t0[3], t1[3];
	for (int i = 0; i < 3; i++) {
		const int a = pix(img, w, h, x + i - 1, y - 1), b = pix(img, w, h, x + i - 1, y), c = pix(img, w, h, x + i - 1, y + 1);
		t0[i] = 3 * (a + c) + 10 * b;
		t1[i] = c - a;
	}
	*ix = t0[2] - t0[0];
	*iy = 3 * (t1[0] + t1[2]) + 10 * t1[1];
}

/* ---------------------------------------------------------------- LK (App. A.3) */

#define OLK_MAXWIN 21
static inline int descale(int x, int n) { return (x + (1 << (n - 1))) >> n; }

static inline void q14_weights(float a, float b, int *w00, int *w01, int *w10, int *w11)
{
	*w00 = (int)lrintf((1.f - a) * (1.f - b) * 16384.f); /* cvRound: round half to even */
	*w01 = (int)lrintf(a * (1.f - b) * 16384.f);
	*w10 = (int)lrintf((1.f - a) * b * 16384.f);
	*w11 = 16384 - *w00 - *w01 - *w10;
}

/*
 * One pyramidal LK call, flags = 0, err not computed.
 * pyr_prev/pyr_next[l] : level l images (no border; the border is implied), ws/hs sizes.
 * pts_prev[2n] in, pts_next[2n] out, status[n] out.
 */
void olk_track(const u8 *const *pyr_prev, const u8 *const *pyr_next, const int *ws, const int *hs, int top_level,
	       const float *pts_prev, float *pts_next, u8 *status, int n, int win, int max_iter, double eps,
	       float min_eig_thr)
{
	const float half = (win - 1) * 0.5f;
	const double eps2 = eps * eps;
	static __thread int Iw[OLK_MAXWIN * OLK_MAXWIN], Ixw[OLK_MAXWIN * OLK_MAXWIN], Iyw[OLK_MAXWIN * OLK_MAXWIN];
	for (int i = 0; i < n; i++) status[i] = 1;
	for (int level = top_level; level >= 0; level--) {
		const u8 *I = pyr_prev[level], *J = pyr_next[level];
		const int w = ws[level], h = hs[level];
		const float scale = (float)(1. / (1 << level));
		for (int p = 0; p < n; p++) {
			float px = pts_prev[2 * p] * scale, py = pts_prev[2 * p + 1] * scale;
			float nx, ny;
			if (level == top_level) { nx = px; ny = py; }
			else { nx = pts_next[2 * p] * 2.f; ny = pts_next[2 * p + 1] * 2.f; }
			pts_next[2 * p] = nx; pts_next[2 * p + 1] = ny;
			px -= half; py -= half;
			int ipx = (int)floorf(px), ipy = (int)floorf(py);
			if (ipx < -win || ipx >= w || ipy < -win || ipy >= h) {
				if (level == 0) status[p] = 0;
				continue;
			}
			int w00, w01, w10, w11;
			q14_weights(px - ipx, py - ipy, &w00, &w01, &w10, &w11);
			int64_t s11 = 0, s12 = 0, s22 = 0;
			for (int y = 0; y < win; y++)
				for (int x = 0; x < win; x++) {
					const int X = ipx + x, Y = ipy + y;
					int dx[4], dy[4];
					deriv_at(I, w, h, X, Y, &dx[0], &dy[0]);
					deriv_at(I, w, h, X + 1, Y, &dx[1], &dy[1]);
					deriv_at(I, w, h, X, Y + 1, &dx[2], &dy[2]);
					deriv_at(I, w, h, X + 1, Y + 1, &dx[3], &dy[3]);
					const int iv = descale(pix(I, w, h, X, Y) * w00 + pix(I, w, h, X + 1, Y) * w01 +
								       pix(I, w, h, X, Y + 1) * w10 + pix(I, w, h, X + 1, Y + 1) * w11, 9);
					const int ixv = (int16_t)descale(dx[0] * w00 + dx[1] * w01 + dx[2] * w10 + dx[3] * w11, 14);
					const int iyv = (int16_t)descale(dy[0] * w00 + dy[1] * w01 + dy[2] * w10 + dy[3] * w11, 14);
					Iw[y * win + x] = (int16_t)iv; Ixw[y * win + x] = ixv; Iyw[y * win + x] = iyv;
					s11 += (int64_t)ixv * ixv; s12 += (int64_t)ixv * iyv; s22 += (int64_t)iyv * iyv;
				}
			const float FLT_SCALE = 1.f / (1 << 20);
			const float A11 = (float)s11 * FLT_SCALE, A12 = (float)s12 * FLT_SCALE, A22 = (float)s22 * FLT_SCALE;
			float D = A11 * A22 - A12 * A12;
			const float min_eig = (A22 + A11 - sqrtf((A11 - A22) * (A11 - A22) + 4.f * A12 * A12)) / (float)(2 * win * win);
			if (min_eig < min_eig_thr || D < FLT_EPSILON) {
				if (level == 0) status[p] = 0;
				continue;
			}
			D = 1.f / D;
			nx -= half; ny -= half;
			float pdx = 0.f, pdy = 0.f;
			for (int j = 0; j < max_iter; j++) {
				const int inx = (int)floorf(nx), iny = (int)floorf(ny);
				if (inx < -win || inx >= w || iny < -win || iny >= h) {
					if (level == 0) status[p] = 0;
					break;
				}
				q14_weights(nx - inx, ny - iny, &w00, &w01, &w10, &w11);
				int64_t sb1 = 0, sb2 = 0;
				for (int y = 0; y < win; y++)
					for (int x = 0; x < win; x++) {
						const int X = inx + x, Y = iny + y;
						const int diff = descale(pix(J, w, h, X, Y) * w00 + pix(J, w, h, X + 1, Y) * w01 +
									 pix(J, w, h, X, Y + 1) * w10 + pix(J, w, h, X + 1, Y + 1) * w11, 9) -
								 Iw[y * win + x];
						sb1 += (int64_t)diff * Ixw[y * win + x];
						sb2 += (int64_t)diff * Iyw[y * win + x];
					}
				const float b1 = (float)sb1 * FLT_SCALE, b2 = (float)sb2 * FLT_SCALE;
				const float ddx = (float)((A12 * b2 - A22 * b1) * D), ddy = (float)((A12 * b1 - A11 * b2) * D);
				nx += ddx; ny += ddy;
				pts_next[2 * p] = nx + half; pts_next[2 * p + 1] = ny + half;
				if ((double)ddx * ddx + (double)ddy * ddy <= eps2) break;
				if (j > 0 && (double)fabsf(ddx + pdx) < 0.01 && (double)fabsf(ddy + pdy) < 0.01) {
					pts_next[2 * p] -= ddx * 0.5f; pts_next[2 * p + 1] -= ddy * 0.5f;
					break;
				}
				pdx = ddx; pdy = ddy;
			}
			/* OpenCV re-tests the window origin of the FINAL position at level 0 when the caller asks for the
			 * error vector (the reference does: trackFeatures.cpp:94,98; cv2 always does) and clears the status
			 * if the last update pushed it out of range; the error itself is never read by the reference. */
			if (level == 0 && status[p]) {
				const int fx_ = (int)floorf(pts_next[2 * p] - half), fy_ = (int)floorf(pts_next[2 * p + 1] - half);
				if (fx_ < -win || fx_ >= w || fy_ < -win || fy_ >= h) status[p] = 0;
			}
		}
	}
}

/* ---------------------------------------------------------------- FAST-9/16 (App. A.4) */

static const int ring_dx[16] = {0, 1, 2, 3, 3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1};
static const int ring_dy[16] = {3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1, 0, 1, 2, 3};

/* largest m such that 9 contiguous ring pixels are all brighter than c + (m-1) or all darker than c - (m-1) */
static int fast_m(const u8 *p, const int *off, int thr)
{
	int d[25];
	const int c = p[0];
	{ /* a 9-arc always contains two of the four compass points (ring 0, 4, 8, 12) */
		const int a = c - p[off[0]], b = c - p[off[4]], e = c - p[off[8]], f = c - p[off[12]];
		if ((a > thr) + (b > thr) + (e > thr) + (f > thr) < 2 && (a < -thr) + (b < -thr) + (e < -thr) + (f < -thr) < 2) return thr;
	}
	for (int k = 0; k < 16; k++) d[k] = c - p[off[k]];
	/* cheap exact reject (what makes cv::FAST fast): a corner needs 9 contiguous ring pixels beyond +-thr */
	unsigned mp = 0, mn = 0;
	for (int k = 0; k < 16; k++) { mp |= (unsigned)(d[k] > thr) << k; mn |= (unsigned)(d[k] < -thr) << k; }
	int corner = 0;
	for (int pass = 0; pass < 2 && !corner; pass++) {
		unsigned m = pass ? mn : mp;
		m |= m << 16;
		const unsigned a = m & (m >> 1), b = a & (a >> 2), e = b & (b >> 4);
		corner = (e & (m >> 8) & 0xFFFFu) != 0;
	}
	if (!corner) return thr; /* not a corner: any value <= thr */
	for (int k = 16; k < 25; k++) d[k] = d[k - 16];
	int m = -1000;
	for (int k = 0; k < 16; k++) {
		int mn9 = d[k], mx9 = d[k];
		for (int i = 1; i < 9; i++) {
			if (d[k + i] < mn9) mn9 = d[k + i];
			if (d[k + i] > mx9) mx9 = d[k + i];
		}
		if (mn9 > m) m = mn9;
		if (-mx9 > m) m = -mx9;
	}
	return m;
}

/*
 * cv::FAST(roi, kps, thr) with non-max suppression on a w x h ROI (row pitch `stride`)
 * treated as its own image.  score[w*h] gets the keypoint response (m - 1) at
 * keypoints and 0 elsewhere; returns the keypoint count.  Raster order of the
 * non-zero entries = OpenCV's output order.
 */
int olk_fast(const u8 *roi, int stride, int w, int h, int thr, u8 *score)
{
	u8 *raw = (u8 *)calloc((size_t)w * h, 1);
	int off[16];
	for (int k = 0; k < 16; k++) off[k] = ring_dy[k] * stride + ring_dx[k];
	memset(score, 0, (size_t)w * h);
	for (int y = 3; y < h - 3; y++)
		for (int x = 3; x < w - 3; x++) {
			const int m = fast_m(roi + (size_t)y * stride + x, off, thr);
			if (m > thr) raw[y * w + x] = (u8)(m - 1);
		}
	int n = 0;
	for (int y = 3; y < h - 3; y++)
		for (int x = 3; x < w - 3; x++) {
			const int s = raw[y * w + x];
			if (!s) continue;
			int keep = 1;
			for (int j = -1; j <= 1 && keep; j++)
				for (int i = -1; i <= 1; i++)
					if ((i || j) && raw[(y + j) * w + x + i] >= s) { keep = 0; break; }
			if (keep) { score[y * w + x] = (u8)s; n++; }
		}
	free(raw);
	return n;
}

/* ---------------------------------------------------------------- tracker state */

#define OLK_MAXF 100 /* trackFeatures.cpp:56 */

typedef struct olk_kp { int x, y, resp; } olk_kp;

typedef struct olk_tracker {
	/* parameters */
	int w, h, n_feat, output_rate;
	float focal_x, focal_y, conf;
	int win, max_level, fast_thr;
	/* trackFeatures.cpp:53-56, per object here */
	int have_prev;
	u8 *pyr[2][8]; /* [slot][level]; slot `cur` holds the newest frame */
	int ws[8], hs[8], top;
	int cur;
	int n_prev_corners;                 /* prev_corners.size(): 0 until the first successful init */
	float prev_corners[2 * OLK_MAXF];
	u8 prev_status[OLK_MAXF];
	/* flow_opencv.hpp:66-67 */
	int update_vec[OLK_MAXF];
	int have_update_vec;
	float feat_cur[2 * OLK_MAXF], feat_prev[2 * OLK_MAXF];
	int have_feat_prev;
	/* limiter, per object (optical_flow.cpp:55 static made per stream) */
	float sum_x, sum_y; int sum_q, n_valid; uint32_t t_last;
	/* flow_opencv.cpp:72-87: camera model, used only when both setters were called */
	int have_cam, have_dist;
	float cam[4];  /* fx, fy, cx, cy */
	float dc[5];   /* k1, k2, p1, p2, k3 */
	/* diagnostics of the last call */
	int last_fast_bins, last_inliers, last_active;
	float last_px_x, last_px_y;
} olk_tracker;

olk_tracker *olk_new(float fx, float fy, int rate, int w, int h, int n_feat, float conf)
{
	if (n_feat < 1 || n_feat > OLK_MAXF) return NULL;
	olk_tracker *t = (olk_tracker *)calloc(1, sizeof *t);
	t->w = w; t->h = h; t->n_feat = n_feat; t->output_rate = rate; t->focal_x = fx; t->focal_y = fy; t->conf = conf;
	t->win = 9; t->max_level = 3; t->fast_thr = 10;
	t->top = olk_pyr_top_level(w, h, t->win, t->max_level);
	int lw = w, lh = h;
	for (int l = 0; l <= t->top; l++) {
		t->ws[l] = lw; t->hs[l] = lh;
		t->pyr[0][l] = (u8 *)malloc((size_t)lw * lh);
		t->pyr[1][l] = (u8 *)malloc((size_t)lw * lh);
		lw = (lw + 1) / 2; lh = (lh + 1) / 2;
	}
	return t;
}

void olk_free(olk_tracker *t)
{
	if (!t) return;
	for (int l = 0; l <= t->top; l++) { free(t->pyr[0][l]); free(t->pyr[1][l]); }
	free(t);
}

/* Test hook: replaces the deterministic sort below by a caller-supplied one.  tests/test_ref_lk.py installs
 * libstdc++'s own std::sort with the reference's response-only comparator (oracle/stdsort_shim.cpp), which makes
 * this restatement reproduce the reference objects bit for bit on inputs full of response ties as well. */
typedef void (*olk_sort_fn)(olk_kp *kps, size_t n, int W);
static olk_sort_fn sort_hook;
void olk_set_sort_hook(olk_sort_fn fn) { sort_hook = fn; }

/* descending response, ties by raster index ascending (documented divergence from std::sort) */
static int kp_cmp(const void *a, const void *b, void *wp)
{
	const olk_kp *p = (const olk_kp *)a, *q = (const olk_kp *)b;
	const int W = *(const int *)wp;
	if (p->resp != q->resp) return q->resp - p->resp;
	const long ip = (long)p->y * W + p->x, iq = (long)q->y * W + q->x;
	return ip < iq ? -1 : ip > iq;
}

/* Number of sort calls since the last reset whose result depends on how ties are ordered by an UNSTABLE sort:
 * more than 16 elements (libstdc++ sorts <= 16 by a stable insertion sort) and two DIFFERENT keypoints of equal
 * response among them. */
static long tie_sensitive_sorts;
long olk_tie_sensitive_sorts(int reset)
{
	const long v = tie_sensitive_sorts;
	if (reset) tie_sensitive_sorts = 0;
	return v;
}

static void sort_kps(olk_kp *kps, size_t n, int W)
{
	if (n > 16) {
		int first[256]; /* index + 1 of the first keypoint with this response; the same keypoint listed twice
				   (trackFeatures.cpp:212,245-247 can do that) is interchangeable with itself */
		memset(first, 0, sizeof first);
		for (size_t i = 0; i < n; i++) {
			const int r = kps[i].resp & 255;
			if (first[r] && (kps[first[r] - 1].x != kps[i].x || kps[first[r] - 1].y != kps[i].y)) { tie_sensitive_sorts++; break; }
			if (!first[r]) first[r] = (int)i + 1;
		}
	}
	if (sort_hook) sort_hook(kps, n, W);
	else qsort_r(kps, n, sizeof(olk_kp), kp_cmp, (void *)&W);
}

static int too_close_existing(const olk_tracker *t, int x, int y, unsigned dist)
{
	for (int j = 0; j < t->n_prev_corners; j++) {
		if (!t->prev_status[j]) continue;
		const int ex = (int)t->prev_corners[2 * j], ey = (int)t->prev_corners[2 * j + 1];
		if ((unsigned)abs(ex - x) < dist && (unsigned)abs(ey - y) < dist) return 1;
	}
	return 0;
}

static int too_close_list(const olk_kp *l, int n, int x, int y, unsigned dist)
{
	for (int j = 0; j < n; j++)
		if ((unsigned)abs(l[j].x - x) < dist && (unsigned)abs(l[j].y - y) < dist) return 1;
	return 0;
}

/* trackFeatures.cpp:136-363, stereo = 0 */
static void init_more_points(olk_tracker *t, const u8 *img, int *status, float *feat)
{
	const int N = t->n_feat, W = t->w, H = t->h;
	unsigned target = 0;
	for (int i = 0; i < N; i++) target += status[i] == 2;
	t->last_fast_bins = 0;
	if (!target) return;
	const int bw = W / 4, bh = H / 4, tpb = (N - 1) / 16 + 1;
	int per_bin[4][4] = {{0}};
	for (int i = 0; i < t->n_prev_corners; i++)
		if (status[i] == 1) {
			const int bx = (int)(t->prev_corners[2 * i] / bw), by = (int)(t->prev_corners[2 * i + 1] / bh);
			if (bx >= 4 || by >= 4) continue;
			per_bin[bx][by]++;
		}
	unsigned dist = bw / tpb;
	olk_kp *good = (olk_kp *)malloc(sizeof(olk_kp) * ((size_t)16 * tpb + N + 16));
	olk_kp *unused = (olk_kp *)malloc(sizeof(olk_kp) * ((size_t)W * H + 16));
	olk_kp *kps = (olk_kp *)malloc(sizeof(olk_kp) * ((size_t)bw * bh + 16));
	u8 *score = (u8 *)malloc((size_t)bw * bh + 16);
	size_t n_good = 0, n_unused = 0;
	for (int x = 0; x < 4; x++)
		for (int y = 0; y < 4; y++) {
			const int needed = tpb - per_bin[x][y] > 0 ? tpb - per_bin[x][y] : 0;
			if (!needed) continue;
			t->last_fast_bins++;
			const int c0 = x * bw, r0 = y * bh; /* (x+1)*bw <= W always */
			olk_fast(img + (size_t)r0 * W + c0, W, bw, bh, t->fast_thr, score);
			size_t nk = 0;
			for (int yy = 0; yy < bh; yy++)
				for (int xx = 0; xx < bw; xx++)
					if (score[yy * bw + xx]) { kps[nk].x = xx + c0; kps[nk].y = yy + r0; kps[nk].resp = score[yy * bw + xx]; nk++; }
			sort_kps(kps, nk, W);
			olk_kp *bin_good = good + n_good;
			int n_bin = 0;
			size_t idx = 0;
			for (; idx < nk; idx++) {
				const int nx = kps[idx].x, ny = kps[idx].y;
				if (too_close_existing(t, nx, ny, dist) || too_close_list(bin_good, n_bin, nx, ny, dist)) {
					unused[n_unused++] = kps[idx];
					continue;
				}
				bin_good[n_bin++] = kps[idx];
				if (n_bin == needed) break;
			}
			n_good += n_bin;
			if (idx < nk - 1) /* size_t wrap for nk == 0 as in the reference: an empty range */
				for (size_t k = idx; k < nk; k++) unused[n_unused++] = kps[k];
		}
	if (n_good > target) {
		const int n_remove = (int)(n_good - target), step = (int)target / n_remove + 2;
		size_t m = 0;
		for (size_t i = 0; i < n_good; i++)
			if (i % step) good[m++] = good[i];
		n_good = m;
	}
	if (n_good < target) {
		sort_kps(unused, n_unused, W);
		dist /= 2;
		for (size_t i = 0; i < n_unused; i++) {
			const int nx = unused[i].x, ny = unused[i].y;
			if (too_close_existing(t, nx, ny, dist) || too_close_list(good, (int)n_good, nx, ny, dist)) continue;
			good[n_good++] = unused[i];
			if (n_good == target) break;
		}
	}
	if (!n_good) {
		for (int i = 0; i < N; i++)
			if (status[i] == 2) status[i] = 0;
	} else {
		if (t->n_prev_corners < N) {
			for (int i = t->n_prev_corners; i < N; i++) t->prev_corners[2 * i] = t->prev_corners[2 * i + 1] = 0.f;
			t->n_prev_corners = N;
		}
		size_t m = 0;
		for (int i = 0; i < N; i++) {
			if (status[i] != 2) continue;
			if (m < n_good) {
				t->prev_corners[2 * i] = feat[2 * i] = (float)good[m].x;
				t->prev_corners[2 * i + 1] = feat[2 * i + 1] = (float)good[m].y;
				t->prev_status[i] = 1;
				m++;
			} else status[i] = 0;
		}
	}
	free(good); free(unused); free(kps); free(score);
}

/* optional: make the next track call use these points instead of detecting (corner injection for LK KATs) */

/* trackFeatures.cpp:70-132, stereo = 0 */
static void track_features(olk_tracker *t, const u8 *img, float *feat, int *status)
{
	const int N = t->n_feat;
	for (int i = 0; i < N; i++) { feat[2 * i] = feat[2 * i + 1] = -100.f; }
	for (int i = 0; i < N; i++) t->prev_status[i] = status[i] == 1;
	/* new pyramid into the other slot */
	const int nxt = t->cur ^ 1;
	memcpy(t->pyr[nxt][0], img, (size_t)t->w * t->h);
	for (int l = 1; l <= t->top; l++) olk_pyr_down(t->pyr[nxt][l - 1], t->ws[l - 1], t->hs[l - 1], t->pyr[nxt][l]);
	if (t->have_prev && t->n_prev_corners) {
		float nextp[2 * OLK_MAXF];
		u8 st[OLK_MAXF];
		olk_track((const u8 *const *)t->pyr[t->cur], (const u8 *const *)t->pyr[nxt], t->ws, t->hs, t->top, t->prev_corners,
			  nextp, st, t->n_prev_corners, t->win, 30, 0.01, 1e-4f);
		memcpy(t->prev_corners, nextp, sizeof(float) * 2 * t->n_prev_corners);
		for (int i = 0; i < t->n_prev_corners && i < N; i++) {
			if (!(t->prev_status[i] && st[i])) t->prev_status[i] = 0;
			if (t->prev_status[i] == 1) {
				const float x = t->prev_corners[2 * i], y = t->prev_corners[2 * i + 1];
				if (x < 0 || x > t->w || y < 0 || y > t->h) status[i] = 0;
				else { feat[2 * i] = x; feat[2 * i + 1] = y; status[i] = 1; }
			} else if (status[i] == 1) status[i] = 0;
		}
	}
	t->cur = nxt;
	t->have_prev = 1;
	init_more_points(t, img, status, feat);
}

/* optical_flow.cpp:51-96 with a per-object time_last_pub */
static int limit_rate(olk_tracker *t, int q, uint32_t t_us, int *dt_us, float *fx, float *fy)
{
	if (t->output_rate <= 0) {
		*dt_us = (int)(t_us - t->t_last);
		t->t_last = t_us;
		return q;
	}
	if (q > 0) { t->sum_x += *fx; t->sum_y += *fy; t->sum_q += q; t->n_valid++; }
	if ((float)(uint32_t)(t_us - t->t_last) > 1.0e6f / t->output_rate) {
		int avg = 0;
		if (t->n_valid > 0) avg = t->sum_q / t->n_valid; /* int / int, then std::floor of an int */
		*fx = t->sum_x; *fy = t->sum_y;
		t->sum_x = t->sum_y = 0.f; t->sum_q = 0; t->n_valid = 0;
		*dt_us = (int)(t_us - t->t_last);
		t->t_last = t_us;
		return avg;
	}
	return -1;
}

void olk_set_camera_matrix(olk_tracker *t, float fx, float fy, float cx, float cy)
{
	t->cam[0] = fx; t->cam[1] = fy; t->cam[2] = cx; t->cam[3] = cy; t->have_cam = 1;
}

/* setFocalLengthX/Y, setOutputRate (optical_flow.hpp:76-78), setConfMultiplier (flow_opencv.hpp:73): plain
 * assignments, all tracker / limiter state is kept */
void olk_set_params(olk_tracker *t, float fx, float fy, int rate, float conf)
{
	t->focal_x = fx; t->focal_y = fy; t->output_rate = rate; t->conf = conf;
}

/* argument order of OpticalFlowOpenCV::setCameraDistortion(k1, k2, k3, p1, p2) */
void olk_set_camera_distortion(olk_tracker *t, float k1, float k2, float k3, float p1, float p2)
{
	t->dc[0] = k1; t->dc[1] = k2; t->dc[2] = p1; t->dc[3] = p2; t->dc[4] = k3; t->have_dist = 1;
}

/* cv::undistortPoints(src, dst, K, D) (no R, no P; 5 fixed-point iterations, in double): normalised coordinates,
 * rounded to f32 as cv::Point2f stores them */
void olk_undistort_norm(const float cam[4], const float dc[5], const float *pt, float *out)
{
	const double fx = cam[0], fy = cam[1], cx = cam[2], cy = cam[3];
	const double k1 = dc[0], k2 = dc[1], p1 = dc[2], p2 = dc[3], k3 = dc[4];
	const double ifx = 1. / fx, ify = 1. / fy;
	const double u = pt[0], v = pt[1];
	double x = (u - cx) * ifx, y = (v - cy) * ify;
	const double x0 = x, y0 = y;
	for (int j = 0; j < 5; j++) {
		const double r2 = x * x + y * y;
		const double icdist = (1 + ((0 * r2 + 0) * r2 + 0) * r2) / (1 + ((k3 * r2 + k2) * r2 + k1) * r2);
		if (icdist < 0) { x = (u - cx) * ifx; y = (v - cy) * ify; break; }
		const double dx = 2 * p1 * x * y + p2 * (r2 + 2 * x * x) + 0 * r2 + 0 * r2 * r2;
		const double dy = p1 * (r2 + 2 * y * y) + 2 * p2 * x * y + 0 * r2 + 0 * r2 * r2;
		x = (x0 - dx) * icdist; y = (y0 - dy) * icdist;
	}
	out[0] = (float)x; out[1] = (float)y;
}

/* olk_undistort_norm followed by the re-projection of flow_opencv.cpp:113-118 (in float) */
void olk_undistort(const float cam[4], const float dc[5], float *pt)
{
	float n[2];
	olk_undistort_norm(cam, dc, pt, n);
	pt[0] = n[0] * cam[0] + cam[2];
	pt[1] = n[1] * cam[1] + cam[3];
}

/* flow_opencv.cpp:89-210 */
int olk_calc_flow(olk_tracker *t, const u8 *img, uint32_t t_us, int *dt_us, float *flow_x, float *flow_y)
{
	const int N = t->n_feat;
	if (!t->have_update_vec) {
		for (int i = 0; i < N; i++) t->update_vec[i] = 2;
		t->have_update_vec = 1;
	}
	int meancount = 0;
	float mx = 0.f, my = 0.f, sx = 0.f, sy = 0.f;
	track_features(t, img, t->feat_cur, t->update_vec);
	if (t->have_cam && t->have_dist)
		for (int i = 0; i < N; i++) olk_undistort(t->cam, t->dc, &t->feat_cur[2 * i]);
	int *uv = t->update_vec;
	const float *c = t->feat_cur, *p = t->feat_prev;
	t->last_active = 0;
	if (t->have_feat_prev) {
		for (int i = 0; i < N; i++)
			if (uv[i] == 1) { mx += c[2 * i] - p[2 * i]; my += c[2 * i + 1] - p[2 * i + 1]; meancount++; }
		t->last_active = meancount;
		if (meancount) {
			mx /= meancount; my /= meancount;
			for (int i = 0; i < N; i++)
				if (uv[i] == 1) {
					const float ex = c[2 * i] - p[2 * i] - mx, ey = c[2 * i + 1] - p[2 * i + 1] - my;
					sx = (float)((double)sx + (double)ex * (double)ex); /* pow(float, 2) is evaluated in double */
					sy = (float)((double)sy + (double)ey * (double)ey);
				}
			sx = sqrtf(sx / meancount); sy = sqrtf(sy / meancount);
			float tx = 0.f, ty = 0.f;
			meancount = 0;
			for (int i = 0; i < N; i++)
				if (uv[i] == 1) {
					const float fx = c[2 * i] - p[2 * i], fy = c[2 * i + 1] - p[2 * i + 1];
					if (fabsf(fx - mx) < sx * t->conf && fabsf(fy - my) < sy * t->conf) { tx += fx; ty += fy; meancount++; }
					else uv[i] = 0;
				}
			if (meancount) { mx = tx / meancount; my = ty / meancount; }
		}
	}
	memcpy(t->feat_prev, t->feat_cur, sizeof(float) * 2 * N);
	t->have_feat_prev = 1;
	for (int i = 0; i < N; i++) {
		if (uv[i] == 2) uv[i] = 1;
		if (uv[i] == 0) uv[i] = 2;
	}
	t->last_inliers = meancount; t->last_px_x = mx; t->last_px_y = my;
	float fx = mx, fy = my;
	int q = (int)round(255.0 * meancount / N);
	q = limit_rate(t, q, t_us, dt_us, &fx, &fy);
	*flow_x = (float)atan2((double)fx, (double)t->focal_x);
	*flow_y = (float)atan2((double)fy, (double)t->focal_y);
	return q;
}

/* state access for tests */
void olk_get_state(const olk_tracker *t, int *update_vec, float *feat_cur, float *prev_corners, u8 *prev_status, int *diag)
{
	memcpy(update_vec, t->update_vec, sizeof(int) * t->n_feat);
	memcpy(feat_cur, t->feat_cur, sizeof(float) * 2 * t->n_feat);
	memcpy(prev_corners, t->prev_corners, sizeof(float) * 2 * t->n_feat);
	memcpy(prev_status, t->prev_status, t->n_feat);
	diag[0] = t->last_fast_bins; diag[1] = t->last_inliers; diag[2] = t->last_active; diag[3] = t->n_prev_corners;
	memcpy(&diag[4], &t->last_px_x, 4); memcpy(&diag[5], &t->last_px_y, 4);
}

/* ---------------------------------------------------------------- Shi-Tomasi / goodFeaturesToTrack (App. A.6)
 * named by north_star, not called by the reference.  Restates cv::cornerMinEigenVal(blockSize 3, Sobel 3) and
 * cv::goodFeaturesToTrack as cv2 4.13 evaluates them for u8 input:
 *   Dx = sepFilter(row [-1 0 1], column [s 2s s]),  Dy = sepFilter(row [s 2s s], column [-1 0 1]),  s = (float)(1/3060),
 *   both f32 with BORDER_REFLECT_101;  cov = (DxDx, DxDy, DyDy) box-summed 3x3 in double (reflect-101), cast to f32;
 *   lambda_min = (a/2 + c/2) - sqrt((a/2 - c/2)^2 + b^2).
 * cv2's vectorised column filter may fuse multiply-adds, so parity with cv2 is a tolerance (tests/test_oracle_lk.py). */
static void sobel_dxdy(const u8 *img, int w, int h, int x, int y, float *dx, float *dy)
{
	const float s = (float)(1. / 3060.), s2 = 2.f * s;
	float rx[3], ry[3]; /* row-filtered values of rows y-1, y, y+1 */
	for (int j = 0; j < 3; j++) {
		const u8 *r = img + (size_t)reflect101(y + j - 1, h) * w;
		const int a = r[reflect101(x - 1, w)], b = r[x], c = r[reflect101(x + 1, w)];
		rx[j] = (float)(c - a);
		ry[j] = (float)(a + c) * s + (float)b * s2;
	}
	*dx = (rx[0] + rx[2]) * s + rx[1] * s2;
	*dy = ry[2] - ry[0];
}

void olk_min_eig(const u8 *img, int w, int h, float *eig)
{
	float *cov = (float *)malloc(sizeof(float) * 3 * (size_t)w * h);
	for (int y = 0; y < h; y++)
		for (int x = 0; x < w; x++) {
			float dx, dy;
			sobel_dxdy(img, w, h, x, y, &dx, &dy);
			float *c = cov + 3 * ((size_t)y * w + x);
			c[0] = dx * dx; c[1] = dx * dy; c[2] = dy * dy;
		}
	for (int y = 0; y < h; y++)
		for (int x = 0; x < w; x++) {
			double s[3] = {0, 0, 0};
			for (int j = -1; j <= 1; j++)
				for (int i = -1; i <= 1; i++) {
					const float *c = cov + 3 * ((size_t)reflect101(y + j, h) * w + reflect101(x + i, w));
					s[0] += c[0]; s[1] += c[1]; s[2] += c[2];
				}
			const float a = (float)s[0] * 0.5f, b = (float)s[1], c = (float)s[2] * 0.5f;
			eig[(size_t)y * w + x] = (a + c) - sqrtf((a - c) * (a - c) + b * b);
		}
	free(cov);
}

typedef struct { float v; int idx; } gftt_cand;
static int gftt_cmp(const void *a, const void *b)
{
	const gftt_cand *p = (const gftt_cand *)a, *q = (const gftt_cand *)b;
	if (p->v != q->v) return p->v > q->v ? -1 : 1;
	return q->idx - p->idx; /* ties: higher address first */
}

/* -> number of corners written to xy[2 * max_corners] */
int olk_gftt(const u8 *img, int w, int h, int max_corners, double quality, double min_dist, float *xy)
{
	float *eig = (float *)malloc(sizeof(float) * (size_t)w * h);
	olk_min_eig(img, w, h, eig);
	float mx = eig[0];
	for (size_t i = 1; i < (size_t)w * h; i++) if (eig[i] > mx) mx = eig[i];
	const float thr = (float)((double)mx * quality);
	for (size_t i = 0; i < (size_t)w * h; i++) if (!(eig[i] > thr)) eig[i] = 0.f;
	gftt_cand *cand = (gftt_cand *)malloc(sizeof(gftt_cand) * (size_t)w * h);
	int nc = 0;
	for (int y = 1; y < h - 1; y++)
		for (int x = 1; x < w - 1; x++) {
			const float v = eig[(size_t)y * w + x];
			if (v == 0.f) continue;
			float m = v;
			for (int j = -1; j <= 1; j++)
				for (int i = -1; i <= 1; i++) { const float u = eig[(size_t)(y + j) * w + x + i]; if (u > m) m = u; }
			if (v == m) { cand[nc].v = v; cand[nc].idx = y * w + x; nc++; }
		}
	qsort(cand, nc, sizeof(gftt_cand), gftt_cmp);
	int n = 0;
	const float md2 = (float)(min_dist * min_dist);
	for (int i = 0; i < nc && (max_corners <= 0 || n < max_corners); i++) {
		const int x = cand[i].idx % w, y = cand[i].idx / w;
		int ok = 1;
		if (min_dist >= 1)
			for (int j = 0; j < n && ok; j++) {
				const float dx = (float)x - xy[2 * j], dy = (float)y - xy[2 * j + 1];
				if (dx * dx + dy * dy < md2) ok = 0;
			}
		if (ok) { xy[2 * n] = (float)x; xy[2 * n + 1] = (float)y; n++; }
	}
	free(cand); free(eig);
	return n;
}
