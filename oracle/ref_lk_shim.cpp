// oracle/ref_lk_shim.cpp -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
//
// extern "C" access to the UNMODIFIED reference LK-path objects:
//   modules/OpticalFlow/src/flow_opencv.cpp, src/optical_flow.cpp and
//   modules/OpticalFlow/external/klt_feature_tracker/src/trackFeatures.cpp
// are compiled where they lie under /root/reference by oracle/Makefile against the stand-in
// OpenCV headers of oracle/cv_compat/ (no reference source is copied).  This TU includes the
// reference HEADERS only; `private` is opened for this TU alone so tests can read updateVector
// and features_current.  It also owns the backend table behind cv::FAST /
// cv::calcOpticalFlowPyrLK / cv::undistortPoints (cv_compat.hpp), whose default entries are the
// cv2-pinned primitives of oracle/lk_oracle.c linked into the same library.
//
// The tracker keeps its state in file statics (trackFeatures.cpp:53-56) and limitRate in a
// function static (optical_flow.cpp:55): ONE live OpticalFlowOpenCV per loaded copy of this
// library.  oracle/ref_lk.py loads a private copy of the .so per tracker for that reason.
//
// Output: oracle/_ref/libaof_ref_lk.so (git-ignored, travels to the GPU box).
#define private public
#include "flow_opencv.hpp"
#undef private

#include <cstdint>
#include <cstring>
#include <vector>

extern "C" {
// oracle/lk_oracle.c (C linkage)
int olk_pyr_top_level(int w, int h, int win, int max_level);
void olk_pyr_down(const uint8_t *src, int w, int h, uint8_t *dst);
void olk_track(const uint8_t *const *pyr_prev, const uint8_t *const *pyr_next, const int *ws, const int *hs, int top_level,
	       const float *pts_prev, float *pts_next, uint8_t *status, int n, int win, int max_iter, double eps,
	       float min_eig_thr);
int olk_fast(const uint8_t *roi, int stride, int w, int h, int thr, uint8_t *score);
void olk_undistort_norm(const float cam[4], const float dc[5], const float *pt, float *out);
}

static int dflt_fast(const uchar *roi, int stride, int w, int h, int thr, float *xyr, int cap)
{
	std::vector<uint8_t> score((size_t)w * h + 16);
	olk_fast(roi, stride, w, h, thr, score.data());
	int n = 0;
	for (int y = 0; y < h; y++)
		for (int x = 0; x < w; x++)
			if (score[(size_t)y * w + x] && n < cap) {
				xyr[3 * n] = (float)x;
				xyr[3 * n + 1] = (float)y;
				xyr[3 * n + 2] = (float)score[(size_t)y * w + x];
				n++;
			}
	return n;
}

static void dflt_pyrlk(const uchar *prev, const uchar *next, int w, int h, const float *pts, int n, float *out, uchar *status,
		       int win, int max_level)
{
	const int top = olk_pyr_top_level(w, h, win, max_level);
	std::vector<std::vector<uint8_t> > lv[2];
	const uint8_t *pp[8], *pn[8];
	int ws[8], hs[8];
	ws[0] = w;
	hs[0] = h;
	pp[0] = prev;
	pn[0] = next;
	lv[0].resize(top + 1);
	lv[1].resize(top + 1);
	for (int l = 1; l <= top; l++) {
		ws[l] = (ws[l - 1] + 1) / 2;
		hs[l] = (hs[l - 1] + 1) / 2;
		lv[0][l].resize((size_t)ws[l] * hs[l]);
		lv[1][l].resize((size_t)ws[l] * hs[l]);
		olk_pyr_down(pp[l - 1], ws[l - 1], hs[l - 1], lv[0][l].data());
		olk_pyr_down(pn[l - 1], ws[l - 1], hs[l - 1], lv[1][l].data());
		pp[l] = lv[0][l].data();
		pn[l] = lv[1][l].data();
	}
	olk_track(pp, pn, ws, hs, top, pts, out, status, n, win, 30, 0.01, 1e-4f);
}

static void dflt_undistort(const float *pts, int n, const float *K9, const float *D5, float *out)
{
	const float cam[4] = {K9[0], K9[4], K9[2], K9[5]};
	for (int i = 0; i < n; i++) olk_undistort_norm(cam, D5, pts + 2 * i, out + 2 * i);
}

static cvc_backend g_backend = {dflt_fast, dflt_pyrlk, dflt_undistort, 0, 0, 0};

extern "C" {

cvc_backend *cvc_backend_get(void) { return &g_backend; }

// NULL keeps / restores the default (oracle primitive) for that call
void ref_lk_set_backend(cvc_fast_fn f, cvc_pyrlk_fn p, cvc_undistort_fn u)
{
	g_backend.fast = f ? f : dflt_fast;
	g_backend.pyrlk = p ? p : dflt_pyrlk;
	g_backend.undistort = u ? u : dflt_undistort;
}

void ref_lk_call_counts(long *out3)
{
	out3[0] = g_backend.n_fast;
	out3[1] = g_backend.n_pyrlk;
	out3[2] = g_backend.n_undistort;
}

void *ref_lk_new(float fx, float fy, int rate, int w, int h, int n_feat, float conf)
{
	return new OpticalFlowOpenCV(fx, fy, rate, w, h, n_feat, conf);
}

void ref_lk_set_camera_matrix(void *h, float fx, float fy, float cx, float cy)
{
	static_cast<OpticalFlowOpenCV *>(h)->setCameraMatrix(fx, fy, cx, cy);
}

void ref_lk_set_camera_distortion(void *h, float k1, float k2, float k3, float p1, float p2)
{
	static_cast<OpticalFlowOpenCV *>(h)->setCameraDistortion(k1, k2, k3, p1, p2);
}

// setters of the base class, for the "setters keep state" tests (optical_flow.hpp:74-78)
void ref_lk_set_params(void *h, float fx, float fy, int rate, float conf)
{
	OpticalFlowOpenCV *o = static_cast<OpticalFlowOpenCV *>(h);
	o->setFocalLengthX(fx);
	o->setFocalLengthY(fy);
	o->setOutputRate(rate);
	o->setConfMultiplier(conf);
}

int ref_lk_calc_flow(void *h, uint8_t *img, uint32_t t_us, int *dt_us, float *fx, float *fy)
{
	return static_cast<OpticalFlowOpenCV *>(h)->calcFlow(img, t_us, *dt_us, *fx, *fy);
}

// updateVector (after the call's 2->1 / 0->2 rewrite) and features_current
void ref_lk_get_state(void *h, int *update_vec, float *feat_cur)
{
	OpticalFlowOpenCV *o = static_cast<OpticalFlowOpenCV *>(h);
	for (size_t i = 0; i < o->updateVector.size(); i++) update_vec[i] = o->updateVector[i];
	for (size_t i = 0; i < o->features_current.size(); i++) {
		feat_cur[2 * i] = o->features_current[i].x;
		feat_cur[2 * i + 1] = o->features_current[i].y;
	}
}

void ref_lk_free(void *h) { delete static_cast<OpticalFlowOpenCV *>(h); }

} // extern "C"
