// oracle/cv_compat/cv_compat.hpp -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
//
// The smallest stand-in for the OpenCV C++ API that lets the reference's LK-path sources
//   modules/OpticalFlow/external/klt_feature_tracker/src/trackFeatures.cpp
//   modules/OpticalFlow/src/flow_opencv.cpp   (+ include/flow_opencv.hpp, which wants <cv.h>)
// compile UNMODIFIED, where they lie under /root/reference, into oracle/_ref/libaof_ref_lk.so
// (oracle/Makefile, target `ref`).  No OpenCV C++ SDK exists in this image; only these two files'
// view of it is provided:
//   types   : Mat (header over external or owned data, rowRange/colRange/copyTo), Mat_<T> with the
//             `m << a, b, c` initialiser, Point2f, KeyPoint, DMatch, Size, TermCriteria
//   calls   : FAST (trackFeatures.cpp:196), calcOpticalFlowPyrLK (:98,:101,:386), undistortPoints
//             (flow_opencv.cpp:110), cornerSubPix (:382, stereo only: aborts, never reached with stereo = 0)
// The three live calls go through a backend table.  Its default entries are the cv2-4.13-pinned
// primitives of oracle/lk_oracle.c (olk_fast, olk_pyr_down + olk_track, olk_undistort_norm:
// tests/test_oracle_lk.py checks each against cv2 executing); tests and bench.py may install
// callbacks that run the real OpenCV through cv2 instead (oracle/ref_lk.py).
// Everything the reference AUTHORED on this path -- the status machine, the bin bookkeeping,
// std::sort(compareKeypoints), the distance gating, thinning, second chance, slot assignment,
// the f32 statistics, limitRate -- is the reference's own object code.
//
// Nothing under aero-optical-flow_b200/ includes or links this.
#pragma once
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <vector>

typedef unsigned char uchar;

#define CV_8UC1 0
#define CV_32FC1 5
#define CV_TERMCRIT_ITER 1
#define CV_TERMCRIT_EPS 2

extern "C" {
// backend of the three OpenCV calls (plain C types so that ctypes callbacks fit)
typedef int (*cvc_fast_fn)(const uchar *roi, int stride, int w, int h, int thr, float *xyr, int cap);
typedef void (*cvc_pyrlk_fn)(const uchar *prev, const uchar *next, int w, int h, const float *pts, int n, float *out,
			     uchar *status, int win, int max_level);
typedef void (*cvc_undistort_fn)(const float *pts, int n, const float *K9, const float *D5, float *out);
struct cvc_backend {
	cvc_fast_fn fast;
	cvc_pyrlk_fn pyrlk;
	cvc_undistort_fn undistort;
	long n_fast, n_pyrlk, n_undistort; // call counters (tests assert the reference really went through here)
};
cvc_backend *cvc_backend_get(void); // defined in oracle/ref_lk_shim.cpp
}

namespace cv
{

template <typename T> struct Point_ {
	T x, y;
	Point_() : x(0), y(0) {}
	Point_(T x_, T y_) : x(x_), y(y_) {}
};
typedef Point_<float> Point2f;
typedef Point_<int> Point;

struct Size {
	int width, height;
	Size() : width(0), height(0) {}
	Size(int w, int h) : width(w), height(h) {}
};

struct TermCriteria {
	int type, maxCount;
	double epsilon;
	TermCriteria() : type(0), maxCount(0), epsilon(0) {}
	TermCriteria(int t, int n, double e) : type(t), maxCount(n), epsilon(e) {}
};

struct KeyPoint {
	Point2f pt;
	float size, angle, response;
	int octave, class_id;
	KeyPoint() : size(0), angle(-1), response(0), octave(0), class_id(-1) {}
	KeyPoint(float x, float y, float s, float a = -1, float r = 0, int o = 0, int c = -1)
		: pt(x, y), size(s), angle(a), response(r), octave(o), class_id(c) {}
};

struct DMatch {
	int queryIdx, trainIdx, imgIdx;
	float distance;
	DMatch() : queryIdx(-1), trainIdx(-1), imgIdx(-1), distance(0) {}
};

// 8-bit single-channel matrix header.  `data` is public and assignable (flow_opencv.cpp:103 points it
// at the caller's frame); owned storage is released through `own`, as cv::Mat releases through its
// refcount and not through `data`.
class Mat
{
public:
	int rows, cols;
	size_t step;
	uchar *data;
	std::shared_ptr<std::vector<uchar> > own;

	Mat() : rows(0), cols(0), step(0), data(0) {}
	Mat(int r, int c, int /*type*/) : rows(r), cols(c), step((size_t)c), data(0) { alloc(r, c); }

	bool empty() const { return data == 0 || rows == 0 || cols == 0; }
	bool isContinuous() const { return step == (size_t)cols; }

	Mat rowRange(int a, int b) const
	{
		Mat m(*this);
		m.rows = b - a;
		m.data = data + (size_t)a * step;
		return m;
	}
	Mat colRange(int a, int b) const
	{
		Mat m(*this);
		m.cols = b - a;
		m.data = data + a;
		return m;
	}
	void copyTo(Mat &dst) const
	{
		if (dst.rows != rows || dst.cols != cols || !dst.own || dst.data != dst.own->data()) dst.alloc(rows, cols);
		for (int y = 0; y < rows; y++) std::memcpy(dst.data + (size_t)y * dst.step, data + (size_t)y * step, (size_t)cols);
	}

private:
	void alloc(int r, int c)
	{
		own = std::make_shared<std::vector<uchar> >((size_t)r * c + 1);
		data = own->data();
		rows = r;
		cols = c;
		step = (size_t)c;
	}
};

template <typename T> class Mat_;

template <typename T> class MatCommaInitializer_
{
public:
	explicit MatCommaInitializer_(Mat_<T> *m) : mat(m), idx(0) {}
	template <typename T2> MatCommaInitializer_<T> &operator,(T2 v)
	{
		mat->at_linear(idx++) = (T)v;
		return *this;
	}
	Mat_<T> *mat;
	size_t idx;
};

template <typename T> class Mat_
{
public:
	int rows, cols;
	std::vector<T> v;
	Mat_() : rows(0), cols(0) {}
	void create(int r, int c)
	{
		rows = r;
		cols = c;
		v.assign((size_t)r * c, T());
	}
	T &operator()(int i, int j) { return v[(size_t)i * cols + j]; }
	const T &operator()(int i, int j) const { return v[(size_t)i * cols + j]; }
	T &at_linear(size_t i) { return v.at(i); }
};

template <typename T, typename T2> static inline MatCommaInitializer_<T> operator<<(Mat_<T> &m, T2 val)
{
	MatCommaInitializer_<T> ci(&m);
	return (ci, val);
}

// ---- the three OpenCV calls on the path --------------------------------------------------------

// cv::FAST(image, keypoints, threshold, nonmaxSuppression = true): FAST-9/16, keypoints in raster
// order, pt relative to the ROI, response = score, size 7, angle -1.
static inline void FAST(const Mat &img, std::vector<KeyPoint> &kps, int threshold, bool nms = true)
{
	kps.clear();
	if (!nms) {
		std::fprintf(stderr, "cv_compat: FAST without non-max suppression is not on the path\n");
		std::abort();
	}
	cvc_backend *b = cvc_backend_get();
	b->n_fast++;
	const int cap = ((img.cols + 1) / 2) * ((img.rows + 1) / 2) + 16; // strict 3x3 maxima cannot be denser
	std::vector<float> xyr((size_t)cap * 3);
	const int n = b->fast(img.data, (int)img.step, img.cols, img.rows, threshold, xyr.data(), cap);
	kps.reserve(n);
	for (int i = 0; i < n; i++) kps.push_back(KeyPoint(xyr[3 * i], xyr[3 * i + 1], 7.f, -1.f, xyr[3 * i + 2]));
}

static inline void contiguous(const Mat &m, std::vector<uchar> &buf, const uchar *&p)
{
	if (m.isContinuous()) {
		p = m.data;
		return;
	}
	buf.resize((size_t)m.rows * m.cols);
	for (int y = 0; y < m.rows; y++) std::memcpy(&buf[(size_t)y * m.cols], m.data + (size_t)y * m.step, (size_t)m.cols);
	p = buf.data();
}

// cv::calcOpticalFlowPyrLK(prev, next, prevPts, nextPts, status, err, winSize, maxLevel) with the
// defaults the reference leaves alone: criteria (COUNT + EPS, 30, 0.01), flags 0, minEigThreshold 1e-4.
// `err` is sized but not computed (the reference never reads it, trackFeatures.cpp:94,98).
static inline void calcOpticalFlowPyrLK(const Mat &prev, const Mat &next, const std::vector<Point2f> &prevPts,
					std::vector<Point2f> &nextPts, std::vector<uchar> &status, std::vector<float> &err,
					Size winSize = Size(21, 21), int maxLevel = 3)
{
	const int n = (int)prevPts.size();
	nextPts.resize(n);
	status.resize(n);
	err.assign(n, 0.f);
	if (!n) return;
	if (prev.rows != next.rows || prev.cols != next.cols || winSize.width != winSize.height) {
		std::fprintf(stderr, "cv_compat: calcOpticalFlowPyrLK: unsupported arguments\n");
		std::abort();
	}
	cvc_backend *b = cvc_backend_get();
	b->n_pyrlk++;
	std::vector<uchar> b0, b1;
	const uchar *p0, *p1;
	contiguous(prev, b0, p0);
	contiguous(next, b1, p1);
	b->pyrlk(p0, p1, prev.cols, prev.rows, &prevPts[0].x, n, &nextPts[0].x, status.data(), winSize.width, maxLevel);
}

// cv::undistortPoints(src, dst, cameraMatrix, distCoeffs): normalised coordinates, no R, no P.
static inline void undistortPoints(const std::vector<Point2f> &src, std::vector<Point2f> &dst, const Mat_<float> &K,
				   const Mat_<float> &D)
{
	cvc_backend *b = cvc_backend_get();
	b->n_undistort++;
	std::vector<Point2f> out(src.size());
	if (!src.empty()) b->undistort(&src[0].x, (int)src.size(), K.v.data(), D.v.data(), &out[0].x);
	dst = out;
}

// stereo branch only (trackFeatures.cpp:382); aero always passes stereo = 0 (flow_opencv.cpp:106)
static inline void cornerSubPix(const Mat &, std::vector<Point2f> &, Size, Size, TermCriteria)
{
	std::fprintf(stderr, "cv_compat: cornerSubPix (klt stereo branch) is not on the calcFlow path\n");
	std::abort();
}

} // namespace cv
