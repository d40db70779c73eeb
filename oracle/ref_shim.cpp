// oracle/ref_shim.cpp -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
//
// extern "C" access to the UNMODIFIED reference SAD objects (px4flow.cpp,
// flow_px4.cpp, optical_flow.cpp are compiled where they lie under
// /root/reference by oracle/Makefile; no reference source is copied).  This TU
// only includes the reference headers; `private` is opened for this TU alone so
// the per-block primitives can be called for known-answer tests.
//
// Output: oracle/_ref/libaof_ref.so (git-ignored, travels to the GPU box).
#define private public
#include "px4flow.hpp"
#include "flow_px4.hpp"
#undef private

#include <thread>
#include <vector>
#include <cstdint>
#include <cstring>

extern "C" {

uint32_t ref_compute_diff(uint8_t *img, int x, int y, int pitch)
{
	PX4Flow f(pitch, 4, 30, 3000);
	return f.compute_diff(img, (uint16_t)x, (uint16_t)y, (uint16_t)pitch);
}

uint32_t ref_compute_sad_8x8(uint8_t *a, uint8_t *b, int ax, int ay, int bx, int by, int pitch)
{
	PX4Flow f(pitch, 4, 30, 3000);
	return f.compute_sad_8x8(a, b, (uint16_t)ax, (uint16_t)ay, (uint16_t)bx, (uint16_t)by, (uint16_t)pitch);
}

void ref_compute_subpixel(uint8_t *a, uint8_t *b, int ax, int ay, int bx, int by, int pitch, uint32_t *acc)
{
	PX4Flow f(pitch, 4, 30, 3000);
	f.compute_subpixel(a, b, (uint16_t)ax, (uint16_t)ay, (uint16_t)bx, (uint16_t)by, acc, (uint16_t)pitch);
}

int ref_compute_flow(uint8_t *prev, uint8_t *cur, int width, int search, int feat_thr, int val_thr, float *fx,
		     float *fy)
{
	PX4Flow f(width, search, feat_thr, val_thr);
	return f.compute_flow(prev, cur, 0.f, 0.f, 0.f, fx, fy);
}

// n stateless pairs over `threads` host threads, each with its own PX4Flow
// (PX4Flow has no statics; OpticalFlow::limitRate does, so it stays out of here).
void ref_compute_flow_batch(uint8_t *prev, uint8_t *cur, int n, int width, int height, int search, int feat_thr,
			    int val_thr, float *fx, float *fy, int *quality, int threads)
{
	if (threads < 1) threads = 1;
	size_t fs = (size_t)width * height;
	auto work = [&](int t) {
		PX4Flow f(width, search, feat_thr, val_thr);
		for (int i = t; i < n; i += threads)
			quality[i] = f.compute_flow(prev + i * fs, cur + i * fs, 0.f, 0.f, 0.f, &fx[i], &fy[i]);
	};
	if (threads == 1) { work(0); return; }
	std::vector<std::thread> pool;
	for (int t = 0; t < threads; t++) pool.emplace_back(work, t);
	for (auto &th : pool) th.join();
}

// OpticalFlowPX4 sequences.  NOTE: limitRate's time_last_pub is one static for
// the whole process (optical_flow.cpp:55): use one live instance at a time.
void *ref_px4_new(float fx, float fy, int rate, int w, int h, int search, int feat_thr, int val_thr)
{
	return new OpticalFlowPX4(fx, fy, rate, w, h, search, feat_thr, val_thr);
}

int ref_px4_calc_flow(void *h, uint8_t *img, uint32_t t_us, int *dt_us, float *fx, float *fy)
{
	return static_cast<OpticalFlowPX4 *>(h)->calcFlow(img, t_us, *dt_us, *fx, *fy);
}

// the reference's own setters (include/optical_flow.hpp:76-78): they only assign
void ref_px4_set_params(void *h, float fx, float fy, int rate)
{
	OpticalFlowPX4 *o = static_cast<OpticalFlowPX4 *>(h);
	o->setFocalLengthX(fx);
	o->setFocalLengthY(fy);
	o->setOutputRate(rate);
}

void ref_px4_free(void *h) { delete static_cast<OpticalFlowPX4 *>(h); }

} // extern "C"
