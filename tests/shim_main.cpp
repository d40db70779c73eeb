// Test driver for the header-compatible C++ shims (aero-optical-flow_b200/cpp): uses the classes exactly as
// src/mainloop.cpp:423,322 of the reference does (new OpticalFlowOpenCV(fx, fy, rate, w, h); calcFlow(...)).
//   shim_main <px4|lk|lkcam|px4retune|lkretune> <frames.raw> <w> <h> <n_frames> <rate> <t_step_us>
// The *retune modes call the setters between frames (focal length, output rate, confidence multiplier), which in
// the reference only assign (include/optical_flow.hpp:74-78) and must not disturb tracker or limiter state.
// prints one line per frame: quality dt_us flow_x flow_y (hex floats; "untouched" outputs show the sentinels).
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "flow_opencv.hpp"
#include "flow_px4.hpp"

int main(int argc, char **argv)
{
	if (argc < 8) return 2;
	const int w = atoi(argv[3]), h = atoi(argv[4]), n = atoi(argv[5]), rate = atoi(argv[6]), step = atoi(argv[7]);
	std::vector<uint8_t> buf((size_t)w * h * n);
	FILE *f = fopen(argv[2], "rb");
	if (!f || fread(buf.data(), 1, buf.size(), f) != buf.size()) return 3;
	fclose(f);
	OpticalFlow *flow;
	OpticalFlowOpenCV *lk_obj = NULL;
	const bool retune = strstr(argv[1], "retune") != NULL;
	if (!strncmp(argv[1], "px4", 3)) flow = new OpticalFlowPX4(216.6677f, 216.2457f, rate, w, h, 4);
	else {
		OpticalFlowOpenCV *lk = new OpticalFlowOpenCV(216.6677f, 216.2457f, rate, w, h);
		if (!strcmp(argv[1], "lkcam")) {
			lk->setCameraMatrix(216.6677f, 216.2457f, w / 2.f, h / 2.f);
			lk->setCameraDistortion(-0.2f, 0.05f, 0.01f, 0.001f, -0.002f);
		}
		flow = lk;
		lk_obj = lk;
	}
	if (flow->getImageWidth() != w || flow->getImageHeight() != h) return 4;
	for (int k = 0; k < n; k++) {
		if (retune && (k == 4 || k == 8 || k == 11)) {
			flow->setFocalLengthX(k == 11 ? 216.f : 300.f);
			flow->setFocalLengthY(k == 11 ? 216.f : 310.f);
			flow->setOutputRate(k == 4 ? 15 : k == 8 ? -1 : 20);
			if (lk_obj) lk_obj->setConfMultiplier(k == 4 ? 1.645f : k == 8 ? 1.0f : 2.5f);
		}
		int dt_us = -123456;
		float fx = -77.f, fy = -77.f;
		const int q = flow->calcFlow(buf.data() + (size_t)k * w * h, (uint32_t)(1000 + step * k), dt_us, fx, fy);
		printf("%d %d %a %a\n", q, dt_us, fx, fy);
	}
	delete flow;
	return 0;
}
