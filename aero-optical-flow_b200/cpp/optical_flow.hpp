// Drop-in replacement for modules/OpticalFlow/include/optical_flow.hpp:51-88 of aero-optical-flow:
// the same abstract class (names, argument meaning, defaults), so src/mainloop.cpp:322,423 compiles
// unchanged.  The work happens behind the C ABI of libaof_b200.so (include/aof.h) on a B200; the
// rate limiter (src/optical_flow.cpp:51-96) runs on the device with the rest of calcFlow, so the
// sum_* members below are kept only for source compatibility of subclasses that read them.
#pragma once
#include <cstdint>
#include <cstdio>

#include "aof.h"

#define DEFAULT_OUTPUT_RATE 15
#define DEFAULT_IMAGE_WIDTH 64
#define DEFAULT_IMAGE_HEIGHT 64

class OpticalFlow
{
protected:
	int image_width = DEFAULT_IMAGE_WIDTH;
	int image_height = DEFAULT_IMAGE_HEIGHT;
	float focal_length_x = 0.f; // [pixel]
	float focal_length_y = 0.f;
	int output_rate = DEFAULT_OUTPUT_RATE;
	float sum_flow_x = 0.f;
	float sum_flow_y = 0.f;
	int sum_flow_quality = 0;
	int valid_frame_count = 0;
	// The reference's setters only assign (include/optical_flow.hpp:74-78) and every object keeps its state.
	// Here a change of GEOMETRY (width, height, feature count) rebuilds the device object -- the reference
	// would index its old buffers with the new sizes -- while focal length, output rate and the confidence
	// multiplier are forwarded to the live handle (aof_*_set_params): tracker, limiter sums and the time of
	// the last publish survive, as they do in the reference.
	bool aof_dirty = true;   // geometry changed since the device object was built
	bool aof_retune = false; // focal length / rate / confidence changed since they were last forwarded

	// no exceptions cross this API (the reference has none): failures are reported on stderr
	// and calcFlow answers -1 ("do not publish", src/optical_flow.cpp:92)
	static int aof_fail(const char *what, int rc)
	{
		std::fprintf(stderr, "aof_b200: %s failed (%d): %s\n", what, rc, aof_last_error());
		return -1;
	}

public:
	virtual ~OpticalFlow() {}

	inline void setImageWidth(int img_width) { image_width = img_width; aof_dirty = true; }
	inline void setImageHeight(int img_height) { image_height = img_height; aof_dirty = true; }
	inline void setFocalLengthX(float f_lengh) { focal_length_x = f_lengh; aof_retune = true; }
	inline void setFocalLengthY(float f_lengh) { focal_length_y = f_lengh; aof_retune = true; }
	inline void setOutputRate(int out_rate) { output_rate = out_rate; aof_retune = true; }

	inline int getImageWidth() { return image_width; }
	inline int getImageHeight() { return image_height; }
	inline int getFocalLengthX() { return focal_length_x; }
	inline int getFocalLengthy() { return focal_length_y; } // spelling as in the reference (:83)
	inline int getOutputRate() { return output_rate; }

	virtual int calcFlow(uint8_t *img_current, const uint32_t &img_time_us, int &dt_us, float &flow_x, float &flow_y) = 0;
};
