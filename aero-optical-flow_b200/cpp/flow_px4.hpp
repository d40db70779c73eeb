// Drop-in replacement for modules/OpticalFlow/include/flow_px4.hpp:51-71 (class OpticalFlowPX4)
// over aof_px4_* of libaof_b200.so.  One object = one stream on CUDA device AOF_DEVICE (env, default 0).
#pragma once
#include <cstdlib>

#include "optical_flow.hpp"

#define DEFAULT_SEARCH_SIZE 6
#define DEFAULT_FLOW_FEATURE_THRESHOLD 30
#define DEFAULT_FLOW_VALUE_THRESHOLD 3000

class OpticalFlowPX4 : public OpticalFlow
{
private:
	aof_px4 *handle = nullptr;
	int search_size, flow_feature_threshold, flow_value_threshold;

	int build()
	{
		if (handle) { aof_px4_destroy(handle); handle = nullptr; }
		aof_px4_params p;
		p.focal_length_x = focal_length_x; p.focal_length_y = focal_length_y; p.output_rate = output_rate;
		p.image_width = image_width; p.image_height = image_height; p.search_size = search_size;
		p.flow_feature_threshold = flow_feature_threshold; p.flow_value_threshold = flow_value_threshold;
		const char *dev = std::getenv("AOF_DEVICE");
		const int rc = aof_px4_create(&p, 1, dev ? std::atoi(dev) : 0, &handle);
		aof_dirty = false;
		aof_retune = false;
		return rc;
	}

public:
	OpticalFlowPX4(float f_length_x, float f_length_y, int ouput_rate = DEFAULT_OUTPUT_RATE,
		       int img_width = DEFAULT_IMAGE_WIDTH, int img_height = DEFAULT_IMAGE_HEIGHT,
		       int search_size = DEFAULT_SEARCH_SIZE, int flow_feature_threshold = DEFAULT_FLOW_FEATURE_THRESHOLD,
		       int flow_value_threshold = DEFAULT_FLOW_VALUE_THRESHOLD)
		: search_size(search_size), flow_feature_threshold(flow_feature_threshold), flow_value_threshold(flow_value_threshold)
	{
		setImageWidth(img_width); setImageHeight(img_height);
		setFocalLengthX(f_length_x); setFocalLengthY(f_length_y);
		setOutputRate(ouput_rate);
	}
	~OpticalFlowPX4() { aof_px4_destroy(handle); }
	OpticalFlowPX4(const OpticalFlowPX4 &) = delete;
	OpticalFlowPX4 &operator=(const OpticalFlowPX4 &) = delete;

	// src/flow_px4.cpp:69-98
	int calcFlow(uint8_t *img_current, const uint32_t &img_time_us, int &dt_us, float &flow_x, float &flow_y)
	{
		if (aof_dirty || !handle) {
			const int rc = build();
			if (rc != AOF_OK) return aof_fail("aof_px4_create", rc);
		}
		if (aof_retune) {
			aof_px4_set_params(handle, focal_length_x, focal_length_y, output_rate);
			aof_retune = false;
		}
		const int q = aof_px4_calc_flow(handle, img_current, img_time_us, &dt_us, &flow_x, &flow_y);
		return q <= AOF_ERR_BASE ? aof_fail("aof_px4_calc_flow", q) : q;
	}
};
