// Drop-in replacement for modules/OpticalFlow/include/flow_opencv.hpp:56-89 (class OpticalFlowOpenCV)
// over aof_lk_* of libaof_b200.so.  Needs no OpenCV header (the reference's needs <cv.h>).
// One object = one stream on CUDA device AOF_DEVICE (env, default 0).
#pragma once
#include <cstdlib>

#include "optical_flow.hpp"

#define DEFAULT_NUMBER_OF_FEATURES 20
#define DEFAULT_CONFIDENCE_MULTIPLIER 1.645f // 90% confidence interval

class OpticalFlowOpenCV : public OpticalFlow
{
private:
	aof_lk *handle = nullptr;
	int num_features;
	float confidence_multiplier;
	bool set_camera_matrix = false, set_camera_distortion = false, cam_dirty = false;
	float cam[4] = {0, 0, 0, 0}, dist[5] = {0, 0, 0, 0, 0};

	int build()
	{
		if (handle) { aof_lk_destroy(handle); handle = nullptr; }
		aof_lk_params p;
		p.focal_length_x = focal_length_x; p.focal_length_y = focal_length_y; p.output_rate = output_rate;
		p.image_width = image_width; p.image_height = image_height; p.num_features = num_features;
		p.confidence_multiplier = confidence_multiplier;
		const char *dev = std::getenv("AOF_DEVICE");
		int rc = aof_lk_create(&p, 1, dev ? std::atoi(dev) : 0, &handle);
		if (rc == AOF_OK && set_camera_matrix) rc = aof_lk_set_camera_matrix(handle, cam[0], cam[1], cam[2], cam[3]);
		if (rc == AOF_OK && set_camera_distortion) rc = aof_lk_set_camera_distortion(handle, dist[0], dist[1], dist[2], dist[3], dist[4]);
		aof_dirty = false;
		aof_retune = false;
		cam_dirty = false;
		return rc;
	}

public:
	inline void setNumFeatures(int n_feat) { num_features = n_feat; aof_dirty = true; }
	inline void setConfMultiplier(float conf_multi) { confidence_multiplier = conf_multi; aof_retune = true; }
	// src/flow_opencv.cpp:72-87
	void setCameraMatrix(float focal_len_x, float focal_len_y, float principal_point_x, float principal_point_y)
	{
		cam[0] = focal_len_x; cam[1] = focal_len_y; cam[2] = principal_point_x; cam[3] = principal_point_y;
		set_camera_matrix = true; cam_dirty = true;
	}
	void setCameraDistortion(float k1, float k2, float k3, float p1 = 0.0f, float p2 = 0.0f)
	{
		dist[0] = k1; dist[1] = k2; dist[2] = k3; dist[3] = p1; dist[4] = p2;
		set_camera_distortion = true; cam_dirty = true;
	}
	inline int getNumFeatures() { return num_features; }
	inline int getConfMultiplier() { return confidence_multiplier; }

	OpticalFlowOpenCV(float f_length_x, float f_length_y, int output_rate = DEFAULT_OUTPUT_RATE,
			  int img_width = DEFAULT_IMAGE_WIDTH, int img_height = DEFAULT_IMAGE_HEIGHT,
			  int num_feat = DEFAULT_NUMBER_OF_FEATURES, float conf_multi = DEFAULT_CONFIDENCE_MULTIPLIER)
		: num_features(num_feat), confidence_multiplier(conf_multi)
	{
		setImageWidth(img_width); setImageHeight(img_height);
		setFocalLengthX(f_length_x); setFocalLengthY(f_length_y);
		setOutputRate(output_rate);
	}
	~OpticalFlowOpenCV() { aof_lk_destroy(handle); }
	OpticalFlowOpenCV(const OpticalFlowOpenCV &) = delete;
	OpticalFlowOpenCV &operator=(const OpticalFlowOpenCV &) = delete;

	// src/flow_opencv.cpp:89-210
	int calcFlow(uint8_t *img_current, const uint32_t &img_time_us, int &dt_us, float &flow_x, float &flow_y)
	{
		if (aof_dirty || !handle) {
			const int rc = build();
			if (rc != AOF_OK) return aof_fail("aof_lk_create", rc);
		}
		if (aof_retune) {
			aof_lk_set_params(handle, focal_length_x, focal_length_y, output_rate, confidence_multiplier);
			aof_retune = false;
		}
		if (cam_dirty) { // the camera model is forwarded to the live handle too (src/flow_opencv.cpp:72-87 only assign)
			if (set_camera_matrix) aof_lk_set_camera_matrix(handle, cam[0], cam[1], cam[2], cam[3]);
			if (set_camera_distortion) aof_lk_set_camera_distortion(handle, dist[0], dist[1], dist[2], dist[3], dist[4]);
			cam_dirty = false;
		}
		const int q = aof_lk_calc_flow(handle, img_current, img_time_us, &dt_us, &flow_x, &flow_y);
		return q <= AOF_ERR_BASE ? aof_fail("aof_lk_calc_flow", q) : q;
	}
};
