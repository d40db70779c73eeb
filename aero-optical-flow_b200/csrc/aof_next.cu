// The steps either side of calcFlow that SURVEY.md 8(f) ranks "next": the centre crop + staging before it
// (src/mainloop.cpp:294-298,317-320, src/camera.cpp:97-115), the auto-exposure statistics taken from the same
// cropped frame (src/mainloop.cpp:197-220) and the OPTICAL_FLOW_RAD wire format after it (src/mainloop.cpp:359-371,
// src/mavlink_tcp.cpp:142-149).  Paths relative to /root/reference/.
#include <cstring>

#include "aof_common.cuh"
#include "aof_internal.h"
#include "aof_host.h"

using namespace aof;

namespace aof {

// ---- (f3) 10-bin histogram of the central 128 x 128 + mean sample value, one CTA per frame -----------------
// cv::calcHist(uniform, range [0, 255), 10 bins): bin = floor(v * 10 / 255) for v < 255, 255 itself is out of range.
// Thread = one word (4 pixels) per row step; per-warp private bins in shared memory, then one sum.
constexpr int AE_SIZE = 128, AE_BINS = 10, AE_THREADS = 256;

__global__ void __launch_bounds__(AE_THREADS)
ae_stats_kernel(const uint8_t *__restrict__ frames, int w, int h, aof_ae_stats *__restrict__ out)
{
	__shared__ uint32_t s_hist[AE_THREADS / 32][AE_BINS + 1]; // bin 10 collects the value 255
	const int tid = threadIdx.x, warp = tid >> 5;
	if (tid < (AE_THREADS / 32) * (AE_BINS + 1)) (&s_hist[0][0])[tid] = 0;
	__syncthreads();
	const uint8_t *img = frames + (size_t)blockIdx.x * w * h;
	const int x0 = w / 2 - AE_SIZE / 2, y0 = h / 2 - AE_SIZE / 2;
	for (int i = tid; i < AE_SIZE * AE_SIZE; i += AE_THREADS) {
		const int r = i >> 7, c = i & 127;
		const uint32_t v = img[(size_t)(y0 + r) * w + x0 + c];
		atomicAdd(&s_hist[warp][(v * 10u) / 255u], 1u); // 255 -> bin 10
	}
	__syncthreads();
	if (tid < AE_BINS) {
		uint32_t t = 0;
		for (int k = 0; k < AE_THREADS / 32; k++) t += s_hist[k][tid];
		s_hist[0][tid] = t;
		out[blockIdx.x].hist[tid] = t;
	}
	__syncthreads();
	if (tid == 0) {
		float msv = 0.0f;
		for (int i = 0; i < AE_BINS; i++) msv += (float)(i + 1) * (float)s_hist[0][i] / 16384.0f; // src/mainloop.cpp:217-220
		out[blockIdx.x].msv = msv;
		out[blockIdx.x].pad = 0;
	}
}

} // namespace aof

extern "C" {

int aof_host_register(void *ptr, size_t bytes)
{
	if (!ptr || !bytes) { set_error("aof_host_register: bad argument"); return AOF_ERR_INVALID; }
	AOF_CUDA(cudaHostRegister(ptr, bytes, cudaHostRegisterDefault));
	return AOF_OK;
}

int aof_host_unregister(void *ptr)
{
	if (!ptr) { set_error("aof_host_unregister: bad argument"); return AOF_ERR_INVALID; }
	AOF_CUDA(cudaHostUnregister(ptr));
	return AOF_OK;
}

int aof_crop_to_device(const uint8_t *const *frames, int n, int cam_w, int cam_h, int w, int h, uint8_t *d_batch,
		       void *cuda_stream)
{
	if (!frames || !d_batch || n < 0 || w < 1 || h < 1 || w > cam_w || h > cam_h) {
		set_error("aof_crop_to_device: bad argument (crop %dx%d of %dx%d)", w, h, cam_w, cam_h);
		return AOF_ERR_INVALID;
	}
	// cv::Rect(cam_w / 2 - w / 2, cam_h / 2 - h / 2, w, h)   (src/mainloop.cpp:294-297)
	const size_t x0 = (size_t)(cam_w / 2 - w / 2), y0 = (size_t)(cam_h / 2 - h / 2);
	cudaStream_t st = (cudaStream_t)cuda_stream;
	for (int i = 0; i < n; i++) {
		if (!frames[i]) { set_error("aof_crop_to_device: frame %d is null", i); return AOF_ERR_INVALID; }
		AOF_CUDA(cudaMemcpy2DAsync(d_batch + (size_t)i * w * h, (size_t)w, frames[i] + y0 * cam_w + x0, (size_t)cam_w, (size_t)w,
					   (size_t)h, cudaMemcpyHostToDevice, st));
	}
	return AOF_OK;
}

int aof_ae_stats_device(const uint8_t *d_frames, int n, int w, int h, aof_ae_stats *d_out, void *cuda_stream)
{
	if (!d_frames || !d_out || n < 0) { set_error("aof_ae_stats_device: bad argument"); return AOF_ERR_INVALID; }
	if (w < AE_SIZE || h < AE_SIZE) { set_error("auto-exposure window is 128x128: frame %dx%d is too small", w, h); return AOF_ERR_UNSUPPORTED; }
	if (n == 0) return AOF_OK;
	ae_stats_kernel<<<n, AE_THREADS, 0, (cudaStream_t)cuda_stream>>>(d_frames, w, h, d_out);
	count_launch();
	AOF_CUDA(cudaGetLastError());
	return AOF_OK;
}

int aof_ae_stats_host(const uint8_t *frames, int n, int w, int h, int device, aof_ae_stats *out)
{
	if (!frames || !out || n < 0) { set_error("aof_ae_stats_host: bad argument"); return AOF_ERR_INVALID; }
	if (w < AE_SIZE || h < AE_SIZE) { set_error("auto-exposure window is 128x128: frame %dx%d is too small", w, h); return AOF_ERR_UNSUPPORTED; }
	if (n == 0) return AOF_OK;
	int rc = select_device(device);
	if (rc != AOF_OK) return rc;
	uint8_t *d_f = nullptr;
	aof_ae_stats *d_o = nullptr;
	cudaError_t e;
	rc = AOF_OK;
	if ((e = cudaMalloc(&d_f, (size_t)n * w * h)) != cudaSuccess || (e = cudaMalloc(&d_o, sizeof(aof_ae_stats) * n)) != cudaSuccess ||
	    (e = cudaMemcpy(d_f, frames, (size_t)n * w * h, cudaMemcpyHostToDevice)) != cudaSuccess) {
		set_error("aof_ae_stats_host: %s", cudaGetErrorString(e));
		rc = AOF_ERR_CUDA;
	}
	if (rc == AOF_OK) rc = aof_ae_stats_device(d_f, n, w, h, d_o, nullptr);
	if (rc == AOF_OK && (e = cudaMemcpy(out, d_o, sizeof(aof_ae_stats) * n, cudaMemcpyDeviceToHost)) != cudaSuccess) {
		set_error("aof_ae_stats_host: %s", cudaGetErrorString(e));
		rc = AOF_ERR_CUDA;
	}
	cudaFree(d_f); cudaFree(d_o);
	return rc;
}

} // extern "C"

// ---- (f4) OPTICAL_FLOW_RAD as a MAVLink 2 frame (host side) --------------------------------------------------
namespace {

// X.25 / MCRF4XX accumulate, as modules/mavlink_c/checksum.h
inline void crc_acc(uint8_t data, uint16_t &crc)
{
	uint8_t tmp = data ^ (uint8_t)(crc & 0xff);
	tmp ^= (uint8_t)(tmp << 4);
	crc = (uint16_t)((crc >> 8) ^ ((uint16_t)tmp << 8) ^ ((uint16_t)tmp << 3) ^ (tmp >> 4));
}

template <typename T> inline void put(uint8_t *&p, T v) { memcpy(p, &v, sizeof(T)); p += sizeof(T); } // little-endian host

} // namespace

extern "C" int aof_pack_optical_flow_rad(const aof_flow_out *res, int n, uint64_t time_offset_us, const uint32_t *img_time_us,
					 const double *gyro_xyz, uint8_t system_id, uint8_t component_id, uint8_t *seq,
					 uint8_t *buf, int buf_cap, int *frame_off)
{
	if (!res || !img_time_us || !gyro_xyz || !seq || !buf || n < 0) { set_error("aof_pack_optical_flow_rad: bad argument"); return AOF_ERR_INVALID; }
	const int MSG_ID = 106, PAYLOAD = 44, CRC_EXTRA = 138, HDR = 10;
	int used = 0;
	for (int i = 0; i < n; i++) {
		if (frame_off) frame_off[i] = -1;
		if (res[i].quality < 0) continue; // not published (src/mainloop.cpp:328)
		uint8_t pl[PAYLOAD];
		uint8_t *p = pl;
		put<uint64_t>(p, time_offset_us + img_time_us[i]);
		put<uint32_t>(p, (uint32_t)res[i].dt_us);
		put<float>(p, res[i].flow_x);
		put<float>(p, res[i].flow_y);
		put<float>(p, (float)-gyro_xyz[3 * i + 1]); // switch to match pixel directions (src/mainloop.cpp:363-364)
		put<float>(p, (float)gyro_xyz[3 * i]);
		put<float>(p, (float)gyro_xyz[3 * i + 2]);
		put<uint32_t>(p, 0u);                        // time_delta_distance_us
		put<float>(p, -1.0f);                        // distance unknown
		put<int16_t>(p, 0);                          // temperature
		put<uint8_t>(p, 0);                          // sensor_id
		put<uint8_t>(p, (uint8_t)res[i].quality);
		int len = PAYLOAD;                           // MAVLink 2 trims trailing zero bytes, keeping at least one
		while (len > 1 && pl[len - 1] == 0) len--;
		if (used + HDR + len + 2 > buf_cap) { set_error("aof_pack_optical_flow_rad: buffer too small"); return AOF_ERR_INVALID; }
		uint8_t *f = buf + used;
		f[0] = 0xFD; f[1] = (uint8_t)len; f[2] = 0; f[3] = 0; f[4] = (*seq)++; f[5] = system_id; f[6] = component_id;
		f[7] = (uint8_t)(MSG_ID & 0xFF); f[8] = (uint8_t)((MSG_ID >> 8) & 0xFF); f[9] = (uint8_t)((MSG_ID >> 16) & 0xFF);
		memcpy(f + HDR, pl, (size_t)len);
		uint16_t crc = 0xFFFF;
		for (int k = 1; k < HDR + len; k++) crc_acc(f[k], crc);
		crc_acc((uint8_t)CRC_EXTRA, crc);
		f[HDR + len] = (uint8_t)(crc & 0xFF);
		f[HDR + len + 1] = (uint8_t)(crc >> 8);
		if (frame_off) frame_off[i] = used;
		used += HDR + len + 2;
	}
	return used;
}
