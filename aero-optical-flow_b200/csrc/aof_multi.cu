// Multi-GPU host side of both paths (include/aof.h, aof_*_multi_*): SURVEY.md section 8(e).
//
// The path shards by whole streams: every stream's state (previous frame or pyramid, feature slots, limiter sums)
// is private, nothing is reduced across streams (src/flow_px4.cpp:69-98, src/flow_opencv.cpp:89-210).  So a
// multi-GPU handle is n_devices single-device handles, device d owning the contiguous stream range
// [first_d, first_d + count_d), plus ONE host worker thread per device.  A batch call hands every worker its
// slice of the caller's host frames; the worker runs the single-device batch entry point on its own device
// (pinned staging, chunked async H2D on a copy stream under the kernels of the previous chunk, 16 B / stream
// D2H) with its slice of ONE pinned result array as the D2H target: that array is the "host gather".
// No NCCL, no peer access, no collective: NVLink is unused by design.
//
// Worker threads are pinned to the CPUs local to their GPU when the kernel exposes them
// (/sys/bus/pci/devices/<bdf>/local_cpulist), so that the staging memcpy of pageable frames and the pinned
// buffers they first touch stay on the GPU's NUMA node.
#include <atomic>
#include <condition_variable>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <new>
#include <string>
#include <thread>
#include <vector>

#include <sched.h>

#include "aof_internal.h"
#include "aof_host.h"

using namespace aof;

namespace {

// The multi entry points touch every device from the CALLER's thread (create, destroy, state read-back): the caller's
// current device is put back when they return, so that a caller (torch, for one) that tracks the current device itself
// does not find it changed behind its back.
struct DeviceGuard {
	int dev = -1;
	DeviceGuard() { if (cudaGetDevice(&dev) != cudaSuccess) { dev = -1; cudaGetLastError(); } }
	~DeviceGuard() { if (dev >= 0) cudaSetDevice(dev); }
};

// one job = one batch call on one device
struct Job {
	const uint8_t *imgs = nullptr;
	const uint32_t *t_us = nullptr;
	aof_flow_out *out = nullptr;
	aof_lk_raw *raw = nullptr;
};

struct Worker {
	int device = 0, first = 0, count = 0;
	aof_px4 *px4 = nullptr;
	aof_lk *lk = nullptr;
	std::thread th;
	std::mutex mu;
	std::condition_variable cv;
	Job job;
	int state = 0;   // 0 idle, 1 job posted, 2 done, 3 quit
	int rc = AOF_OK;
	char err[256] = "";
};

// CPUs local to the GPU (same PCIe root / NUMA node), as the kernel reports them; false if unknown
bool pin_to_gpu_cpus(int device)
{
	char bdf[32] = "";
	if (cudaDeviceGetPCIBusId(bdf, sizeof bdf, device) != cudaSuccess) { cudaGetLastError(); return false; }
	for (char *c = bdf; *c; c++) *c = (char)tolower(*c);
	const std::string path = std::string("/sys/bus/pci/devices/") + bdf + "/local_cpulist";
	FILE *f = fopen(path.c_str(), "r");
	if (!f) return false;
	char line[1024] = "";
	const bool got = fgets(line, sizeof line, f) != nullptr;
	fclose(f);
	if (!got) return false;
	cpu_set_t set;
	CPU_ZERO(&set);
	int n = 0;
	for (char *p = line; *p && *p != '\n';) { // "0-31,64-95"
		char *end;
		long a = strtol(p, &end, 10), b = a;
		if (end == p) break;
		if (*end == '-') b = strtol(end + 1, &end, 10);
		for (long c = a; c <= b && c < CPU_SETSIZE; c++) { CPU_SET((int)c, &set); n++; }
		p = *end == ',' ? end + 1 : end;
	}
	return n > 0 && sched_setaffinity(0, sizeof set, &set) == 0;
}

void worker_main(Worker *w)
{
	pin_to_gpu_cpus(w->device);
	cudaSetDevice(w->device);
	for (;;) {
		Job j;
		{
			std::unique_lock<std::mutex> lk(w->mu);
			w->cv.wait(lk, [&] { return w->state == 1 || w->state == 3; });
			if (w->state == 3) return;
			j = w->job;
		}
		int rc;
		if (w->px4) rc = aof_px4_calc_flow_batch(w->px4, j.imgs, j.t_us, j.out);
		else rc = aof_lk_calc_flow_batch(w->lk, j.imgs, j.t_us, j.out, j.raw);
		{
			std::lock_guard<std::mutex> lk(w->mu);
			w->rc = rc;
			if (rc != AOF_OK) snprintf(w->err, sizeof w->err, "device %d: %s", w->device, last_error()); // thread-local text
			w->state = 2;
		}
		w->cv.notify_all();
	}
}

} // namespace

struct aof_multi {
	int n_total = 0, n_devices = 0;
	size_t frame_bytes = 0;
	bool is_lk = false;
	std::vector<Worker *> workers;
	aof_flow_out *h_out = nullptr;   // ONE pinned array, n_total entries: every device's D2H lands in its slice
	aof_lk_raw *h_raw = nullptr;
};

static void multi_destroy(aof_multi *m)
{
	if (!m) return;
	DeviceGuard guard;
	for (Worker *w : m->workers) {
		if (w->th.joinable()) {
			{ std::lock_guard<std::mutex> lk(w->mu); w->state = 3; }
			w->cv.notify_all();
			w->th.join();
		}
		if (w->px4) aof_px4_destroy(w->px4);
		if (w->lk) aof_lk_destroy(w->lk);
		delete w;
	}
	cudaFreeHost(m->h_out);
	cudaFreeHost(m->h_raw);
	delete m;
}

// contiguous whole-stream ranges, the first (n_total % n_devices) devices own one stream more
static int multi_create(const aof_px4_params *pp, const aof_lk_params *lp, int n_total, const int *devices, int n_devices,
			aof_multi **out)
{
	if ((!pp && !lp) || !out || !devices || n_devices < 1 || n_total < n_devices) {
		set_error("create_multi: bad argument (need n_streams_total >= n_devices >= 1)");
		return AOF_ERR_INVALID;
	}
	*out = nullptr;
	DeviceGuard guard;
	for (int i = 0; i < n_devices; i++)
		for (int j = 0; j < i; j++)
			if (devices[i] == devices[j]) { set_error("create_multi: device %d listed twice", devices[i]); return AOF_ERR_INVALID; }
	aof_multi *m = new (std::nothrow) aof_multi();
	if (!m) return AOF_ERR_INVALID;
	m->n_total = n_total; m->n_devices = n_devices; m->is_lk = lp != nullptr;
	m->frame_bytes = pp ? (size_t)pp->image_width * pp->image_height : (size_t)lp->image_width * lp->image_height;
	const int base = n_total / n_devices, extra = n_total % n_devices;
	int first = 0, rc = AOF_OK;
	for (int d = 0; d < n_devices && rc == AOF_OK; d++) {
		Worker *w = new (std::nothrow) Worker();
		if (!w) { rc = AOF_ERR_INVALID; break; }
		m->workers.push_back(w);
		w->device = devices[d]; w->first = first; w->count = base + (d < extra ? 1 : 0);
		first += w->count;
		rc = pp ? aof_px4_create(pp, w->count, w->device, &w->px4) : aof_lk_create(lp, w->count, w->device, &w->lk);
	}
	if (rc == AOF_OK) {
		// portable: page-locked for every device's context, so each worker's D2H can target its slice directly
		if (cudaHostAlloc(&m->h_out, sizeof(aof_flow_out) * n_total, cudaHostAllocPortable) != cudaSuccess ||
		    (m->is_lk && cudaHostAlloc(&m->h_raw, sizeof(aof_lk_raw) * n_total, cudaHostAllocPortable) != cudaSuccess)) {
			set_error("create_multi: pinned result array: %s", cudaGetErrorString(cudaGetLastError()));
			rc = AOF_ERR_CUDA;
		}
	}
	if (rc != AOF_OK) { multi_destroy(m); return rc; }
	for (Worker *w : m->workers) w->th = std::thread(worker_main, w);
	*out = m;
	return AOF_OK;
}

static int multi_run(aof_multi *m, const uint8_t *imgs, const uint32_t *t_us, aof_flow_out *out, aof_lk_raw *raw)
{
	if (!m || !imgs || !t_us) { set_error("multi calc_flow_batch: null argument"); return AOF_ERR_INVALID; }
	for (Worker *w : m->workers) {
		{
			std::lock_guard<std::mutex> lk(w->mu);
			w->job.imgs = imgs + (size_t)w->first * m->frame_bytes;
			w->job.t_us = t_us + w->first;
			w->job.out = m->h_out + w->first;
			w->job.raw = (m->is_lk && raw) ? m->h_raw + w->first : nullptr;
			w->state = 1;
		}
		w->cv.notify_all();
	}
	int rc = AOF_OK;
	for (Worker *w : m->workers) {
		std::unique_lock<std::mutex> lk(w->mu);
		w->cv.wait(lk, [&] { return w->state == 2; });
		w->state = 0;
		if (w->rc != AOF_OK && rc == AOF_OK) { rc = w->rc; set_error("%s", w->err); }
	}
	if (rc != AOF_OK) return rc;
	// the gather is complete when the last worker returns: every slice of the pinned array has been written by its
	// device's DMA.  A caller that passes its own array gets a copy; NULL = read aof_*_multi_results().
	if (out && out != m->h_out) memcpy(out, m->h_out, sizeof(aof_flow_out) * m->n_total);
	if (raw && m->is_lk && raw != m->h_raw) memcpy(raw, m->h_raw, sizeof(aof_lk_raw) * m->n_total);
	return AOF_OK;
}

static int multi_range(const aof_multi *m, int i, int *device, int *first, int *count)
{
	if (!m || i < 0 || i >= m->n_devices) return AOF_ERR_INVALID;
	if (device) *device = m->workers[i]->device;
	if (first) *first = m->workers[i]->first;
	if (count) *count = m->workers[i]->count;
	return AOF_OK;
}

extern "C" {

int aof_px4_create_multi(const aof_px4_params *p, int n_streams_total, const int *devices, int n_devices, aof_px4_multi **out)
{
	return multi_create(p, nullptr, n_streams_total, devices, n_devices, reinterpret_cast<aof_multi **>(out));
}
void aof_px4_destroy_multi(aof_px4_multi *m) { multi_destroy(reinterpret_cast<aof_multi *>(m)); }
int aof_px4_multi_calc_flow_batch(aof_px4_multi *m, const uint8_t *imgs, const uint32_t *t_us, aof_flow_out *out)
{
	return multi_run(reinterpret_cast<aof_multi *>(m), imgs, t_us, out, nullptr);
}
const aof_flow_out *aof_px4_multi_results(const aof_px4_multi *m) { return m ? reinterpret_cast<const aof_multi *>(m)->h_out : nullptr; }
int aof_px4_multi_range(const aof_px4_multi *m, int i, int *device, int *first, int *count)
{
	return multi_range(reinterpret_cast<const aof_multi *>(m), i, device, first, count);
}
int aof_px4_multi_set_params(aof_px4_multi *m, float fx, float fy, int rate)
{
	if (!m) return AOF_ERR_INVALID;
	for (Worker *w : reinterpret_cast<aof_multi *>(m)->workers) aof_px4_set_params(w->px4, fx, fy, rate);
	return AOF_OK;
}

int aof_lk_create_multi(const aof_lk_params *p, int n_streams_total, const int *devices, int n_devices, aof_lk_multi **out)
{
	return multi_create(nullptr, p, n_streams_total, devices, n_devices, reinterpret_cast<aof_multi **>(out));
}
void aof_lk_destroy_multi(aof_lk_multi *m) { multi_destroy(reinterpret_cast<aof_multi *>(m)); }
int aof_lk_multi_calc_flow_batch(aof_lk_multi *m, const uint8_t *imgs, const uint32_t *t_us, aof_flow_out *out, aof_lk_raw *raw)
{
	// raw != NULL asks for the diagnostics as well; pass aof_lk_multi_raw() itself to avoid the copy
	aof_multi *mm = reinterpret_cast<aof_multi *>(m);
	return multi_run(mm, imgs, t_us, out, raw);
}
const aof_flow_out *aof_lk_multi_results(const aof_lk_multi *m) { return m ? reinterpret_cast<const aof_multi *>(m)->h_out : nullptr; }
const aof_lk_raw *aof_lk_multi_raw(const aof_lk_multi *m) { return m ? reinterpret_cast<const aof_multi *>(m)->h_raw : nullptr; }
int aof_lk_multi_range(const aof_lk_multi *m, int i, int *device, int *first, int *count)
{
	return multi_range(reinterpret_cast<const aof_multi *>(m), i, device, first, count);
}
int aof_lk_multi_set_params(aof_lk_multi *m, float fx, float fy, int rate, float conf)
{
	if (!m) return AOF_ERR_INVALID;
	for (Worker *w : reinterpret_cast<aof_multi *>(m)->workers) aof_lk_set_params(w->lk, fx, fy, rate, conf);
	return AOF_OK;
}
int aof_lk_multi_get_state(aof_lk_multi *m, int32_t *update_vec, float *feat_cur, float *prev_corners, uint8_t *prev_status)
{
	if (!m) return AOF_ERR_INVALID;
	DeviceGuard guard;
	aof_multi *mm = reinterpret_cast<aof_multi *>(m);
	for (Worker *w : mm->workers) {
		const int N = aof_lk_num_features(w->lk);
		const size_t o = (size_t)w->first * N;
		const int rc = aof_lk_get_state(w->lk, update_vec ? update_vec + o : nullptr, feat_cur ? feat_cur + 2 * o : nullptr,
						prev_corners ? prev_corners + 2 * o : nullptr, prev_status ? prev_status + o : nullptr);
		if (rc != AOF_OK) return rc;
	}
	return AOF_OK;
}

} // extern "C"
