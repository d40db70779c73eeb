// Host helpers shared by the C-ABI translation units: error text, device selection, pinned
// staging, launch counting and the optional per-kernel CUDA-event timing (aof_timing_*).
#include <atomic>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <vector>

#include "aof_internal.h"
#include "aof_host.h"

namespace aof {

static std::atomic<uint64_t> g_launches{0};
void count_launch(uint64_t n) { g_launches.fetch_add(n, std::memory_order_relaxed); }

static thread_local char g_err[512] = "";
void set_error(const char *fmt, ...)
{
	va_list ap;
	va_start(ap, fmt);
	vsnprintf(g_err, sizeof g_err, fmt, ap);
	va_end(ap);
}
const char *last_error() { return g_err; }

int env_int(const char *name, int dflt)
{
	const char *v = getenv(name);
	return v && *v ? atoi(v) : dflt;
}

int select_device(int device)
{
	int n = 0;
	cudaError_t e = cudaGetDeviceCount(&n);
	if (e != cudaSuccess) { set_error("cudaGetDeviceCount: %s", cudaGetErrorString(e)); return AOF_ERR_CUDA; }
	if (n == 0) { set_error("no CUDA device (this library has no CPU fallback)"); return AOF_ERR_NO_DEVICE; }
	if (device < 0 || device >= n) { set_error("device %d out of range (%d present)", device, n); return AOF_ERR_INVALID; }
	AOF_CUDA(cudaSetDevice(device));
	return AOF_OK;
}

bool host_pinned(const void *p)
{
	cudaPointerAttributes at{};
	const bool pinned = cudaPointerGetAttributes(&at, p) == cudaSuccess && at.type == cudaMemoryTypeHost;
	cudaGetLastError();
	return pinned;
}

int stage_h2d(void *d_dst, const void *src, size_t bytes, void *stage, cudaStream_t st)
{
	const bool pinned = host_pinned(src);
	if (pinned) {
		AOF_CUDA(cudaMemcpyAsync(d_dst, src, bytes, cudaMemcpyHostToDevice, st));
		return AOF_OK;
	}
	const size_t chunk = 4u << 20;
	for (size_t o = 0; o < bytes; o += chunk) {
		const size_t len = bytes - o < chunk ? bytes - o : chunk;
		memcpy((char *)stage + o, (const char *)src + o, len);
		AOF_CUDA(cudaMemcpyAsync((char *)d_dst + o, (char *)stage + o, len, cudaMemcpyHostToDevice, st));
	}
	return AOF_OK;
}

} // namespace aof

// ---------------------------------------------------------------------------
// Per-kernel timing: when enabled every launch wrapper brackets its kernel with two CUDA
// events on the launching stream; aof_timing_read() synchronises and sums them.  Off by
// default (two extra event records per launch), used by bench.py to attribute a step's
// time to its kernels and by whoever wants a trace without a profiler.
namespace aof {

// slot -> kernel symbol.  The entries below are the defaults; a launch wrapper that picks among several kernels passes
// the symbol of the one it launches to KernelTimer, which records it here, so that the timing table names what ran.
static const char *g_kernel_names[KID_COUNT] = {
	"sad_flow_kernel", "px4_epilogue_kernel", "lk_pyr_down_kernel", "lk_track_update_kernel", "lk_fast_kernel",
	"lk_select_epilogue_kernel", "lk_track_points_kernel", "sad_dense_kernel", "gftt_min_eig_kernel", "gftt_select_kernel",
	"lk_pyr_fused_kernel", "lk_fast_pick_kernel", "lk_fast_band_flagged_kernel", "lk_fast_plan_kernel",
};

struct TimingState {
	std::mutex mu;
	bool on = false;
	struct Pair { cudaEvent_t a, b; int kid; };
	std::vector<Pair> pending, pool;
	double ms[KID_COUNT] = {};
	uint64_t count[KID_COUNT] = {};
};
static TimingState g_t;
static std::atomic<bool> g_timing_on{false};

bool timing_on() { return g_timing_on.load(std::memory_order_relaxed); }

KernelTimer::KernelTimer(int kid, cudaStream_t st, const char *symbol) : kid_(kid), st_(st)
{
	if (!g_timing_on.load(std::memory_order_relaxed)) return;
	std::lock_guard<std::mutex> lk(g_t.mu);
	if (symbol) g_kernel_names[kid] = symbol;
	TimingState::Pair p;
	if (!g_t.pool.empty()) { p = g_t.pool.back(); g_t.pool.pop_back(); }
	else if (cudaEventCreate(&p.a) != cudaSuccess || cudaEventCreate(&p.b) != cudaSuccess) return;
	p.kid = kid;
	a_ = p.a; b_ = p.b;
	cudaEventRecord(p.a, st);
	active_ = true;
}

KernelTimer::~KernelTimer()
{
	if (!active_) return;
	cudaEventRecord((cudaEvent_t)b_, st_);
	std::lock_guard<std::mutex> lk(g_t.mu);
	g_t.pending.push_back({(cudaEvent_t)a_, (cudaEvent_t)b_, kid_});
}

static void timing_drain()
{
	for (auto &p : g_t.pending) {
		float ms = 0.f;
		if (cudaEventSynchronize(p.b) == cudaSuccess && cudaEventElapsedTime(&ms, p.a, p.b) == cudaSuccess) {
			g_t.ms[p.kid] += ms;
			g_t.count[p.kid]++;
		}
		g_t.pool.push_back(p);
	}
	g_t.pending.clear();
}

} // namespace aof

extern "C" {

const char *aof_version(void) { return "aof_b200 0.2 (sm_100a)"; }
const char *aof_last_error(void) { return aof::last_error(); }
uint64_t aof_launch_count(void) { return aof::g_launches.load(); }

int aof_device_count(void)
{
	int n = 0;
	if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
	return n;
}

void aof_timing_enable(int on)
{
	std::lock_guard<std::mutex> lk(aof::g_t.mu);
	aof::timing_drain();
	aof::g_timing_on.store(on != 0);
}

void aof_timing_reset(void)
{
	std::lock_guard<std::mutex> lk(aof::g_t.mu);
	aof::timing_drain();
	for (int k = 0; k < aof::KID_COUNT; k++) { aof::g_t.ms[k] = 0; aof::g_t.count[k] = 0; }
}

int aof_timing_kernels(void) { return aof::KID_COUNT; }

int aof_timing_read(int kernel, const char **name, double *total_ms, uint64_t *launches)
{
	if (kernel < 0 || kernel >= aof::KID_COUNT) return AOF_ERR_INVALID;
	std::lock_guard<std::mutex> lk(aof::g_t.mu);
	aof::timing_drain();
	if (name) *name = aof::g_kernel_names[kernel];
	if (total_ms) *total_ms = aof::g_t.ms[kernel];
	if (launches) *launches = aof::g_t.count[kernel];
	return AOF_OK;
}

} // extern "C"
