// aof_px4_calc_flow_pairs_static: OpticalFlowPX4::calcFlow of n independent (previous, current) frame pairs WITHOUT any
// allocation -- no handle, no cudaMalloc / cudaFree, nothing from an allocating library.  For callers that cannot
// allocate: code under CUDA graph capture, tight serving loops, harnesses that forbid it.  No counterpart in the
// reference.
//
// The handle-based path keeps two per-stream scratch arrays between its kernels (aof_px4.cu: d_raw, the result of the
// SAD kernel that the epilogue kernel turns into flow angles, and d_lim, the rate limiter's sums).  Here both live in ONE
// file-scope __device__ array of fixed size: AOF_STATIC_MAX_PAIRS pairs, (16 + 24) B each = 320 KiB, held by every GPU
// that loads the library until the process exits.
//
//   * In host code the array's name is only a stand-in for the symbol, not a device address.  Its address is taken with
//     cudaGetSymbolAddress, once per device with that device current, and THAT pointer goes to the kernels.
//   * Every call uses the same array and finds the previous call's contents in it: the limiter part, which must start at
//     zero (a freshly constructed object, src/optical_flow.cpp:55-60), is cleared on the caller's stream before the
//     kernels; the raw part is fully overwritten by the SAD kernel.
//   * Calls that overlap on one GPU (two streams, two threads) race on the array: the caller serialises them.
//
// The arithmetic is the library's own: sad_flow_launch + px4_epilogue_launch, the launches aof_px4_calc_flow_device makes.
#include <mutex>

#include "aof_internal.h"
#include "aof_host.h"

using namespace aof;

namespace {

struct StaticScratch {
	aof_px4_raw raw[AOF_STATIC_MAX_PAIRS];
	LimiterState lim[AOF_STATIC_MAX_PAIRS];
};

static __device__ __align__(256) unsigned char g_px4_static_scratch[sizeof(StaticScratch)];

constexpr int MAX_DEVICES = 64;
std::mutex g_mu;
StaticScratch *g_scratch_dev[MAX_DEVICES] = {};   // device address of the array on each device, filled on first use

// the array's device address on `device` (current); nullptr + error text on failure
StaticScratch *scratch_of(int device)
{
	std::lock_guard<std::mutex> lk(g_mu);
	if (!g_scratch_dev[device]) {
		void *p = nullptr;
		const cudaError_t e = cudaGetSymbolAddress(&p, g_px4_static_scratch);
		if (e != cudaSuccess || !p) {
			set_error("aof_px4_calc_flow_pairs_static: cudaGetSymbolAddress on device %d: %s", device, cudaGetErrorString(e));
			cudaGetLastError();
			return nullptr;
		}
		g_scratch_dev[device] = static_cast<StaticScratch *>(p);
	}
	return g_scratch_dev[device];
}

} // namespace

extern "C" {

int aof_px4_static_max_pairs(void) { return AOF_STATIC_MAX_PAIRS; }

int aof_px4_calc_flow_pairs_static(const aof_px4_params *p, int device, const uint8_t *d_prev, const uint8_t *d_cur,
				   const uint32_t *d_t_us, int n, aof_flow_out *d_out, void *cuda_stream)
{
	if (!p || !d_prev || !d_cur || !d_t_us || !d_out) { set_error("aof_px4_calc_flow_pairs_static: null argument"); return AOF_ERR_INVALID; }
	if (n < 1 || n > AOF_STATIC_MAX_PAIRS) {
		set_error("aof_px4_calc_flow_pairs_static: n = %d outside 1..%d (the fixed scratch array holds AOF_STATIC_MAX_PAIRS pairs; "
			  "split the batch or use a handle)", n, AOF_STATIC_MAX_PAIRS);
		return AOF_ERR_INVALID;
	}
	if (((uintptr_t)d_prev | (uintptr_t)d_cur) & 15) { // the kernels stage frames with cp.async.bulk
		set_error("aof_px4_calc_flow_pairs_static: d_prev / d_cur must be 16-byte aligned");
		return AOF_ERR_INVALID;
	}
	// the geometry rules of aof_px4_create (aof_px4.cu), restated here so that no existing code path changes
	const int W = p->image_width, H = p->image_height, S = p->search_size;
	if (S < 1 || S > 127) { set_error("search_size %d out of range (1..127)", S); return AOF_ERR_UNSUPPORTED; }
	if (W < 2 * (S + 1) + 8 + 1 || H < W || (long)W * H > 65535 || (W & 3) || ((W * H) & 15)) {
		set_error("image %dx%d unsupported (need W%%4==0, W*H%%16==0, H>=W>=%d, W*H<=65535)", W, H, 2 * (S + 1) + 9);
		return AOF_ERR_UNSUPPORTED;
	}
	if (device < 0 || device >= MAX_DEVICES) { set_error("device %d out of range", device); return AOF_ERR_INVALID; }
	const int rc = select_device(device);
	if (rc != AOF_OK) return rc;
	StaticScratch *sc = scratch_of(device);
	if (!sc) return AOF_ERR_CUDA;

	SadGeom g{};
	g.width = W;
	g.stage_bytes = (uint32_t)((size_t)W * H);
	g.frame_stride = (size_t)W * H;
	const int lo = S + 1, hi = W - (S + 1) - 8;
	g.lo = lo;
	g.step = (hi - lo) / 5 + 1;
	g.nb = 0;
	for (int v = lo; v < hi; v += g.step) g.nb++;
	g.tun_v1 = g.tun_v2 = g.tun_ctas = 0;             // (no environment switches on this entry point)
	g.feat_thr = (uint32_t)p->flow_feature_threshold;
	g.val_thr = (uint32_t)p->flow_value_threshold;

	cudaStream_t st = (cudaStream_t)cuda_stream;
	// the limiter of a fresh object starts at zero; the array still holds the previous call's sums
	AOF_CUDA(cudaMemsetAsync(sc->lim, 0, sizeof(LimiterState) * (size_t)n, st));
	AOF_CUDA(sad_flow_launch(S, d_prev, d_cur, g, n, sc->raw, nullptr, st));
	AOF_CUDA(px4_epilogue_launch(sc->raw, d_t_us, sc->lim, d_out, n, p->output_rate, p->focal_length_x, p->focal_length_y, st));
	return AOF_OK;
}

} // extern "C"
