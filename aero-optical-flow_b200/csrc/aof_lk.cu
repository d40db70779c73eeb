// C-ABI host side of the LK path (include/aof.h): per-stream device state, pinned
// staging, async copies, launches.  Replaces the host logic of OpticalFlowOpenCV
// (src/flow_opencv.cpp:47-210) and of trackFeatures' file statics
// (external/klt_feature_tracker/src/trackFeatures.cpp:53-56); the arithmetic is in
// lk_kernels.cu.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <vector>

#include "aof_internal.h"
#include "aof_host.h"

using namespace aof;

constexpr int LK_MAX_PARTS = 8;

struct aof_lk {
	aof_lk_params p;
	int n = 0, device = 0;
	LkGeom g{};
	cudaStream_t st = nullptr, st_copy = nullptr;
	cudaStream_t st_hi = nullptr, st_lo = nullptr; // device path pipeline: latency-bound kernels / FAST (see aof_lk_calc_flow_device)
	cudaEvent_t ev_chunk[8] = {};        // host-batch path: chunk c of the frames has landed
	cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
	cudaEvent_t ev_trk[LK_MAX_PARTS] = {}, ev_fast[LK_MAX_PARTS] = {};
	int tun_split = 0;                   // AOF_LK_SPLIT: 1 = never pipeline the parts, 2 = always (read once at create)
	int tun_parts = 0;                   // AOF_LK_PARTS: parts a batch is cut into (0 = by batch size)
	size_t lvl_px[LK_MAX_LEVELS] = {};   // pixels of one frame at level l
	size_t upper_off[LK_MAX_LEVELS] = {}; // offset of the level-l block (l >= 1) inside d_upper[slot]
	size_t upper_bytes = 0;
	uint8_t *d_upper[2] = {nullptr, nullptr};
	const uint8_t *l0[2] = {nullptr, nullptr};
	int cur = 0;
	bool have_prev = false, have_feat_prev = false;
	bool set_cam = false, set_dist = false;
	LkCamera cam{};
	LkState s{};
	// host-batch path (allocated on first use)
	uint8_t *d_frames[2] = {nullptr, nullptr};
	uint8_t *h_stage = nullptr;
	uint32_t *h_t = nullptr, *d_t = nullptr;
	aof_flow_out *h_out = nullptr, *d_out = nullptr;
	aof_lk_raw *h_raw = nullptr, *d_raw = nullptr;
};

// level base pointers of slot `slot`, seen from stream `first` on
static LkPyr pyr_of(const aof_lk *h, int slot, size_t first = 0)
{
	LkPyr p{};
	p.lvl[0] = h->l0[slot] + first * h->lvl_px[0];
	for (int l = 1; l <= h->g.top; l++) p.lvl[l] = h->d_upper[slot] + h->upper_off[l] + first * h->lvl_px[l];
	return p;
}

// the per-stream state arrays seen from stream `first` on
static LkState state_from(const LkState &s, const LkGeom &g, size_t first)
{
	LkState v = s;
	const size_t N = (size_t)g.n_feat;
	v.prev_corners += first * N; v.prev_status += first * N; v.n_prev_corners += first; v.update_vec += first * N;
	v.feat_cur += first * N; v.feat_prev += first * N; v.lim += first; v.bin_needed += first * LK_BINS; v.target += first;
	v.kp_count += first * LK_BINS; v.kp += first * LK_BINS * (size_t)g.kp_cap;
	v.lk_next += first * N; v.lk_stat += first * N; v.ticket += first;
	v.lz_pick += first * LK_BINS * LK_MAX_TPB; v.lz_npick += first * LK_BINS; v.lz_floor += first * LK_BINS;
	v.need_full += first; v.n_full += first; v.full_list += first;
	const size_t nb = (size_t)LK_BINS * g.fb_bands;
	v.plan_units += first * nb * g.fb_ucap; v.plan_hdr += first * nb; v.work_list += first * nb; v.n_work += first;
	return v;
}

// levels 1..top of n frames whose level 0 is at l0; level blocks at upper + off[l]
static cudaError_t build_upper(const LkGeom &g, const uint8_t *l0, uint8_t *upper, const size_t *off, int n, cudaStream_t st)
{
	if (lk_pyr_stream_supported(g, l0)) return lk_pyr_stream_launch(g, l0, upper + off[1], upper + off[2], upper + off[3], n, st);
	if (g.top == 3) return lk_pyr_fused_launch(g, l0, upper + off[1], upper + off[2], upper + off[3], n, st);
	for (int l = 1; l <= g.top; l++) {
		const uint8_t *src = l == 1 ? l0 : upper + off[l - 1];
		cudaError_t e = lk_pyr_down_launch(src, upper + off[l], g.w[l - 1], g.h[l - 1], n, st);
		if (e != cudaSuccess) return e;
	}
	return cudaSuccess;
}

static int reset_state(aof_lk *h)
{
	const size_t nN = (size_t)h->n * h->g.n_feat;
	std::vector<int> twos(nN, 2);
	AOF_CUDA(cudaMemcpy(h->s.update_vec, twos.data(), sizeof(int) * nN, cudaMemcpyHostToDevice));
	AOF_CUDA(cudaMemset(h->s.prev_corners, 0, sizeof(float2) * nN));
	AOF_CUDA(cudaMemset(h->s.prev_status, 0, nN));
	AOF_CUDA(cudaMemset(h->s.n_prev_corners, 0, sizeof(int) * h->n));
	AOF_CUDA(cudaMemset(h->s.feat_cur, 0, sizeof(float2) * nN));
	AOF_CUDA(cudaMemset(h->s.feat_prev, 0, sizeof(float2) * nN));
	AOF_CUDA(cudaMemset(h->s.lim, 0, sizeof(LimiterState) * h->n));
	AOF_CUDA(cudaMemset(h->s.bin_needed, 0, sizeof(int) * LK_BINS * h->n));
	AOF_CUDA(cudaMemset(h->s.target, 0, sizeof(int) * h->n));
	AOF_CUDA(cudaMemset(h->s.kp_count, 0, sizeof(int) * LK_BINS * h->n));
	AOF_CUDA(cudaMemset(h->s.ticket, 0, sizeof(int) * h->n));
	AOF_CUDA(cudaMemset(h->s.lz_npick, 0, sizeof(int) * LK_BINS * h->n));
	AOF_CUDA(cudaMemset(h->s.lz_floor, 0, sizeof(int) * LK_BINS * h->n));
	AOF_CUDA(cudaMemset(h->s.need_full, 0, sizeof(int) * h->n));
	AOF_CUDA(cudaMemset(h->s.n_full, 0, sizeof(int) * h->n));
	AOF_CUDA(cudaMemset(h->s.n_work, 0, sizeof(int) * h->n));
	AOF_CUDA(cudaMemset(h->s.lz_stats, 0, sizeof(unsigned long long) * 4));
	h->have_prev = h->have_feat_prev = false;
	h->cur = 0;
	h->l0[0] = h->l0[1] = nullptr;
	return AOF_OK;
}

extern "C" {

void aof_lk_default_params(aof_lk_params *p)
{
	if (!p) return;
	p->focal_length_x = p->focal_length_y = 0.f;
	p->output_rate = 15;              // include/optical_flow.hpp:47
	p->image_width = 64;              // :48
	p->image_height = 64;             // :49
	p->num_features = 20;             // include/flow_opencv.hpp:53
	p->confidence_multiplier = 1.645f; // :54
}

int aof_lk_create(const aof_lk_params *p, int n_streams, int device, aof_lk **out)
{
	if (!p || !out || n_streams < 1) { set_error("aof_lk_create: bad argument"); return AOF_ERR_INVALID; }
	*out = nullptr;
	const int W = p->image_width, H = p->image_height, N = p->num_features;
	if (N < 1 || N > LK_MAX_FEATURES) { set_error("num_features %d outside 1..%d (trackFeatures.cpp:56)", N, LK_MAX_FEATURES); return AOF_ERR_UNSUPPORTED; }
	const int tpb = (N - 1) / LK_BINS + 1;
	if (W < 16 || H < 16 || (W / 4) / tpb < 2 || n_streams > 65535) {
		set_error("image %dx%d / %d features / %d streams unsupported (need W,H >= 16, (W/4)/ceil(N/16) >= 2, streams <= 65535)", W, H, N, n_streams);
		return AOF_ERR_UNSUPPORTED;
	}
	int rc = select_device(device);
	if (rc != AOF_OK) return rc;
	aof_lk *h = new (std::nothrow) aof_lk();
	if (!h) return AOF_ERR_INVALID;
	h->p = *p; h->n = n_streams; h->device = device;
	LkGeom &g = h->g;
	g.n_feat = N; g.bw = W / 4; g.bh = H / 4; g.tpb = tpb; g.dist = g.bw / tpb; g.fast_thr = 10;
	g.kp_cap = ((g.bw + 1) / 2) * ((g.bh + 1) / 2);
	g.fb_bands = g.bh > 6 ? (g.bh - 6 + FB_ROWS - 1) / FB_ROWS : 1;
	g.fb_ucap = (FB_ROWS + 2) * ((g.bw / 2 + 15) / 16);
	if (g.bw > 4096 || g.bh > 4096) { // keypoint entries carry 12-bit bin-local coordinates
		delete h;
		set_error("frame %dx%d too large: bins above 4096 px", W, H);
		return AOF_ERR_UNSUPPORTED;
	}
	// cv::buildOpticalFlowPyramid: stop when the next level would be <= the window
	int lw = W, lh = H, level = 0;
	g.w[0] = W; g.h[0] = H;
	for (; level < LK_MAX_LEVELS - 1; level++) {
		const int nw = (lw + 1) / 2, nh = (lh + 1) / 2;
		if (nw <= LK_WIN || nh <= LK_WIN) break;
		lw = nw; lh = nh;
		g.w[level + 1] = lw; g.h[level + 1] = lh;
	}
	g.top = level;
	size_t off = 0;
	for (int l = 0; l <= g.top; l++) {
		h->lvl_px[l] = (size_t)g.w[l] * g.h[l];
		if (l >= 1) { h->upper_off[l] = off; off += (h->lvl_px[l] * n_streams + 255) & ~(size_t)255; }
	}
	h->upper_bytes = off ? off : 256;
	const size_t nN = (size_t)n_streams * N;
	// every environment switch of this path is read here, once (never on a launch path)
	g.tun_pyr_bands = env_int("AOF_PYR_BANDS", 0); g.tun_pyr_occ = env_int("AOF_PYR_OCC", 0);
	g.tun_fast_lazy = env_int("AOF_FAST_LAZY", 2); g.tun_pdl = env_int("AOF_LK_PDL", 1);
	g.tun_pyr_v1 = env_int("AOF_PYR_V1", 0); g.tun_fast_v1 = env_int("AOF_FAST_V1", 0); g.tun_track_v1 = env_int("AOF_TRACK_V1", 0);
	h->tun_split = env_int("AOF_LK_SPLIT", 0); h->tun_parts = env_int("AOF_LK_PARTS", 0);
	cudaError_t e;
	int prio_lo = 0, prio_hi = 0;
	// measured (r2c): with the latency-bound kernels on a HIGH-priority stream 4 parts take 0.65 ms per 512 streams
	// against 0.57 for two equal-priority halves -- the small kernels then run one after the other at their latency
	// floors while FAST waits.  AOF_LK_PRIO=1 turns the priorities on for A/B runs.
	if (env_int("AOF_LK_PRIO", 0) == 1) cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi); // (least, greatest)
	for (int c = 0; c < 8; c++)
		if (cudaEventCreateWithFlags(&h->ev_chunk[c], cudaEventDisableTiming) != cudaSuccess) h->ev_chunk[c] = nullptr;
	for (int c = 0; c < LK_MAX_PARTS; c++)
		if (cudaEventCreateWithFlags(&h->ev_trk[c], cudaEventDisableTiming) != cudaSuccess ||
		    cudaEventCreateWithFlags(&h->ev_fast[c], cudaEventDisableTiming) != cudaSuccess) h->ev_trk[c] = h->ev_fast[c] = nullptr;
	if ((e = cudaStreamCreateWithFlags(&h->st, cudaStreamNonBlocking)) != cudaSuccess ||
	    (e = cudaStreamCreateWithFlags(&h->st_copy, cudaStreamNonBlocking)) != cudaSuccess ||
	    (e = cudaStreamCreateWithPriority(&h->st_hi, cudaStreamNonBlocking, prio_hi)) != cudaSuccess ||
	    (e = cudaStreamCreateWithPriority(&h->st_lo, cudaStreamNonBlocking, prio_lo)) != cudaSuccess ||
	    (e = cudaEventCreateWithFlags(&h->ev_fork, cudaEventDisableTiming)) != cudaSuccess ||
	    (e = cudaEventCreateWithFlags(&h->ev_join, cudaEventDisableTiming)) != cudaSuccess ||
	    (e = cudaMalloc(&h->d_upper[0], h->upper_bytes)) != cudaSuccess ||
	    (e = cudaMalloc(&h->d_upper[1], h->upper_bytes)) != cudaSuccess ||
	    (e = cudaMalloc(&h->s.prev_corners, sizeof(float2) * nN)) != cudaSuccess ||
	    (e = cudaMalloc(&h->s.prev_status, nN)) != cudaSuccess ||
	    (e = cudaMalloc(&h->s.n_prev_corners, sizeof(int) * n_streams)) != cudaSuccess ||
	    (e = cudaMalloc(&h->s.update_vec, sizeof(int) * nN)) != cudaSuccess ||
	    (e = cudaMalloc(&h->s.feat_cur, sizeof(float2) * nN)) != cudaSuccess ||
	    (e = cudaMalloc(&h->s.feat_prev, sizeof(float2) * nN)) != cudaSuccess ||
	    (e = cudaMalloc(&h->s.lim, sizeof(LimiterState) * n_streams)) != cudaSuccess ||
	    (e = cudaMalloc(&h->s.bin_needed, sizeof(int) * LK_BINS * n_streams)) != cudaSuccess ||
	    (e = cudaMalloc(&h->s.target, sizeof(int) * n_streams)) != cudaSuccess ||
	    (e = cudaMalloc(&h->s.kp_count, sizeof(int) * LK_BINS * n_streams)) != cudaSuccess ||
	    (e = cudaMalloc(&h->s.lk_next, sizeof(float2) * nN)) != cudaSuccess ||
	    (e = cudaMalloc(&h->s.lk_stat, nN)) != cudaSuccess ||
	    (e = cudaMalloc(&h->s.ticket, sizeof(int) * n_streams)) != cudaSuccess ||
	    (e = cudaMalloc(&h->s.lz_pick, sizeof(uint32_t) * LK_BINS * LK_MAX_TPB * n_streams)) != cudaSuccess ||
	    (e = cudaMalloc(&h->s.lz_npick, sizeof(int) * LK_BINS * n_streams)) != cudaSuccess ||
	    (e = cudaMalloc(&h->s.lz_floor, sizeof(int) * LK_BINS * n_streams)) != cudaSuccess ||
	    (e = cudaMalloc(&h->s.need_full, sizeof(int) * n_streams)) != cudaSuccess ||
	    (e = cudaMalloc(&h->s.n_full, sizeof(int) * n_streams)) != cudaSuccess ||
	    (e = cudaMalloc(&h->s.full_list, sizeof(int) * n_streams)) != cudaSuccess ||
	    (e = cudaMalloc(&h->s.lz_stats, sizeof(unsigned long long) * 4)) != cudaSuccess ||
	    (e = cudaMalloc(&h->s.plan_units, sizeof(uint16_t) * LK_BINS * g.fb_bands * (size_t)g.fb_ucap * n_streams)) != cudaSuccess ||
	    (e = cudaMalloc(&h->s.plan_hdr, sizeof(uint32_t) * LK_BINS * g.fb_bands * (size_t)n_streams)) != cudaSuccess ||
	    (e = cudaMalloc(&h->s.work_list, sizeof(uint32_t) * LK_BINS * g.fb_bands * (size_t)n_streams)) != cudaSuccess ||
	    (e = cudaMalloc(&h->s.n_work, sizeof(int) * n_streams)) != cudaSuccess ||
	    (e = cudaMalloc(&h->s.kp, sizeof(uint32_t) * LK_BINS * (size_t)g.kp_cap * n_streams)) != cudaSuccess) {
		set_error("aof_lk_create: %s", cudaGetErrorString(e));
		aof_lk_destroy(h);
		return AOF_ERR_CUDA;
	}
	rc = reset_state(h);
	if (rc != AOF_OK) { aof_lk_destroy(h); return rc; }
	*out = h;
	return AOF_OK;
}

void aof_lk_destroy(aof_lk *h)
{
	if (!h) return;
	cudaSetDevice(h->device);
	if (h->st) cudaStreamSynchronize(h->st);
	cudaFree(h->d_upper[0]); cudaFree(h->d_upper[1]);
	cudaFree(h->s.prev_corners); cudaFree(h->s.prev_status); cudaFree(h->s.n_prev_corners); cudaFree(h->s.update_vec);
	cudaFree(h->s.feat_cur); cudaFree(h->s.feat_prev); cudaFree(h->s.lim); cudaFree(h->s.bin_needed);
	cudaFree(h->s.target); cudaFree(h->s.kp_count); cudaFree(h->s.kp);
	cudaFree(h->s.lk_next); cudaFree(h->s.lk_stat); cudaFree(h->s.ticket);
	cudaFree(h->s.lz_pick); cudaFree(h->s.lz_npick); cudaFree(h->s.lz_floor);
	cudaFree(h->s.need_full); cudaFree(h->s.n_full); cudaFree(h->s.full_list); cudaFree(h->s.lz_stats);
	cudaFree(h->s.plan_units); cudaFree(h->s.plan_hdr); cudaFree(h->s.work_list); cudaFree(h->s.n_work);
	cudaFree(h->d_frames[0]); cudaFree(h->d_frames[1]);
	cudaFreeHost(h->h_stage); cudaFreeHost(h->h_t); cudaFreeHost(h->h_out); cudaFreeHost(h->h_raw);
	cudaFree(h->d_t); cudaFree(h->d_out); cudaFree(h->d_raw);
	for (int c = 0; c < 8; c++)
		if (h->ev_chunk[c]) cudaEventDestroy(h->ev_chunk[c]);
	if (h->ev_fork) cudaEventDestroy(h->ev_fork);
	if (h->ev_join) cudaEventDestroy(h->ev_join);
	for (int c = 0; c < LK_MAX_PARTS; c++) {
		if (h->ev_trk[c]) cudaEventDestroy(h->ev_trk[c]);
		if (h->ev_fast[c]) cudaEventDestroy(h->ev_fast[c]);
	}
	if (h->st_hi) cudaStreamDestroy(h->st_hi);
	if (h->st_lo) cudaStreamDestroy(h->st_lo);
	if (h->st_copy) cudaStreamDestroy(h->st_copy);
	if (h->st) cudaStreamDestroy(h->st);
	delete h;
}

int aof_lk_num_streams(const aof_lk *h) { return h ? h->n : AOF_ERR_INVALID; }
int aof_lk_num_features(const aof_lk *h) { return h ? h->g.n_feat : AOF_ERR_INVALID; }

int aof_lk_lazy_stats(aof_lk *h, uint64_t out[4])
{
	if (!h || !out) return AOF_ERR_INVALID;
	AOF_CUDA(cudaSetDevice(h->device));
	AOF_CUDA(cudaDeviceSynchronize());
	unsigned long long v[4];
	AOF_CUDA(cudaMemcpy(v, h->s.lz_stats, sizeof v, cudaMemcpyDeviceToHost));
	for (int i = 0; i < 4; i++) out[i] = v[i];
	return AOF_OK;
}

int aof_lk_reset(aof_lk *h)
{
	if (!h) return AOF_ERR_INVALID;
	AOF_CUDA(cudaSetDevice(h->device));
	AOF_CUDA(cudaStreamSynchronize(h->st));
	AOF_CUDA(cudaDeviceSynchronize());
	return reset_state(h);
}

// the four stages of one calcFlow pass over streams [first, first + count) of the batch whose level 0 is h->l0[nxt]
static int run_pyr_track(aof_lk *h, int nxt, int first, int count, cudaStream_t st)
{
	const size_t f = (size_t)first;
	const LkPyr next = pyr_of(h, nxt, f);
	const LkPyr prev = h->have_prev ? pyr_of(h, h->cur, f) : next;
	const LkState sv = state_from(h->s, h->g, f);
	if (lk_pyr_stream_supported(h->g, next.lvl[0]))
		AOF_CUDA(lk_pyr_stream_launch(h->g, next.lvl[0], const_cast<uint8_t *>(next.lvl[1]), const_cast<uint8_t *>(next.lvl[2]),
					      const_cast<uint8_t *>(next.lvl[3]), count, st));
	else if (h->g.top == 3)
		AOF_CUDA(lk_pyr_fused_launch(h->g, next.lvl[0], const_cast<uint8_t *>(next.lvl[1]), const_cast<uint8_t *>(next.lvl[2]),
					     const_cast<uint8_t *>(next.lvl[3]), count, st));
	else
		for (int l = 1; l <= h->g.top; l++)
			AOF_CUDA(lk_pyr_down_launch(next.lvl[l - 1], const_cast<uint8_t *>(next.lvl[l]), h->g.w[l - 1], h->g.h[l - 1], count, st));
	AOF_CUDA(lk_track_update_launch(h->g, prev, next, sv, count, h->have_prev ? 1 : 0, st));
	return AOF_OK;
}

// How the keypoints of the bins that need features are found (AOF_FAST_LAZY, read at create):
//   LK_SEL_MASKED (2, default)  the band kernel scoring only the pixels the first selection stage can accept
//   LK_SEL_LAZY (1)             the lazy arg-max kernel (lk_fast_lazy.cu)
//   LK_SEL_COMPLETE (0, and every shape the band kernel does not take)  every pixel scored
// MASKED and LAZY hand the streams whose selection needs the second-chance pass to the complete path.
static int fast_mode(const aof_lk *h, const uint8_t *l0)
{
	if (h->g.tun_fast_lazy == 2 && lk_fast_masked_supported(h->g, l0)) return LK_SEL_MASKED;
	if (h->g.tun_fast_lazy == 1 && lk_fast_lazy_supported(h->g, l0)) return LK_SEL_LAZY;
	return LK_SEL_COMPLETE;
}

static int run_fast(aof_lk *h, int nxt, int first, int count, cudaStream_t st)
{
	const size_t f = (size_t)first;
	const uint8_t *l0 = pyr_of(h, nxt, f).lvl[0];
	const int mode = fast_mode(h, l0);
	if (mode == LK_SEL_LAZY) AOF_CUDA(lk_fast_pick_launch(h->g, l0, state_from(h->s, h->g, f), count, st));
	else if (mode == LK_SEL_MASKED) AOF_CUDA(lk_fast_masked_launch(h->g, l0, state_from(h->s, h->g, f), count, st));
	else AOF_CUDA(lk_fast_launch(h->g, l0, state_from(h->s, h->g, f), count, nullptr, st));
	return AOF_OK;
}

static int run_select(aof_lk *h, const uint32_t *d_t_us, aof_flow_out *d_out, aof_lk_raw *d_raw, int first, int count,
		      cudaStream_t st)
{
	const size_t f = (size_t)first;
	const LkState sv = state_from(h->s, h->g, f);
	const uint8_t *l0 = pyr_of(h, h->cur ^ 1, f).lvl[0];
	const int mode = fast_mode(h, l0);
	const bool lazy = mode != LK_SEL_COMPLETE;
	// masked / lazy: the selection flags the streams whose second-chance pass needs every keypoint; those (usually none)
	// then go through the complete FAST and the complete selection
	for (int pass = 0; pass < (lazy ? 2 : 1); pass++) {
		if (pass == 1) AOF_CUDA(lk_fast_band_flagged_launch(h->g, l0, sv, count, st));
		AOF_CUDA(lk_select_epilogue_launch(h->g, sv, d_t_us + f, d_out + f, d_raw ? d_raw + f : nullptr, count,
						   h->have_feat_prev ? 1 : 0, h->p.output_rate, h->p.confidence_multiplier,
						   h->p.focal_length_x, h->p.focal_length_y, h->cam, st,
						   pass == 0 ? mode : LK_SEL_FLAGGED));
	}
	return AOF_OK;
}

static int run_range(aof_lk *h, int nxt, const uint32_t *d_t_us, aof_flow_out *d_out, aof_lk_raw *d_raw, int first, int count,
		     cudaStream_t st)
{
	int rc = run_pyr_track(h, nxt, first, count, st);
	if (rc == AOF_OK) rc = run_fast(h, nxt, first, count, st);
	if (rc == AOF_OK) rc = run_select(h, d_t_us, d_out, d_raw, first, count, st);
	return rc;
}

int aof_lk_calc_flow_device(aof_lk *h, const uint8_t *d_imgs, const uint32_t *d_t_us, aof_flow_out *d_out,
			    aof_lk_raw *d_raw, void *cuda_stream)
{
	if (!h || !d_imgs || !d_t_us || !d_out) { set_error("lk calc_flow_device: null argument"); return AOF_ERR_INVALID; }
	AOF_CUDA(cudaSetDevice(h->device));
	const int nxt = h->cur ^ 1;
	h->l0[nxt] = d_imgs;
	// FAST is bound by the ALU pipe (and is two thirds of the step); the pyramid, the tracker and the selection are
	// bound by latency.  The batch is cut into parts of whole streams and software-pipelined over two internal
	// streams: the latency-bound kernels of every part go to a HIGH-priority stream, FAST to a LOW-priority one, with
	// events for the per-part dependencies (track -> FAST -> select).  While FAST of part p fills the SMs, the
	// pyramid / tracker of part p + 1 and the selection of part p - 1 take the CTA slots that free up first, so only the
	// head (pyramid + tracker of the first part) and the tail (selection of the last part) are exposed.
	// Off while per-kernel timing is on (the timed kernels must not overlap); AOF_LK_SPLIT=1 turns it off, =2 forces it,
	// AOF_LK_PARTS sets the number of parts (both read once, in aof_lk_create).
	int parts = h->tun_parts > 0 ? h->tun_parts : 2;
	if (parts > LK_MAX_PARTS) parts = LK_MAX_PARTS;
	if (parts > h->n) parts = h->n;
	// Measured with the masked detector (r2r, 512 streams): one pass 0.499 ms, 2 parts 0.563, 3 parts 0.562, 4 parts 0.601 --
	// its band kernel is a persistent grid that holds every CTA slot it can get, so a second part's latency-bound
	// kernels find no room beside it and the parts only add launches.  The pipeline stays the default for the complete
	// and the lazy detector (short CTAs that leave slots as they finish).
	const bool auto_split = h->n >= 256 && fast_mode(h, d_imgs) != LK_SEL_MASKED;
	const bool pipelined = (h->tun_split == 0 ? auto_split : h->tun_split >= 2) && parts >= 2 && !timing_on() && h->ev_trk[LK_MAX_PARTS - 1];
	cudaStream_t st = (cudaStream_t)cuda_stream;
	int rc = AOF_OK;
	if (!pipelined) rc = run_range(h, nxt, d_t_us, d_out, d_raw, 0, h->n, st);
	else {
		const int per = (h->n + parts - 1) / parts;
		AOF_CUDA(cudaEventRecord(h->ev_fork, st));
		AOF_CUDA(cudaStreamWaitEvent(h->st_hi, h->ev_fork, 0));
		int np = 0;
		for (int first = 0; first < h->n && rc == AOF_OK; first += per, np++) {
			const int count = first + per <= h->n ? per : h->n - first;
			rc = run_pyr_track(h, nxt, first, count, h->st_hi);
			if (rc != AOF_OK) break;
			AOF_CUDA(cudaEventRecord(h->ev_trk[np], h->st_hi));
			AOF_CUDA(cudaStreamWaitEvent(h->st_lo, h->ev_trk[np], 0));
			rc = run_fast(h, nxt, first, count, h->st_lo);
			AOF_CUDA(cudaEventRecord(h->ev_fast[np], h->st_lo));
		}
		for (int first = 0, q = 0; first < h->n && rc == AOF_OK; first += per, q++) {
			const int count = first + per <= h->n ? per : h->n - first;
			AOF_CUDA(cudaStreamWaitEvent(h->st_hi, h->ev_fast[q], 0));
			rc = run_select(h, d_t_us, d_out, d_raw, first, count, h->st_hi);
		}
		AOF_CUDA(cudaEventRecord(h->ev_join, h->st_hi));
		AOF_CUDA(cudaStreamWaitEvent(st, h->ev_join, 0));
	}
	if (rc != AOF_OK) return rc;
	h->cur = nxt;
	h->have_prev = h->have_feat_prev = true;
	return AOF_OK;
}

static int ensure_batch_buffers(aof_lk *h)
{
	if (h->h_stage) return AOF_OK;
	const size_t fb = h->lvl_px[0] * h->n;
	AOF_CUDA(cudaMalloc(&h->d_frames[0], fb));
	AOF_CUDA(cudaMalloc(&h->d_frames[1], fb));
	AOF_CUDA(cudaMallocHost(&h->h_t, sizeof(uint32_t) * h->n));
	AOF_CUDA(cudaMallocHost(&h->h_out, sizeof(aof_flow_out) * h->n));
	AOF_CUDA(cudaMallocHost(&h->h_raw, sizeof(aof_lk_raw) * h->n));
	AOF_CUDA(cudaMalloc(&h->d_t, sizeof(uint32_t) * h->n));
	AOF_CUDA(cudaMalloc(&h->d_out, sizeof(aof_flow_out) * h->n));
	AOF_CUDA(cudaMalloc(&h->d_raw, sizeof(aof_lk_raw) * h->n));
	AOF_CUDA(cudaMallocHost(&h->h_stage, fb));
	return AOF_OK;
}

// Host frames in, results out.  The batch is cut into up to 8 chunks of whole streams: chunk c + 1 crosses PCIe on the
// copy stream while the kernels of chunk c run (streams are independent, so a chunk is a complete calcFlow pass).
int aof_lk_calc_flow_batch(aof_lk *h, const uint8_t *imgs, const uint32_t *t_us, aof_flow_out *out, aof_lk_raw *raw)
{
	if (!h || !imgs || !t_us || !out) { set_error("lk calc_flow_batch: null argument"); return AOF_ERR_INVALID; }
	AOF_CUDA(cudaSetDevice(h->device));
	int rc = ensure_batch_buffers(h);
	if (rc != AOF_OK) return rc;
	const int nxt = h->cur ^ 1;
	uint8_t *d_cur = h->d_frames[nxt];
	h->l0[nxt] = d_cur;
	memcpy(h->h_t, t_us, sizeof(uint32_t) * h->n);
	AOF_CUDA(cudaMemcpyAsync(h->d_t, h->h_t, sizeof(uint32_t) * h->n, cudaMemcpyHostToDevice, h->st));
	// chunks of >= 16 streams and >= 1 MiB, 8-byte aligned starts (the pyramid kernel loads 8 bytes per lane)
	int chunks = h->n / 16;
	if ((size_t)chunks > h->lvl_px[0] * h->n >> 20) chunks = (int)(h->lvl_px[0] * h->n >> 20);
	if (chunks > 8) chunks = 8;
	if (chunks < 1 || h->lvl_px[0] % 8) chunks = 1;
	const int per = (h->n + chunks - 1) / chunks;
	for (int c = 0, first = 0; first < h->n; c++, first += per) {
		const int count = first + per <= h->n ? per : h->n - first;
		const size_t off = (size_t)first * h->lvl_px[0];
		rc = stage_h2d(d_cur + off, imgs + off, (size_t)count * h->lvl_px[0], h->h_stage + off, h->st_copy);
		if (rc != AOF_OK) return rc;
		AOF_CUDA(cudaEventRecord(h->ev_chunk[c], h->st_copy));
		AOF_CUDA(cudaStreamWaitEvent(h->st, h->ev_chunk[c], 0));
		rc = run_range(h, nxt, h->d_t, h->d_out, h->d_raw, first, count, h->st);
		if (rc != AOF_OK) return rc;
	}
	h->cur = nxt;
	h->have_prev = h->have_feat_prev = true;
	// the "host gather": straight into the caller's arrays when they are pinned (see aof_px4_calc_flow_batch)
	const bool direct = host_pinned(out), direct_raw = raw && host_pinned(raw);
	AOF_CUDA(cudaMemcpyAsync(direct ? out : h->h_out, h->d_out, sizeof(aof_flow_out) * h->n, cudaMemcpyDeviceToHost, h->st));
	if (raw) AOF_CUDA(cudaMemcpyAsync(direct_raw ? raw : h->h_raw, h->d_raw, sizeof(aof_lk_raw) * h->n, cudaMemcpyDeviceToHost, h->st));
	AOF_CUDA(cudaStreamSynchronize(h->st));
	if (!direct) memcpy(out, h->h_out, sizeof(aof_flow_out) * h->n);
	if (raw && !direct_raw) memcpy(raw, h->h_raw, sizeof(aof_lk_raw) * h->n);
	return AOF_OK;
}

int aof_lk_calc_flow(aof_lk *h, const uint8_t *img, uint32_t img_time_us, int *dt_us, float *flow_x, float *flow_y)
{
	if (!h || !img || !dt_us || !flow_x || !flow_y) { set_error("lk calc_flow: null argument"); return AOF_ERR_INVALID; }
	if (h->n != 1) { set_error("calc_flow needs a 1-stream handle (this one has %d)", h->n); return AOF_ERR_STATE; }
	aof_flow_out o;
	int rc = aof_lk_calc_flow_batch(h, img, &img_time_us, &o, nullptr);
	if (rc != AOF_OK) return rc;
	*flow_x = o.flow_x;
	*flow_y = o.flow_y;
	if (o.quality >= 0 || h->p.output_rate <= 0) *dt_us = o.dt_us;
	return o.quality;
}

/* setFocalLengthX/Y, setOutputRate (include/optical_flow.hpp:76-78), setConfMultiplier (include/flow_opencv.hpp:73) */
int aof_lk_set_params(aof_lk *h, float focal_length_x, float focal_length_y, int output_rate, float confidence_multiplier)
{
	if (!h) { set_error("aof_lk_set_params: null handle"); return AOF_ERR_INVALID; }
	// passed to the selection / epilogue kernel by value on every call: nothing on the device changes
	h->p.focal_length_x = focal_length_x; h->p.focal_length_y = focal_length_y; h->p.output_rate = output_rate;
	h->p.confidence_multiplier = confidence_multiplier;
	return AOF_OK;
}

/* OpticalFlowOpenCV::setCameraMatrix (src/flow_opencv.cpp:72-80) */
int aof_lk_set_camera_matrix(aof_lk *h, float focal_len_x, float focal_len_y, float principal_point_x, float principal_point_y)
{
	if (!h) return AOF_ERR_INVALID;
	h->cam.fx = focal_len_x; h->cam.fy = focal_len_y; h->cam.cx = principal_point_x; h->cam.cy = principal_point_y;
	h->set_cam = true;
	h->cam.on = h->set_cam && h->set_dist;
	return AOF_OK;
}

/* OpticalFlowOpenCV::setCameraDistortion(k1, k2, k3, p1, p2) (src/flow_opencv.cpp:82-87) */
int aof_lk_set_camera_distortion(aof_lk *h, float k1, float k2, float k3, float p1, float p2)
{
	if (!h) return AOF_ERR_INVALID;
	h->cam.k1 = k1; h->cam.k2 = k2; h->cam.k3 = k3; h->cam.p1 = p1; h->cam.p2 = p2;
	h->set_dist = true;
	h->cam.on = h->set_cam && h->set_dist;
	return AOF_OK;
}

int aof_lk_get_state(aof_lk *h, int32_t *update_vec, float *feat_cur, float *prev_corners, uint8_t *prev_status)
{
	if (!h) return AOF_ERR_INVALID;
	AOF_CUDA(cudaSetDevice(h->device));
	AOF_CUDA(cudaDeviceSynchronize());
	const size_t nN = (size_t)h->n * h->g.n_feat;
	if (update_vec) AOF_CUDA(cudaMemcpy(update_vec, h->s.update_vec, sizeof(int) * nN, cudaMemcpyDeviceToHost));
	if (feat_cur) AOF_CUDA(cudaMemcpy(feat_cur, h->s.feat_cur, sizeof(float2) * nN, cudaMemcpyDeviceToHost));
	if (prev_corners) AOF_CUDA(cudaMemcpy(prev_corners, h->s.prev_corners, sizeof(float2) * nN, cudaMemcpyDeviceToHost));
	if (prev_status) AOF_CUDA(cudaMemcpy(prev_status, h->s.prev_status, nN, cudaMemcpyDeviceToHost));
	return AOF_OK;
}

} // extern "C"

// ---- kernel-level entry points for known-answer tests (host pointers, synchronous) ----

namespace {
struct TmpPyr { // pyramids of n frames in scratch device memory
	uint8_t *l0 = nullptr, *upper = nullptr;
	size_t off[LK_MAX_LEVELS] = {};
	~TmpPyr() { cudaFree(l0); cudaFree(upper); }
};

int make_tmp_pyr(aof_lk *h, const uint8_t *imgs, int n, TmpPyr &t)
{
	const LkGeom &g = h->g;
	size_t o = 0;
	for (int l = 1; l <= g.top; l++) { t.off[l] = o; o += (h->lvl_px[l] * n + 255) & ~(size_t)255; }
	AOF_CUDA(cudaMalloc(&t.l0, h->lvl_px[0] * n));
	AOF_CUDA(cudaMalloc(&t.upper, o ? o : 256));
	AOF_CUDA(cudaMemcpyAsync(t.l0, imgs, h->lvl_px[0] * n, cudaMemcpyHostToDevice, h->st));
	AOF_CUDA(build_upper(g, t.l0, t.upper, t.off, n, h->st));
	return AOF_OK;
}

LkPyr pyr_of_tmp(const LkGeom &g, const TmpPyr &t)
{
	LkPyr p{};
	p.lvl[0] = t.l0;
	for (int l = 1; l <= g.top; l++) p.lvl[l] = t.upper + t.off[l];
	return p;
}
} // namespace

extern "C" {

int aof_lk_pyramid(aof_lk *h, const uint8_t *imgs, int n, int *lw, int *lh, uint8_t *levels)
{
	if (!h || n < 0 || n > 65535) { set_error("aof_lk_pyramid: bad argument"); return AOF_ERR_INVALID; }
	const LkGeom &g = h->g;
	for (int l = 0; l <= g.top; l++) { if (lw) lw[l] = g.w[l]; if (lh) lh[l] = g.h[l]; }
	if (!levels || !imgs || n == 0) return g.top + 1;
	AOF_CUDA(cudaSetDevice(h->device));
	TmpPyr t;
	int rc = make_tmp_pyr(h, imgs, n, t);
	if (rc != AOF_OK) return rc;
	size_t per_frame = 0;
	for (int l = 0; l <= g.top; l++) per_frame += h->lvl_px[l];
	size_t lo = 0;
	for (int l = 0; l <= g.top; l++) { // strided copy: level l of frame i -> levels + i * per_frame + lo
		const uint8_t *src = l == 0 ? t.l0 : t.upper + t.off[l];
		AOF_CUDA(cudaMemcpy2DAsync(levels + lo, per_frame, src, h->lvl_px[l], h->lvl_px[l], n, cudaMemcpyDeviceToHost, h->st));
		lo += h->lvl_px[l];
	}
	AOF_CUDA(cudaStreamSynchronize(h->st));
	return g.top + 1;
}

int aof_lk_track_points(aof_lk *h, const uint8_t *prev, const uint8_t *cur, int n, const float *pts, int n_pts,
			float *next_pts, uint8_t *status)
{
	if (!h || !prev || !cur || !pts || !next_pts || !status || n < 0 || n > 65535 || n_pts < 0) {
		set_error("aof_lk_track_points: bad argument");
		return AOF_ERR_INVALID;
	}
	if (n == 0 || n_pts == 0) return AOF_OK;
	AOF_CUDA(cudaSetDevice(h->device));
	TmpPyr a, b;
	int rc = make_tmp_pyr(h, prev, n, a);
	if (rc != AOF_OK) return rc;
	rc = make_tmp_pyr(h, cur, n, b);
	if (rc != AOF_OK) return rc;
	const size_t np = (size_t)n * n_pts;
	float2 *d_pts = nullptr, *d_out = nullptr;
	uint8_t *d_st = nullptr;
	cudaError_t e;
	if ((e = cudaMalloc(&d_pts, sizeof(float2) * np)) != cudaSuccess || (e = cudaMalloc(&d_out, sizeof(float2) * np)) != cudaSuccess ||
	    (e = cudaMalloc(&d_st, np)) != cudaSuccess ||
	    (e = cudaMemcpyAsync(d_pts, pts, sizeof(float2) * np, cudaMemcpyHostToDevice, h->st)) != cudaSuccess ||
	    (e = lk_track_points_launch(h->g, pyr_of_tmp(h->g, a), pyr_of_tmp(h->g, b), d_pts, n_pts, d_out, d_st, n, h->st)) != cudaSuccess ||
	    (e = cudaMemcpyAsync(next_pts, d_out, sizeof(float2) * np, cudaMemcpyDeviceToHost, h->st)) != cudaSuccess ||
	    (e = cudaMemcpyAsync(status, d_st, np, cudaMemcpyDeviceToHost, h->st)) != cudaSuccess ||
	    (e = cudaStreamSynchronize(h->st)) != cudaSuccess) {
		set_error("aof_lk_track_points: %s", cudaGetErrorString(e));
		rc = AOF_ERR_CUDA;
	}
	cudaFree(d_pts); cudaFree(d_out); cudaFree(d_st);
	return rc;
}

int aof_lk_fast_scores(aof_lk *h, const uint8_t *imgs, int n, uint8_t *score)
{
	if (!h || !imgs || !score || n < 0 || n > 65535) { set_error("aof_lk_fast_scores: bad argument"); return AOF_ERR_INVALID; }
	if (n == 0) return AOF_OK;
	AOF_CUDA(cudaSetDevice(h->device));
	const size_t bytes = h->lvl_px[0] * n;
	uint8_t *d_img = nullptr, *d_sc = nullptr;
	int rc = AOF_OK;
	cudaError_t e;
	if ((e = cudaMalloc(&d_img, bytes)) != cudaSuccess || (e = cudaMalloc(&d_sc, bytes)) != cudaSuccess ||
	    (e = cudaMemcpyAsync(d_img, imgs, bytes, cudaMemcpyHostToDevice, h->st)) != cudaSuccess ||
	    (e = cudaMemsetAsync(d_sc, 0, bytes, h->st)) != cudaSuccess ||
	    (e = lk_fast_launch(h->g, d_img, h->s, n, d_sc, h->st)) != cudaSuccess ||
	    (e = cudaMemcpyAsync(score, d_sc, bytes, cudaMemcpyDeviceToHost, h->st)) != cudaSuccess ||
	    (e = cudaStreamSynchronize(h->st)) != cudaSuccess) {
		set_error("aof_lk_fast_scores: %s", cudaGetErrorString(e));
		rc = AOF_ERR_CUDA;
	}
	cudaFree(d_img); cudaFree(d_sc);
	return rc;
}

} // extern "C"
