// C-ABI host side of the SAD path (include/aof.h): per-stream device state,
// pinned staging, async copies, launches.  Replaces the host logic of
// OpticalFlowPX4 (src/flow_px4.cpp:47-98); the arithmetic is in sad_kernels.cu.
#include <cstdio>
#include <cstring>
#include <new>

#include "aof_internal.h"
#include "aof_host.h"

using namespace aof;

struct aof_px4 {
	aof_px4_params p;
	int n = 0, device = 0;
	SadGeom g{};
	size_t frame_bytes = 0;
	cudaStream_t st = nullptr, st_copy = nullptr;
	cudaEvent_t ev_chunk[8] = {};    // host-batch path: chunk c of the frames has landed
	bool primed = false;
	const uint8_t *d_prev = nullptr; // previous frames of all streams (own buffer or the caller's)
	// state + results
	LimiterState *d_lim = nullptr;
	aof_px4_raw *d_raw = nullptr;
	// host-batch path (allocated on first use)
	uint8_t *d_frames[2] = {nullptr, nullptr};
	int slot = 0;
	uint8_t *h_stage = nullptr;
	uint32_t *h_t = nullptr, *d_t = nullptr;
	aof_flow_out *h_out = nullptr, *d_out = nullptr;
	// stateless pairs path scratch
	uint8_t *d_pp = nullptr, *d_pc = nullptr;
	aof_px4_raw *d_praw = nullptr;
	aof_px4_block *d_pblk = nullptr;
	int pairs_cap = 0;
	bool pairs_blocks = false;
};

extern "C" {



void aof_px4_default_params(aof_px4_params *p)
{
	if (!p) return;
	p->focal_length_x = p->focal_length_y = 0.f;
	p->output_rate = 15;          // include/optical_flow.hpp:47
	p->image_width = 64;          // :48
	p->image_height = 64;         // :49
	p->search_size = 6;           // include/flow_px4.hpp:47
	p->flow_feature_threshold = 30;
	p->flow_value_threshold = 3000;
}

int aof_px4_create(const aof_px4_params *p, int n_streams, int device, aof_px4 **out)
{
	if (!p || !out || n_streams < 1) { set_error("aof_px4_create: bad argument"); return AOF_ERR_INVALID; }
	*out = nullptr;
	const int W = p->image_width, H = p->image_height, S = p->search_size;
	// any search size the reference's int8_t displacement loop can express (px4flow.cpp:393-396); 4, 6 and 8 have
	// specialised kernels, the rest run sad_flow_any_kernel
	if (S < 1 || S > 127) { set_error("search_size %d out of range (1..127)", S); return AOF_ERR_UNSUPPORTED; }
	// the reference grid uses image_width on both axes (px4flow.cpp:360,380) and 16-bit
	// pixel offsets (:117,304): its valid domain is H >= W, W*H <= 65535.
	if (W < 2 * (S + 1) + 8 + 1 || H < W || (long)W * H > 65535 || (W & 3) || ((W * H) & 15)) {
		set_error("image %dx%d unsupported (need W%%4==0, W*H%%16==0, H>=W>=%d, W*H<=65535)", W, H, 2 * (S + 1) + 9);
		return AOF_ERR_UNSUPPORTED;
	}
	int rc = select_device(device);
	if (rc != AOF_OK) return rc;
	aof_px4 *h = new (std::nothrow) aof_px4();
	if (!h) return AOF_ERR_INVALID;
	h->p = *p; h->n = n_streams; h->device = device;
	h->frame_bytes = (size_t)W * H;
	SadGeom &g = h->g;
	g.width = W;
	g.stage_bytes = (uint32_t)h->frame_bytes;
	g.frame_stride = h->frame_bytes;
	const int lo = S + 1, hi = W - (S + 1) - 8;
	g.lo = lo;
	g.step = (hi - lo) / 5 + 1;
	g.nb = 0;
	for (int v = lo; v < hi; v += g.step) g.nb++;
	g.tun_v1 = env_int("AOF_SAD_V1", 0); g.tun_v2 = env_int("AOF_SAD_V2", 0); g.tun_ctas = env_int("AOF_SAD_CTAS", 0);
	g.feat_thr = (uint32_t)p->flow_feature_threshold;
	g.val_thr = (uint32_t)p->flow_value_threshold;
	cudaError_t e;
	for (int c = 0; c < 8; c++)
		if (cudaEventCreateWithFlags(&h->ev_chunk[c], cudaEventDisableTiming) != cudaSuccess) h->ev_chunk[c] = nullptr;
	if ((e = cudaStreamCreateWithFlags(&h->st, cudaStreamNonBlocking)) != cudaSuccess ||
	    (e = cudaStreamCreateWithFlags(&h->st_copy, cudaStreamNonBlocking)) != cudaSuccess ||
	    (e = cudaMalloc(&h->d_lim, sizeof(LimiterState) * n_streams)) != cudaSuccess ||
	    (e = cudaMalloc(&h->d_raw, sizeof(aof_px4_raw) * n_streams)) != cudaSuccess ||
	    (e = cudaMemset(h->d_lim, 0, sizeof(LimiterState) * n_streams)) != cudaSuccess) {
		set_error("aof_px4_create: %s", cudaGetErrorString(e));
		aof_px4_destroy(h);
		return AOF_ERR_CUDA;
	}
	*out = h;
	return AOF_OK;
}

int aof_px4_set_params(aof_px4 *h, float focal_length_x, float focal_length_y, int output_rate)
{
	if (!h) { set_error("aof_px4_set_params: null handle"); return AOF_ERR_INVALID; }
	// passed to the epilogue kernel by value on every call: nothing on the device changes
	h->p.focal_length_x = focal_length_x; h->p.focal_length_y = focal_length_y; h->p.output_rate = output_rate;
	return AOF_OK;
}

void aof_px4_destroy(aof_px4 *h)
{
	if (!h) return;
	cudaSetDevice(h->device);
	if (h->st) cudaStreamSynchronize(h->st);
	cudaFree(h->d_lim); cudaFree(h->d_raw);
	cudaFree(h->d_frames[0]); cudaFree(h->d_frames[1]);
	cudaFreeHost(h->h_stage); cudaFreeHost(h->h_t); cudaFreeHost(h->h_out);
	cudaFree(h->d_t); cudaFree(h->d_out);
	cudaFree(h->d_pp); cudaFree(h->d_pc); cudaFree(h->d_praw); cudaFree(h->d_pblk);
	for (int c = 0; c < 8; c++)
		if (h->ev_chunk[c]) cudaEventDestroy(h->ev_chunk[c]);
	if (h->st_copy) cudaStreamDestroy(h->st_copy);
	if (h->st) cudaStreamDestroy(h->st);
	delete h;
}

int aof_px4_num_streams(const aof_px4 *h) { return h ? h->n : AOF_ERR_INVALID; }

int aof_px4_reset(aof_px4 *h)
{
	if (!h) return AOF_ERR_INVALID;
	AOF_CUDA(cudaSetDevice(h->device));
	AOF_CUDA(cudaStreamSynchronize(h->st));
	AOF_CUDA(cudaMemset(h->d_lim, 0, sizeof(LimiterState) * h->n));
	h->primed = false;
	h->d_prev = nullptr;
	return AOF_OK;
}

int aof_px4_calc_flow_device(aof_px4 *h, const uint8_t *d_imgs, const uint32_t *d_t_us, aof_flow_out *d_out,
			     aof_px4_raw *d_raw, aof_px4_block *d_blocks, void *cuda_stream)
{
	if (!h || !d_imgs || !d_t_us || !d_out) { set_error("calc_flow_device: null argument"); return AOF_ERR_INVALID; }
	if ((uintptr_t)d_imgs & 15) { // the kernels stage frames with cp.async.bulk: a misaligned source faults the context
		set_error("calc_flow_device: d_imgs must be 16-byte aligned");
		return AOF_ERR_INVALID;
	}
	AOF_CUDA(cudaSetDevice(h->device));
	cudaStream_t st = (cudaStream_t)cuda_stream;
	if (!h->primed) { // first frame: remember it, report quality 0 (flow_px4.cpp:73-78)
		AOF_CUDA(cudaMemsetAsync(d_out, 0, sizeof(aof_flow_out) * h->n, st));
		if (d_raw) AOF_CUDA(cudaMemsetAsync(d_raw, 0, sizeof(aof_px4_raw) * h->n, st));
		h->primed = true;
		h->d_prev = d_imgs;
		return AOF_OK;
	}
	aof_px4_raw *raw = d_raw ? d_raw : h->d_raw;
	AOF_CUDA(sad_flow_launch(h->p.search_size, h->d_prev, d_imgs, h->g, h->n, raw, d_blocks, st));
	AOF_CUDA(px4_epilogue_launch(raw, d_t_us, h->d_lim, d_out, h->n, h->p.output_rate, h->p.focal_length_x,
				     h->p.focal_length_y, st));
	h->d_prev = d_imgs;
	return AOF_OK;
}

static int ensure_batch_buffers(aof_px4 *h)
{
	if (h->h_stage) return AOF_OK;
	const size_t fb = h->frame_bytes * h->n;
	AOF_CUDA(cudaMalloc(&h->d_frames[0], fb + 16));
	AOF_CUDA(cudaMalloc(&h->d_frames[1], fb + 16));
	AOF_CUDA(cudaMallocHost(&h->h_stage, fb));
	AOF_CUDA(cudaMallocHost(&h->h_t, sizeof(uint32_t) * h->n));
	AOF_CUDA(cudaMallocHost(&h->h_out, sizeof(aof_flow_out) * h->n));
	AOF_CUDA(cudaMalloc(&h->d_t, sizeof(uint32_t) * h->n));
	AOF_CUDA(cudaMalloc(&h->d_out, sizeof(aof_flow_out) * h->n));
	return AOF_OK;
}

// Host frames in, results out.  Up to 8 chunks of whole streams: chunk c + 1 crosses PCIe on the copy stream while the
// kernels of chunk c run.
int aof_px4_calc_flow_batch(aof_px4 *h, const uint8_t *imgs, const uint32_t *t_us, aof_flow_out *out)
{
	if (!h || !imgs || !t_us || !out) { set_error("calc_flow_batch: null argument"); return AOF_ERR_INVALID; }
	AOF_CUDA(cudaSetDevice(h->device));
	int rc = ensure_batch_buffers(h);
	if (rc != AOF_OK) return rc;
	uint8_t *d_cur = h->d_frames[h->slot];
	memcpy(h->h_t, t_us, sizeof(uint32_t) * h->n);
	AOF_CUDA(cudaMemcpyAsync(h->d_t, h->h_t, sizeof(uint32_t) * h->n, cudaMemcpyHostToDevice, h->st));
	if (!h->primed) { // first frame: remember it, report quality 0 (flow_px4.cpp:73-78)
		rc = stage_h2d(d_cur, imgs, h->frame_bytes * h->n, h->h_stage, h->st);
		if (rc != AOF_OK) return rc;
		rc = aof_px4_calc_flow_device(h, d_cur, h->d_t, h->d_out, nullptr, nullptr, h->st);
		if (rc != AOF_OK) return rc;
	} else {
		int chunks = (int)(h->frame_bytes * h->n >> 21);   // >= 2 MiB per chunk
		if (chunks > 8) chunks = 8;
		if (chunks < 1) chunks = 1;
		const int per = (h->n + chunks - 1) / chunks;
		for (int c = 0, first = 0; first < h->n; c++, first += per) {
			const int count = first + per <= h->n ? per : h->n - first;
			const size_t off = (size_t)first * h->frame_bytes;
			rc = stage_h2d(d_cur + off, imgs + off, (size_t)count * h->frame_bytes, h->h_stage + off, h->st_copy);
			if (rc != AOF_OK) return rc;
			AOF_CUDA(cudaEventRecord(h->ev_chunk[c], h->st_copy));
			AOF_CUDA(cudaStreamWaitEvent(h->st, h->ev_chunk[c], 0));
			AOF_CUDA(sad_flow_launch(h->p.search_size, h->d_prev + off, d_cur + off, h->g, count, h->d_raw + first, nullptr, h->st));
			AOF_CUDA(px4_epilogue_launch(h->d_raw + first, h->d_t + first, h->d_lim + first, h->d_out + first, count,
						     h->p.output_rate, h->p.focal_length_x, h->p.focal_length_y, h->st));
		}
		h->d_prev = d_cur;
	}
	// the "host gather": 16 B per stream, straight into the caller's array when it is pinned (the multi-GPU entry
	// points hand every device its slice of ONE pinned array), else through the handle's pinned buffer
	const bool direct = host_pinned(out);
	AOF_CUDA(cudaMemcpyAsync(direct ? out : h->h_out, h->d_out, sizeof(aof_flow_out) * h->n, cudaMemcpyDeviceToHost, h->st));
	AOF_CUDA(cudaStreamSynchronize(h->st));
	if (!direct) memcpy(out, h->h_out, sizeof(aof_flow_out) * h->n);
	h->slot ^= 1;
	return AOF_OK;
}

int aof_px4_calc_flow(aof_px4 *h, const uint8_t *img, uint32_t img_time_us, int *dt_us, float *flow_x, float *flow_y)
{
	if (!h || !img || !dt_us || !flow_x || !flow_y) { set_error("calc_flow: null argument"); return AOF_ERR_INVALID; }
	if (h->n != 1) { set_error("calc_flow needs a 1-stream handle (this one has %d)", h->n); return AOF_ERR_STATE; }
	const bool first = !h->primed;
	aof_flow_out o;
	int rc = aof_px4_calc_flow_batch(h, img, &img_time_us, &o);
	if (rc != AOF_OK) return rc;
	if (first) return 0; // outputs untouched (flow_px4.cpp:73-78)
	*flow_x = o.flow_x;
	*flow_y = o.flow_y;
	if (o.quality >= 0 || h->p.output_rate <= 0) *dt_us = o.dt_us;
	return o.quality;
}

static int ensure_pairs(aof_px4 *h, int n, bool blocks)
{
	if (n > h->pairs_cap) {
		cudaFree(h->d_pp); cudaFree(h->d_pc); cudaFree(h->d_praw); cudaFree(h->d_pblk);
		h->d_pp = h->d_pc = nullptr; h->d_praw = nullptr; h->d_pblk = nullptr; h->pairs_cap = 0; h->pairs_blocks = false;
		AOF_CUDA(cudaMalloc(&h->d_pp, h->frame_bytes * n + 16));
		AOF_CUDA(cudaMalloc(&h->d_pc, h->frame_bytes * n + 16));
		AOF_CUDA(cudaMalloc(&h->d_praw, sizeof(aof_px4_raw) * n));
		h->pairs_cap = n;
	}
	if (blocks && !h->pairs_blocks) {
		AOF_CUDA(cudaMalloc(&h->d_pblk, sizeof(aof_px4_block) * AOF_PX4_MAX_BLOCKS * h->pairs_cap));
		h->pairs_blocks = true;
	}
	return AOF_OK;
}

int aof_px4_compute_flow_pairs_device(aof_px4 *h, const uint8_t *d_prev, const uint8_t *d_cur, int n,
				      aof_px4_raw *d_raw, aof_px4_block *d_blocks, void *cuda_stream)
{
	if (!h || !d_prev || !d_cur || !d_raw || n < 0) { set_error("compute_flow_pairs_device: bad argument"); return AOF_ERR_INVALID; }
	if (((uintptr_t)d_prev | (uintptr_t)d_cur) & 15) {
		set_error("compute_flow_pairs_device: d_prev and d_cur must be 16-byte aligned");
		return AOF_ERR_INVALID;
	}
	AOF_CUDA(cudaSetDevice(h->device));
	AOF_CUDA(sad_flow_launch(h->p.search_size, d_prev, d_cur, h->g, n, d_raw, d_blocks, (cudaStream_t)cuda_stream));
	return AOF_OK;
}

int aof_px4_compute_flow_pairs(aof_px4 *h, const uint8_t *prev, const uint8_t *cur, int n, aof_px4_raw *raw,
			       aof_px4_block *blocks)
{
	if (!h || !prev || !cur || !raw || n < 0) { set_error("compute_flow_pairs: bad argument"); return AOF_ERR_INVALID; }
	if (n == 0) return AOF_OK;
	AOF_CUDA(cudaSetDevice(h->device));
	int rc = ensure_pairs(h, n, blocks != nullptr);
	if (rc != AOF_OK) return rc;
	const size_t total = h->frame_bytes * n;
	AOF_CUDA(cudaMemcpyAsync(h->d_pp, prev, total, cudaMemcpyHostToDevice, h->st));
	AOF_CUDA(cudaMemcpyAsync(h->d_pc, cur, total, cudaMemcpyHostToDevice, h->st));
	rc = aof_px4_compute_flow_pairs_device(h, h->d_pp, h->d_pc, n, h->d_praw, blocks ? h->d_pblk : nullptr, h->st);
	if (rc != AOF_OK) return rc;
	AOF_CUDA(cudaMemcpyAsync(raw, h->d_praw, sizeof(aof_px4_raw) * n, cudaMemcpyDeviceToHost, h->st));
	if (blocks)
		AOF_CUDA(cudaMemcpyAsync(blocks, h->d_pblk, sizeof(aof_px4_block) * AOF_PX4_MAX_BLOCKS * n,
					 cudaMemcpyDeviceToHost, h->st));
	AOF_CUDA(cudaStreamSynchronize(h->st));
	return AOF_OK;
}

} // extern "C"
