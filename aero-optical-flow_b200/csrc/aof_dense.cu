// C-ABI host side of the dense SAD field (include/aof.h, aof_sad_dense_*).
#include <new>

#include "aof_internal.h"
#include "aof_host.h"

using namespace aof;

struct aof_sad_dense {
	aof_dense_params p;
	int device = 0;
	DenseGeom g{};
	cudaStream_t st = nullptr;
	uint8_t *d_prev = nullptr, *d_cur = nullptr;
	aof_dense_block *d_field = nullptr;
	int cap = 0;
};

extern "C" {

int aof_sad_dense_create(const aof_dense_params *p, int device, aof_sad_dense **out)
{
	if (!p || !out) { set_error("aof_sad_dense_create: bad argument"); return AOF_ERR_INVALID; }
	*out = nullptr;
	const int W = p->image_width, H = p->image_height, S = p->search_size;
	if (S != 4 && S != 6 && S != 8) { set_error("search_size %d not built (4, 6, 8)", S); return AOF_ERR_UNSUPPORTED; }
	if ((W & 3) || W < 8 + 2 * (S + 1) + 8 || H < 8 + 2 * (S + 1) + 8 || ((size_t)W * H & 3)) {
		set_error("image %dx%d unsupported (need W %% 4 == 0 and room for one block with its search window)", W, H);
		return AOF_ERR_UNSUPPORTED;
	}
	int rc = select_device(device);
	if (rc != AOF_OK) return rc;
	aof_sad_dense *h = new (std::nothrow) aof_sad_dense();
	if (!h) return AOF_ERR_INVALID;
	h->p = *p; h->device = device;
	DenseGeom &g = h->g;
	g.width = W; g.height = H; g.frame_stride = (size_t)W * H;
	g.i_lo = (S + 1 + 7) / 8; g.i_hi = (W - S - 9) / 8;
	g.j_lo = (S + 1 + 7) / 8; g.j_hi = (H - S - 9) / 8;
	g.nbx = g.i_hi - g.i_lo + 1; g.nby = g.j_hi - g.j_lo + 1;
	g.feat_thr = (uint32_t)p->flow_feature_threshold; g.val_thr = (uint32_t)p->flow_value_threshold;
	if (g.nbx < 1 || g.nby < 1) { delete h; set_error("no block fits"); return AOF_ERR_UNSUPPORTED; }
	cudaError_t e = cudaStreamCreateWithFlags(&h->st, cudaStreamNonBlocking);
	if (e != cudaSuccess) { set_error("aof_sad_dense_create: %s", cudaGetErrorString(e)); delete h; return AOF_ERR_CUDA; }
	*out = h;
	return AOF_OK;
}

void aof_sad_dense_destroy(aof_sad_dense *h)
{
	if (!h) return;
	cudaSetDevice(h->device);
	if (h->st) cudaStreamSynchronize(h->st);
	cudaFree(h->d_prev); cudaFree(h->d_cur); cudaFree(h->d_field);
	if (h->st) cudaStreamDestroy(h->st);
	delete h;
}

int aof_sad_dense_grid(const aof_sad_dense *h, int *nbx, int *nby, int *x0, int *y0)
{
	if (!h) return AOF_ERR_INVALID;
	if (nbx) *nbx = h->g.nbx;
	if (nby) *nby = h->g.nby;
	if (x0) *x0 = 8 * h->g.i_lo;
	if (y0) *y0 = 8 * h->g.j_lo;
	return AOF_OK;
}

int aof_sad_dense_compute_device(aof_sad_dense *h, const uint8_t *d_prev, const uint8_t *d_cur, int n_pairs,
				 aof_dense_block *d_field, void *cuda_stream)
{
	if (!h || !d_prev || !d_cur || !d_field || n_pairs < 0 || n_pairs > 65535) { set_error("aof_sad_dense_compute_device: bad argument"); return AOF_ERR_INVALID; }
	if ((reinterpret_cast<uintptr_t>(d_prev) | reinterpret_cast<uintptr_t>(d_cur)) & 3) { set_error("frames must be 4-byte aligned"); return AOF_ERR_INVALID; }
	AOF_CUDA(cudaSetDevice(h->device));
	AOF_CUDA(sad_dense_launch(h->p.search_size, d_prev, d_cur, h->g, n_pairs, d_field, (cudaStream_t)cuda_stream));
	return AOF_OK;
}

int aof_sad_dense_compute(aof_sad_dense *h, const uint8_t *prev, const uint8_t *cur, int n_pairs, aof_dense_block *field)
{
	if (!h || !prev || !cur || !field || n_pairs < 0 || n_pairs > 65535) { set_error("aof_sad_dense_compute: bad argument"); return AOF_ERR_INVALID; }
	if (n_pairs == 0) return AOF_OK;
	AOF_CUDA(cudaSetDevice(h->device));
	const size_t fb = h->g.frame_stride * n_pairs, nb = (size_t)h->g.nbx * h->g.nby * n_pairs;
	if (n_pairs > h->cap) {
		cudaFree(h->d_prev); cudaFree(h->d_cur); cudaFree(h->d_field);
		h->d_prev = h->d_cur = nullptr; h->d_field = nullptr; h->cap = 0;
		AOF_CUDA(cudaMalloc(&h->d_prev, fb));
		AOF_CUDA(cudaMalloc(&h->d_cur, fb));
		AOF_CUDA(cudaMalloc(&h->d_field, sizeof(aof_dense_block) * nb));
		h->cap = n_pairs;
	}
	AOF_CUDA(cudaMemcpyAsync(h->d_prev, prev, fb, cudaMemcpyHostToDevice, h->st));
	AOF_CUDA(cudaMemcpyAsync(h->d_cur, cur, fb, cudaMemcpyHostToDevice, h->st));
	int rc = aof_sad_dense_compute_device(h, h->d_prev, h->d_cur, n_pairs, h->d_field, h->st);
	if (rc != AOF_OK) return rc;
	AOF_CUDA(cudaMemcpyAsync(field, h->d_field, sizeof(aof_dense_block) * nb, cudaMemcpyDeviceToHost, h->st));
	AOF_CUDA(cudaStreamSynchronize(h->st));
	return AOF_OK;
}

} // extern "C"
