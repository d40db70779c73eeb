// Reproducer: nvcc/ptxas 12.9 for sm_100a folds max(m, -b) into VIMNMX3 and loses the negation (variant 0 and 1
// both return max(a, b)).  Found while bringing up lk_fast_kernel; the product code materialises e = -d instead.
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cstdint>
__device__ __forceinline__ bool has_arc9(unsigned m){ m |= m << 16; const unsigned a = m & (m >> 1), b = a & (a >> 2), c = b & (b >> 4); return (c & (m >> 8) & 0xFFFFu) != 0; }
template <int V> __device__ int fs(const uint8_t *p, int pitch, int thr, int *dbg)
{
	const int c = p[0];
	int d[16];
	d[0] = c - p[3 * pitch];      d[1] = c - p[3 * pitch + 1];  d[2] = c - p[2 * pitch + 2];  d[3] = c - p[pitch + 3];
	d[4] = c - p[3];              d[5] = c - p[-pitch + 3];     d[6] = c - p[-2 * pitch + 2]; d[7] = c - p[-3 * pitch + 1];
	d[8] = c - p[-3 * pitch];     d[9] = c - p[-3 * pitch - 1]; d[10] = c - p[-2 * pitch - 2]; d[11] = c - p[-pitch - 3];
	d[12] = c - p[-3];            d[13] = c - p[pitch - 3];     d[14] = c - p[2 * pitch - 2]; d[15] = c - p[3 * pitch - 1];
	if (dbg) for (int k = 0; k < 16; k++) dbg[k] = d[k];
	unsigned mp = 0, mn = 0;
#pragma unroll
	for (int k = 0; k < 16; k++) { mp |= (unsigned)(d[k] > thr) << k; mn |= (unsigned)(d[k] < -thr) << k; }
	if (dbg) { dbg[16] = mp; dbg[17] = mn; }
	if (!has_arc9(mp) && !has_arc9(mn)) return 0;
	int m = -1000;
	if (V == 0) {
		int lo2[16], hi2[16], lo4[16], hi4[16];
#pragma unroll
		for (int k = 0; k < 16; k++) { lo2[k] = min(d[k], d[(k + 1) & 15]); hi2[k] = max(d[k], d[(k + 1) & 15]); }
#pragma unroll
		for (int k = 0; k < 16; k++) { lo4[k] = min(lo2[k], lo2[(k + 2) & 15]); hi4[k] = max(hi2[k], hi2[(k + 2) & 15]); }
#pragma unroll
		for (int k = 0; k < 16; k++) {
			const int lo9 = min(min(lo4[k], lo4[(k + 4) & 15]), d[(k + 8) & 15]);
			const int hi9 = max(max(hi4[k], hi4[(k + 4) & 15]), d[(k + 8) & 15]);
			m = max(m, max(lo9, -hi9));
		}
	} else {
#pragma unroll
		for (int k = 0; k < 16; k++) {
			int a = d[k], b = d[k];
#pragma unroll
			for (int i = 1; i < 9; i++) { const int v = d[(k + i) & 15]; a = v < a ? v : a; b = v > b ? v : b; }
			m = a > m ? a : m; m = -b > m ? -b : m;
		}
	}
	return m - 1;
}
template <int V> __global__ void raw_kernel(const uint8_t *img, int W, int H, uint8_t *out, int *dbg, int dy, int dx)
{
	int x = blockIdx.x * 32 + threadIdx.x, y = blockIdx.y;
	if (x < 3 || x >= W - 3 || y < 3 || y >= H - 3) return;
	out[y * W + x] = (uint8_t)fs<V>(img + y * W + x, W, 10, (y == dy && x == dx) ? dbg : nullptr);
}
static const int rdx[16] = {0, 1, 2, 3, 3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1};
static const int rdy[16] = {3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1, 0, 1, 2, 3};
static int fast_m(const uint8_t *p, int stride, int *dd){int d[25];const int c=p[0];for(int k=0;k<16;k++){d[k]=c-p[rdy[k]*stride+rdx[k]]; if(dd)dd[k]=d[k];}for(int k=16;k<25;k++)d[k]=d[k-16];int m=-1000;for(int k=0;k<16;k++){int a=d[k],b=d[k];for(int i=1;i<9;i++){if(d[k+i]<a)a=d[k+i];if(d[k+i]>b)b=d[k+i];}if(a>m)m=a;if(-b>m)m=-b;}return m;}
template <int V> void run(const std::vector<uint8_t> &img, int W, int H, uint8_t *d_img, uint8_t *d_raw, int *d_dbg)
{
	std::vector<uint8_t> got(W * H, 0);
	cudaMemset(d_raw, 0, W * H);
	raw_kernel<V><<<dim3(W / 32, H), 32>>>(d_img, W, H, d_raw, d_dbg, 3, 8);
	cudaMemcpy(got.data(), d_raw, W * H, cudaMemcpyDeviceToHost);
	int dbg[18]; cudaMemcpy(dbg, d_dbg, sizeof dbg, cudaMemcpyDeviceToHost);
	int bad = 0;
	for (int y = 3; y < H - 3; y++) for (int x = 3; x < W - 3; x++) {
		int m = fast_m(&img[y * W + x], W, nullptr); int want = m > 10 ? m - 1 : 0;
		if (want != got[y * W + x]) bad++;
	}
	int hd[16]; int m = fast_m(&img[3 * W + 8], W, hd);
	printf("variant %d: %d mismatches; at (3,8) host m=%d dev=%d\n  host d:", V, bad, m, got[3 * W + 8]);
	for (int k = 0; k < 16; k++) printf(" %d", hd[k]);
	printf("\n  dev  d:");
	for (int k = 0; k < 16; k++) printf(" %d", dbg[k]);
	printf("  mp %x mn %x\n", dbg[16], dbg[17]);
}
int main()
{
	const int W = 128, H = 128;
	std::vector<uint8_t> img(W * H);
	srand(3);
	for (auto &v : img) v = 100 + rand() % 60;
	uint8_t *d_img, *d_raw; int *d_dbg;
	cudaMalloc(&d_img, W * H); cudaMalloc(&d_raw, W * H); cudaMalloc(&d_dbg, 18 * 4);
	cudaMemcpy(d_img, img.data(), W * H, cudaMemcpyHostToDevice);
	run<0>(img, W, H, d_img, d_raw, d_dbg);
	run<1>(img, W, H, d_img, d_raw, d_dbg);
	return 0;
}
