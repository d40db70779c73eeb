// tools/ubench.cu -- issue-rate micro-benchmark for the packed-byte instructions the SAD
// kernels live on (VABSDIFF4.U8.ACC, PRMT, SHF, LOP3, IMAD) on sm_100a.  Sets the
// integer-ALU roofline quoted in DESIGN.md.  Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#define ITERS 4096
#define UNROLL 12 // independent chains per thread

template <int OP> __global__ void __launch_bounds__(1024) k(uint32_t *out, uint32_t seed)
{
	uint32_t a[UNROLL], b = seed * 2654435761u + threadIdx.x, c = seed ^ 0x9e3779b9u;
#pragma unroll
	for (int i = 0; i < UNROLL; i++) a[i] = seed + i * 977 + threadIdx.x;
	for (int it = 0; it < ITERS; it++) {
#pragma unroll
		for (int i = 0; i < UNROLL; i++) {
			if (OP == 0) asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(a[i]) : "r"(b), "r"(c));
			if (OP == 1) asm volatile("prmt.b32 %0, %0, %1, %2;" : "+r"(a[i]) : "r"(b), "r"(c & 0x7777));
			if (OP == 2) asm volatile("shf.r.wrap.b32 %0, %0, %1, %2;" : "+r"(a[i]) : "r"(b), "r"(c));
			if (OP == 3) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(a[i]) : "r"(b), "r"(c));
			if (OP == 4) asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(a[i]) : "r"(b), "r"(c));
			if (OP == 5) { // 1 SAD : 1 IMAD mix (different pipes?)
				if (i & 1) asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(a[i]) : "r"(b), "r"(c));
				else asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(a[i]) : "r"(b), "r"(c));
			}
			if (OP == 6) { // 3 SAD : 1 PRMT mix (same pipe?)
				if ((i & 3) == 3) asm volatile("prmt.b32 %0, %0, %1, %2;" : "+r"(a[i]) : "r"(b), "r"(c & 0x7777));
				else asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(a[i]) : "r"(b), "r"(c));
			}
			if (OP == 7) asm volatile("add.u32 %0, %0, %1;" : "+r"(a[i]) : "r"(b));
			if (OP == 8) asm volatile("dp4a.u32.u32 %0, %1, %2, %0;" : "+r"(a[i]) : "r"(b), "r"(c));
			if (OP == 9) { // 1 DP4A : 1 PRMT (different pipes?)
				if (i & 1) asm volatile("prmt.b32 %0, %0, %1, %2;" : "+r"(a[i]) : "r"(b), "r"(c & 0x7777));
				else asm volatile("dp4a.u32.u32 %0, %1, %2, %0;" : "+r"(a[i]) : "r"(b), "r"(c));
			}
			if (OP == 10) { // 1 DP4A : 1 IMAD (same pipe?)
				if (i & 1) asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(a[i]) : "r"(b), "r"(c));
				else asm volatile("dp4a.u32.u32 %0, %1, %2, %0;" : "+r"(a[i]) : "r"(b), "r"(c));
			}
			// max then min: a chain of max alone is fused by ptxas into VIMNMX3 (two source ops per instruction), which made
			// the first version of this line read 125 lanes/clk; min(max(a, b), c) cannot be fused
			if (OP == 11) {
				asm volatile("max.u16x2 %0, %0, %1;" : "+r"(a[i]) : "r"(b));
				asm volatile("min.u16x2 %0, %0, %1;" : "+r"(a[i]) : "r"(c));
			}
			if (OP == 13) a[i] = __vimax3_u16x2(a[i], b, c + i);
			if (OP == 14) { // 2 VIMNMX : 1 VIMNMX3 (the FAST network's mix), unfusable order
				asm volatile("max.u16x2 %0, %0, %1;" : "+r"(a[i]) : "r"(b));
				asm volatile("min.u16x2 %0, %0, %1;" : "+r"(a[i]) : "r"(c));
				a[i] = __vimax3_u16x2(a[i], b, c + i);
			}
			if (OP == 15) { // 1 VIMNMX : 1 IMAD
				asm volatile("max.u16x2 %0, %0, %1;" : "+r"(a[i]) : "r"(b));
				asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(a[i]) : "r"(b), "r"(c));
			}
			if (OP == 12) asm volatile("mad.hi.u32 %0, %0, %1, %2;" : "+r"(a[i]) : "r"(b), "r"(c));
			if (OP == 16) { // packed fp16 min / max (HMNMX2): which pipe, what rate?
				asm volatile("max.f16x2 %0, %0, %1;" : "+r"(a[i]) : "r"(b));
				asm volatile("min.f16x2 %0, %0, %1;" : "+r"(a[i]) : "r"(c));
			}
			if (OP == 17) { // HMNMX2 : VIMNMX3 1:1 -- do they dual-issue?
				asm volatile("max.f16x2 %0, %0, %1;" : "+r"(a[i]) : "r"(b));
				a[i] = __vimin3_u16x2(a[i], c, b + i);
			}
			if (OP == 18) { // HMNMX2 : VIMNMX : VIMNMX3 2:1:1
				asm volatile("max.f16x2 %0, %0, %1;" : "+r"(a[i]) : "r"(b));
				asm volatile("min.u16x2 %0, %0, %1;" : "+r"(a[i]) : "r"(c));
				asm volatile("min.f16x2 %0, %0, %1;" : "+r"(a[i]) : "r"(c));
				a[i] = __vimax3_u16x2(a[i], b, c + i);
			}
		}
	}
	uint32_t s = 0;
#pragma unroll
	for (int i = 0; i < UNROLL; i++) s += a[i];
	out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int OP> void run(const char *name, int sms, int clk_khz, uint32_t *d)
{
	cudaEvent_t e0, e1;
	cudaEventCreate(&e0); cudaEventCreate(&e1);
	const int blocks = sms * 2, threads = 1024;
	k<OP><<<blocks, threads>>>(d, 1);
	cudaDeviceSynchronize();
	float best = 1e30f;
	for (int r = 0; r < 5; r++) {
		cudaEventRecord(e0);
		k<OP><<<blocks, threads>>>(d, r + 2);
		cudaEventRecord(e1);
		cudaEventSynchronize(e1);
		float ms; cudaEventElapsedTime(&ms, e0, e1);
		if (ms < best) best = ms;
	}
	double ops = (double)blocks * threads * ITERS * UNROLL * (OP == 11 || OP == 15 || OP == 16 || OP == 17 ? 2 : OP == 14 ? 3 : OP == 18 ? 4 : 1);
	double per_s = ops / (best * 1e-3);
	printf("{\"op\": \"%s\", \"ms\": %.4f, \"thread_ops_per_s\": %.4e, \"lanes_per_clk_per_sm_at_max_clk\": %.2f}\n", name, best,
	       per_s, per_s / sms / (clk_khz * 1e3));
}

int main()
{
	cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
	int clk = 0; cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
	printf("{\"gpu\": \"%s\", \"sms\": %d, \"max_clock_khz\": %d}\n", p.name, p.multiProcessorCount, clk);
	uint32_t *d; cudaMalloc(&d, 4u * 1024 * 2 * p.multiProcessorCount);
	run<0>("vabsdiff4.acc", p.multiProcessorCount, clk, d);
	run<1>("prmt", p.multiProcessorCount, clk, d);
	run<2>("shf.r.wrap", p.multiProcessorCount, clk, d);
	run<3>("lop3", p.multiProcessorCount, clk, d);
	run<4>("imad", p.multiProcessorCount, clk, d);
	run<5>("vabsdiff4+imad 1:1", p.multiProcessorCount, clk, d);
	run<6>("vabsdiff4+prmt 3:1", p.multiProcessorCount, clk, d);
	run<7>("iadd", p.multiProcessorCount, clk, d);
	run<8>("dp4a", p.multiProcessorCount, clk, d);
	run<9>("dp4a+prmt 1:1", p.multiProcessorCount, clk, d);
	run<10>("dp4a+imad 1:1", p.multiProcessorCount, clk, d);
	run<11>("vimnmx.u16x2 (max,min alternating)", p.multiProcessorCount, clk, d);
	run<16>("hmnmx2 (max,min f16x2)", p.multiProcessorCount, clk, d);
	run<17>("hmnmx2+vimnmx3 1:1", p.multiProcessorCount, clk, d);
	run<18>("hmnmx2+vimnmx+vimnmx3 2:1:1", p.multiProcessorCount, clk, d);
	run<12>("imad.hi", p.multiProcessorCount, clk, d);
	run<13>("vimnmx3.u16x2", p.multiProcessorCount, clk, d);
	run<14>("vimnmx+vimnmx3 2:1", p.multiProcessorCount, clk, d);
	run<15>("vimnmx+imad 1:1", p.multiProcessorCount, clk, d);
	return 0;
}
