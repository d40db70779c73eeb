// Host-only helpers for the C-ABI translation units.
#pragma once
#include <cstdarg>
#include <cuda_runtime.h>

#include "../../include/aof.h"

namespace aof {
void set_error(const char *fmt, ...);
const char *last_error();
int select_device(int device);
// host -> device copy of `bytes` on `st`: directly if `src` is pinned (cudaHostAlloc / cudaHostRegister),
// else through the pinned staging buffer `stage` (>= bytes) in chunks so the CPU copy overlaps the DMA
// true when p is page-locked host memory (cudaHostAlloc / cudaHostRegister): DMA can target it directly
bool host_pinned(const void *p);
int stage_h2d(void *d_dst, const void *src, size_t bytes, void *stage, cudaStream_t st);
} // namespace aof

// CUDA errors never cross the ABI as exceptions: record the text, return a code
#define AOF_CUDA(expr)                                                                         \
	do {                                                                                   \
		cudaError_t _e = (expr);                                                       \
		if (_e != cudaSuccess) {                                                       \
			aof::set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #expr, cudaGetErrorString(_e)); \
			return AOF_ERR_CUDA;                                                   \
		}                                                                              \
	} while (0)
