// Shared device helpers for the sm_100a kernels of libaof_b200.so.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace aof {

// ---- mbarrier + 1-D bulk (TMA engine) global->shared copy -----------------
__device__ __forceinline__ uint32_t smem_u32(const void *p)
{
	return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

// Programmatic dependent launch (launch_k(pdl = true, ...) in aof_internal.h): the kernels of an LK step are launched with
// programmatic stream serialisation, so the NEXT kernel is set up on the device while this one runs and starts the moment
// this one's last CTA exits.  Every kernel of such a chain calls pdl_enter() before it touches global memory: it blocks
// until the kernel before it has COMPLETED and its writes are visible -- nothing is read or written early, and completion
// stays transitive along the chain.  Without the launch attribute the instruction does nothing.
// Measured (r2P, 512 LK streams, 7 launches per step): plain launches 0.4835 ms per step, this 0.4664 ms.  Letting the
// dependents become resident early as well (griddepcontrol.launch_dependents at the top of every kernel,
// AOF_PDL_TRIGGER=1) was SLOWER, 0.502-0.505 ms: the blocked CTAs of the next kernel hold registers and warp slots in the
// tail of the running one.
#ifndef AOF_PDL_TRIGGER
#define AOF_PDL_TRIGGER 0
#endif
__device__ __forceinline__ void pdl_enter()
{
#if AOF_PDL_TRIGGER
	asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
#endif
	asm volatile("griddepcontrol.wait;" ::: "memory");
}

__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count)
{
	asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
	asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}

__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes)
{
	asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
		     : "memory");
}

__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity)
{
	asm volatile(
		"{\n"
		".reg .pred p;\n"
		"WAIT_%=:\n"
		"mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
		"@p bra DONE_%=;\n"
		"bra WAIT_%=;\n"
		"DONE_%=:\n"
		"}\n" ::"r"(smem_u32(bar)),
		"r"(parity)
		: "memory");
}

// size and both addresses must be multiples of 16 B
__device__ __forceinline__ void bulk_g2s(void *dst_smem, const void *src_gmem, uint32_t bytes, uint64_t *bar)
{
	asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
			     smem_u32(dst_smem)),
		     "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
		     : "memory");
}

// ---- packed-byte arithmetic ------------------------------------------------
// USADA8: acc + sum_i |a_i - b_i|   (one VABSDIFF4.U8.ACC on sm_100a)
// (inline PTX: with the __vsadu4 intrinsic nvcc splits the accumulate into a separate IADD3)
__device__ __forceinline__ uint32_t sad4(uint32_t a, uint32_t b, uint32_t acc)
{
	uint32_t d;
	asm("vabsdiff4.u32.u32.u32.add %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(acc));
	return d;
}

// UHADD8: per-byte (a + b) >> 1, truncating
__device__ __forceinline__ uint32_t havg4(uint32_t a, uint32_t b)
{
	return (a & b) + (((a ^ b) & 0xFEFEFEFEu) >> 1);
}

// bytes [o, o+4) of the little-endian 8-byte string lo:hi, 0 <= o <= 4
__device__ __forceinline__ uint32_t bytes_at(uint32_t lo, uint32_t hi, uint32_t o)
{
	return __byte_perm(lo, hi, 0x3210u + 0x1111u * o);
}

} // namespace aof
