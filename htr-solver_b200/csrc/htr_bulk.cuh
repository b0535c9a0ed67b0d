// TMA (1-D bulk async copy) staging for the tiles of the fast path.
//
// A tile row of an AoS field ([x][component], x unit-stride) is ONE contiguous span of global memory, so it
// is moved with one `cp.async.bulk.shared::cluster.global` instruction issued by one lane; completion is
// counted in bytes on an mbarrier in shared memory.  No per-element address arithmetic, no LSU
// wavefronts, and all rows of all fields of the tile are in flight at once.  The tile keeps the field's
// own layout in shared memory ([row][x][component]); with an odd component count (3, 5, 9; 1) the strided
// 8-byte reads of the flux stencils are bank-conflict free.
//
// Requirements checked on the host (htr_kernels.cu: bulkEligible): field base and y/z strides multiples of
// 16 bytes, x stride = 8*ncomp.  Checked per CTA on the device: the row span [xs, xs+W) starts on an even
// cell (16-byte alignment for odd ncomp), lies inside the allocation and is not cut by the periodic wrap.
// CTAs that fail use cooperative per-element loads into the same layout.
#pragma once
#include "htr_access.cuh"

namespace htr {
namespace bulk {

__device__ __forceinline__ uint32_t sAddr(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbarInit(uint64_t *bar, int count) {
   asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(sAddr(bar)), "r"(count) : "memory");
   asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbarExpectTx(uint64_t *bar, uint32_t bytes) {
   asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(sAddr(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void copyRow(void *dst, const void *src, uint32_t bytes, uint64_t *bar) {
   asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                :: "r"(sAddr(dst)), "l"(src), "r"(bytes), "r"(sAddr(bar)) : "memory");
}
__device__ __forceinline__ bool mbarTry(uint64_t *bar, uint32_t parity) {
   uint32_t ok;
   asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                : "=r"(ok) : "r"(sAddr(bar)), "r"(parity) : "memory");
   return ok != 0;
}
__device__ __forceinline__ void mbarWait(uint64_t *bar, uint32_t parity) {
   while (!mbarTry(bar, parity)) { }
}

// Cooperative per-element copy of one field's box into the raw layout (fallback of copyRow):
// element n of the box = (row rr, column cc, component k); rows/columns map to global coordinates through
// the caller's functor `coord(rr, cc, Pt&) -> bool` (false: cell not needed / does not exist).
template<int NCOMP, int W, int ROWS, int NT, class Coord>
__device__ __forceinline__ void loadBoxElements(double *dst, const htr_acc &a, const int tid, const Coord &coord) {
   constexpr int ROWE = W*NCOMP;
   for (int n = tid; n < ROWS*ROWE; n += NT) {
      const int rr = n/ROWE, rem = n - rr*ROWE;
      const int cc = rem/NCOMP, k = rem - cc*NCOMP;
      Pt q;
      if (coord(rr, cc, q)) dst[n] = ldc<double>(a, q, k);
   }
}

} // namespace bulk
} // namespace htr
