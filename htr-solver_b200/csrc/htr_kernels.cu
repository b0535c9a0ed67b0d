// Kernels and C ABI (include/htr_b200.h) of the B200-native HTR right-hand-side engine.
// Built once per mixture for sm_100a (see htr-solver_b200/build.py).
//
// Kernel shapes (x is the unit-stride direction of every field):
//   flux kernels   one thread per FACE of a (BX,BY,BZ) block of faces; the face fluxes are
//                  exchanged through shared memory and each thread then updates the cell on
//                  the minus side of its face:  Conserved_t[p] += m[p]*(F[p+1/2] - F[p-1/2])
//                  (src/prometeo_rhs.inl:884, :968, :1047, :1277).  A block of B faces along the
//                  flux direction updates B-1 cells.
//   point kernels  one thread per cell, x fastest.
#include <cuda_runtime.h>
#include <stdio.h>
#include <string.h>
#include <math.h>
#include <initializer_list>
#include <utility>
#include "htr_euler.cuh"
#include "htr_diffusion.cuh"
#include "htr_pointwise.cuh"
#include "htr_fast.cuh"
#include "htr_fast_bulk.cuh"
#include "htr_march.cuh"
#include "htr_diff_async.cuh"

namespace htr {

// ---------------------------------------------------------------------------------------------
// flux kernels
// ---------------------------------------------------------------------------------------------
enum { FLUX_EULER = 0, FLUX_DIFFUSION = 1 };

struct EulerParams { double RK0, RK1, dt; };

template<int dir> struct FaceBlock;            // faces per block in (x, y, z)
template<> struct FaceBlock<0> { static constexpr int X = 64, Y = 4, Z = 1; };
template<> struct FaceBlock<1> { static constexpr int X = 32, Y = 8, Z = 1; };
template<> struct FaceBlock<2> { static constexpr int X = 32, Y = 1, Z = 8; };

template<int dir, int SCHEME>
HTR_DEV void eulerFace(double *Flux, const htr_fluid &f, const Pt &p, int t, const htr_rect &B, const EulerParams &ep) {
   if (SCHEME == HTR_SCHEME_SKEWSYMMETRIC) {
      kgFaceFlux<dir>(Flux, f, p, t, B);
      return;
   }
   double lim_f = 0.0;
   if (nSpec > 1)
      lim_f = ep.dt*(ld<double>(m_eOf<dir>(f), p) + ld<double>(m_eOf<dir>(f), shift<dir>(p, offP1(t), B)));
   if (SCHEME == HTR_SCHEME_HYBRID) {
      const htr_acc &ss = sensorOf<dir>(f);
      const bool smooth = ld<uint8_t>(ss, shift<dir>(p, offM1(t), B)) && ld<uint8_t>(ss, p) &&
                          ld<uint8_t>(ss, shift<dir>(p, offP1(t), B));
      if (smooth) kgFaceFlux<dir>(Flux, f, p, t, B);
      else        tenoFaceFlux<dir, OP_TENO>(Flux, f, p, t, B, ep.RK0, ep.RK1, lim_f);
   }
   else if (SCHEME == HTR_SCHEME_TENO)    tenoFaceFlux<dir, OP_TENO>(Flux, f, p, t, B, ep.RK0, ep.RK1, lim_f);
   else if (SCHEME == HTR_SCHEME_TENOA)   tenoFaceFlux<dir, OP_TENOA>(Flux, f, p, t, B, ep.RK0, ep.RK1, lim_f);
   else if (SCHEME == HTR_SCHEME_TENOLAD) tenoFaceFlux<dir, OP_TENOLAD>(Flux, f, p, t, B, ep.RK0, ep.RK1, lim_f);
}

} // namespace htr
#include "htr_teno_coop.cuh"     // eight lanes per TENO face (uses EulerParams, kgFaceFlux and the accessors above)
namespace htr {
// htr_set_option("coop", 1): the lane-cooperative TENO kernels (eight lanes per face) instead of the one-thread-per-face ones.
// Off by default: bit-identical (tests/test_coop_gpu.py) but SLOWER on B200 (profiles/r2_v21_*): 128^3 ConstPropMix TENO 8.35 vs
// 4.14 ms per RHS, air256_front 29.6 vs 10.6 ms, air512_front 152.0 vs 71.6 ms -- see htr_teno_coop.cuh for why.
static bool g_coop = false;
static int numSMs();

// KIND = FLUX_EULER (SCHEME selects the reconstruction) or FLUX_DIFFUSION
// MASKED: only the cells flagged in `mask` (dense over the tile, see htr_fast.cuh) are updated and only
// the faces they need are evaluated; a block without flagged cells returns at once.
// dense index of global cell c in the tile's mask (same enumeration as fast::maskIndex on relative coordinates)
HTR_DEV int64_t maskIndexG(const Pt &c, const htr_rect &tile) {
   const int64_t nx = tile.hi[0] - tile.lo[0] + 1, ny = tile.hi[1] - tile.lo[1] + 1;
   return ((c.z - tile.lo[2])*ny + (c.y - tile.lo[1]))*nx + (c.x - tile.lo[0]);
}

template<int dir, int KIND, int SCHEME, bool MASKED>
HTR_DEV void fluxBlock(const unsigned bx, const unsigned by, const unsigned bz, const htr_fluid &f, const htr_rect &tile,
                       const htr_rect &B, const EulerParams &ep, const uint8_t *__restrict__ mask) {
   using FB = FaceBlock<dir>;
   constexpr int NT = FB::X*FB::Y*FB::Z;
   __shared__ double sF[nEq][NT];

   const int tid = threadIdx.x + FB::X*(threadIdx.y + FB::Y*threadIdx.z);
   const int fd = (dir == 0) ? threadIdx.x : ((dir == 1) ? threadIdx.y : threadIdx.z);   // face index along dir
   constexpr int CX = FB::X - (dir == 0), CY = FB::Y - (dir == 1), CZ = FB::Z - (dir == 2); // cells per block
   // cell on the minus side of this thread's face; fd == 0 is the face below the block's first cell
   Pt c;
   c.x = tile.lo[0] + (int64_t)bx*CX + threadIdx.x - (dir == 0);
   c.y = tile.lo[1] + (int64_t)by*CY + threadIdx.y - (dir == 1);
   c.z = tile.lo[2] + (int64_t)bz*CZ + threadIdx.z - (dir == 2);
   const bool inCross = (dir == 0 || c.x <= tile.hi[0]) && (dir == 1 || c.y <= tile.hi[1]) && (dir == 2 || c.z <= tile.hi[2]);
   const int64_t cd = (dir == 0) ? c.x : ((dir == 1) ? c.y : c.z);
   const bool active = inCross && (cd <= tile.hi[dir]);
   bool need = active, mine = true;
   if (MASKED) {
      mine = false;
      bool next = false;
      if (active) {
         if (cd >= tile.lo[dir]) mine = mask[maskIndexG(c, tile)] != 0;
         if (cd < tile.hi[dir]) {
            Pt n = c;
            ((dir == 0) ? n.x : ((dir == 1) ? n.y : n.z)) += 1;
            next = mask[maskIndexG(n, tile)] != 0;
         }
      }
      need = mine || next;
      if (!__syncthreads_or(need)) return;
   }

   if (need) {
      Pt p = c;
      if (fd == 0) {
         // first minus face of the block: the face owned by the neighbour below the first cell
         Pt p0 = c;
         ((dir == 0) ? p0.x : ((dir == 1) ? p0.y : p0.z)) += 1;
         p = shift<dir>(p0, offM1(ld<int>(nTypeOf<dir>(f), p0)), B);
      }
      const int t = ld<int>(nTypeOf<dir>(f), p);
      double Flux[nEq];
      if (KIND == FLUX_EULER) eulerFace<dir, SCHEME>(Flux, f, p, t, B, ep);
      else                    diffusionFaceFlux<dir>(Flux, f, p, B, fd != 0 && cd == tile.lo[dir]);
      HTR_UNROLL
      for (int i = 0; i < nEq; i++) sF[i][tid] = Flux[i];
   }
   __syncthreads();
   if (active && fd > 0 && mine) {
      constexpr int below = (dir == 0) ? 1 : ((dir == 1) ? FB::X : FB::X*FB::Y);
      const double m = ld<double>((KIND == FLUX_EULER) ? m_eOf<dir>(f) : m_dOf<dir>(f), c);
      double *Ct = at<double>(f.Conserved_t, c);
      HTR_UNROLL
      for (int i = 0; i < nEq; i++) Ct[i] += m*(sF[i][tid] - sF[i][tid - below]);
   }
}

// resident CTAs per SM the register allocation of the generic per-face kernels aims at (1 = up to 255 registers)
#ifndef HTR_FLUX_MINB
#define HTR_FLUX_MINB 1
#endif
#ifndef HTR_LIST_MINB
#define HTR_LIST_MINB 1
#endif
template<int dir, int KIND, int SCHEME>
__global__ void __launch_bounds__(256, HTR_FLUX_MINB)
flux_divergence_kernel(const htr_fluid f, const htr_rect tile, const htr_rect B, const EulerParams ep) {
   fluxBlock<dir, KIND, SCHEME, false>(blockIdx.x, blockIdx.y, blockIdx.z, f, tile, B, ep, nullptr);
}

// Generic flux pass over the cells the staged kernels deferred.  Persistent grid: nothing to do (the usual
// case of a smooth periodic box) costs one flag load per CTA; otherwise the CTAs stride over the blocks of the
// tile and skip those without flagged cells.
template<int dir, int KIND, int SCHEME>
__global__ void __launch_bounds__(256)
masked_flux_kernel(const htr_fluid f, const htr_rect tile, const htr_rect B, const EulerParams ep,
                    const uint8_t *__restrict__ mask, const int *__restrict__ flag, const unsigned gx, const unsigned gy, const unsigned gz) {
   if (*flag == 0) return;
   const unsigned total = gx*gy*gz;
   for (unsigned vb = blockIdx.x; vb < total; vb += gridDim.x) {
      fluxBlock<dir, KIND, SCHEME, true>(vb % gx, (vb/gx) % gy, vb/(gx*gy), f, tile, B, ep, mask);
      __syncthreads();
   }
}

template<int dir, int KIND, int SCHEME>
static cudaError_t launchFlux(const htr_fluid &f, const htr_rect &tile, const htr_rect &B, const EulerParams &ep, cudaStream_t st) {
   using FB = FaceBlock<dir>;
   const int64_t n[3] = {tile.hi[0] - tile.lo[0] + 1, tile.hi[1] - tile.lo[1] + 1, tile.hi[2] - tile.lo[2] + 1};
   if (n[0] <= 0 || n[1] <= 0 || n[2] <= 0) return cudaSuccess;
   const int C[3] = {FB::X - (dir == 0), FB::Y - (dir == 1), FB::Z - (dir == 2)};
   dim3 grid((unsigned)((n[0] + C[0] - 1)/C[0]), (unsigned)((n[1] + C[1] - 1)/C[1]), (unsigned)((n[2] + C[2] - 1)/C[2]));
   dim3 block(FB::X, FB::Y, FB::Z);
   if constexpr (KIND == FLUX_EULER && (SCHEME == HTR_SCHEME_TENO || SCHEME == HTR_SCHEME_TENOA || SCHEME == HTR_SCHEME_TENOLAD)) {
      if (g_coop) {       // runs of 8 cells along dir, one group of eight lanes each (htr_teno_coop.cuh)
         constexpr int RUN = 8;
         const int64_t runs = ((n[dir] + RUN - 1)/RUN)*(n[0]*n[1]*n[2]/n[dir]);
         int perSM = 1;
         if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSM, coop::euler_coop_kernel<dir, SCHEME, false>, coop::GL*coop::GROUPS, 0) != cudaSuccess || perSM < 1) perSM = 1;
         const int64_t want = (runs + coop::GROUPS - 1)/coop::GROUPS, cap = (int64_t)numSMs()*perSM*8;
         coop::euler_coop_kernel<dir, SCHEME, false><<<(unsigned)(want < cap ? want : cap), coop::GL*coop::GROUPS, 0, st>>>(f, tile, B, ep, nullptr, nullptr, RUN);
         return cudaGetLastError();
      }
   }
   flux_divergence_kernel<dir, KIND, SCHEME><<<grid, block, 0, st>>>(f, tile, B, ep);
   return cudaGetLastError();
}

// ---- per-device library state -------------------------------------------------------------------------------
// One Legion rank drives several GPUs (one GPU processor per device, src/prometeo_mapper.cc:231-262), so everything the
// library caches is keyed on the CURRENT device: the "mixture loaded" flag (the __constant__ block itself is per
// device already), the SM count, the cudaFuncSetAttribute guards, the scratch buffers and the host-pipeline streams.
constexpr int HTR_MAX_DEVICES = 64;
static inline int curDev() { int d = 0; cudaGetDevice(&d); return (d >= 0 && d < HTR_MAX_DEVICES) ? d : 0; }
static int g_num_sms[HTR_MAX_DEVICES] = {};
static int numSMs() {
   const int d = curDev();
   if (g_num_sms[d] == 0 &&
       (cudaDeviceGetAttribute(&g_num_sms[d], cudaDevAttrMultiProcessorCount, d) != cudaSuccess || g_num_sms[d] <= 0)) g_num_sms[d] = 148;
   return g_num_sms[d];
}
// true once per (device, slot): the caller then raises the dynamic shared-memory limits of its kernels on that device
static bool firstUseOnDevice(int slot) {
   static bool done[8][HTR_MAX_DEVICES] = {};
   const int d = curDev();
   if (done[slot][d]) return false;
   done[slot][d] = true;
   return true;
}

// ---- deferred cells as a compacted list --------------------------------------------------------------------------
// The block scan (masked_flux_kernel) evaluates a 255-register TENO face for every lane of a block that holds one flagged
// cell; flagged cells of a shock or a front form thin sheets, so most of those lanes idle (air256_front: 13 of 21.7 ms for
// 1.8 % of the cells).  Instead the mask is compacted once into a list of cell indices and one thread takes one flagged
// cell: both its faces, then the reference's  Conserved_t[p] += m[p]*(F[p+1/2] - F[p-1/2]).  Same face functions, same bits.
__global__ void __launch_bounds__(256) compact_mask_kernel(const uint8_t *__restrict__ mask, const int64_t cells, const int *__restrict__ flag,
                                                           int *__restrict__ list, int *__restrict__ count) {
   if (*flag == 0) return;
   const int lane = threadIdx.x & 31;
   for (int64_t base = ((int64_t)blockIdx.x*blockDim.x + threadIdx.x) - lane; base < cells; base += (int64_t)gridDim.x*blockDim.x) {
      const int64_t c = base + lane;
      const bool on = c < cells && mask[c] != 0;
      const unsigned b = __ballot_sync(0xffffffffu, on);
      if (b == 0) continue;
      int pos = 0;
      if (lane == 0) pos = atomicAdd(count, __popc(b));
      pos = __shfl_sync(0xffffffffu, pos, 0);
      if (on) list[pos + __popc(b & ((1u << lane) - 1))] = (int)c;
   }
}

template<int dir, int KIND, int SCHEME>
__global__ void __launch_bounds__(128, HTR_LIST_MINB)
listed_flux_kernel(const htr_fluid f, const htr_rect tile, const htr_rect B, const EulerParams ep,
                   const int *__restrict__ list, const int *__restrict__ count) {
   const int n = *count;
   const int64_t nx = tile.hi[0] - tile.lo[0] + 1, ny = tile.hi[1] - tile.lo[1] + 1;
   for (int t = blockIdx.x*blockDim.x + threadIdx.x; t < n; t += gridDim.x*blockDim.x) {
      const int64_t idx = list[t];
      Pt c;
      c.x = tile.lo[0] + idx % nx; c.y = tile.lo[1] + (idx/nx) % ny; c.z = tile.lo[2] + idx/(nx*ny);
      const int tc = ld<int>(nTypeOf<dir>(f), c);
      const Pt pm = shift<dir>(c, offM1(tc), B);             // owner of the cell's minus face (fluxBlock: fd == 0)
      const int64_t cd = (dir == 0) ? c.x : ((dir == 1) ? c.y : c.z);
      double Fp[nEq], Fm[nEq];
      if (KIND == FLUX_EULER) {
         eulerFace<dir, SCHEME>(Fm, f, pm, ld<int>(nTypeOf<dir>(f), pm), B, ep);
         eulerFace<dir, SCHEME>(Fp, f, c, tc, B, ep);
      } else {
         // the face below the tile's second cell is the span's first plus face (htr_diffusion.cuh: spanFirst)
         diffusionFaceFlux<dir>(Fm, f, pm, B, cd == tile.lo[dir] + 1);
         diffusionFaceFlux<dir>(Fp, f, c, B, cd == tile.lo[dir]);
      }
      const double m = ld<double>((KIND == FLUX_EULER) ? m_eOf<dir>(f) : m_dOf<dir>(f), c);
      double *Ct = at<double>(f.Conserved_t, c);
      HTR_UNROLL
      for (int i = 0; i < nEq; i++) Ct[i] += m*(Fp[i] - Fm[i]);
   }
}

template<int dir, int KIND, int SCHEME>
static cudaError_t launchMaskedScan(const htr_fluid &f, const htr_rect &tile, const htr_rect &B, const EulerParams &ep,
                                     const uint8_t *mask, const int *flag, cudaStream_t st);
static thread_local int *g_list = nullptr;      // list + counter of the running htr_rhs call (library scratch), or null
static bool g_no_list = false;                  // htr_set_option("list", 0): the block scan (masked_flux_kernel) instead

// `fresh`: the mask was (re)written since the last compaction; false = same mask as the previous call, reuse its list
template<int dir, int KIND, int SCHEME>
static cudaError_t launchMasked(const htr_fluid &f, const htr_rect &tile, const htr_rect &B, const EulerParams &ep,
                                     const uint8_t *mask, const int *flag, cudaStream_t st, bool fresh = true) {
   const int64_t cells = (tile.hi[0] - tile.lo[0] + 1)*(tile.hi[1] - tile.lo[1] + 1)*(tile.hi[2] - tile.lo[2] + 1);
   if (g_list && !g_no_list && cells < ((int64_t)1 << 31)) {
      int *count = g_list, *list = g_list + 4;
      if (fresh) {
         cudaError_t e = cudaMemsetAsync(count, 0, sizeof(int), st);
         if (e != cudaSuccess) return e;
         const int64_t g = (cells + 255)/256;
         compact_mask_kernel<<<(unsigned)(g < (int64_t)numSMs()*8 ? g : (int64_t)numSMs()*8), 256, 0, st>>>(mask, cells, flag, list, count);
      }
      int perSM = 1;
      if constexpr (KIND == FLUX_EULER && SCHEME != HTR_SCHEME_SKEWSYMMETRIC) {
         if (g_coop) {    // a deferred Euler cell = a run of one cell for a group of eight lanes (htr_teno_coop.cuh)
            if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSM, coop::euler_coop_kernel<dir, SCHEME, true>, coop::GL*coop::GROUPS, 0) != cudaSuccess || perSM < 1) perSM = 1;
            coop::euler_coop_kernel<dir, SCHEME, true><<<(unsigned)(numSMs()*perSM), coop::GL*coop::GROUPS, 0, st>>>(f, tile, B, ep, list, count, 1);
            return cudaGetLastError();
         }
      }
      if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSM, listed_flux_kernel<dir, KIND, SCHEME>, 128, 0) != cudaSuccess || perSM < 1) perSM = 1;
      listed_flux_kernel<dir, KIND, SCHEME><<<(unsigned)(numSMs()*perSM), 128, 0, st>>>(f, tile, B, ep, list, count);
      return cudaGetLastError();
   }
   return launchMaskedScan<dir, KIND, SCHEME>(f, tile, B, ep, mask, flag, st);
}

template<int dir, int KIND, int SCHEME>
static cudaError_t launchMaskedScan(const htr_fluid &f, const htr_rect &tile, const htr_rect &B, const EulerParams &ep,
                                     const uint8_t *mask, const int *flag, cudaStream_t st) {
   using FB = FaceBlock<dir>;
   const int64_t n[3] = {tile.hi[0] - tile.lo[0] + 1, tile.hi[1] - tile.lo[1] + 1, tile.hi[2] - tile.lo[2] + 1};
   const int C[3] = {FB::X - (dir == 0), FB::Y - (dir == 1), FB::Z - (dir == 2)};
   const unsigned gx = (unsigned)((n[0] + C[0] - 1)/C[0]), gy = (unsigned)((n[1] + C[1] - 1)/C[1]), gz = (unsigned)((n[2] + C[2] - 1)/C[2]);
   int perSM = 1;          // resident CTAs of this instantiation (1 at 255 registers, 2-3 for the KG / diffusion forms)
   if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSM, masked_flux_kernel<dir, KIND, SCHEME>, FB::X*FB::Y*FB::Z, 0) != cudaSuccess || perSM < 1) perSM = 1;
   const uint64_t total = (uint64_t)gx*gy*gz, slots = (uint64_t)numSMs()*(uint64_t)perSM;
   const unsigned grid = (unsigned)(total < slots ? total : slots);
   dim3 block(FB::X, FB::Y, FB::Z);
   masked_flux_kernel<dir, KIND, SCHEME><<<grid, block, 0, st>>>(f, tile, B, ep, mask, flag, gx, gy, gz);
   return cudaGetLastError();
}

// Fast path of  Conserved_t = 0 (+chem) + Euler z,y,x  for the Hybrid and SkewSymmetric schemes
// (htr_fast.cuh) followed by the masked generic kernels for the cells it deferred.
// `scratch` holds two mask bytes per cell of the tile.
// htr_set_option("diff3", 1): diffusion z, y, x in one pass (htr_march.cuh: diffusion3_staged_kernel) instead of the two staged
// kernels.  Off by default: measured on B200 at 512^3 AirMix (profiles/r2_v8_*) 51.1 ms (45.6 ms at 2 CTAs per SM) against
// 13.7 + 23.1 ms -- the operands of the planes p-1 / p+1 are read from global memory inside the flux phase (L2 hits, but
// ~80 dependent, component-strided loads per cell on the critical path of every warp), which costs more than the pass it saves.
static bool g_diff3 = false;
// htr_set_option("dasync", 1) (default): the diffusion tiles reach shared memory by cp.async row copies (htr_diff_async.cuh:
// diffusion_async_kernel) instead of the cooperative load/store staging of fast::diffusion_staged_kernel; same bits.
// Measured on B200, 512^3 AirMix (profiles/r2_v15_*): diffusion z 13.53 -> 12.38 ms, y+x 23.39 -> 21.29 ms.
static bool g_dasync = true;
static int g_slab_tiles = 0;            // htr_set_option("slab", k): z-slab thickness of htr_host_rhs in z tiles of 7 planes (0 = 14 planes)
static bool g_no_pipe = false;          // htr_set_option("pipeline", 0): htr_host_rhs without the z-slab copy/compute overlap
static bool g_no_bulk = false;          // htr_set_option("bulk", 0): cooperative loads instead of TMA bulk copies

// Tile-relative view of the fields for the staged kernels (htr_access.cuh).  false: a stride or an extent does not
// fit the 32-bit form -> the caller uses the per-task kernels.
static bool makeRelative(const htr_fluid &f, const htr_rect &tile, const htr_rect &B, rfluid &rf, rbox &rt, rbox &rb) {
   const htr_acc *src = reinterpret_cast<const htr_acc*>(&f);
   racc *dst = reinterpret_cast<racc*>(&rf);
   const int64_t LIM = (int64_t)1 << 30;
   for (int d = 0; d < 3; d++) {
      const int64_t n = tile.hi[d] - tile.lo[d] + 1;
      if (n >= LIM) return false;
      rt.lo[d] = 0; rt.hi[d] = (int)(n - 1);
      const int64_t lo = B.lo[d] - tile.lo[d], hi = B.hi[d] - tile.lo[d];
      rb.lo[d] = (int)(lo < -LIM ? -LIM : (lo > LIM ? LIM : lo));
      rb.hi[d] = (int)(hi < -LIM ? -LIM : (hi > LIM ? LIM : hi));
   }
   for (size_t k = 0; k < sizeof(htr_fluid)/sizeof(htr_acc); k++) {
      const htr_acc &a = src[k];
      for (int d = 0; d < 3; d++) {
         if (a.stride[d] >= 2*LIM || a.stride[d] <= -2*LIM) return false;
         dst[k].s[d] = (int)a.stride[d];
      }
      dst[k].base = a.base ? static_cast<char*>(a.base) + tile.lo[0]*a.stride[0] + tile.lo[1]*a.stride[1] + tile.lo[2]*a.stride[2] : nullptr;
   }
   return true;
}

// Can rows of these fields be moved by cp.async.bulk?  Returns 0 (no) or 1 | parity<<1 (htr_bulk.cuh).
static int bulkEligible(std::initializer_list<std::pair<const racc*, int>> fields) {
   if (g_no_bulk) return 0;
   int parity = -1;
   for (auto &fc : fields) {
      const racc &a = *fc.first;
      const uintptr_t b = (uintptr_t)a.base;
      if (a.s[0] != 8*fc.second || (a.s[1] & 15) || (a.s[2] & 15) || (b & 7)) return 0;
      const int p = (int)((b >> 3) & 1);
      // a field with an EVEN component count (e.g. MassFracs of the 4-species CH41StMix) has 16-byte cells: shifting the row start
      // by one cell cannot repair a base that is 8 mod 16, so it must be aligned as it is
      if ((fc.second & 1) == 0 && p != 0) return 0;
      if (parity < 0) parity = p; else if (p != parity) return 0;
   }
   return 1 | (parity << 1);
}

// ---- TMA tensor maps of the Euler-xy kernel (htr_fast_bulk.cuh: XYMaps) ---------------------------------------------
// cuTensorMapEncodeTiled is a driver entry point: fetched through the runtime so that the library does not link libcuda.
static bool g_no_tensor = false;        // htr_set_option("tensor", 0): row-wise bulk copies instead of tensor-map copies
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn encodeTiled() {
   static EncodeTiledFn fn = nullptr;
   static bool tried = false;
   if (!tried) {
      tried = true;
      void *p = nullptr;
      cudaDriverEntryPointQueryResult q;
      if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess)
         fn = reinterpret_cast<EncodeTiledFn>(p);
   }
   return fn;
}
// Maps rho, pressure, velocity, MassFracs over the cells this call may read: [max(tile.lo-3, bounds.lo), min(tile.hi+3,
// bounds.hi)] in x and y, the tile's planes in z.  Requires what bulkEligible already checked (x dense, y/z strides
// multiples of 16 bytes, common 8-byte parity); the x origin is moved one cell down if that makes the base 16-byte aligned
// (that cell is never part of a box).
template<int TY>
static void makeTensorMaps(const rfluid &rf, const rbox &rt, const rbox &rb, int bulkOK, fast::XYMaps &maps) {
   using T = fast::XYB<TY>;
   memset(&maps, 0, sizeof(maps));
   EncodeTiledFn enc = encodeTiled();
   if (!enc || !(bulkOK & 1) || g_no_tensor) return;
   int lo[3], hi[3];
   for (int d = 0; d < 2; d++) {
      lo[d] = (-3 > rb.lo[d]) ? -3 : rb.lo[d];
      hi[d] = (rt.hi[d] + 3 < rb.hi[d]) ? rt.hi[d] + 3 : rb.hi[d];
   }
   lo[2] = 0; hi[2] = rt.hi[2];
   const racc *fld[4] = {&rf.rho, &rf.pressure, &rf.velocity, &rf.MassFracs};
   const int nc[4] = {1, 1, 3, nSpec};
   // common x origin: aligned for every field (odd component counts share the parity; even ones are aligned anywhere)
   int x0 = lo[0];
   for (int k = 0; k < 4; k++) {
      const uintptr_t a = (uintptr_t)(fld[k]->base + (int64_t)lo[0]*fld[k]->s[0] + (int64_t)lo[1]*fld[k]->s[1]);
      if (a & 15) { if (nc[k] & 1) x0 = lo[0] - 1; else return; }
   }
   for (int k = 0; k < 4; k++) {
      const racc &a = *fld[k];
      char *base = a.base + (int64_t)x0*a.s[0] + (int64_t)lo[1]*a.s[1];
      if (((uintptr_t)base & 15) || (a.s[1] & 15) || (a.s[2] & 15) || a.s[0] != 8*nc[k]) return;
      const cuuint64_t dims[3] = {(cuuint64_t)(hi[0] - x0 + 1)*nc[k], (cuuint64_t)(hi[1] - lo[1] + 1), (cuuint64_t)(hi[2] + 1)};
      const cuuint64_t strides[2] = {(cuuint64_t)a.s[1], (cuuint64_t)a.s[2]};
      const cuuint32_t box[3] = {(cuuint32_t)(T::W*nc[k]), (cuuint32_t)T::ROWS, 1};
      const cuuint32_t es[3] = {1, 1, 1};
      if (box[0] > 256 || dims[0] < 1 || dims[1] < 1) return;
      if (enc(&maps.m[k], CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 3, base, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
              CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS) return;
   }
   maps.x0 = x0; maps.y0 = lo[1]; maps.z0 = 0; maps.ok = 1;
}

enum { PART_EZ = 1, PART_EXY = 2, PART_DZ = 4, PART_DXY = 8, PART_ALL = 15 };

template<int SCHEME>
static cudaError_t launchFastEuler(bool chem, const htr_fluid &f, const htr_rect &tile, const htr_rect &B,
                                   const rfluid &rf, const rbox &rt, const rbox &rb,
                                   const EulerParams &ep, uint8_t *scratch, cudaStream_t st, int parts = PART_ALL) {
   constexpr bool HY = (SCHEME == HTR_SCHEME_HYBRID);
   constexpr int TY = HTR_K2_TY;
   const int64_t n[3] = {tile.hi[0] - tile.lo[0] + 1, tile.hi[1] - tile.lo[1] + 1, tile.hi[2] - tile.lo[2] + 1};
   int *flags = reinterpret_cast<int*>(scratch);      // cells deferred by: [0] Euler z, [1] Euler xy, [2] diffusion z, [3] diffusion xy
   uint8_t *maskZ = scratch + 16, *maskXY = maskZ + n[0]*n[1]*n[2];
   { cudaError_t e = cudaMemsetAsync(flags, 0, 16, st); if (e != cudaSuccess) return e; }
   if (firstUseOnDevice(HY ? 0 : 1)) {
      cudaFuncSetAttribute(fast::euler_z_chem_kernel<HY, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)fast::ZT_SMEM);
      cudaFuncSetAttribute(fast::euler_z_chem_kernel<HY, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)fast::ZT_SMEM);
      cudaFuncSetAttribute(fast::euler_xy_bulk_kernel<TY, HY>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)fast::XYB<TY>::SMEM);
   }
   if (parts & PART_EZ) {
      dim3 grid((unsigned)((n[2] + fast::ZT_Z - 2)/(fast::ZT_Z - 1)), (unsigned)((n[0] + fast::ZT_X - 1)/fast::ZT_X), (unsigned)n[1]);
      dim3 block(fast::ZT_X, fast::ZT_Z, 1);
      if (chem) fast::euler_z_chem_kernel<HY, true><<<grid, block, fast::ZT_SMEM, st>>>(rf, rt, rb, maskZ, flags);
      else      fast::euler_z_chem_kernel<HY, false><<<grid, block, fast::ZT_SMEM, st>>>(rf, rt, rb, maskZ, flags);
      cudaError_t e = cudaGetLastError();
      if (e != cudaSuccess) return e;
      if ((e = launchMasked<2, FLUX_EULER, SCHEME>(f, tile, B, ep, maskZ, flags, st)) != cudaSuccess) return e;
   }
   if (parts & PART_EXY) {
      dim3 grid((unsigned)((n[0] + 30)/31), (unsigned)((n[1] + TY - 2)/(TY - 1)), (unsigned)n[2]);
      dim3 block(32, TY, 1);
      const int bulkOK = bulkEligible({{&rf.rho, 1}, {&rf.pressure, 1}, {&rf.velocity, 3}, {&rf.MassFracs, nSpec}});
      fast::XYMaps maps;
      makeTensorMaps<TY>(rf, rt, rb, bulkOK, maps);
      fast::euler_xy_bulk_kernel<TY, HY><<<grid, block, fast::XYB<TY>::SMEM, st>>>(rf, rt, rb, maskXY, flags + 1, bulkOK, maps);
      cudaError_t e = cudaGetLastError();
      if (e != cudaSuccess) return e;
      if ((e = launchMasked<1, FLUX_EULER, SCHEME>(f, tile, B, ep, maskXY, flags + 1, st)) != cudaSuccess) return e;
      if ((e = launchMasked<0, FLUX_EULER, SCHEME>(f, tile, B, ep, maskXY, flags + 1, st, false)) != cudaSuccess) return e;
   }
   return cudaSuccess;
}

static bool makeUniform(const rfluid &rf, const rbox &rt, const rbox &rb, bool hybrid, bool chem, bool diffusion, march::UF &u);

// Conserved_t += diffusion z; += y; += x through the staged kernels + masked generic passes (same scratch,
// after launchFastEuler has consumed it in stream order; flags [2], [3] were cleared there).
static cudaError_t launchFastDiffusion(const htr_fluid &f, const htr_rect &tile, const htr_rect &B,
                                       const rfluid &rf, const rbox &rt, const rbox &rb, uint8_t *scratch, cudaStream_t st,
                                       int parts = PART_ALL) {
   constexpr int TJ = HTR_D_TJ;
   const int64_t n[3] = {tile.hi[0] - tile.lo[0] + 1, tile.hi[1] - tile.lo[1] + 1, tile.hi[2] - tile.lo[2] + 1};
   int *flags = reinterpret_cast<int*>(scratch);
   uint8_t *maskZ = scratch + 16, *maskXY = maskZ + n[0]*n[1]*n[2];
   const EulerParams ep = {0, 0, 0};
   cudaError_t e;
   if (g_diff3) {
      // all three directions in one pass over Conserved_t (htr_march.cuh: diffusion3_staged_kernel); PART_DXY stands for it
      if (!(parts & PART_DXY)) return cudaSuccess;
      if (firstUseOnDevice(6))
         cudaFuncSetAttribute(march::diffusion3_staged_kernel<TJ>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)fast::DTile<1, TJ>::SMEM);
      dim3 g3((unsigned)((n[0] + 30)/31), (unsigned)((n[1] + TJ - 2)/(TJ - 1)), (unsigned)n[2]);
      march::diffusion3_staged_kernel<TJ><<<g3, dim3(32, TJ, 1), fast::DTile<1, TJ>::SMEM, st>>>(rf, rt, rb, maskXY, flags + 3);
      if ((e = cudaGetLastError()) != cudaSuccess) return e;
      if ((e = launchMasked<2, FLUX_DIFFUSION, 0>(f, tile, B, ep, maskXY, flags + 3, st)) != cudaSuccess) return e;
      if ((e = launchMasked<1, FLUX_DIFFUSION, 0>(f, tile, B, ep, maskXY, flags + 3, st, false)) != cudaSuccess) return e;
      return launchMasked<0, FLUX_DIFFUSION, 0>(f, tile, B, ep, maskXY, flags + 3, st, false);
   }
   // (a variant with 1-D TMA row copies existed in round 1: ~90 copies of 0.3-2.6 KB per CTA made it TMA-issue bound,
   //  Dz 3.2 vs 1.9 ms and Dxy 4.2 vs 3.3 ms at 256^3, profiles/r1_v8_*; removed in round 2)
   if (firstUseOnDevice(3)) {
      cudaFuncSetAttribute(fast::diffusion_staged_kernel<2, TJ>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)fast::DTile<2, TJ>::SMEM);
      cudaFuncSetAttribute(fast::diffusion_staged_kernel<1, TJ>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)fast::DTile<1, TJ>::SMEM);
   }
   // cp.async staging on the uniform view of the fields (htr_diff_async.cuh); any other layout takes the cooperative kernels
   march::UF u;
   const bool dasync = g_dasync && makeUniform(rf, rt, rb, false, false, true, u);
   if (dasync && firstUseOnDevice(7)) {
      cudaFuncSetAttribute(fast::diffusion_async_kernel<2, TJ>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)fast::ATile<2, TJ>::SMEM);
      cudaFuncSetAttribute(fast::diffusion_async_kernel<1, TJ>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)fast::ATile<1, TJ>::SMEM);
   }
   dim3 gz((unsigned)((n[0] + 31)/32), (unsigned)((n[2] + TJ - 2)/(TJ - 1)), (unsigned)n[1]);
   if (parts & PART_DZ) {
      if (dasync) fast::diffusion_async_kernel<2, TJ><<<gz, dim3(32, TJ, 1), fast::ATile<2, TJ>::SMEM, st>>>(u, maskZ, flags + 2);
      else        fast::diffusion_staged_kernel<2, TJ><<<gz, dim3(32, TJ, 1), fast::DTile<2, TJ>::SMEM, st>>>(rf, rt, rb, maskZ, flags + 2);
      if ((e = cudaGetLastError()) != cudaSuccess) return e;
      if ((e = launchMasked<2, FLUX_DIFFUSION, 0>(f, tile, B, ep, maskZ, flags + 2, st)) != cudaSuccess) return e;
   }
   if (!(parts & PART_DXY)) return cudaSuccess;
   dim3 gxy((unsigned)((n[0] + 30)/31), (unsigned)((n[1] + TJ - 2)/(TJ - 1)), (unsigned)n[2]);
   if (dasync) fast::diffusion_async_kernel<1, TJ><<<gxy, dim3(32, TJ, 1), fast::ATile<1, TJ>::SMEM, st>>>(u, maskXY, flags + 3);
   else        fast::diffusion_staged_kernel<1, TJ><<<gxy, dim3(32, TJ, 1), fast::DTile<1, TJ>::SMEM, st>>>(rf, rt, rb, maskXY, flags + 3);
   if ((e = cudaGetLastError()) != cudaSuccess) return e;
   if ((e = launchMasked<1, FLUX_DIFFUSION, 0>(f, tile, B, ep, maskXY, flags + 3, st)) != cudaSuccess) return e;
   return launchMasked<0, FLUX_DIFFUSION, 0>(f, tile, B, ep, maskXY, flags + 3, st, false);
}

// ---- z-marching two-pass form (htr_march.cuh) ---------------------------------------------------
// htr_set_option("march", m): bit 0 = euler_z_march_kernel instead of fast::euler_z_chem_kernel, bit 1 = plane_march_kernel
// instead of the three staged kernels euler_xy / diffusion_z / diffusion_xy, bit 2 (without bit 1) = plane_march_kernel
// without its diffusion phase instead of euler_xy only.  Default 1: measured on B200 at 512^3 AirMix
// (profiles/r2_*): Euler-z march 14.4 ms vs 15.7 ms staged; plane march 60.5 ms vs 13.5 + 13.6 + 23.0 ms staged (16 warps per
// SM at 112 KB of shared memory per CTA leave its load latency exposed: issue slots 25 % busy vs 36-45 %).
static int g_march = 1;
static int g_march_kz = 64;             // planes per CTA of euler_z_march_kernel   ("march_kz")
static int g_march_kp = 32;             // planes per CTA of plane_march_kernel     ("march_kp")

// Uniform-layout view: every field the kernels read is x-dense and all share one cell pitch in y and in z
// (what one Legion instance per tile gives, src/prometeo_mapper.cc:1142-1204).  false -> htr_fast.cuh path.
static bool makeUniform(const rfluid &rf, const rbox &rt, const rbox &rb, bool hybrid, bool chem, bool diffusion, march::UF &u) {
   int64_t SY = -1, SZ = -1;
   auto ok = [&](const racc &a, int elem, int ncomp) -> bool {
      if (!a.base || ((uintptr_t)a.base & (uintptr_t)(elem - 1))) return false;
      const int64_t cs = (int64_t)elem*ncomp;
      if (a.s[0] != cs || a.s[1] % cs || a.s[2] % cs) return false;
      const int64_t sy = a.s[1]/cs, sz = a.s[2]/cs;
      if (SY < 0) { SY = sy; SZ = sz; }
      return sy == SY && sz == SZ && sy > 0 && sz > 0;
   };
   #define HTR_UD(dst, src, nc) if (!ok(rf.src, 8, nc)) return false; u.dst = reinterpret_cast<double*>(rf.src.base);
   HTR_UD(rho, rho, 1) HTR_UD(p, pressure, 1) HTR_UD(Y, MassFracs, nSpec) HTR_UD(vel, velocity, 3) HTR_UD(C, Conserved, nEq)
   if (!ok(rf.Conserved_t, 8, nEq)) return false;
   u.Ct = reinterpret_cast<double*>(rf.Conserved_t.base);
   HTR_UD(me[0], dcsi_e, 1) HTR_UD(me[1], deta_e, 1) HTR_UD(me[2], dzet_e, 1)
   u.T = u.mu = u.lam = u.X = u.Di = nullptr;
   for (int d = 0; d < 3; d++) { u.md[d] = u.ms[d] = nullptr; u.sens[d] = nullptr; }
   if (chem || diffusion) { HTR_UD(T, temperature, 1) }
   if (diffusion) {
      HTR_UD(mu, mu, 1) HTR_UD(lam, lam, 1) HTR_UD(X, MolarFracs, nSpec) HTR_UD(Di, Di, nSpec)
      HTR_UD(md[0], dcsi_d, 1) HTR_UD(md[1], deta_d, 1) HTR_UD(md[2], dzet_d, 1)
      HTR_UD(ms[0], dcsi_s, 1) HTR_UD(ms[1], deta_s, 1) HTR_UD(ms[2], dzet_s, 1)
   }
   #undef HTR_UD
   const racc *nt[3] = {&rf.nType_x, &rf.nType_y, &rf.nType_z};
   const racc *ss[3] = {&rf.shockSensorX, &rf.shockSensorY, &rf.shockSensorZ};
   for (int d = 0; d < 3; d++) {
      if (!ok(*nt[d], 4, 1)) return false;
      u.nT[d] = reinterpret_cast<const int*>(nt[d]->base);
      if (hybrid) {
         if (!ok(*ss[d], 1, 1)) return false;
         u.sens[d] = reinterpret_cast<const uint8_t*>(ss[d]->base);
      }
   }
   if (SY >= ((int64_t)1 << 24) || SZ >= ((int64_t)1 << 40)) return false;
   u.SY = (int)SY;
   if (SZ >= ((int64_t)1 << 31)) return false;
   u.SZ = (int)SZ;
   int64_t ext[3];                        // largest |coordinate| a (wrapped) stencil point can have, per direction
   for (int d = 0; d < 3; d++) {
      u.n[d] = rt.hi[d] + 1;
      u.blo[d] = rb.lo[d]; u.bhi[d] = rb.hi[d];
      // a stencil point overshoots the box by at most 3 cells: the single-step wrap needs a box at least that wide
      if ((int64_t)rb.hi[d] - rb.lo[d] + 1 < 8) return false;
      const int64_t BIG = (int64_t)1 << 29;
      const int64_t lo = (rb.lo[d] > -BIG && rb.lo[d] < -3) ? rb.lo[d] : -3;
      const int64_t hi = (rb.hi[d] < BIG && rb.hi[d] > rt.hi[d] + 3) ? rb.hi[d] : rt.hi[d] + 3;
      ext[d] = (-lo > hi ? -lo : hi) + 64;
   }
   // the kernels index cells with 32 bits
   if (ext[0] + ext[1]*SY + ext[2]*SZ >= ((int64_t)1 << 31)) return false;
   return true;
}

template<int SCHEME>
static cudaError_t launchMarch(bool chem, bool diffusion, const htr_fluid &f, const htr_rect &tile, const htr_rect &B,
                               const march::UF &u, const EulerParams &ep, uint8_t *scratch, cudaStream_t st, int parts) {
   constexpr bool HY = (SCHEME == HTR_SCHEME_HYBRID);
   const int64_t n[3] = {u.n[0], u.n[1], u.n[2]};
   int *flags = reinterpret_cast<int*>(scratch);      // cells deferred by: [0] Euler z pass, [1] plane pass
   uint8_t *maskZ = scratch + 16, *maskP = maskZ + n[0]*n[1]*n[2];
   cudaError_t e;
   if ((e = cudaMemsetAsync(flags, 0, 16, st)) != cudaSuccess) return e;
   if (firstUseOnDevice(HY ? 4 : 5)) {
      cudaFuncSetAttribute(march::plane_march_kernel<HY, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)march::PLay<true>::SMEM);
      cudaFuncSetAttribute(march::plane_march_kernel<HY, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)march::PLay<false>::SMEM);
   }
   if (parts & PART_EZ) {
      int kch = g_march_kz < 6 ? 6 : g_march_kz;
      dim3 grid((unsigned)((n[0] + 31)/32), (unsigned)((n[1] + march::MZ_TY - 1)/march::MZ_TY), (unsigned)((n[2] + kch - 1)/kch));
      dim3 block(32, march::MZ_TY, 1);
      if (chem) march::euler_z_march_kernel<HY, true><<<grid, block, 0, st>>>(u, maskZ, flags, kch);
      else      march::euler_z_march_kernel<HY, false><<<grid, block, 0, st>>>(u, maskZ, flags, kch);
      if ((e = cudaGetLastError()) != cudaSuccess) return e;
      if ((e = launchMasked<2, FLUX_EULER, SCHEME>(f, tile, B, ep, maskZ, flags, st)) != cudaSuccess) return e;
   }
   if (parts & PART_EXY) {
      int kch = g_march_kp < 1 ? 1 : g_march_kp;
      dim3 grid((unsigned)((n[0] + 30)/31), (unsigned)((n[1] + 6)/7), (unsigned)((n[2] + kch - 1)/kch));
      dim3 block(32, 8, 1);
      if (diffusion) march::plane_march_kernel<HY, true><<<grid, block, march::PLay<true>::SMEM, st>>>(u, maskP, flags + 1, kch);
      else           march::plane_march_kernel<HY, false><<<grid, block, march::PLay<false>::SMEM, st>>>(u, maskP, flags + 1, kch);
      if ((e = cudaGetLastError()) != cudaSuccess) return e;
      if ((e = launchMasked<1, FLUX_EULER, SCHEME>(f, tile, B, ep, maskP, flags + 1, st)) != cudaSuccess) return e;
      if ((e = launchMasked<0, FLUX_EULER, SCHEME>(f, tile, B, ep, maskP, flags + 1, st, false)) != cudaSuccess) return e;
      if (diffusion) {
         const EulerParams e0 = {0, 0, 0};
         if ((e = launchMasked<2, FLUX_DIFFUSION, 0>(f, tile, B, e0, maskP, flags + 1, st, false)) != cudaSuccess) return e;
         if ((e = launchMasked<1, FLUX_DIFFUSION, 0>(f, tile, B, e0, maskP, flags + 1, st, false)) != cudaSuccess) return e;
         if ((e = launchMasked<0, FLUX_DIFFUSION, 0>(f, tile, B, e0, maskP, flags + 1, st, false)) != cudaSuccess) return e;
      }
   }
   return cudaSuccess;
}

// ---------------------------------------------------------------------------------------------
// point kernels
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ bool cellOf(Pt &p, const htr_rect &tile) {
   const int64_t nx = tile.hi[0] - tile.lo[0] + 1, ny = tile.hi[1] - tile.lo[1] + 1, nz = tile.hi[2] - tile.lo[2] + 1;
   const int64_t g = (int64_t)blockIdx.x*blockDim.x + threadIdx.x;
   if (g >= nx*ny*nz) return false;
   p.x = tile.lo[0] + g % nx;
   p.y = tile.lo[1] + (g/nx) % ny;
   p.z = tile.lo[2] + g/(nx*ny);
   return true;
}

__global__ void __launch_bounds__(128) ducros_kernel(const htr_fluid f, const htr_rect tile, const htr_rect B, const double eps) {
   Pt p; if (!cellOf(p, tile)) return;
   *at<double>(f.DucrosSensor, p) = ducrosSensor(f, p, B, eps);
}

template<int dir>
__global__ void __launch_bounds__(128) shock_sensor_kernel(const htr_fluid f, const htr_rect tile, const htr_rect B) {
   Pt p; if (!cellOf(p, tile)) return;
   *at<uint8_t>(sensorOf<dir>(f), p) = shockSensor<dir>(f, p, B) ? 1 : 0;
}

__global__ void __launch_bounds__(128) chem_source_kernel(const htr_fluid f, const htr_rect tile) {
   Pt p; if (!cellOf(p, tile)) return;
   double Yi[nSpec], w[nSpec];
   HTR_UNROLL
   for (int i = 0; i < nSpec; i++) Yi[i] = ldc<double>(f.MassFracs, p, i);
   mix::GetProductionRates(w, ld<double>(f.rho, p), ld<double>(f.pressure, p), ld<double>(f.temperature, p), Yi);
   double *Ct = at<double>(f.Conserved_t, p);
   HTR_UNROLL
   for (int i = 0; i < nSpec; i++) Ct[i] += w[i];
}

__device__ int d_chem_failures;

__device__ int *d_chem_calls;      // optional diagnostic output (htr_chemistry_rhs_counts): rhs() evaluations per cell of the tile

// resident CTAs per SM the register allocation aims at: the integration is a long chain of dependent fp64 operations
// (top stall: fixed-latency `wait`), so it wants warps more than registers; its state lives in local memory anyway
#ifndef HTR_CHEM_MINB
#define HTR_CHEM_MINB 6      // B200, 96^3 hot AirMix box: 695 ms at 2, 588 ms at 4, 541 ms at 6 (902 ms before rhs() went out of line)
#endif
__global__ void __launch_bounds__(128, HTR_CHEM_MINB) update_chemistry_kernel(const htr_fluid f, const htr_rect tile, const double dt) {
   Pt p; if (!cellOf(p, tile)) return;
   double S[nEq], x[nEq];
   const double *Ct = at<const double>(f.Conserved_t, p);
   double *Cto = at<double>(f.Conserved_t_old, p);
   double *C = at<double>(f.Conserved, p);
   for (int i = 0; i < nEq; i++) { S[i] = Ct[i]; Cto[i] = S[i]; x[i] = C[i]; }
   ChemProblem prob; prob.S = S; prob.T = *at<double>(f.temperature, p); prob.calls = 0;
   if (!rosenbrockSolve<nEq>(prob, x, dt, dt)) atomicAdd(&d_chem_failures, 1);
   for (int i = 0; i < nEq; i++) C[i] = x[i];
   *at<double>(f.temperature, p) = prob.T;
   if (d_chem_calls) d_chem_calls[maskIndexG(p, tile)] = prob.calls;
}

// NEWTON = false: the temperature field already holds the converged value of these cells (ghost copies of a
// neighbour's cells after the exchange of Conserved and temperature): every other primitive follows from it with the
// owner's own operations, hence the owner's bits; iterating again from the converged value would move T within the
// Newton tolerance and leave the two ranks with different ghost and owner states.
template<bool NEWTON>
__global__ void __launch_bounds__(128) primitive_from_conserved_kernel(const htr_fluid f, const htr_rect tile) {
   Pt p; if (!cellOf(p, tile)) return;
   double U[nEq], rho, Yi[nSpec], MixW, u[3], P;
   HTR_UNROLL
   for (int i = 0; i < nEq; i++) U[i] = ldc<double>(f.Conserved, p, i);
   double T = *at<double>(f.temperature, p);
   primitivesFromConserved<NEWTON>(U, rho, Yi, MixW, u, T, P);
   double Xi[nSpec];
   mix::GetMolarFractions(Xi, MixW, Yi);
   double *X = at<double>(f.MolarFracs, p), *v = at<double>(f.velocity, p);
   HTR_UNROLL
   for (int i = 0; i < nSpec; i++) X[i] = Xi[i];
   HTR_UNROLL
   for (int i = 0; i < 3; i++) v[i] = u[i];
   *at<double>(f.temperature, p) = T;
   *at<double>(f.pressure, p) = P;
}

// rho, mu, lam, Di, SoS, MassFracs of one cell from (Xi, P, T) (src/prometeo_variables.hpp:62-89)
__device__ __forceinline__ void storeProperties(const htr_fluid &f, const Pt &p, const double *Xi, const double P, const double T) {
   double Yi[nSpec], Di[nSpec];
   const double MixW = mix::GetMolarWeightFromXi(Xi);
   *at<double>(f.rho, p) = mix::GetRho(P, T, MixW);
#if HTR_MULTICOMP && !defined(HTR_EXACT_POW)
   double mu, lam;
   mix::GetTransport(mu, lam, Di, P, T, MixW, Xi);
   *at<double>(f.mu, p) = mu;
   *at<double>(f.lam, p) = lam;
#else
   *at<double>(f.mu, p) = mix::GetViscosity(T, Xi);
   *at<double>(f.lam, p) = mix::GetHeatConductivity(T, Xi);
   mix::GetDiffusivity(Di, P, T, MixW, Xi);
#endif
   mix::GetMassFractions(Yi, MixW, Xi);
   const double gamma = mix::GetGamma(T, MixW, Yi);
   *at<double>(f.SoS, p) = mix::GetSpeedOfSound(T, gamma, MixW);
   double *oD = at<double>(f.Di, p), *oY = at<double>(f.MassFracs, p);
   HTR_UNROLL
   for (int i = 0; i < nSpec; i++) { oD[i] = Di[i]; oY[i] = Yi[i]; }
}

// ---- Runge-Kutta glue (src/prometeo_rk.rg:46-111, coefficients src/prometeo_const.rg:54-58) -------------
__constant__ double c_rk[3][3] = {{1.0, 0.0, 1.0}, {3.0/4.0, 1.0/4.0, 1.0/4.0}, {1.0/3.0, 2.0/3.0, 2.0/3.0}};

__global__ void __launch_bounds__(128) rk_initialize_temporaries_kernel(const htr_fluid f, const htr_rect tile) {
   Pt p; if (!cellOf(p, tile)) return;
   const double *C = at<const double>(f.Conserved, p);
   double *Co = at<double>(f.Conserved_old, p);
   HTR_UNROLL
   for (int i = 0; i < nEq; i++) Co[i] = C[i];
}

__global__ void __launch_bounds__(128) rk_initialize_time_derivatives_kernel(const htr_fluid f, const htr_rect tile, const int implicitChemistry) {
   Pt p; if (!cellOf(p, tile)) return;
   double *Ct = at<double>(f.Conserved_t, p);
   if (implicitChemistry) {
      const double *Cto = at<const double>(f.Conserved_t_old, p);
      HTR_UNROLL
      for (int i = 0; i < nEq; i++) Ct[i] = -Cto[i];
   } else {
      HTR_UNROLL
      for (int i = 0; i < nEq; i++) Ct[i] = 0.0;
   }
}

// The same three operations when the nEq doubles of consecutive x cells are contiguous (stride[0] == 8*nEq, every
// layout the library allocates): a row of the tile is a flat array of nx*nEq doubles and one thread moves one double,
// so every warp access is a full 256-byte line instead of 32 spans 72 bytes apart.
// OP 0: Conserved_old = Conserved   1: Conserved_t = 0   2: Conserved_t = -Conserved_t_old   3: UpdateVars(stage)
//    4: Conserved_t -= Conserved_t_old (htr_time_step with implicit chemistry)
template<int OP>
__global__ void __launch_bounds__(256) rk_rows_kernel(const htr_fluid f, const htr_rect tile, const int stage, const double dt) {
   const int64_t e = (int64_t)blockIdx.x*256 + threadIdx.x;
   if (e >= (tile.hi[0] - tile.lo[0] + 1)*(int64_t)nEq) return;
   Pt p; p.x = tile.lo[0]; p.y = tile.lo[1] + blockIdx.y; p.z = tile.lo[2] + blockIdx.z;
   if (OP == 0) at<double>(f.Conserved_old, p)[e] = at<const double>(f.Conserved, p)[e];
   if (OP == 1) at<double>(f.Conserved_t, p)[e] = 0.0;
   if (OP == 2) at<double>(f.Conserved_t, p)[e] = -at<const double>(f.Conserved_t_old, p)[e];
   if (OP == 4) at<double>(f.Conserved_t, p)[e] -= at<const double>(f.Conserved_t_old, p)[e];
   if (OP == 3) {
      const double c1 = c_rk[stage][0], c2 = c_rk[stage][1], c3 = c_rk[stage][2];
      double *C = at<double>(f.Conserved, p) + e;
      *C = (c1*at<const double>(f.Conserved_old, p)[e] + c2*(*C)) + (c3*at<const double>(f.Conserved_t, p)[e])*dt;
   }
}
static bool rowsContiguous(std::initializer_list<const htr_acc*> fields, const htr_rect &t) {
   for (const htr_acc *a : fields) if (a->stride[0] != 8*nEq) return false;
   return (t.hi[1] - t.lo[1]) < 65535 && (t.hi[2] - t.lo[2]) < 65535;
}
static dim3 rowsGrid(const htr_rect &t) {
   return dim3((unsigned)(((t.hi[0] - t.lo[0] + 1)*(int64_t)nEq + 255)/256), (unsigned)(t.hi[1] - t.lo[1] + 1), (unsigned)(t.hi[2] - t.lo[2] + 1));
}

// UpdateVars(STAGE); with RECOVER also UpdatePrimitiveFromConserved and UpdatePropertiesFromPrimitive of the
// same cell from the registers (the three tasks the reference launches back to back after every stage,
// src/prometeo.rg:1427, :452, :470).  The operation order of each task is kept, so the fused pass writes the
// same bits as the three kernels in sequence (strict build; FMA contraction may differ by an ulp otherwise).
template<bool RECOVER>
__global__ void __launch_bounds__(128) rk_update_vars_kernel(const htr_fluid f, const htr_rect tile, const int stage, const double dt) {
   Pt p; if (!cellOf(p, tile)) return;
   const double c1 = c_rk[stage][0], c2 = c_rk[stage][1], c3 = c_rk[stage][2];
   const double *Co = at<const double>(f.Conserved_old, p), *Ct = at<const double>(f.Conserved_t, p);
   double *C = at<double>(f.Conserved, p);
   double U[nEq];
   HTR_UNROLL
   for (int i = 0; i < nEq; i++) U[i] = (c1*Co[i] + c2*C[i]) + (c3*Ct[i])*dt;
   HTR_UNROLL
   for (int i = 0; i < nEq; i++) C[i] = U[i];
   if (RECOVER) {
      double rho, Yi[nSpec], MixW, u[3], P, Xi[nSpec];
      double T = *at<double>(f.temperature, p);
      primitivesFromConserved(U, rho, Yi, MixW, u, T, P);
      mix::GetMolarFractions(Xi, MixW, Yi);
      double *X = at<double>(f.MolarFracs, p), *v = at<double>(f.velocity, p);
      HTR_UNROLL
      for (int i = 0; i < nSpec; i++) X[i] = Xi[i];
      HTR_UNROLL
      for (int i = 0; i < 3; i++) v[i] = u[i];
      *at<double>(f.temperature, p) = T;
      *at<double>(f.pressure, p) = P;
      storeProperties(f, p, Xi, P, T);
   }
}

// The fused stage with 128 consecutive x cells per block and every multi-component field moved as flat row spans
// through shared memory (coalesced 8-byte accesses; the per-cell view in shared memory has an odd stride in doubles,
// hence no bank conflicts).  Needs x-contiguous cells in all AoS fields (rkStageRowsOK); same arithmetic and bits
// as rk_update_vars_kernel<true>.
#ifndef HTR_RKS_MINB
#define HTR_RKS_MINB 6
#endif
constexpr int RKS_NT = 128;
constexpr int RKS_POOL = RKS_NT*((3*nEq > nEq + 3*nSpec + 3) ? 3*nEq : nEq + 3*nSpec + 3);
__global__ void __launch_bounds__(RKS_NT, HTR_RKS_MINB) rk_stage_rows_kernel(const htr_fluid f, const htr_rect tile, const int stage, const double dt) {
   __shared__ double pool[RKS_POOL];
   const int tid = threadIdx.x;
   const int x0 = blockIdx.x*RKS_NT;
   const int nx = (int)(tile.hi[0] - tile.lo[0] + 1);
   const int nc = min(RKS_NT, nx - x0);                      // cells of this block
   Pt p0; p0.x = tile.lo[0] + x0; p0.y = tile.lo[1] + blockIdx.y; p0.z = tile.lo[2] + blockIdx.z;
   {
      const double *gCo = at<const double>(f.Conserved_old, p0), *gC = at<const double>(f.Conserved, p0), *gCt = at<const double>(f.Conserved_t, p0);
      // global -> shared without a register stop (LDGSTS): all 3 nEq copies of a thread are in flight together; ncu had 11.6 % of
      // the kernel's stall samples on the load-then-store form of this loop (profiles/r2_v26_hot_source_lines.txt)
      for (int e = tid; e < nc*nEq; e += RKS_NT) {
         fast::cpAsync8(pool + e, gCo + e); fast::cpAsync8(pool + RKS_NT*nEq + e, gC + e); fast::cpAsync8(pool + 2*RKS_NT*nEq + e, gCt + e);
      }
   }
   const bool valid = tid < nc;
   Pt p = p0; p.x += tid;
   double T0 = 0.0;
   if (valid) T0 = *at<double>(f.temperature, p);            // the Newton guess: requested while the rows arrive
   fast::cpAsyncWaitAll();
   __syncthreads();
   const double c1 = c_rk[stage][0], c2 = c_rk[stage][1], c3 = c_rk[stage][2];
   double U[nEq], Xi[nSpec], u[3], Yi[nSpec], Di[nSpec];
   HTR_UNROLL
   for (int i = 0; i < nEq; i++)
      U[i] = (c1*pool[tid*nEq + i] + c2*pool[RKS_NT*nEq + tid*nEq + i]) + (c3*pool[2*RKS_NT*nEq + tid*nEq + i])*dt;
   __syncthreads();
   if (valid) {
      double rho, MixW, P;
      double T = T0;
      primitivesFromConserved(U, rho, Yi, MixW, u, T, P);
      mix::GetMolarFractions(Xi, MixW, Yi);
      *at<double>(f.temperature, p) = T;
      *at<double>(f.pressure, p) = P;
      // storeProperties, with the vector outputs kept in registers
      const double MixW2 = mix::GetMolarWeightFromXi(Xi);
      *at<double>(f.rho, p) = mix::GetRho(P, T, MixW2);
#if HTR_MULTICOMP && !defined(HTR_EXACT_POW)
      double mu, lam;
      mix::GetTransport(mu, lam, Di, P, T, MixW2, Xi);
      *at<double>(f.mu, p) = mu;
      *at<double>(f.lam, p) = lam;
#else
      *at<double>(f.mu, p) = mix::GetViscosity(T, Xi);
      *at<double>(f.lam, p) = mix::GetHeatConductivity(T, Xi);
      mix::GetDiffusivity(Di, P, T, MixW2, Xi);
#endif
      mix::GetMassFractions(Yi, MixW2, Xi);
      const double gamma = mix::GetGamma(T, MixW2, Yi);
      *at<double>(f.SoS, p) = mix::GetSpeedOfSound(T, gamma, MixW2);
   }
   double *sC = pool, *sX = sC + RKS_NT*nEq, *sY = sX + RKS_NT*nSpec, *sD = sY + RKS_NT*nSpec, *sV = sD + RKS_NT*nSpec;
   HTR_UNROLL
   for (int i = 0; i < nEq; i++) sC[tid*nEq + i] = U[i];
   HTR_UNROLL
   for (int i = 0; i < nSpec; i++) { sX[tid*nSpec + i] = Xi[i]; sY[tid*nSpec + i] = Yi[i]; sD[tid*nSpec + i] = Di[i]; }
   HTR_UNROLL
   for (int i = 0; i < 3; i++) sV[tid*3 + i] = u[i];
   __syncthreads();
   {
      double *gC = at<double>(f.Conserved, p0), *gX = at<double>(f.MolarFracs, p0), *gY = at<double>(f.MassFracs, p0);
      double *gD = at<double>(f.Di, p0), *gV = at<double>(f.velocity, p0);
      for (int e = tid; e < nc*nEq; e += RKS_NT) gC[e] = sC[e];
      for (int e = tid; e < nc*nSpec; e += RKS_NT) { gX[e] = sX[e]; gY[e] = sY[e]; gD[e] = sD[e]; }
      for (int e = tid; e < nc*3; e += RKS_NT) gV[e] = sV[e];
   }
}
static bool rkStageRowsOK(const htr_fluid &f, const htr_rect &t) {
   return rowsContiguous({&f.Conserved, &f.Conserved_old, &f.Conserved_t}, t) && f.MolarFracs.stride[0] == 8*nSpec &&
          f.MassFracs.stride[0] == 8*nSpec && f.Di.stride[0] == 8*nSpec && f.velocity.stride[0] == 24;
}

#ifndef HTR_PROP_MINB
#define HTR_PROP_MINB 6
#endif
__global__ void __launch_bounds__(128, HTR_PROP_MINB) properties_from_primitive_kernel(const htr_fluid f, const htr_rect tile) {
   Pt p; if (!cellOf(p, tile)) return;
   double Xi[nSpec];
   HTR_UNROLL
   for (int i = 0; i < nSpec; i++) Xi[i] = ldc<double>(f.MolarFracs, p, i);
   const double P = ld<double>(f.pressure, p), T = ld<double>(f.temperature, p);
   storeProperties(f, p, Xi, P, T);
}

__global__ void __launch_bounds__(128) conserved_from_primitive_kernel(const htr_fluid f, const htr_rect tile) {
   Pt p; if (!cellOf(p, tile)) return;
   double Yi[nSpec], u[3];
   HTR_UNROLL
   for (int i = 0; i < nSpec; i++) Yi[i] = ldc<double>(f.MassFracs, p, i);
   HTR_UNROLL
   for (int i = 0; i < 3; i++) u[i] = ldc<double>(f.velocity, p, i);
   const double rho = ld<double>(f.rho, p);
   double *C = at<double>(f.Conserved, p);
   HTR_UNROLL
   for (int i = 0; i < nSpec; i++) C[i] = rho*Yi[i];
   HTR_UNROLL
   for (int i = 0; i < 3; i++) C[i + irU] = rho*u[i];
   C[irE] = rho*(0.5*mod2(u) + mix::GetInternalEnergy(ld<double>(f.temperature, p), Yi));
}

__global__ void __launch_bounds__(128) zero_rhs_kernel(const htr_fluid f, const htr_rect tile) {
   Pt p; if (!cellOf(p, tile)) return;
   double *Ct = at<double>(f.Conserved_t, p);
   HTR_UNROLL
   for (int i = 0; i < nEq; i++) Ct[i] = 0.0;
}

// max over the tile: warp shuffle, one shared slot per warp, then one atomic per block on the
// bit pattern (radii are >= 0, so integer order == floating order; max is order independent)
__global__ void __launch_bounds__(256) spectral_radius_kernel(const htr_fluid f, const htr_rect tile, double *result) {
   Pt p;
   double r = 0.0;
   if (cellOf(p, tile)) r = spectralRadius(f, p);
   HTR_UNROLL
   for (int o = 16; o > 0; o >>= 1) r = fmax(r, __shfl_xor_sync(0xffffffffu, r, o));
   __shared__ double sm[8];
   if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = r;
   __syncthreads();
   if (threadIdx.x < 32) {
      r = (threadIdx.x < (blockDim.x >> 5)) ? sm[threadIdx.x] : 0.0;
      HTR_UNROLL
      for (int o = 4; o > 0; o >>= 1) r = fmax(r, __shfl_xor_sync(0xffffffffu, r, o));
      if (threadIdx.x == 0) atomicMax(reinterpret_cast<unsigned long long*>(result), (unsigned long long)__double_as_longlong(r));
   }
}

} // namespace htr

// =============================================================================================
// C ABI
// =============================================================================================
using namespace htr;

#if HTR_MULTICOMP
#if defined(HTR_MIX_AirMix)
#include "generated/AirMix_tables.inc"
#define HTR_TABLES htr_AirMix_tables
#define HTR_TABLE_BYTES HTR_AIRMIX_TABLE_BYTES
#elif defined(HTR_MIX_CH41StMix)
#include "generated/CH41StMix_tables.inc"
#define HTR_TABLES htr_CH41StMix_tables
#define HTR_TABLE_BYTES HTR_CH41STMIX_TABLE_BYTES
#endif
#endif

static bool g_mix_loaded_dev[HTR_MAX_DEVICES] = {};
#define g_mix_loaded g_mix_loaded_dev[curDev()]
static bool g_no_fast = false;          // htr_set_option("fast", 0): per-task kernels only (A/B testing)

// Device scratch owned by the library (mask bytes of the fast path).  One buffer per stream that ever
// asked, grown on demand; growth synchronises the device once.
struct Scratch { int dev; cudaStream_t st; void *p; size_t bytes; };
constexpr int HTR_NSCRATCH = 64;
static Scratch g_scratch[HTR_NSCRATCH];
static int g_nscratch = 0;
static void *scratchFor(size_t bytes, cudaStream_t st) {
   const int dev = curDev();
   int k = -1;
   for (int i = 0; i < g_nscratch; i++) if (g_scratch[i].dev == dev && g_scratch[i].st == st) { k = i; break; }
   if (k < 0) {
      if (g_nscratch < HTR_NSCRATCH) { k = g_nscratch++; g_scratch[k] = {dev, st, nullptr, 0}; }
      else {
         // table full: recycle an entry of THIS device (its buffer may still be in use by queued work: drain that stream first)
         for (int i = 0; i < g_nscratch; i++) if (g_scratch[i].dev == dev) { k = i; break; }
         if (k < 0) return nullptr;
         cudaStreamSynchronize(g_scratch[k].st);
         g_scratch[k].st = st;
      }
   }
   Scratch &s = g_scratch[k];
   if (s.bytes < bytes) {
      // only work queued on this stream can still read the old buffer
      if (s.p) { cudaStreamSynchronize(st); cudaFree(s.p); s.p = nullptr; s.bytes = 0; }
      if (cudaMalloc(&s.p, bytes) != cudaSuccess) { s.p = nullptr; return nullptr; }
      s.bytes = bytes;
   }
   return s.p;
}

static inline int cuerr(cudaError_t e) { return e == cudaSuccess ? HTR_OK : 1000 + (int)e; }
static inline bool rectEmpty(const htr_rect &r) { return r.hi[0] < r.lo[0] || r.hi[1] < r.lo[1] || r.hi[2] < r.lo[2]; }
static inline unsigned blocksFor(const htr_rect &t, int threads) {
   const int64_t n = (t.hi[0] - t.lo[0] + 1)*(t.hi[1] - t.lo[1] + 1)*(t.hi[2] - t.lo[2] + 1);
   return (unsigned)((n + threads - 1)/threads);
}

template<int KIND, int SCHEME>
static int dispatchDir(int dir, const htr_fluid &f, const htr_rect &tile, const htr_rect &B, const EulerParams &ep, cudaStream_t st) {
   switch (dir) {
   case 0: return cuerr(launchFlux<0, KIND, SCHEME>(f, tile, B, ep, st));
   case 1: return cuerr(launchFlux<1, KIND, SCHEME>(f, tile, B, ep, st));
   case 2: return cuerr(launchFlux<2, KIND, SCHEME>(f, tile, B, ep, st));
   }
   return HTR_EINVAL;
}

extern "C" {

int htr_nspec(void) { return nSpec; }
int htr_neq(void) { return nEq; }
const char *htr_mixture_name(void) { return HTR_MIX_NAME; }
int htr_sizeof_mix(void) { return (int)sizeof(MixData); }
const char *htr_build_info(void) {
   return "htr_b200 " HTR_MIX_NAME " sm_100a"
#ifdef HTR_EXACT_POW
      " exact-pow"
#endif
#ifdef HTR_FMAD_OFF
      " fmad=false"
#endif
      ;
}

int htr_mix_init_constprop(void *out, double gasConstant, double gamma, double prandtl,
                           int visc_model, double p0, double p1, double p2) {
#if HTR_MULTICOMP
   (void)out; (void)gasConstant; (void)gamma; (void)prandtl; (void)visc_model; (void)p0; (void)p1; (void)p2;
   return HTR_EMIXTYPE;
#else
   if (!out || visc_model < 0 || visc_model > 2) return HTR_EINVAL;
   MixData m;
   memset(&m, 0, sizeof(m));
   m.R = gasConstant; m.gamma = gamma; m.viscosityModel = visc_model; m.Prandtl = prandtl;
   if (visc_model == 0) m.constantVisc = p0;
   else if (visc_model == 1) { m.powerlawViscRef = p0; m.powerlawTempRef = p1; }
   else { m.sutherlandViscRef = p0; m.sutherlandTempRef = p1; m.sutherlandSRef = p2; }
   memcpy(out, &m, sizeof(m));
   return HTR_OK;
#endif
}

int htr_mix_init_multicomp(void *out, double LRef, double PRef, double TRef,
                           int n, const char *names, const double *molarFracs) {
#if !HTR_MULTICOMP
   (void)out; (void)LRef; (void)PRef; (void)TRef; (void)n; (void)names; (void)molarFracs;
   return HTR_EMIXTYPE;
#else
   if (!out || !(PRef > 0) || !(LRef > 0) || !(TRef > 0) || n < 0 || (n > 0 && (!names || !molarFracs))) return HTR_EINVAL;
   static_assert(HTR_TABLE_BYTES == offsetof(MixData, PRef), "species/reaction tables must end where PRef starts");
   MixData m;
   memset(&m, 0, sizeof(m));
   memcpy(&m, HTR_TABLES, HTR_TABLE_BYTES);
   // reference quantities (src/Mixtures/MultiComponent.inl:438-488)
   m.PRef = PRef; m.TRef = TRef;
   for (int i = 0; i < nSpec; i++) m.XiRef[i] = 1.0e-60;
   double check = 0.0;
   for (int k = 0; k < n; k++) {
      char nm[11]; memcpy(nm, names + 10*k, 10); nm[10] = 0;
      int is = -1;
      for (int i = 0; i < nSpec; i++) if (strcmp(m.species[i].Name, nm) == 0) { is = i; break; }
      if (is < 0) return HTR_EINVAL;
      m.XiRef[is] = molarFracs[k];
      check += molarFracs[k];
   }
   if (!(fabs(check - 1.0) < 1e-3)) return HTR_EINVAL;
   double W = 0.0;
   for (int i = 0; i < nSpec; i++) W += m.XiRef[i]*m.species[i].W;
   m.MixWRef = W;
   m.iMixWRef = 1.0/m.MixWRef;
   m.rhoRef = PRef*m.MixWRef/(RGAS*TRef);
   m.ieRef = m.rhoRef/PRef;
   m.iuRef = sqrt(m.ieRef);
   m.iCpRef = m.MixWRef/RGAS;
   m.imuRef = 1.0/(LRef*sqrt(PRef*m.rhoRef));
   m.ilamRef = m.MixWRef/(LRef*sqrt(PRef*m.rhoRef)*RGAS);
   m.iDiRef = sqrt(m.rhoRef/PRef)/(LRef);
   m.iwiRef = LRef/sqrt(m.rhoRef*PRef);
   m.iKiRef = sqrt(PRef/m.rhoRef)*m.MixWRef/(Na*LRef*eCrg);
   m.Eps0 = m.MixWRef/(m.rhoRef*Na*eCrg*LRef);
   m.Eps0 = eps_0*PRef*m.Eps0*m.Eps0;
   m.TMax = 1e20; m.TMin = 0.0;
   for (int i = 0; i < nSpec; i++) {
      m.TMax = fmin(m.TMax, m.species[i].cp.TMax/TRef);
      m.TMin = fmax(m.TMin, m.species[i].cp.TMin/TRef);
   }
   memcpy(out, &m, sizeof(m));
   return HTR_OK;
#endif
}

int htr_set_option(const char *name, int value) {
   if (!name) return HTR_EINVAL;
   if (strcmp(name, "fast") == 0) { g_no_fast = (value == 0); return HTR_OK; }
   if (strcmp(name, "bulk") == 0) { g_no_bulk = (value == 0); return HTR_OK; }
   if (strcmp(name, "tensor") == 0) { g_no_tensor = (value == 0); return HTR_OK; }
   if (strcmp(name, "list") == 0) { g_no_list = (value == 0); return HTR_OK; }
   if (strcmp(name, "diff3") == 0) { g_diff3 = (value != 0); return HTR_OK; }
   if (strcmp(name, "dasync") == 0) { g_dasync = (value != 0); return HTR_OK; }
   if (strcmp(name, "coop") == 0) { g_coop = (value != 0); return HTR_OK; }
   if (strcmp(name, "pipeline") == 0) { g_no_pipe = (value == 0); return HTR_OK; }
   if (strcmp(name, "slab") == 0) { if (value < 0 || value > 64) return HTR_EINVAL; g_slab_tiles = value; return HTR_OK; }
   if (strcmp(name, "march") == 0) { if (value < 0 || value > 7) return HTR_EINVAL; g_march = value; return HTR_OK; }
   if (strcmp(name, "march_kz") == 0) { if (value < 1) return HTR_EINVAL; g_march_kz = value; return HTR_OK; }
   if (strcmp(name, "march_kp") == 0) { if (value < 1) return HTR_EINVAL; g_march_kp = value; return HTR_OK; }
   return HTR_EINVAL;
}

int htr_set_mixture(const void *mix, size_t bytes) {
   if (!mix || bytes != sizeof(MixData)) return HTR_EINVAL;
#if HTR_MULTICOMP
   // the production kinetics are specialised on the structure of this library's mechanism
   // (compared field by field: padding bytes of a caller-constructed struct are not defined)
   {
      MixData a, b;
      memcpy(&a, mix, sizeof(a));
      memset(&b, 0, sizeof(b));
      memcpy(&b, HTR_TABLES, HTR_TABLE_BYTES);
      auto sameTerms = [](const StoichTerm *x, const StoichTerm *y, int n) {
         for (int i = 0; i < n; i++) if (x[i].ind != y[i].ind || x[i].nu != y[i].nu) return false;
         return true;
      };
#ifdef HTR_FWD_ORDERS
      auto sameReact = [](const ReactTerm *x, const ReactTerm *y, int n) {
         for (int i = 0; i < n; i++) if (x[i].ind != y[i].ind || x[i].nu != y[i].nu || x[i].ord != y[i].ord) return false;
         return true;
      };
#else
      auto sameReact = sameTerms;
#endif
      auto sameK = [](const Arrhenius &x, const Arrhenius &y) {
         return (x.n != 0.0f) == (y.n != 0.0f) && (fabsf(x.EovR) > 1e-5f) == (fabsf(y.EovR) > 1e-5f);
      };
#if HTR_NREAC > 0
      for (int r = 0; r < HTR_NREAC; r++) {
         const PlainReaction &x = a.reactions[r], &y = b.reactions[r];
         if (x.has_backward != y.has_backward || x.Neducts != y.Neducts || x.Npducts != y.Npducts || !sameK(x.k, y.k) ||
             !sameReact(x.educts, y.educts, y.Neducts) || !sameTerms(x.pducts, y.pducts, y.Npducts)) return HTR_EMIXTYPE;
      }
#endif
#if HTR_NTBREAC > 0
      for (int r = 0; r < HTR_NTBREAC; r++) {
         const CollisionReaction &x = a.tbReactions[r], &y = b.tbReactions[r];
         if (x.has_backward != y.has_backward || x.Neducts != y.Neducts || x.Npducts != y.Npducts || x.Nthirdb != y.Nthirdb ||
             !sameK(x.k, y.k) || !sameReact(x.educts, y.educts, y.Neducts) || !sameTerms(x.pducts, y.pducts, y.Npducts) ||
             !sameTerms(x.thirdb, y.thirdb, y.Nthirdb)) return HTR_EMIXTYPE;
      }
#endif
   }
#endif
   cudaError_t e = cudaMemcpyToSymbol(c_mix, mix, sizeof(MixData));
#if HTR_MULTICOMP
   if (e == cudaSuccess) {
      MixData m;
      memcpy(&m, mix, sizeof(m));
      SpecDerived der[nSpec];
      for (int i = 0; i < nSpec; i++) {
         der[i].rOvW = RGAS/m.species[i].W;
         const double *cp[3] = {m.species[i].cp.cpH, m.species[i].cp.cpM, m.species[i].cp.cpL};
         for (int r = 0; r < 3; r++) {
            const double *c = cp[r];
            der[i].q[r][0] = c[3]*0.50; der[i].q[r][1] = c[4]/3; der[i].q[r][2] = c[5]*0.25;
            der[i].q[r][3] = c[6]/5;    der[i].q[r][4] = c[5]/6;
         }
      }
      e = cudaMemcpyToSymbol(c_der, der, sizeof(der));
      // transport constants (reference expressions: Species.inl GetMu/GetDif/GetSelfDiffusion/GetFZrot, MultiComponent.inl GetViscosity)
      TransDerived tr;
      memset(&tr, 0, sizeof(tr));
      tr.pi15 = pow(PI, 1.5); tr.kb3 = pow(kb, 3);
      for (int i = 0; i < nSpec; i++) {
         const SpeciesData &a = m.species[i];
         tr.sigma2[i] = pow(a.diff.sigma, 2);
         const double tmp = 1.0/(a.diff.kbOveps*298.0);
         tr.fz298[i] = 1 + 0.5*pow(PI, 1.5)*sqrt(tmp) + (2 + 0.25*PI*PI)*tmp + pow(PI, 1.5)*pow(tmp, 1.5);
         for (int j = 0; j < nSpec; j++) {
            const SpeciesData &b = m.species[j];
            tr.w25[i][j] = pow(b.W/a.W, 0.25);
            tr.s8[i][j] = sqrt(8*(1 + a.W/b.W));
            double xi = 1.0;
            if ((a.diff.mu*b.diff.mu == 0.0) && (a.diff.mu + b.diff.mu != 0.0)) {
               double mup, alp, epr;
               if (a.diff.mu != 0.0) {
                  mup = a.diff.mu/sqrt(4*PI*eps_0*kb*pow(a.diff.sigma, 3)/a.diff.kbOveps);
                  alp = b.diff.alpha/pow(b.diff.sigma, 3);
                  epr = sqrt(b.diff.kbOveps/a.diff.kbOveps);
               } else {
                  mup = b.diff.mu/sqrt(4*PI*eps_0*kb*pow(b.diff.sigma, 3)/b.diff.kbOveps);
                  alp = a.diff.alpha/pow(a.diff.sigma, 3);
                  epr = sqrt(a.diff.kbOveps/b.diff.kbOveps);
               }
               xi = 1 + 0.25*mup*alp*epr;
            }
            tr.sigmaij[i][j] = 0.5*(a.diff.sigma + b.diff.sigma)*pow(xi, -1./6);
            tr.kboEpsij[i][j] = sqrt(a.diff.kbOveps*b.diff.kbOveps)/(xi*xi);
            tr.invWij[i][j] = (a.W + b.W)/(a.W*b.W);
            tr.is8[i][j] = 1.0/tr.s8[i][j];
            tr.cdif[i][j] = 16*PI*tr.sigmaij[i][j]*tr.sigmaij[i][j]/(3*sqrt(2*PI*tr.kb3*Na*tr.invWij[i][j]));
         }
         tr.cmu[i] = 5*sqrt(PI*a.W/Na*kb)/(16*PI*tr.sigma2[i]);
         tr.zf[i] = a.diff.Z298*tr.fz298[i];
      }
      if (e == cudaSuccess) e = cudaMemcpyToSymbol(c_tr, &tr, sizeof(tr));
   }
#endif
   if (e == cudaSuccess) g_mix_loaded = true;
   return cuerr(e);
}

#define HTR_CHECK_COMMON(f, tile) \
   if (!(f) || !(tile)) return HTR_EINVAL; \
   if (!g_mix_loaded) return HTR_ENOMIX; \
   if (rectEmpty(*(tile))) return HTR_OK;

int htr_euler_flux(int scheme, int dir, const htr_fluid *f, const htr_rect *tile, const htr_rect *bounds,
                   double RK0, double RK1, double deltaTime, void *stream) {
   HTR_CHECK_COMMON(f, tile)
   if (!bounds || dir < 0 || dir > 2) return HTR_EINVAL;
   const EulerParams ep = {RK0, RK1, deltaTime};
   cudaStream_t st = (cudaStream_t)stream;
   switch (scheme) {
   case HTR_SCHEME_HYBRID:        return dispatchDir<FLUX_EULER, HTR_SCHEME_HYBRID>(dir, *f, *tile, *bounds, ep, st);
   case HTR_SCHEME_TENO:          return dispatchDir<FLUX_EULER, HTR_SCHEME_TENO>(dir, *f, *tile, *bounds, ep, st);
   case HTR_SCHEME_TENOA:         return dispatchDir<FLUX_EULER, HTR_SCHEME_TENOA>(dir, *f, *tile, *bounds, ep, st);
   case HTR_SCHEME_TENOLAD:       return dispatchDir<FLUX_EULER, HTR_SCHEME_TENOLAD>(dir, *f, *tile, *bounds, ep, st);
   case HTR_SCHEME_SKEWSYMMETRIC: return dispatchDir<FLUX_EULER, HTR_SCHEME_SKEWSYMMETRIC>(dir, *f, *tile, *bounds, ep, st);
   }
   return HTR_EINVAL;
}

int htr_diffusion_flux(int dir, const htr_fluid *f, const htr_rect *tile, const htr_rect *bounds, void *stream) {
   HTR_CHECK_COMMON(f, tile)
   if (!bounds || dir < 0 || dir > 2) return HTR_EINVAL;
   const EulerParams ep = {0, 0, 0};
   return dispatchDir<FLUX_DIFFUSION, 0>(dir, *f, *tile, *bounds, ep, (cudaStream_t)stream);
}

int htr_ducros_sensor(const htr_fluid *f, const htr_rect *tile, const htr_rect *bounds, double vorticityScale, void *stream) {
   HTR_CHECK_COMMON(f, tile)
   if (!bounds) return HTR_EINVAL;
   const double eps = fmax(vorticityScale*vorticityScale, 1e-6);      // src/prometeo_sensor.cc:62
   ducros_kernel<<<blocksFor(*tile, 128), 128, 0, (cudaStream_t)stream>>>(*f, *tile, *bounds, eps);
   return cuerr(cudaGetLastError());
}

int htr_shock_sensor(int dir, const htr_fluid *f, const htr_rect *tile, const htr_rect *bounds, void *stream) {
   HTR_CHECK_COMMON(f, tile)
   if (!bounds || dir < 0 || dir > 2) return HTR_EINVAL;
   cudaStream_t st = (cudaStream_t)stream;
   const unsigned nb = blocksFor(*tile, 128);
   if (dir == 0) shock_sensor_kernel<0><<<nb, 128, 0, st>>>(*f, *tile, *bounds);
   else if (dir == 1) shock_sensor_kernel<1><<<nb, 128, 0, st>>>(*f, *tile, *bounds);
   else shock_sensor_kernel<2><<<nb, 128, 0, st>>>(*f, *tile, *bounds);
   return cuerr(cudaGetLastError());
}

int htr_add_chemistry_sources(const htr_fluid *f, const htr_rect *tile, void *stream) {
   HTR_CHECK_COMMON(f, tile)
   chem_source_kernel<<<blocksFor(*tile, 128), 128, 0, (cudaStream_t)stream>>>(*f, *tile);
   return cuerr(cudaGetLastError());
}

int htr_update_chemistry(const htr_fluid *f, const htr_rect *tile, double dt, void *stream) {
   HTR_CHECK_COMMON(f, tile)
   update_chemistry_kernel<<<blocksFor(*tile, 128), 128, 0, (cudaStream_t)stream>>>(*f, *tile, dt);
   return cuerr(cudaGetLastError());
}

/* number of cells whose implicit integration hit the iteration caps since the last call
 * (the reference asserts, src/Utils/math_utils.hpp:233,240); synchronises the device */
int htr_chemistry_failures(void);
/* diagnostic: subsequent htr_update_chemistry calls write the number of rhs() evaluations of every cell of their tile
 * (dense, x fastest) to the DEVICE buffer `d_counts`; NULL switches it off.  (2 nEq + 1) evaluations per integration step
 * plus 3 per attempt: equal counts = same adaptive step sequence as the reference (oracle: ..._counted). */
int htr_chemistry_rhs_counts(int *d_counts) {
   return cuerr(cudaMemcpyToSymbol(d_chem_calls, &d_counts, sizeof(int*)));
}
int htr_mixture_loaded(void) { return g_mix_loaded ? 1 : 0; }

/* the same, ordered after the work queued on `stream` (cudaMemcpyFromSymbol alone only orders against the legacy default
 * stream, not against non-blocking streams such as PyTorch side streams or Realm task streams) */
int htr_chemistry_failures_on(void *stream) {
   if (cudaStreamSynchronize((cudaStream_t)stream) != cudaSuccess) return -1;
   return htr_chemistry_failures();
}

int htr_chemistry_failures(void) {
   int n = 0, zero = 0;
   if (cudaMemcpyFromSymbol(&n, d_chem_failures, sizeof(int)) != cudaSuccess) return -1;
   cudaMemcpyToSymbol(d_chem_failures, &zero, sizeof(int));
   return n;
}

// ---- fp64 pipe probe: the denominator of the fp64 roofline -----------------------
// Every thread runs 16 independent DFMA chains; 2 flops per DFMA.  Timed with CUDA events on `stream`.
__global__ void __launch_bounds__(256) fp64_probe_kernel(double *out, const int iters, const double a, const double b) {
   double v[16];
   HTR_UNROLL
   for (int k = 0; k < 16; k++) v[k] = a + k + threadIdx.x*1e-3;
   for (int it = 0; it < iters; it++) {
      HTR_UNROLL
      for (int k = 0; k < 16; k++) v[k] = fma(v[k], a, b);
   }
   double sacc = 0.0;
   HTR_UNROLL
   for (int k = 0; k < 16; k++) sacc += v[k];
   if (sacc == 123.456) out[0] = sacc;       // never true: keeps the chains alive
}
int htr_probe_fp64(double seconds, double *tflops, void *stream) {
   if (!tflops) return HTR_EINVAL;
   cudaStream_t st = (cudaStream_t)stream;
   int dev = 0, sms = 148;
   cudaGetDevice(&dev);
   cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
   double *d = nullptr;
   if (cudaMalloc(&d, 8) != cudaSuccess) return 1000 + (int)cudaErrorMemoryAllocation;
   cudaEvent_t e0, e1;
   cudaEventCreate(&e0); cudaEventCreate(&e1);
   const int blocks = sms*8, iters = 4096;
   const double flopsPerLaunch = 2.0*16*iters*256.0*blocks;
   fp64_probe_kernel<<<blocks, 256, 0, st>>>(d, iters, 0.999999, 1e-9);      // warm-up
   cudaStreamSynchronize(st);
   double best = 0.0, elapsed = 0.0;
   int launches = 0;
   while (elapsed < seconds*1e3 || launches < 3) {
      cudaEventRecord(e0, st);
      for (int r = 0; r < 20; r++) fp64_probe_kernel<<<blocks, 256, 0, st>>>(d, iters, 0.999999, 1e-9);
      cudaEventRecord(e1, st);
      if (cudaEventSynchronize(e1) != cudaSuccess) break;
      float ms = 0; cudaEventElapsedTime(&ms, e0, e1);
      elapsed += ms; launches++;
      const double tf = 20*flopsPerLaunch/(ms*1e-3)/1e12;
      tflops[0] = tf;                           // sustained: the last batch
      if (tf > best) best = tf;
      if (launches > 2000) break;
   }
   tflops[1] = best;                            // burst: the best batch
   cudaEventDestroy(e0); cudaEventDestroy(e1);
   cudaFree(d);
   return cuerr(cudaGetLastError());
}

int htr_update_primitive_from_conserved(const htr_fluid *f, const htr_rect *tile, void *stream) {
   HTR_CHECK_COMMON(f, tile)
   primitive_from_conserved_kernel<true><<<blocksFor(*tile, 128), 128, 0, (cudaStream_t)stream>>>(*f, *tile);
   return cuerr(cudaGetLastError());
}
int htr_update_ghost_primitive_from_conserved(const htr_fluid *f, const htr_rect *tile, void *stream) {
   HTR_CHECK_COMMON(f, tile)
   primitive_from_conserved_kernel<false><<<blocksFor(*tile, 128), 128, 0, (cudaStream_t)stream>>>(*f, *tile);
   return cuerr(cudaGetLastError());
}
int htr_update_properties_from_primitive(const htr_fluid *f, const htr_rect *tile, void *stream) {
   HTR_CHECK_COMMON(f, tile)
   properties_from_primitive_kernel<<<blocksFor(*tile, 128), 128, 0, (cudaStream_t)stream>>>(*f, *tile);
   return cuerr(cudaGetLastError());
}
int htr_rk_initialize_temporaries(const htr_fluid *f, const htr_rect *tile, void *stream) {
   HTR_CHECK_COMMON(f, tile)
   if (rowsContiguous({&f->Conserved, &f->Conserved_old}, *tile))
      rk_rows_kernel<0><<<rowsGrid(*tile), 256, 0, (cudaStream_t)stream>>>(*f, *tile, 0, 0.0);
   else
      rk_initialize_temporaries_kernel<<<blocksFor(*tile, 128), 128, 0, (cudaStream_t)stream>>>(*f, *tile);
   return cuerr(cudaGetLastError());
}
int htr_rk_initialize_time_derivatives(const htr_fluid *f, const htr_rect *tile, int implicitChemistry, void *stream) {
   HTR_CHECK_COMMON(f, tile)
   if (rowsContiguous({&f->Conserved_t, implicitChemistry ? &f->Conserved_t_old : &f->Conserved_t}, *tile)) {
      if (implicitChemistry) rk_rows_kernel<2><<<rowsGrid(*tile), 256, 0, (cudaStream_t)stream>>>(*f, *tile, 0, 0.0);
      else rk_rows_kernel<1><<<rowsGrid(*tile), 256, 0, (cudaStream_t)stream>>>(*f, *tile, 0, 0.0);
   } else
      rk_initialize_time_derivatives_kernel<<<blocksFor(*tile, 128), 128, 0, (cudaStream_t)stream>>>(*f, *tile, implicitChemistry);
   return cuerr(cudaGetLastError());
}
static int rkUpdate(bool recover, int stage, double dt, int implicitChemistry, const htr_fluid *f, const htr_rect *tile, void *stream) {
   HTR_CHECK_COMMON(f, tile)
   if (stage < 1 || stage > 3) return HTR_EINVAL;
   if (implicitChemistry) dt *= 0.5;
   if (recover && rkStageRowsOK(*f, *tile) && !g_no_fast)
      rk_stage_rows_kernel<<<dim3((unsigned)((tile->hi[0] - tile->lo[0] + RKS_NT)/RKS_NT), (unsigned)(tile->hi[1] - tile->lo[1] + 1),
                                  (unsigned)(tile->hi[2] - tile->lo[2] + 1)), RKS_NT, 0, (cudaStream_t)stream>>>(*f, *tile, stage - 1, dt);
   else if (recover) rk_update_vars_kernel<true><<<blocksFor(*tile, 128), 128, 0, (cudaStream_t)stream>>>(*f, *tile, stage - 1, dt);
   else if (rowsContiguous({&f->Conserved, &f->Conserved_old, &f->Conserved_t}, *tile))
      rk_rows_kernel<3><<<rowsGrid(*tile), 256, 0, (cudaStream_t)stream>>>(*f, *tile, stage - 1, dt);
   else rk_update_vars_kernel<false><<<blocksFor(*tile, 128), 128, 0, (cudaStream_t)stream>>>(*f, *tile, stage - 1, dt);
   return cuerr(cudaGetLastError());
}
int htr_rk_update_vars(int stage, double Integrator_deltaTime, int implicitChemistry, const htr_fluid *f, const htr_rect *tile, void *stream) {
   return rkUpdate(false, stage, Integrator_deltaTime, implicitChemistry, f, tile, stream);
}
int htr_rk_stage(int stage, double Integrator_deltaTime, int implicitChemistry, const htr_fluid *f, const htr_rect *tile, void *stream) {
   return rkUpdate(true, stage, Integrator_deltaTime, implicitChemistry, f, tile, stream);
}
int htr_update_conserved_from_primitive(const htr_fluid *f, const htr_rect *tile, void *stream) {
   HTR_CHECK_COMMON(f, tile)
   conserved_from_primitive_kernel<<<blocksFor(*tile, 128), 128, 0, (cudaStream_t)stream>>>(*f, *tile);
   return cuerr(cudaGetLastError());
}

int htr_max_spectral_radius(const htr_fluid *f, const htr_rect *tile, double *d_result, void *stream) {
   if (!d_result) return HTR_EINVAL;
   if (!f || !tile) return HTR_EINVAL;
   if (!g_mix_loaded) return HTR_ENOMIX;
   cudaStream_t st = (cudaStream_t)stream;
   cudaError_t e = cudaMemsetAsync(d_result, 0, sizeof(double), st);
   if (e != cudaSuccess) return cuerr(e);
   if (rectEmpty(*tile)) return HTR_OK;
   spectral_radius_kernel<<<blocksFor(*tile, 256), 256, 0, st>>>(*f, *tile, d_result);
   return cuerr(cudaGetLastError());
}

int htr_rhs(int scheme, int diffusion, int chemistry, const htr_fluid *f, const htr_rect *tile,
            const htr_rect *bounds, double RK0, double RK1, double deltaTime, void *stream) {
   return htr_rhs_parts(PART_ALL, scheme, diffusion, chemistry, f, tile, bounds, RK0, RK1, deltaTime, stream);
}

int htr_rhs_parts(int parts, int scheme, int diffusion, int chemistry, const htr_fluid *f, const htr_rect *tile,
                  const htr_rect *bounds, double RK0, double RK1, double deltaTime, void *stream) {
   HTR_CHECK_COMMON(f, tile)
   if (!bounds) return HTR_EINVAL;
   // Reference task order (src/prometeo.rg:502-578): 0, +chem, +Euler z,y,x, +diffusion z,y,x.
   cudaStream_t st = (cudaStream_t)stream;
   if (scheme < 0 || scheme > HTR_SCHEME_SKEWSYMMETRIC) return HTR_EINVAL;
   rfluid rf; rbox rt, rb;
   if ((scheme == HTR_SCHEME_HYBRID || scheme == HTR_SCHEME_SKEWSYMMETRIC) && !g_no_fast && makeRelative(*f, *tile, *bounds, rf, rt, rb)) {
      const int64_t cells = (tile->hi[0] - tile->lo[0] + 1)*(tile->hi[1] - tile->lo[1] + 1)*(tile->hi[2] - tile->lo[2] + 1);
      // 16 bytes of flags, two mask bytes per cell, then (16-byte aligned) the counter and the list of deferred cells
      const size_t listOff = ((size_t)(2*cells + 16) + 15)/16*16;
      uint8_t *scratch = (uint8_t*)scratchFor(listOff + 16 + 4*(size_t)cells, st);
      if (!scratch) return 1000 + (int)cudaErrorMemoryAllocation;
      struct ListScope { ListScope(int *p) { g_list = p; } ~ListScope() { g_list = nullptr; } } listScope(reinterpret_cast<int*>(scratch + listOff));
      const EulerParams ep = {RK0, RK1, deltaTime};
      march::UF u;
      if (g_march && makeUniform(rf, rt, rb, scheme == HTR_SCHEME_HYBRID, chemistry != 0, diffusion != 0, u)) {
         // z-marching kernels (htr_march.cuh); parts: PART_EZ = Euler-z pass, PART_EXY = plane pass (Euler y, x + diffusion)
         const bool planeAll = (g_march & 2) != 0, planeEuler = !planeAll && (g_march & 4) != 0;
         const int mparts = (planeAll || planeEuler) ? (parts & (PART_EZ | PART_EXY)) : (parts & PART_EZ);
         const bool mdiff = diffusion != 0 && planeAll;
         const cudaError_t em = (scheme == HTR_SCHEME_HYBRID)
            ? launchMarch<HTR_SCHEME_HYBRID>(chemistry != 0, mdiff, *f, *tile, *bounds, u, ep, scratch, st, mparts)
            : launchMarch<HTR_SCHEME_SKEWSYMMETRIC>(chemistry != 0, mdiff, *f, *tile, *bounds, u, ep, scratch, st, mparts);
         if (em != cudaSuccess || planeAll) return cuerr(em);
         if (planeEuler) {                // diffusion through the staged kernels (flags [2], [3] were cleared by launchMarch)
            if (diffusion) return cuerr(launchFastDiffusion(*f, *tile, *bounds, rf, rt, rb, scratch, st, parts));
            return HTR_OK;
         }
         parts &= ~PART_EZ;               // the staged kernels do the rest
      }
      cudaError_t e = (scheme == HTR_SCHEME_HYBRID)
         ? launchFastEuler<HTR_SCHEME_HYBRID>(chemistry != 0, *f, *tile, *bounds, rf, rt, rb, ep, scratch, st, parts)
         : launchFastEuler<HTR_SCHEME_SKEWSYMMETRIC>(chemistry != 0, *f, *tile, *bounds, rf, rt, rb, ep, scratch, st, parts);
      if (e != cudaSuccess) return cuerr(e);
      if (diffusion && (e = launchFastDiffusion(*f, *tile, *bounds, rf, rt, rb, scratch, st, parts)) != cudaSuccess) return cuerr(e);
      return HTR_OK;
   }
   if (parts != PART_ALL) return HTR_EINVAL;      // the per-task sequence below is not split
   {  // Conserved_t = 0 over the tile (RK.InitializeTimeDerivatives, src/prometeo_rk.rg:61-87)
      zero_rhs_kernel<<<blocksFor(*tile, 128), 128, 0, st>>>(*f, *tile);
      cudaError_t e = cudaGetLastError();
      if (e != cudaSuccess) return cuerr(e);
   }
   int rc = HTR_OK;
   if (chemistry && (rc = htr_add_chemistry_sources(f, tile, stream)) != HTR_OK) return rc;
   for (int d = 2; d >= 0; d--)
      if ((rc = htr_euler_flux(scheme, d, f, tile, bounds, RK0, RK1, deltaTime, stream)) != HTR_OK) return rc;
   if (diffusion)
      for (int d = 2; d >= 0; d--)
         if ((rc = htr_diffusion_flux(d, f, tile, bounds, stream)) != HTR_OK) return rc;
   return rc;
}

// ---- htr_rhs without any allocation (no counterpart in the reference) ----
// Callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it) get the staged
// htr_rhs with its scratch (16 flag bytes, two mask bytes per cell, the counter and the list of deferred cells: 6 bytes per
// cell + 48) in a file-scope device array instead of the cudaMalloc'ed, per-stream buffers of scratchFor().  Every GPU that
// loads the library holds the array until exit; every call on a device shares it, so calls that overlap on the GPU (two
// streams) race.  Nothing in it has to survive a call or start at zero: the flags and the list counter are cleared on the
// stream by the launchers, every mask byte of the tile is written by the staged kernels before it is read.
constexpr int64_t HTR_NOALLOC_MAX_CELLS = (int64_t)128*128*128;
constexpr size_t HTR_NOALLOC_BYTES = 6*(size_t)HTR_NOALLOC_MAX_CELLS + 64;        // 12 MiB + 64 B
static __device__ __align__(256) unsigned char g_noalloc_scratch[HTR_NOALLOC_BYTES];
// the array's name is a host-side stand-in in host code: the device address is fetched once per device
static uint8_t *g_noalloc_ptr[HTR_MAX_DEVICES] = {};

int64_t htr_rhs_no_alloc_max_cells(void) { return HTR_NOALLOC_MAX_CELLS; }

int htr_rhs_no_alloc(int scheme, int diffusion, int chemistry, const htr_fluid *f, const htr_rect *tile,
                     const htr_rect *bounds, double RK0, double RK1, double deltaTime, void *stream) {
   HTR_CHECK_COMMON(f, tile)
   if (!bounds) return HTR_EINVAL;
   if (scheme < 0 || scheme > HTR_SCHEME_SKEWSYMMETRIC) return HTR_EINVAL;
   cudaStream_t st = (cudaStream_t)stream;
   const int64_t cells = (tile->hi[0] - tile->lo[0] + 1)*(tile->hi[1] - tile->lo[1] + 1)*(tile->hi[2] - tile->lo[2] + 1);
   if (cells > HTR_NOALLOC_MAX_CELLS) {
      fprintf(stderr, "htr_rhs_no_alloc: tile of %lld cells exceeds the static scratch (at most %lld cells, 128^3); use htr_rhs\n",
              (long long)cells, (long long)HTR_NOALLOC_MAX_CELLS);
      return HTR_ETOOBIG;
   }
   rfluid rf; rbox rt, rb;
   // the per-task sequence of htr_rhs (other schemes, "fast" = 0, strides beyond 32 bits) uses no scratch at all
   if (!((scheme == HTR_SCHEME_HYBRID || scheme == HTR_SCHEME_SKEWSYMMETRIC) && !g_no_fast && makeRelative(*f, *tile, *bounds, rf, rt, rb)))
      return htr_rhs_parts(PART_ALL, scheme, diffusion, chemistry, f, tile, bounds, RK0, RK1, deltaTime, stream);
   const int dev = curDev();
   if (!g_noalloc_ptr[dev]) {
      void *p = nullptr;
      const cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);      // with this device current
      if (e != cudaSuccess || !p) return cuerr(e != cudaSuccess ? e : cudaErrorInvalidSymbol);
      g_noalloc_ptr[dev] = static_cast<uint8_t*>(p);
   }
   uint8_t *scratch = g_noalloc_ptr[dev];
   const size_t listOff = ((size_t)(2*cells + 16) + 15)/16*16;
   struct ListScope { ListScope(int *p) { g_list = p; } ~ListScope() { g_list = nullptr; } } listScope(reinterpret_cast<int*>(scratch + listOff));
   const EulerParams ep = {RK0, RK1, deltaTime};
   int parts = PART_ALL;
   march::UF u;
   if (g_march && makeUniform(rf, rt, rb, scheme == HTR_SCHEME_HYBRID, chemistry != 0, diffusion != 0, u)) {
      const bool planeAll = (g_march & 2) != 0, planeEuler = !planeAll && (g_march & 4) != 0;
      const int mparts = (planeAll || planeEuler) ? (PART_EZ | PART_EXY) : PART_EZ;
      const bool mdiff = diffusion != 0 && planeAll;
      const cudaError_t em = (scheme == HTR_SCHEME_HYBRID)
         ? launchMarch<HTR_SCHEME_HYBRID>(chemistry != 0, mdiff, *f, *tile, *bounds, u, ep, scratch, st, mparts)
         : launchMarch<HTR_SCHEME_SKEWSYMMETRIC>(chemistry != 0, mdiff, *f, *tile, *bounds, u, ep, scratch, st, mparts);
      if (em != cudaSuccess || planeAll) return cuerr(em);
      if (planeEuler) return diffusion ? cuerr(launchFastDiffusion(*f, *tile, *bounds, rf, rt, rb, scratch, st, parts)) : HTR_OK;
      parts &= ~PART_EZ;
   }
   cudaError_t e = (scheme == HTR_SCHEME_HYBRID)
      ? launchFastEuler<HTR_SCHEME_HYBRID>(chemistry != 0, *f, *tile, *bounds, rf, rt, rb, ep, scratch, st, parts)
      : launchFastEuler<HTR_SCHEME_SKEWSYMMETRIC>(chemistry != 0, *f, *tile, *bounds, rf, rt, rb, ep, scratch, st, parts);
   if (e != cudaSuccess) return cuerr(e);
   if (diffusion && (e = launchFastDiffusion(*f, *tile, *bounds, rf, rt, rb, scratch, st, parts)) != cudaSuccess) return cuerr(e);
   return HTR_OK;
}

// ---- one RK3 time step of a block without boundary nodes (MainLoopBody, src/prometeo.rg:1348-1470) ---------------
// shock sensors once (Hybrid only, src/prometeo_sensor.rg:104-122); with implicit chemistry the operator-split
// half step (:1360-1395: UpdateDerivatives(RK_coeffs = (0, 1), dt/2) without chemistry sources, UpdateChemistry(dt),
// primitive/property recovery); InitializeTemporaries; then per stage UpdateDerivatives(RK_C[STAGE][1..2],
// dt*RK_C[STAGE][3] [/2]) and UpdateVars + primitive + property recovery.  With implicit chemistry every stage's
// Conserved_t starts from -Conserved_t_old (mkInitializeTimeDerivatives(true)); the staged kernels start from 0, so
// Conserved_t_old is subtracted after them: same terms, the old derivative added last instead of first (rounding only).
__global__ void __launch_bounds__(128) subtract_old_rhs_kernel(const htr_fluid f, const htr_rect tile) {
   Pt p; if (!cellOf(p, tile)) return;
   double *Ct = at<double>(f.Conserved_t, p);
   const double *Cto = at<const double>(f.Conserved_t_old, p);
   HTR_UNROLL
   for (int i = 0; i < nEq; i++) Ct[i] -= Cto[i];
}
int htr_time_step(int scheme, int diffusion, int chemistry, int implicitChemistry, double vorticityScale, const htr_fluid *f,
                  const htr_rect *tile, const htr_rect *bounds, double Integrator_deltaTime, void *stream) {
   HTR_CHECK_COMMON(f, tile)
   if (!bounds) return HTR_EINVAL;
   static const double RK_C[3][3] = {{1.0, 0.0, 1.0}, {3.0/4.0, 1.0/4.0, 1.0/4.0}, {1.0/3.0, 2.0/3.0, 2.0/3.0}};
   const double dt = Integrator_deltaTime;
   int rc;
   if (scheme == HTR_SCHEME_HYBRID) {
      if ((rc = htr_ducros_sensor(f, tile, bounds, vorticityScale, stream)) != HTR_OK) return rc;
      for (int d = 0; d < 3; d++)
         if ((rc = htr_shock_sensor(d, f, tile, bounds, stream)) != HTR_OK) return rc;
   }
   if (implicitChemistry) {
      if ((rc = htr_rhs(scheme, diffusion, 0, f, tile, bounds, 0.0, 1.0, dt*0.5, stream)) != HTR_OK) return rc;
      if ((rc = htr_update_chemistry(f, tile, dt, stream)) != HTR_OK) return rc;
      if ((rc = htr_update_primitive_from_conserved(f, tile, stream)) != HTR_OK) return rc;
      if ((rc = htr_update_properties_from_primitive(f, tile, stream)) != HTR_OK) return rc;
   }
   if ((rc = htr_rk_initialize_temporaries(f, tile, stream)) != HTR_OK) return rc;
   for (int stage = 1; stage <= 3; stage++) {
      const double *c = RK_C[stage - 1];
      const double dts = implicitChemistry ? dt*c[2]*0.5 : dt*c[2];
      if ((rc = htr_rhs(scheme, diffusion, chemistry && !implicitChemistry, f, tile, bounds, c[0], c[1], dts, stream)) != HTR_OK) return rc;
      if (implicitChemistry) {
         if (rowsContiguous({&f->Conserved_t, &f->Conserved_t_old}, *tile))
            rk_rows_kernel<4><<<rowsGrid(*tile), 256, 0, (cudaStream_t)stream>>>(*f, *tile, 0, 0.0);
         else
            subtract_old_rhs_kernel<<<blocksFor(*tile, 128), 128, 0, (cudaStream_t)stream>>>(*f, *tile);
         if ((rc = cuerr(cudaGetLastError())) != HTR_OK) return rc;
      }
      if ((rc = htr_rk_stage(stage, dt, implicitChemistry, f, tile, stream)) != HTR_OK) return rc;
   }
   return HTR_OK;
}

// ---- host-buffer entry point (what a host-resident caller binds) ------------------------------
// The solver state a host owns is Conserved (+ the previous temperature, the Newton guess); every
// other field the RHS reads is derived from it on the device.  One call = H2D of the state,
// UpdatePrimitiveFromConserved + UpdatePropertiesFromPrimitive (src/prometeo.rg:450-471),
// [sensors], UpdateDerivatives, D2H of Conserved_t.  `f` is a device workspace whose tile is dense
// ([z][y][x][comp], no halo), metrics and node types already resident.
static bool denseField(const htr_acc &a, const htr_rect &t, int64_t elemBytes) {
   const int64_t nx = t.hi[0] - t.lo[0] + 1, ny = t.hi[1] - t.lo[1] + 1;
   return a.stride[0] == elemBytes && a.stride[1] == elemBytes*nx && a.stride[2] == elemBytes*nx*ny;
}
static void *cellPtr(const htr_acc &a, const htr_rect &t) {
   return static_cast<char*>(a.base) + t.lo[0]*a.stride[0] + t.lo[1]*a.stride[1] + t.lo[2]*a.stride[2];
}

// z-slab pipeline of htr_host_rhs: uploads (copy stream), device work (caller's stream) and downloads (second copy
// stream) of different slabs overlap.  The stages of a slab are skewed in z by the reach of their stencils, so every
// stage only needs planes whose previous stage has already been issued on the same stream:
//   P  primitives+properties of the uploaded planes [z0, z1]
//   D  Ducros sensor of        [z0 - 2, z1 - 2]      (velocity gradients: 1 plane, 2 on one-sided nodes)
//   S  shock sensors of        [z0 - 5, z1 - 5]      (6-point stencil: 3 planes of Conserved and Ducros)
//   R  htr_rhs + download of   [z0 - 8, z1 - 8]      (3 planes of state, properties and sensors)
// The first planes of each stage (whose stencils reach back across the periodic wrap to planes that are uploaded last)
// and the last ones are done after the last upload: 2 x 8 planes.  A download therefore trails its upload by one slab
// and 8 planes, and the call takes about max(H2D, D2H) of the whole state plus that lag.
constexpr int PIPE_MAXSLABS = 128;
struct HostPipe { cudaStream_t up, down; cudaEvent_t evUp[PIPE_MAXSLABS], evR[PIPE_MAXSLABS + 2], evStart, evDone; bool ok; };
static HostPipe g_pipe_dev[HTR_MAX_DEVICES] = {};
#define g_pipe g_pipe_dev[curDev()]
static bool pipeInit() {
   if (g_pipe.ok) return true;
   if (cudaStreamCreateWithFlags(&g_pipe.up, cudaStreamNonBlocking) != cudaSuccess) return false;
   if (cudaStreamCreateWithFlags(&g_pipe.down, cudaStreamNonBlocking) != cudaSuccess) return false;
   for (int i = 0; i < PIPE_MAXSLABS; i++)
      if (cudaEventCreateWithFlags(&g_pipe.evUp[i], cudaEventDisableTiming) != cudaSuccess) return false;
   for (int i = 0; i < PIPE_MAXSLABS + 2; i++)
      if (cudaEventCreateWithFlags(&g_pipe.evR[i], cudaEventDisableTiming) != cudaSuccess) return false;
   if (cudaEventCreateWithFlags(&g_pipe.evStart, cudaEventDisableTiming) != cudaSuccess) return false;
   if (cudaEventCreateWithFlags(&g_pipe.evDone, cudaEventDisableTiming) != cudaSuccess) return false;
   g_pipe.ok = true;
   return true;
}

int htr_host_rhs(int scheme, int diffusion, int chemistry, int update_sensors, double vorticityScale,
                 const htr_fluid *f, const htr_rect *tile, const htr_rect *bounds,
                 const double *h_Conserved, const double *h_Conserved_old, double *h_temperature,
                 double *h_Conserved_t, double RK0, double RK1, double deltaTime, void *stream) {
   HTR_CHECK_COMMON(f, tile)
   if (!bounds || !h_Conserved || !h_temperature || !h_Conserved_t) return HTR_EINVAL;
   if (!denseField(f->Conserved, *tile, 8*nEq) || !denseField(f->Conserved_t, *tile, 8*nEq) ||
       !denseField(f->temperature, *tile, 8) || !denseField(f->Conserved_old, *tile, 8*nEq)) return HTR_EINVAL;
   cudaStream_t st = (cudaStream_t)stream;
   const int64_t nx = tile->hi[0] - tile->lo[0] + 1, ny = tile->hi[1] - tile->lo[1] + 1, nz = tile->hi[2] - tile->lo[2] + 1;
   const bool needOld = (nSpec > 1 && scheme != HTR_SCHEME_SKEWSYMMETRIC);
   const bool sensors = update_sensors && scheme == HTR_SCHEME_HYBRID;
   // slab thickness: a multiple of the planes one CTA of the z kernels updates (Euler-z and diffusion-z: 7)
   constexpr int64_t TZ = fast::ZT_Z - 1;
   // 14 planes: the download of a slab trails its upload by one slab + 8 planes, and that lag is what the call costs beyond
   // max(H2D, D2H); measured on B200 at 512^3 (profiles/r2_v18_*): 28-plane slabs 236.3 ms, 14 planes 230.7, 7 planes 230.7
   int64_t T = 2*TZ;
   if (g_slab_tiles > 0) T = g_slab_tiles*TZ;        // htr_set_option("slab", k): slabs of 7k planes
   while ((nz + T - 1)/T > PIPE_MAXSLABS) T += TZ;
   const bool pipelined = nz >= 60 && !g_no_pipe && pipeInit();
   const int ns = pipelined ? (int)((nz + T - 1)/T) : 1;
   cudaError_t e;
   int rc;
   // all ranges below are planes relative to the tile, inclusive; empty ranges are skipped
   auto box = [&](int64_t zlo, int64_t zhi) { htr_rect r = *tile; r.lo[2] = tile->lo[2] + zlo; r.hi[2] = tile->lo[2] + zhi; return r; };
   auto upload = [&](int64_t zlo, int64_t zhi, cudaStream_t s) -> cudaError_t {
      if (zlo > zhi) return cudaSuccess;
      const htr_rect r = box(zlo, zhi);
      const size_t off = (size_t)zlo*nx*ny, cells = (size_t)(zhi - zlo + 1)*nx*ny;
      cudaError_t q;
      if ((q = cudaMemcpyAsync(cellPtr(f->Conserved, r), h_Conserved + off*nEq, cells*8*nEq, cudaMemcpyHostToDevice, s)) != cudaSuccess) return q;
      if (needOld && h_Conserved_old &&
          (q = cudaMemcpyAsync(cellPtr(f->Conserved_old, r), h_Conserved_old + off*nEq, cells*8*nEq, cudaMemcpyHostToDevice, s)) != cudaSuccess) return q;
      return cudaMemcpyAsync(cellPtr(f->temperature, r), h_temperature + off, cells*8, cudaMemcpyHostToDevice, s);
   };
   auto download = [&](int64_t zlo, int64_t zhi, cudaStream_t s) -> cudaError_t {
      const htr_rect r = box(zlo, zhi);
      const size_t off = (size_t)zlo*nx*ny, cells = (size_t)(zhi - zlo + 1)*nx*ny;
      cudaError_t q;
      if ((q = cudaMemcpyAsync(h_Conserved_t + off*nEq, cellPtr(f->Conserved_t, r), cells*8*nEq, cudaMemcpyDeviceToHost, s)) != cudaSuccess) return q;
      return cudaMemcpyAsync(h_temperature + off, cellPtr(f->temperature, r), cells*8, cudaMemcpyDeviceToHost, s);
   };
   auto P = [&](int64_t zlo, int64_t zhi) -> int {
      if (zlo > zhi) return HTR_OK;
      const htr_rect r = box(zlo, zhi);
      int q;
      if ((q = htr_update_primitive_from_conserved(f, &r, stream)) != HTR_OK) return q;
      if ((q = htr_update_properties_from_primitive(f, &r, stream)) != HTR_OK) return q;
      if (needOld && !h_Conserved_old) {       // no separate old state: device copy instead of a second upload
         const size_t cells = (size_t)(zhi - zlo + 1)*nx*ny;
         if (cudaMemcpyAsync(cellPtr(f->Conserved_old, r), cellPtr(f->Conserved, r), cells*8*nEq, cudaMemcpyDeviceToDevice, st) != cudaSuccess) return 1000;
      }
      return HTR_OK;
   };
   auto D = [&](int64_t zlo, int64_t zhi) -> int {
      if (!sensors || zlo > zhi) return HTR_OK;
      const htr_rect r = box(zlo, zhi);
      return htr_ducros_sensor(f, &r, bounds, vorticityScale, stream);
   };
   auto S = [&](int64_t zlo, int64_t zhi) -> int {
      if (!sensors || zlo > zhi) return HTR_OK;
      const htr_rect r = box(zlo, zhi);
      for (int d = 0; d < 3; d++) { const int q = htr_shock_sensor(d, f, &r, bounds, stream); if (q != HTR_OK) return q; }
      return HTR_OK;
   };
   auto R = [&](int64_t zlo, int64_t zhi) -> int {
      const htr_rect r = box(zlo, zhi);
      return htr_rhs(scheme, diffusion, chemistry, f, &r, bounds, RK0, RK1, deltaTime, stream);
   };

   if (ns == 1) {
      if ((e = upload(0, nz - 1, st)) != cudaSuccess) return cuerr(e);
      if ((rc = P(0, nz - 1)) != HTR_OK || (rc = D(0, nz - 1)) != HTR_OK || (rc = S(0, nz - 1)) != HTR_OK || (rc = R(0, nz - 1)) != HTR_OK) return rc;
      return cuerr(download(0, nz - 1, st));
   }
   // the copy streams start after whatever the caller queued before this call
   if ((e = cudaEventRecord(g_pipe.evStart, st)) != cudaSuccess) return cuerr(e);
   cudaStreamWaitEvent(g_pipe.up, g_pipe.evStart, 0);
   cudaStreamWaitEvent(g_pipe.down, g_pipe.evStart, 0);
   for (int k = 0; k < ns; k++) {
      const int64_t z0 = k*T, z1 = (z0 + T < nz ? z0 + T : nz) - 1;
      if ((e = upload(z0, z1, g_pipe.up)) != cudaSuccess) return cuerr(e);
      cudaEventRecord(g_pipe.evUp[k], g_pipe.up);
   }
   int nfin = 0;
   auto finish = [&](int64_t zlo, int64_t zhi) -> int {          // R then its download
      if (zlo > zhi) return HTR_OK;
      int q = R(zlo, zhi);
      if (q != HTR_OK) return q;
      cudaEventRecord(g_pipe.evR[nfin], st);
      cudaStreamWaitEvent(g_pipe.down, g_pipe.evR[nfin], 0);
      nfin++;
      return cuerr(download(zlo, zhi, g_pipe.down));
   };
   const int64_t dD = sensors ? 2 : 0, dS = sensors ? dD + 3 : 0, dR = dS + 3;
   auto lo = [](int64_t z, int64_t d) { return z - d > d ? z - d : d; };
   for (int k = 0; k < ns; k++) {
      const int64_t z0 = k*T, z1 = (z0 + T < nz ? z0 + T : nz) - 1;
      cudaStreamWaitEvent(st, g_pipe.evUp[k], 0);
      if ((rc = P(z0, z1)) != HTR_OK) return rc;
      if ((rc = D(lo(z0, dD), z1 - dD)) != HTR_OK) return rc;
      if ((rc = S(lo(z0, dS), z1 - dS)) != HTR_OK) return rc;
      if ((rc = finish(lo(z0, dR), z1 - dR)) != HTR_OK) return rc;
   }
   // the planes next to the wrap (all uploads are in)
   if ((rc = D(nz - dD, nz - 1)) != HTR_OK || (rc = D(0, dD - 1)) != HTR_OK) return rc;
   if ((rc = S(nz - dS, nz - 1)) != HTR_OK || (rc = S(0, dS - 1)) != HTR_OK) return rc;
   if ((rc = finish(nz - dR, nz - 1)) != HTR_OK || (rc = finish(0, dR - 1)) != HTR_OK) return rc;
   cudaEventRecord(g_pipe.evDone, g_pipe.down);
   cudaStreamWaitEvent(st, g_pipe.evDone, 0);
   return HTR_OK;
}

} // extern "C"

// non-periodic boundaries: NSCBC right-hand-side tasks and boundary-cell primitive updates (htr_nscbc_*, htr_set_*_bc)
#include "htr_nscbc.cuh"

// ghost-plane exchange over NCCL + scalar max-reduction (htr_comm_*, htr_halo_exchange, htr_allreduce_max)
#include "htr_halo.cuh"
