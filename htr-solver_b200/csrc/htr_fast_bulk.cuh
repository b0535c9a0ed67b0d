// Euler-y + Euler-x kernel of the fast path with TMA bulk staging (see htr_bulk.cuh for the copy
// mechanics and htr_fast.cuh for the algorithm: same tiles, same arithmetic, same deferral rules as
// fast::euler_xy_kernel -- only the way the tile reaches shared memory and its layout there differ).
#pragma once
#include <cuda.h>            // CUtensorMap (type only; the encoder is fetched with cudaGetDriverEntryPoint)
#include "htr_fast.cuh"
#include "htr_bulk.cuh"

namespace htr {
namespace fast {

constexpr int pad16(int n) { return (n + 15)/16*16; }     // doubles: 128-byte sections (TMA tensor destinations)
template<int TY> struct XYB {
   static constexpr int W = 38, ROWS = TY + 5, NC = W*ROWS, NT = 32*TY;
   // tiles in doubles, each in the field's own [row][x][comp] order (= one TMA box per field); H is derived after arrival
   static constexpr int O_RHO = 0, O_P = pad16(NC), O_U = O_P + pad16(NC), O_Y = O_U + pad16(3*NC), O_H = O_Y + pad16(nSpec*NC),
                        ND = O_H + pad16(NC);
   static constexpr uint32_t ROW_BYTES = 8u*W*(1 + 1 + 3 + nSpec);        // one row of the four copied fields
   static constexpr size_t SMEM = 128 + sizeof(double)*ND + 2*NC + NT;     // mbarrier in the first 128 bytes
};

// Tensor maps of the four copied fields over the part of the allocation this call may read (host: makeTensorMaps in
// htr_kernels.cu): field k is a 3-D tensor (ncomp*nx, ny, nz) of doubles whose element (0,0,0) is cell (x0, y0, z0)
// (tile-relative); one cp.async.bulk.tensor per field moves the whole W x ROWS tile, AoS fields as a merged (comp, x)
// inner dimension.
struct XYMaps { CUtensorMap m[4]; int x0, y0, z0, ok; };

__device__ __forceinline__ void tensorLoad3D(void *dst, const CUtensorMap *map, int c0, int c1, int c2, uint64_t *bar) {
   asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
                :: "r"(bulk::sAddr(dst)), "l"(map), "r"(c0), "r"(c1), "r"(c2), "r"(bulk::sAddr(bar)) : "memory");
}

template<int dir, int NC>
HTR_DEV void ldsPointRaw(KGPoint &c, const double *sRho, const double *sP, const double *sU, const double *sY, const double *sH, const int idx) {
   c.rho = sRho[idx];
   c.p = sP[idx];
   c.H = sH[idx];
   HTR_UNROLL
   for (int i = 0; i < 3; i++) c.u[i] = sU[3*idx + i];
   HTR_UNROLL
   for (int i = 0; i < nSpec; i++) c.Y[i] = sY[nSpec*idx + i];
   c.un = c.u[dir];
}

template<int dir, class T>
HTR_DEV void kgStdFluxRaw(double *Flux, const double *d, const int i0, const int step) {
   using namespace fast;
   KGPoint c[6];
   HTR_UNROLL
   for (int k = 0; k < 6; k++) ldsPointRaw<dir, T::NC>(c[k], d + T::O_RHO, d + T::O_P, d + T::O_U, d + T::O_Y, d + T::O_H, i0 + k*step);
   double fl[nEq + 1];
   HTR_UNROLL
   for (int i = 0; i < nEq + 1; i++) fl[i] = 0.0;
   HTR_UNROLL
   for (int l = 0; l < 3; l++) {
      double acc[nEq + 1];
      HTR_UNROLL
      for (int i = 0; i < nEq + 1; i++) acc[i] = 0.0;
      HTR_UNROLL
      for (int m = 0; m <= l; m++) {
         double S[nEq + 1];
         kgPair(S, c[2 - m], c[2 - m + l + 1]);
         HTR_UNROLL
         for (int i = 0; i < nEq + 1; i++) acc[i] += S[i];
      }
      const double cl = c_KennedyCoeff[Std_node][l];
      HTR_UNROLL
      for (int i = 0; i < nEq + 1; i++) fl[i] += cl*acc[i];
   }
   HTR_UNROLL
   for (int i = 0; i < nEq; i++) Flux[i] = 0.25*fl[i];
   Flux[irU + dir] -= fl[nEq];
}

template<int TY, bool HYBRID>
__global__ void __launch_bounds__(32*TY, HTR_K2_MINB)
euler_xy_bulk_kernel(const rfluid f, const rbox tile, const rbox B, uint8_t *__restrict__ mask, int *__restrict__ flag,
                     const int bulkOK, const __grid_constant__ XYMaps maps) {
   using T = XYB<TY>;
   constexpr int W = T::W, NC = T::NC, NT = T::NT, ROWS = T::ROWS;
   extern __shared__ __align__(128) unsigned char smraw[];
   uint64_t *bar = reinterpret_cast<uint64_t*>(smraw);
   double *d = reinterpret_cast<double*>(smraw + 128);
   uint8_t *sSensX = reinterpret_cast<uint8_t*>(d + T::ND);
   uint8_t *sSensY = sSensX + NC;
   uint8_t *sFlag = sSensY + NC;
   static_assert(T::ND >= nEq*NT, "flux exchange reuses the staging buffer");

   const int i = threadIdx.x, j = threadIdx.y;
   const int tid = j*32 + i;
   const int cx0 = tile.lo[0] - 1 + (int)blockIdx.x*31;
   const int cy0 = tile.lo[1] - 1 + (int)blockIdx.y*(TY - 1);
   // first column of the tile: cx0-2 or cx0-3, whichever makes the row start 16-byte aligned
   // (bit 1 of bulkOK = parity of base/8, common to the copied fields)
   const int xs = (cx0 - 2) - (((cx0 - 2) + ((bulkOK >> 1) & 1)) & 1);
   const int co = (int)((cx0 - 2) - xs);                    // column of stencil offset -2 of thread i = 0
   RPt c;
   c.x = cx0 + i; c.y = cy0 + j; c.z = tile.lo[2] + blockIdx.z;

   // can whole rows be copied?  (inside the allocation, not cut by the periodic wrap)
   const int xlo = (tile.lo[0] - 3 > B.lo[0]) ? tile.lo[0] - 3 : B.lo[0];
   const int xhi = (tile.hi[0] + 3 < B.hi[0]) ? tile.hi[0] + 3 : B.hi[0];
   const bool useBulk = (bulkOK & 1) && xs >= xlo && xs + W - 1 <= xhi;

   if (tid == 0) bulk::mbarInit(bar, 1);
   __syncthreads();

   auto rowValid = [&](int rr) { const int y = cy0 - 2 + rr; return y >= tile.lo[1] - 3 && y <= tile.hi[1] + 3; };
   // one tensor copy per field when no needed row of the tile leaves the box in y (those tiles wrap row by row below);
   // rows / columns outside the mapped region are zero-filled by the TMA unit and never read by an updated cell
   const int yFirst = (cy0 - 2 > tile.lo[1] - 3) ? cy0 - 2 : tile.lo[1] - 3;
   const int yLast = (cy0 - 2 + ROWS - 1 < tile.hi[1] + 3) ? cy0 - 2 + ROWS - 1 : tile.hi[1] + 3;
   const bool useTensor = useBulk && maps.ok && yFirst >= B.lo[1] && yLast <= B.hi[1];
   if (useTensor) {
      if (tid == 0) {
         bulk::mbarExpectTx(bar, (uint32_t)ROWS*T::ROW_BYTES);
         const int cx = xs - maps.x0, cy = (cy0 - 2) - maps.y0, cz = c.z - maps.z0;
         tensorLoad3D(d + T::O_RHO, &maps.m[0], cx, cy, cz, bar);
         tensorLoad3D(d + T::O_P, &maps.m[1], cx, cy, cz, bar);
         tensorLoad3D(d + T::O_U, &maps.m[2], 3*cx, cy, cz, bar);
         tensorLoad3D(d + T::O_Y, &maps.m[3], nSpec*cx, cy, cz, bar);
      }
   } else if (useBulk) {
      if (j == 0) {                                          // warp 0 issues the copies
         if (i == 0) {
            int nv = 0;
            for (int rr = 0; rr < ROWS; rr++) nv += rowValid(rr) ? 1 : 0;
            bulk::mbarExpectTx(bar, (uint32_t)nv*T::ROW_BYTES);
         }
         __syncwarp();
         for (int item = i; item < 4*ROWS; item += 32) {
            const int rr = item >> 2, fld = item & 3;
            if (!rowValid(rr)) continue;
            const RPt q = {xs, wrapCoord(cy0 - 2 + rr, B.lo[1], B.hi[1]), c.z};
            if (fld == 0)      bulk::copyRow(d + T::O_RHO + rr*W, at<const double>(f.rho, q), 8u*W, bar);
            else if (fld == 1) bulk::copyRow(d + T::O_P + rr*W, at<const double>(f.pressure, q), 8u*W, bar);
            else if (fld == 2) bulk::copyRow(d + T::O_U + 3*rr*W, at<const double>(f.velocity, q), 24u*W, bar);
            else               bulk::copyRow(d + T::O_Y + nSpec*rr*W, at<const double>(f.MassFracs, q), 8u*nSpec*W, bar);
         }
      }
   } else {
      // cooperative loads, one cell per thread and pass (edge tiles, unaligned layouts)
      for (int n = tid; n < NC; n += NT) {
         const int rr = n/W, cc = n - rr*W;
         const int x = xs + cc, y = cy0 - 2 + rr;
         if (x >= tile.lo[0] - 3 && x <= tile.hi[0] + 3 && y >= tile.lo[1] - 3 && y <= tile.hi[1] + 3) {
            const RPt q = {wrapCoord(x, B.lo[0], B.hi[0]), wrapCoord(y, B.lo[1], B.hi[1]), c.z};
            d[T::O_RHO + n] = ld<double>(f.rho, q);
            d[T::O_P + n] = ld<double>(f.pressure, q);
            HTR_UNROLL
            for (int k = 0; k < 3; k++) d[T::O_U + 3*n + k] = ldc<double>(f.velocity, q, k);
            HTR_UNROLL
            for (int k = 0; k < nSpec; k++) d[T::O_Y + nSpec*n + k] = ldc<double>(f.MassFracs, q, k);
         }
      }
   }

   // what the threads fetch themselves while the rows fly: rhoE (one component of nEq), sensors, node types
   constexpr int PER = (NC + NT - 1)/NT;
   double rhoE[PER];
   HTR_UNROLL
   for (int k = 0; k < PER; k++) {
      const int n = tid + k*NT;
      rhoE[k] = 0.0;
      if (n < NC) {
         const int rr = n/W, cc = n - rr*W;
         const int x = xs + cc, y = cy0 - 2 + rr;
         const int rc = cc - co;                             // column relative to stencil offset -2 of i = 0
         const bool need = ((rr >= 3 && rr <= TY + 1 && rc >= 0 && rc <= 36) || (rc >= 3 && rc <= 33)) &&
                           x >= tile.lo[0] - 3 && x <= tile.hi[0] + 3 && y >= tile.lo[1] - 3 && y <= tile.hi[1] + 3;
         if (need) {
            const RPt q = {wrapCoord(x, B.lo[0], B.hi[0]), wrapCoord(y, B.lo[1], B.hi[1]), c.z};
            rhoE[k] = ldc<double>(f.Conserved, q, irE);
            if (HYBRID) {
               sSensX[n] = ld<uint8_t>(f.shockSensorX, q);
               sSensY[n] = ld<uint8_t>(f.shockSensorY, q);
            }
         }
      }
   }
   const bool inTile = c.x <= tile.hi[0] && c.y <= tile.hi[1];
   const bool xAct = inTile && j >= 1, yAct = inTile && i >= 1;
   // the cell's own Conserved_t and metrics are requested now, so that their latency hides behind the fluxes
   const bool upd = inTile && i >= 1 && j >= 1;
   double *Ct = at<double>(f.Conserved_t, c);
   double ct[nEq], my = 0.0, mx = 0.0;
   if (upd) {
      HTR_UNROLL
      for (int e = 0; e < nEq; e++) ct[e] = Ct[e];
      my = ld<double>(f.deta_e, c); mx = ld<double>(f.dcsi_e, c);
   }
   bool bad = false;
   if (xAct || yAct) {
      const RPt q = {wrapCoord(c.x, B.lo[0], B.hi[0]), wrapCoord(c.y, B.lo[1], B.hi[1]), c.z};
      if (xAct) bad = ld<int>(f.nType_x, q) != Std_node;
      if (yAct) bad = bad || (ld<int>(f.nType_y, q) != Std_node);
   }
   if (useBulk) bulk::mbarWait(bar, 0);       // tensor or row copies: same barrier phase
   else __syncthreads();
   HTR_UNROLL
   for (int k = 0; k < PER; k++) {
      const int n = tid + k*NT;
      if (n < NC) d[T::O_H + n] = HTR_QDIV(rhoE[k] + d[T::O_P + n], d[T::O_RHO + n]);
   }
   const bool anyBad = __syncthreads_or(bad);

   const int own = (j + 2)*W + i + 2 + co;
   double Fx[nEq], Fy[nEq];
   HTR_UNROLL
   for (int e = 0; e < nEq; e++) { Fx[e] = 0.0; Fy[e] = 0.0; }
   bool hx = false, hy = false;
   if (!anyBad) {
      if (yAct) {
         hy = !HYBRID || (sSensY[own - W] && sSensY[own] && sSensY[own + W]);
         if (hy) kgStdFluxRaw<1, T>(Fy, d, own - 2*W, W);
      }
      if (xAct) {
         hx = !HYBRID || (sSensX[own - 1] && sSensX[own] && sSensX[own + 1]);
         if (hx) kgStdFluxRaw<0, T>(Fx, d, own - 2, 1);
      }
   }
   double Fm[nEq];
   HTR_UNROLL
   for (int e = 0; e < nEq; e++) Fm[e] = __shfl_up_sync(0xffffffffu, Fx[e], 1);
   const bool hxm = __shfl_up_sync(0xffffffffu, (int)hx, 1);
   __syncthreads();
   double *sF = d;
   if (hy) {
      HTR_UNROLL
      for (int e = 0; e < nEq; e++) sF[e*NT + tid] = Fy[e];
   }
   sFlag[tid] = hy;
   __syncthreads();
   bool deferred = true;
   if (upd) {
      deferred = !(hx && hxm && hy && sFlag[tid - 32]);
      if (!deferred) {
         HTR_UNROLL
         for (int e = 0; e < nEq; e++) {
            double v = ct[e];
            v += my*(Fy[e] - sF[e*NT + tid - 32]);
            v += mx*(Fx[e] - Fm[e]);
            ct[e] = v;
         }
      }
      mask[maskIndex(c, tile)] = deferred;
      if (deferred) *flag = 1;
   }
   // rows of Conserved_t leave through shared memory as contiguous spans (deferred cells rewrite their old value)
   if (f.Conserved_t.s[0] == 8*nEq) {
      __syncthreads();
      double *sOut = d + j*(nEq*32);
      if (upd) {
         HTR_UNROLL
         for (int e = 0; e < nEq; e++) sOut[(i - 1)*nEq + e] = ct[e];
      }
      __syncwarp();
      if (j >= 1 && c.y <= tile.hi[1]) {
         RPt r0 = c; r0.x = cx0 + 1;
         const int left = tile.hi[0] - r0.x + 1;
         const int nval = nEq*(int)(left < 31 ? left : 31);
         double *row = at<double>(f.Conserved_t, r0);
         HTR_UNROLL
         for (int e = 0; e < nEq; e++) { const int m = e*32 + i; if (m < nval) row[m] = sOut[m]; }
      }
   } else if (upd && !deferred) {
      HTR_UNROLL
      for (int e = 0; e < nEq; e++) Ct[e] = ct[e];
   }
}

} // namespace fast
} // namespace htr
