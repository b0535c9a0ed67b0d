// Shared-memory staged fast path of UpdateDerivatives for the smooth / standard-node part of the
// grid (the whole grid of a periodic box):
//
//   euler_z_chem_kernel   Conserved_t  = 0 + chemistry source + Euler-z   (one write, no read)
//   euler_xy_kernel       Conserved_t += Euler-y; += Euler-x              (one read-modify-write)
//
// i.e. the reference's accumulation order 0 -> +chem -> +Ez -> +Ey -> +Ex (src/prometeo.rg:502-578)
// in two passes over Conserved_t instead of five.  The Euler flux is the 6th-order Kennedy-Gruber
// split form of src/prometeo_rhs.inl:42-135,981-1052 with the Std_node row of the stencil tables as
// compile-time constant-bank operands.
//
// Data movement: every stencil cell of a tile is read from HBM once, by one thread, field by field
// (a warp covers 32 x-consecutive cells, so each AoS field row is one contiguous span), reduced to
// the 6+nSpec numbers the split form needs (rho, p, H = (rhoE+p)/rho, u, Y_i -- one division per
// staged cell instead of one per stencil point per face) and stored struct-of-arrays in shared
// memory; the 6-point stencils of the faces are then conflict-free 8-byte shared loads.  Face
// fluxes are exchanged between neighbouring threads by warp shuffle (x) or through the retired
// staging buffer (y, z).  Tiles overlap by one cell along each flux direction: thread (i,j) owns
// the PLUS face of its cell and updates the cell if it also received the minus face.
//
// What the fast path does not do, it defers: a cell is flagged in `mask` and left for the
// generic per-face kernels (htr_kernels.cu, masked variants) when
//   * any node type on its tile is not Std_node in the flux direction (boundaries), or
//   * the scheme is Hybrid and one of its faces is not smooth (shockSensor, src/prometeo_rhs.inl:767-889:
//     that face needs the TENO flux).
// Deferred cells receive exactly the reference's sequence of += from the generic kernels, so the
// result does not depend on which path a cell took.
#pragma once
#include "htr_euler.cuh"

namespace htr {
namespace fast {

constexpr int NV = nSpec + 6;                 // staged variables per cell
enum { V_RHO = 0, V_P = 1, V_H = 2, V_U = 3, V_Y = 6 };

HTR_DEV int64_t wrapCoord(int64_t c, int64_t lo, int64_t hi) {
   if (c > hi || c < lo) {
      const int64_t size = hi - lo + 1;
      c = ((c - lo) % size + size) % size + lo;
   }
   return c;
}

template<int NC>
HTR_DEV void stageCell(double *s, const int idx, const htr_fluid &f, const Pt &q) {
   const double rho = ld<double>(f.rho, q);
   const double p = ld<double>(f.pressure, q);
   s[V_RHO*NC + idx] = rho;
   s[V_P*NC + idx] = p;
   s[V_H*NC + idx] = (ldc<double>(f.Conserved, q, irE) + p)/rho;
   HTR_UNROLL
   for (int i = 0; i < 3; i++) s[(V_U + i)*NC + idx] = ldc<double>(f.velocity, q, i);
   HTR_UNROLL
   for (int i = 0; i < nSpec; i++) s[(V_Y + i)*NC + idx] = ldc<double>(f.MassFracs, q, i);
}

template<int dir, int NC>
HTR_DEV void ldsPoint(KGPoint &c, const double *s, const int idx) {
   c.rho = s[V_RHO*NC + idx];
   c.p = s[V_P*NC + idx];
   c.H = s[V_H*NC + idx];
   HTR_UNROLL
   for (int i = 0; i < 3; i++) c.u[i] = s[(V_U + i)*NC + idx];
   HTR_UNROLL
   for (int i = 0; i < nSpec; i++) c.Y[i] = s[(V_Y + i)*NC + idx];
   c.un = c.u[dir];
}

// KG flux of the face between staged cells i0+2*step and i0+3*step (stencil i0 .. i0+5*step), Std_node
template<int dir, int NC>
HTR_DEV void kgStdFlux(double *Flux, const double *s, const int i0, const int step) {
   KGPoint c[6];
   HTR_UNROLL
   for (int k = 0; k < 6; k++) ldsPoint<dir, NC>(c[k], s, i0 + k*step);
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

HTR_DEV int64_t maskIndex(const Pt &c, const htr_rect &tile) {
   const int64_t nx = tile.hi[0] - tile.lo[0] + 1, ny = tile.hi[1] - tile.lo[1] + 1;
   return ((c.z - tile.lo[2])*ny + (c.y - tile.lo[1]))*nx + (c.x - tile.lo[0]);
}

// ---------------------------------------------------------------------------------------------
// Conserved_t = 0 (+ chemistry source) + Euler-z
// block = 32 (x) x ZT_Z (z) cells of one y row; z rows overlap by one between blocks
// ---------------------------------------------------------------------------------------------
constexpr int ZT_X = 32, ZT_Z = 16, ZT_ROWS = ZT_Z + 5, ZT_NC = ZT_ROWS*ZT_X, ZT_NT = ZT_X*ZT_Z;
constexpr size_t ZT_SMEM = sizeof(double)*NV*ZT_NC + ZT_NC + ZT_NT;

template<bool HYBRID, bool CHEM>
__global__ void __launch_bounds__(ZT_NT)
euler_z_chem_kernel(const htr_fluid f, const htr_rect tile, const htr_rect B, uint8_t *__restrict__ mask) {
   extern __shared__ double smem[];
   double *s = smem;
   uint8_t *sSens = reinterpret_cast<uint8_t*>(smem + NV*ZT_NC);
   uint8_t *sFlag = sSens + ZT_NC;
   static_assert(NV*ZT_NC >= nEq*ZT_NT, "flux exchange reuses the staging buffer");

   const int i = threadIdx.x, k = threadIdx.y;
   const int tid = k*ZT_X + i;
   Pt c;
   c.x = tile.lo[0] + (int64_t)blockIdx.x*ZT_X + i;
   c.y = tile.lo[1] + blockIdx.y;
   const int64_t cz0 = tile.lo[2] - 1 + (int64_t)blockIdx.z*(ZT_Z - 1);
   c.z = cz0 + k;
   const bool inX = c.x <= tile.hi[0];

   for (int r = k; r < ZT_ROWS; r += ZT_Z) {
      const int64_t z = cz0 - 2 + r;
      if (inX && z >= tile.lo[2] - 3 && z <= tile.hi[2] + 3) {
         const Pt q = {c.x, c.y, wrapCoord(z, B.lo[2], B.hi[2])};
         stageCell<ZT_NC>(s, r*ZT_X + i, f, q);
         if (HYBRID) sSens[r*ZT_X + i] = ld<uint8_t>(f.shockSensorZ, q);
      }
   }
   const bool faceActive = inX && c.z <= tile.hi[2];
   bool bad = false;
   if (faceActive) {
      const Pt q = {c.x, c.y, wrapCoord(c.z, B.lo[2], B.hi[2])};
      bad = ld<int>(f.nType_z, q) != Std_node;
   }
   const bool anyBad = __syncthreads_or(bad);

   double F[nEq];
   bool smooth = true;
   if (!anyBad && faceActive) {
      if (HYBRID) smooth = sSens[(k + 1)*ZT_X + i] && sSens[(k + 2)*ZT_X + i] && sSens[(k + 3)*ZT_X + i];
      if (smooth) kgStdFlux<2, ZT_NC>(F, s, k*ZT_X + i, ZT_X);
   }
   const bool upd = faceActive && k >= 1;
   double base[nEq];
   HTR_UNROLL
   for (int e = 0; e < nEq; e++) base[e] = 0.0;
   if (CHEM && upd) {
      const int own = (k + 2)*ZT_X + i;
      double Yi[nSpec], w[nSpec];
      HTR_UNROLL
      for (int e = 0; e < nSpec; e++) Yi[e] = s[(V_Y + e)*ZT_NC + own];
      mix::GetProductionRates(w, s[V_RHO*ZT_NC + own], s[V_P*ZT_NC + own], ld<double>(f.temperature, c), Yi);
      HTR_UNROLL
      for (int e = 0; e < nSpec; e++) base[e] += w[e];
   }
   __syncthreads();
   double *sF = smem;
   const bool have = faceActive && !anyBad && smooth;
   if (have) {
      HTR_UNROLL
      for (int e = 0; e < nEq; e++) sF[e*ZT_NT + tid] = F[e];
   }
   sFlag[tid] = have;
   __syncthreads();
   if (upd) {
      const bool deferred = !(have && sFlag[tid - ZT_X]);
      if (!deferred) {
         const double m = ld<double>(f.dzet_e, c);
         HTR_UNROLL
         for (int e = 0; e < nEq; e++) base[e] += m*(F[e] - sF[e*ZT_NT + tid - ZT_X]);
      }
      double *Ct = at<double>(f.Conserved_t, c);
      HTR_UNROLL
      for (int e = 0; e < nEq; e++) Ct[e] = base[e];
      mask[maskIndex(c, tile)] = deferred;
   }
}

// ---------------------------------------------------------------------------------------------
// Conserved_t += Euler-y; Conserved_t += Euler-x
// block = 32 (x) x TY (y) cells of one z plane; blocks overlap by one cell in x and y
// ---------------------------------------------------------------------------------------------
template<int TY> struct XYTile {
   static constexpr int W = 37, ROWS = TY + 5, NC = W*ROWS, NT = 32*TY;
   static constexpr size_t SMEM = sizeof(double)*NV*NC + 2*NC + NT;
};

template<int TY, bool HYBRID>
__global__ void __launch_bounds__(32*TY)
euler_xy_kernel(const htr_fluid f, const htr_rect tile, const htr_rect B, uint8_t *__restrict__ mask) {
   using T = XYTile<TY>;
   constexpr int W = T::W, NC = T::NC, NT = T::NT;
   extern __shared__ double smem[];
   double *s = smem;
   uint8_t *sSensX = reinterpret_cast<uint8_t*>(smem + NV*NC);
   uint8_t *sSensY = sSensX + NC;
   uint8_t *sFlag = sSensY + NC;
   static_assert(NV*NC >= nEq*NT, "flux exchange reuses the staging buffer");

   const int i = threadIdx.x, j = threadIdx.y;
   const int tid = j*32 + i;
   const int64_t cx0 = tile.lo[0] - 1 + (int64_t)blockIdx.x*31;
   const int64_t cy0 = tile.lo[1] - 1 + (int64_t)blockIdx.y*(TY - 1);
   Pt c;
   c.x = cx0 + i; c.y = cy0 + j; c.z = tile.lo[2] + blockIdx.z;

   for (int n = tid; n < NC; n += NT) {
      const int rr = n/W, cc = n - rr*W;
      const int64_t x = cx0 - 2 + cc, y = cy0 - 2 + rr;
      const bool need = ((rr >= 3 && rr <= TY + 1) || (cc >= 3 && cc <= 33)) &&
                        x >= tile.lo[0] - 3 && x <= tile.hi[0] + 3 && y >= tile.lo[1] - 3 && y <= tile.hi[1] + 3;
      if (need) {
         const Pt q = {wrapCoord(x, B.lo[0], B.hi[0]), wrapCoord(y, B.lo[1], B.hi[1]), c.z};
         stageCell<NC>(s, n, f, q);
         if (HYBRID) {
            sSensX[n] = ld<uint8_t>(f.shockSensorX, q);
            sSensY[n] = ld<uint8_t>(f.shockSensorY, q);
         }
      }
   }
   const bool inTile = c.x <= tile.hi[0] && c.y <= tile.hi[1];
   const bool xAct = inTile && j >= 1, yAct = inTile && i >= 1;
   bool bad = false;
   if (xAct || yAct) {
      const Pt q = {wrapCoord(c.x, B.lo[0], B.hi[0]), wrapCoord(c.y, B.lo[1], B.hi[1]), c.z};
      if (xAct) bad = ld<int>(f.nType_x, q) != Std_node;
      if (yAct) bad = bad || (ld<int>(f.nType_y, q) != Std_node);
   }
   const bool anyBad = __syncthreads_or(bad);

   const int own = (j + 2)*W + i + 2;
   double Fx[nEq], Fy[nEq];
   HTR_UNROLL
   for (int e = 0; e < nEq; e++) { Fx[e] = 0.0; Fy[e] = 0.0; }
   bool hx = false, hy = false;
   if (!anyBad) {
      if (yAct) {
         hy = !HYBRID || (sSensY[own - W] && sSensY[own] && sSensY[own + W]);
         if (hy) kgStdFlux<1, NC>(Fy, s, j*W + i + 2, W);
      }
      if (xAct) {
         hx = !HYBRID || (sSensX[own - 1] && sSensX[own] && sSensX[own + 1]);
         if (hx) kgStdFlux<0, NC>(Fx, s, (j + 2)*W + i, 1);
      }
   }
   // minus x face: lane i-1 of the same warp (a warp is one row of the block)
   double Fm[nEq];
   HTR_UNROLL
   for (int e = 0; e < nEq; e++) Fm[e] = __shfl_up_sync(0xffffffffu, Fx[e], 1);
   const bool hxm = __shfl_up_sync(0xffffffffu, (int)hx, 1);
   __syncthreads();
   double *sF = smem;
   if (hy) {
      HTR_UNROLL
      for (int e = 0; e < nEq; e++) sF[e*NT + tid] = Fy[e];
   }
   sFlag[tid] = hy;
   __syncthreads();
   if (inTile && i >= 1 && j >= 1) {
      const bool deferred = !(hx && hxm && hy && sFlag[tid - 32]);
      if (!deferred) {
         double *Ct = at<double>(f.Conserved_t, c);
         const double my = ld<double>(f.deta_e, c), mx = ld<double>(f.dcsi_e, c);
         HTR_UNROLL
         for (int e = 0; e < nEq; e++) {
            double v = Ct[e];
            v += my*(Fy[e] - sF[e*NT + tid - 32]);
            v += mx*(Fx[e] - Fm[e]);
            Ct[e] = v;
         }
      }
      mask[maskIndex(c, tile)] = deferred;
   }
}

} // namespace fast
} // namespace htr
