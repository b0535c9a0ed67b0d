// ---------------------------------------------------------------------------------------------------------------
// ATTRIBUTION.  The kernel structure of this file is original, but parts of its arithmetic restate, expression for
// expression (bit parity with the reference requires it), formulas of the HTR solver (stanfordhpccenter/HTR-solver,
// src/prometeo_rhs.inl, prometeo_metric.inl, prometeo_sensor.inl, prometeo_variables.hpp, prometeo_cfl.hpp, Mixtures/*,
// Utils/math_utils.hpp; the file:line citations below say which).  That code is distributed under the following notice,
// retained here as its conditions require:
//
// Copyright (c) "2019, by Stanford University
//               Developer: Mario Di Renzo
//               Affiliation: Center for Turbulence Research, Stanford University
//               URL: https://ctr.stanford.edu
//               Citation: Di Renzo, M., Lin, F., and Urzay, J. (2020).
//                         HTR solver: An open-source exascale-oriented task-based
//                         multi-GPU high-order code for hypersonic aerothermodynamics.
//                         Computer Physics Communications 255, 107262"
// All rights reserved.
//
// Redistribution and use in source and binary forms, with or without
// modification, are permitted provided that the following conditions are met:
//    * Redistributions of source code must retain the above copyright
//      notice, this list of conditions and the following disclaimer.
//    * Redistributions in binary form must reproduce the above copyright
//      notice, this list of conditions and the following disclaimer in the
//      documentation and/or other materials provided with the distribution.
//
// THIS SOFTWARE IS PROVIDED BY THE COPYRIGHT HOLDERS AND CONTRIBUTORS "AS IS" AND
// ANY EXPRESS OR IMPLIED WARRANTIES, INCLUDING, BUT NOT LIMITED TO, THE IMPLIED
// WARRANTIES OF MERCHANTABILITY AND FITNESS FOR A PARTICULAR PURPOSE ARE
// DISCLAIMED. IN NO EVENT SHALL THE COPYRIGHT HOLDER BE LIABLE FOR ANY
// DIRECT, INDIRECT, INCIDENTAL, SPECIAL, EXEMPLARY, OR CONSEQUENTIAL DAMAGES
// (INCLUDING, BUT NOT LIMITED TO, PROCUREMENT OF SUBSTITUTE GOODS OR SERVICES;
// LOSS OF USE, DATA, OR PROFITS; OR BUSINESS INTERRUPTION) HOWEVER CAUSED AND
// ON ANY THEORY OF LIABILITY, WHETHER IN CONTRACT, STRICT LIABILITY, OR TORT
// (INCLUDING NEGLIGENCE OR OTHERWISE) ARISING IN ANY WAY OUT OF THE USE OF THIS
// SOFTWARE, EVEN IF ADVISED OF THE POSSIBILITY OF SUCH DAMAGE.
// ---------------------------------------------------------------------------------------------------------------
// Shared-memory staged fast path of UpdateDerivatives for the smooth / standard-node part of the
// grid (the whole grid of a periodic box):
//
//   euler_z_chem_kernel   Conserved_t  = 0 + chemistry source + Euler-z   (one write, no read)
//   euler_xy_bulk_kernel  Conserved_t += Euler-y; += Euler-x              (one read-modify-write; htr_fast_bulk.cuh)
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

// resident CTAs per SM the register allocation aims at (tuned on B200 with ncu, see profiles/)
#ifndef HTR_K1_MINB
#define HTR_K1_MINB 3
#endif
#ifndef HTR_K2_MINB
#define HTR_K2_MINB 3
#endif
#ifndef HTR_D_MINB
#define HTR_D_MINB 3
#endif

// tile heights (rows of 32 cells) of the in-plane Euler kernel and of the diffusion kernels
#ifndef HTR_K2_TY
#define HTR_K2_TY 8
#endif
#ifndef HTR_D_TJ
#define HTR_D_TJ 8
#endif
#ifndef HTR_D_HOIST
#define HTR_D_HOIST 0
#endif

namespace htr {
namespace fast {

constexpr int NV = nSpec + 6;                 // staged variables per cell
enum { V_RHO = 0, V_P = 1, V_H = 2, V_U = 3, V_Y = 6 };

HTR_DEV int wrapCoord(int c, int lo, int hi) {
   if (c > hi || c < lo) {
      const int size = hi - lo + 1;
      c = ((c - lo) % size + size) % size + lo;
   }
   return c;
}

template<int NC>
HTR_DEV void stageCell(double *s, const int idx, const rfluid &f, const RPt &q) {
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

HTR_DEV int64_t maskIndex(const RPt &c, const rbox &tile) {      // tile.lo = 0
   const int64_t nx = tile.hi[0] + 1, ny = tile.hi[1] + 1;
   return ((int64_t)c.z*ny + c.y)*nx + c.x;
}

// ---------------------------------------------------------------------------------------------
// Conserved_t = 0 (+ chemistry source) + Euler-z
// block = 32 (x) x ZT_Z (z) cells of one y row; z rows overlap by one between blocks
// ---------------------------------------------------------------------------------------------
#ifndef HTR_K1_TZ
#define HTR_K1_TZ 8
#endif
constexpr int ZT_X = 32, ZT_Z = HTR_K1_TZ, ZT_ROWS = ZT_Z + 5, ZT_NC = ZT_ROWS*ZT_X, ZT_NT = ZT_X*ZT_Z;
constexpr size_t ZT_SMEM = sizeof(double)*NV*ZT_NC + ZT_NC + ZT_NT;

template<bool HYBRID, bool CHEM>
__global__ void __launch_bounds__(ZT_NT, HTR_K1_MINB)
euler_z_chem_kernel(const rfluid f, const rbox tile, const rbox B, uint8_t *__restrict__ mask, int *__restrict__ flag) {
   extern __shared__ double smem[];
   double *s = smem;
   uint8_t *sSens = reinterpret_cast<uint8_t*>(smem + NV*ZT_NC);
   uint8_t *sFlag = sSens + ZT_NC;
   static_assert(NV*ZT_NC >= nEq*ZT_NT, "flux exchange reuses the staging buffer");

   const int i = threadIdx.x, k = threadIdx.y;
   const int tid = k*ZT_X + i;
   RPt c;
   // grid = (z blocks, x blocks, y): consecutive CTAs are z-neighbours, whose tiles share 6 of 21 planes in L2
   c.x = tile.lo[0] + (int)blockIdx.y*ZT_X + i;
   c.y = tile.lo[1] + blockIdx.z;
   const int cz0 = tile.lo[2] - 1 + (int)blockIdx.x*(ZT_Z - 1);
   c.z = cz0 + k;
   const bool inX = c.x <= tile.hi[0];

   for (int r = k; r < ZT_ROWS; r += ZT_Z) {
      const int z = cz0 - 2 + r;
      if (inX && z >= tile.lo[2] - 3 && z <= tile.hi[2] + 3) {
         const RPt q = {c.x, c.y, wrapCoord(z, B.lo[2], B.hi[2])};
         stageCell<ZT_NC>(s, r*ZT_X + i, f, q);
         if (HYBRID) sSens[r*ZT_X + i] = ld<uint8_t>(f.shockSensorZ, q);
      }
   }
   const bool faceActive = inX && c.z <= tile.hi[2];
   const bool upd = faceActive && k >= 1;
   bool bad = false;
   double mz = 0.0, Tc = 0.0;
   if (faceActive) {
      const RPt q = {c.x, c.y, wrapCoord(c.z, B.lo[2], B.hi[2])};
      bad = ld<int>(f.nType_z, q) != Std_node;
      if (upd) { mz = ld<double>(f.dzet_e, c); if (CHEM) Tc = ld<double>(f.temperature, c); }
   }
   const bool anyBad = __syncthreads_or(bad);

   double F[nEq];
   bool smooth = true;
   if (!anyBad && faceActive) {
      if (HYBRID) smooth = sSens[(k + 1)*ZT_X + i] && sSens[(k + 2)*ZT_X + i] && sSens[(k + 3)*ZT_X + i];
      if (smooth) kgStdFlux<2, ZT_NC>(F, s, k*ZT_X + i, ZT_X);
   }
   double base[nEq];
   HTR_UNROLL
   for (int e = 0; e < nEq; e++) base[e] = 0.0;
   if (CHEM && upd) {
      const int own = (k + 2)*ZT_X + i;
      double Yi[nSpec], w[nSpec];
      HTR_UNROLL
      for (int e = 0; e < nSpec; e++) Yi[e] = s[(V_Y + e)*ZT_NC + own];
      mix::GetProductionRates(w, s[V_RHO*ZT_NC + own], s[V_P*ZT_NC + own], Tc, Yi);
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
         HTR_UNROLL
         for (int e = 0; e < nEq; e++) base[e] += mz*(F[e] - sF[e*ZT_NT + tid - ZT_X]);
      }
      mask[maskIndex(c, tile)] = deferred;
      if (deferred) *flag = 1;          // all writers store the same value
   }
   // Conserved_t rows leave through shared memory: a warp's 32 cells are nEq*32 consecutive doubles of the
   // field, written as full 256-byte spans instead of nEq stores of 8 bytes at a 8*nEq-byte stride
   if (f.Conserved_t.s[0] == 8*nEq) {
      __syncthreads();                                  // every warp has read its neighbour's flux from sF
      double *sOut = smem + k*(nEq*ZT_X);
      if (upd) {
         HTR_UNROLL
         for (int e = 0; e < nEq; e++) sOut[i*nEq + e] = base[e];
      }
      __syncwarp();
      if (k >= 1 && c.z <= tile.hi[2]) {                // warp-uniform: this row is updated
         RPt r0 = c; r0.x = c.x - i;
         const int left = tile.hi[0] - r0.x + 1;
         const int nval = nEq*(int)(left < ZT_X ? left : ZT_X);
         double *row = at<double>(f.Conserved_t, r0);
         HTR_UNROLL
         for (int e = 0; e < nEq; e++) { const int m = e*ZT_X + i; if (m < nval) row[m] = sOut[m]; }
      }
   } else if (upd) {
      double *Ct = at<double>(f.Conserved_t, c);
      HTR_UNROLL
      for (int e = 0; e < nEq; e++) Ct[e] = base[e];
   }
}

} // namespace fast
} // namespace htr

// =============================================================================================
// Diffusion flux, staged:   diffusion_staged_kernel<2,TJ>  Conserved_t += Dz
//                           diffusion_staged_kernel<1,TJ>  Conserved_t += Dy; += Dx
// (reference order z, y, x: src/prometeo.rg:560-578; arithmetic of src/prometeo_rhs.inl:1058-1215 with the
// Std_node rows of Interp/Grad as constant-bank operands).
//
// Every cell of the (x, J) tile (J = y or z) plus one more column/row on the plus side is staged once:
// its 2-point-stencil operands (rho, mu, lam, T, MixW(Xi), u, Xi, Di, rho*Yi) and the cell-centred
// velocity derivatives the faces interpolate (the reference recomputes them per face: 4 getDeriv
// triples per face, 12 per cell; here 7 resp. 4 scalars per staged cell).  A face is then ~60 shared
// loads and the per-face species enthalpies.  Face exchange / deferral as in the Euler kernels.
// =============================================================================================
#include "htr_diffusion.cuh"

namespace htr {
namespace fast {

template<int JD> struct DLay {
   static constexpr bool DOX = (JD == 1);
   static constexpr int NG = DOX ? 7 : 4;
   static constexpr int O_RHO = 0, O_MU = 1, O_LAM = 2, O_T = 3, O_MW = 4, O_U = 5, O_X = 8, O_DI = 8 + nSpec,
                        O_RY = 8 + 2*nSpec, O_G = 8 + 3*nSpec, NVD = O_G + NG;
};

HTR_DEV double derivStd(double xm1, double x, double xp1, double m) {
   return (c_Grad[Std_node][0]*(x - xm1) + c_Grad[Std_node][1]*(xp1 - x))*m;
}
HTR_DEV double interpStd(double x1, double x2) { return c_Interp[Std_node][0]*x1 + c_Interp[Std_node][1]*x2; }

// Stages cell q (already wrapped) into slot idx, one PART of its data per call so that the (cell, part) items of
// a tile spread evenly over the threads (a 33x9 box on 256 threads would otherwise leave 215 threads waiting at the
// barrier for the 41 that stage a second cell).  Returns true if a node type the fast path does not handle is
// involved.  Parts: 0 rho mu lam T Xi MixW | 1 Di rhoYi | 2 u, d/dx, d/dy | 3 d/dz (xy kernel only).
// JD = 1: derivatives G0..G6 = dUdX dVdX dUdY dVdY dUdZ dVdZ dWdZ;  JD = 2: dUdX dWdX dVdY dWdY.
template<int JD, int NC>
HTR_DEV bool stageDiffPart(const int part, double *s, const int idx, const rfluid &f, const RPt &q, const rbox &B) {
   using L = DLay<JD>;
   bool bad = false;
   if (part == 0) {
      s[L::O_RHO*NC + idx] = ld<double>(f.rho, q);
      s[L::O_MU*NC + idx] = ld<double>(f.mu, q);
      s[L::O_LAM*NC + idx] = ld<double>(f.lam, q);
      s[L::O_T*NC + idx] = ld<double>(f.temperature, q);
      double Xi[nSpec];
      HTR_UNROLL
      for (int i = 0; i < nSpec; i++) { Xi[i] = ldc<double>(f.MolarFracs, q, i); s[(L::O_X + i)*NC + idx] = Xi[i]; }
      s[L::O_MW*NC + idx] = mix::GetMolarWeightFromXi(Xi);
   } else if (part == 1) {
      HTR_UNROLL
      for (int i = 0; i < nSpec; i++) {
         s[(L::O_DI + i)*NC + idx] = ldc<double>(f.Di, q, i);
         s[(L::O_RY + i)*NC + idx] = ldc<double>(f.Conserved, q, i);
      }
   } else if (part == 2) {
      double u[3];
      HTR_UNROLL
      for (int i = 0; i < 3; i++) { u[i] = ldc<double>(f.velocity, q, i); s[(L::O_U + i)*NC + idx] = u[i]; }
      {  // d/dx
         RPt a = q, b = q;
         a.x = wrapCoord(q.x - 1, B.lo[0], B.hi[0]); b.x = wrapCoord(q.x + 1, B.lo[0], B.hi[0]);
         bad = bad || (ld<int>(f.nType_x, q) != Std_node);
         const double m = ld<double>(f.dcsi_d, q);
         constexpr int c1 = (JD == 1) ? 1 : 2;       // second component needed: V (xy kernel) or W (z kernel)
         s[(L::O_G + 0)*NC + idx] = derivStd(ldc<double>(f.velocity, a, 0), u[0], ldc<double>(f.velocity, b, 0), m);
         s[(L::O_G + 1)*NC + idx] = derivStd(ldc<double>(f.velocity, a, c1), u[c1], ldc<double>(f.velocity, b, c1), m);
      }
      {  // d/dy
         RPt a = q, b = q;
         a.y = wrapCoord(q.y - 1, B.lo[1], B.hi[1]); b.y = wrapCoord(q.y + 1, B.lo[1], B.hi[1]);
         bad = bad || (ld<int>(f.nType_y, q) != Std_node);
         const double m = ld<double>(f.deta_d, q);
         constexpr int c0 = (JD == 1) ? 0 : 1, c1 = (JD == 1) ? 1 : 2;   // (U,V) or (V,W)
         s[(L::O_G + 2)*NC + idx] = derivStd(ldc<double>(f.velocity, a, c0), u[c0], ldc<double>(f.velocity, b, c0), m);
         s[(L::O_G + 3)*NC + idx] = derivStd(ldc<double>(f.velocity, a, c1), u[c1], ldc<double>(f.velocity, b, c1), m);
      }
   } else if (JD == 1) {  // d/dz
      RPt a = q, b = q;
      a.z = wrapCoord(q.z - 1, B.lo[2], B.hi[2]); b.z = wrapCoord(q.z + 1, B.lo[2], B.hi[2]);
      bad = bad || (ld<int>(f.nType_z, q) != Std_node);
      const double m = ld<double>(f.dzet_d, q);
      HTR_UNROLL
      for (int i = 0; i < 3; i++)
         s[(L::O_G + 4 + i)*NC + idx] = derivStd(ldc<double>(f.velocity, a, i), ldc<double>(f.velocity, q, i), ldc<double>(f.velocity, b, i), m);
   }
   return bad;
}

// Diffusion flux of the face between staged cells ip and iq = ip + 1 along `dir`; gA..gD are the staged
// derivative slots in the order the stress formulas of `dir` use them:
//   dir 0: dVdY dUdY dUdZ dWdZ    dir 1: dUdX dVdX dVdZ dWdZ    dir 2: dUdX dWdX dVdY dWdY
template<int dir, int JD, int NC>
HTR_DEV void diffFaceStd(double *Flux, const double *s, const int ip, const int iq, const double m_s,
                         const int gA, const int gB, const int gC, const int gD) {
   using L = DLay<JD>;
   #define HTR_S(v, i) s[(v)*NC + (i)]
   const double rho_s = interpStd(HTR_S(L::O_RHO, ip), HTR_S(L::O_RHO, iq));
   const double iMixW_s = 1.0/interpStd(HTR_S(L::O_MW, ip), HTR_S(L::O_MW, iq));
   const double Tp = HTR_S(L::O_T, ip), Tq = HTR_S(L::O_T, iq);
   const double T_s = interpStd(Tp, Tq);
   double heatFlux = interpStd(HTR_S(L::O_LAM, ip), HTR_S(L::O_LAM, iq))*m_s*(Tp - Tq);

   double hi[nSpec];
   mix::GetSpeciesEnthalpies(hi, T_s);

   double ViCorr = 0.0;
   HTR_UNROLL
   for (int i = 0; i < nSpec; i++) {
      const double YiVi = interpStd(HTR_S(L::O_DI + i, ip), HTR_S(L::O_DI + i, iq))*
                          m_s*(HTR_S(L::O_X + i, ip) - HTR_S(L::O_X + i, iq))*mix::GetSpeciesMolarWeight(i)*iMixW_s;
      Flux[i] = -rho_s*YiVi;
      heatFlux += rho_s*YiVi*hi[i];
      ViCorr += YiVi;
   }
   HTR_UNROLL
   for (int i = 0; i < nSpec; i++) {
      const double rhoYiViCorr = interpStd(HTR_S(L::O_RY + i, ip), HTR_S(L::O_RY + i, iq))*ViCorr;
      Flux[i] += rhoYiViCorr;
      heatFlux -= rhoYiViCorr*hi[i];
   }

   double up[3], uq[3], dN[3];
   HTR_UNROLL
   for (int i = 0; i < 3; i++) { up[i] = HTR_S(L::O_U + i, ip); uq[i] = HTR_S(L::O_U + i, iq); dN[i] = m_s*(uq[i] - up[i]); }
   const double mu_s = interpStd(HTR_S(L::O_MU, ip), HTR_S(L::O_MU, iq));
   const double A = interpStd(HTR_S(L::O_G + gA, ip), HTR_S(L::O_G + gA, iq));
   const double Bv = interpStd(HTR_S(L::O_G + gB, ip), HTR_S(L::O_G + gB, iq));
   const double C = interpStd(HTR_S(L::O_G + gC, ip), HTR_S(L::O_G + gC, iq));
   const double D = interpStd(HTR_S(L::O_G + gD, ip), HTR_S(L::O_G + gD, iq));
   double sigma[3];
   if (dir == 0) {
      sigma[0] = mu_s*(4*dN[0] - 2*A - 2*D)/3;
      sigma[1] = mu_s*(dN[1] + Bv);
      sigma[2] = mu_s*(dN[2] + C);
   } else if (dir == 1) {
      sigma[0] = mu_s*(dN[0] + Bv);
      sigma[1] = mu_s*(4*dN[1] - 2*A - 2*D)/3;
      sigma[2] = mu_s*(dN[2] + C);
   } else {
      sigma[0] = mu_s*(dN[0] + Bv);
      sigma[1] = mu_s*(dN[1] + D);
      sigma[2] = mu_s*(4*dN[2] - 2*A - 2*C)/3;
   }
   HTR_UNROLL
   for (int i = 0; i < 3; i++) Flux[irU + i] = sigma[i];
   double uSigma = 0.0;
   HTR_UNROLL
   for (int i = 0; i < 3; i++) uSigma += interpStd(up[i], uq[i])*sigma[i];
   Flux[irE] = (uSigma - heatFlux);
   #undef HTR_S
}

template<int JD, int TJ> struct DTile {
   static constexpr bool DOX = (JD == 1);
   static constexpr int W = DOX ? 33 : 32, ROWS = TJ + 1, NC = W*ROWS, NT = 32*TJ;
   static constexpr size_t SMEM = sizeof(double)*DLay<JD>::NVD*NC + NT;
};

template<int JD, int TJ>
__global__ void __launch_bounds__(32*TJ, HTR_D_MINB)
diffusion_staged_kernel(const rfluid f, const rbox tile, const rbox B, uint8_t *__restrict__ mask, int *__restrict__ flag) {
   using T = DTile<JD, TJ>;
   using L = DLay<JD>;
   constexpr bool DOX = T::DOX;
   constexpr int W = T::W, NC = T::NC, NT = T::NT;
   constexpr int KD = (JD == 1) ? 2 : 1;              // the direction the grid's third index runs along
   extern __shared__ double smem[];
   double *s = smem;
   uint8_t *sFlag = reinterpret_cast<uint8_t*>(smem + L::NVD*NC);
   static_assert(L::NVD*NC >= nEq*NT, "flux exchange reuses the staging buffer");

   const int i = threadIdx.x, j = threadIdx.y;
   const int tid = j*32 + i;
   const int cx0 = DOX ? (tile.lo[0] - 1 + (int)blockIdx.x*31) : (tile.lo[0] + (int)blockIdx.x*32);
   const int cJ0 = tile.lo[JD] - 1 + (int)blockIdx.y*(TJ - 1);
   const int cK = tile.lo[KD] + blockIdx.z;
   RPt c;
   c.x = cx0 + i;
   (JD == 1 ? c.y : c.z) = cJ0 + j;
   (JD == 1 ? c.z : c.y) = cK;

   const int cJ = cJ0 + j;
   const bool inTile = c.x <= tile.hi[0] && cJ <= tile.hi[JD];
   // own Conserved_t and metrics requested before the staging loop (latency hidden behind it)
   const bool upd = inTile && j >= 1 && (!DOX || i >= 1);
   double *Ct = at<double>(f.Conserved_t, c);
   double ct[nEq], mJ = 0.0, mx = 0.0;
   if (HTR_D_HOIST && upd) {
      HTR_UNROLL
      for (int e = 0; e < nEq; e++) ct[e] = Ct[e];
      mJ = ld<double>(m_dOf<JD>(f), c);
      if (DOX) mx = ld<double>(f.dcsi_d, c);
   }
   bool bad = false;
   constexpr int NP = DOX ? 4 : 3;
   for (int item = tid; item < NP*NC; item += NT) {
      const int part = item/NC, n = item - part*NC;
      const int rr = n/W, cc = n - rr*W;
      const int x = cx0 + cc, J = cJ0 + rr;
      if (x <= tile.hi[0] + 1 && J <= tile.hi[JD] + 1) {
         RPt q;
         q.x = wrapCoord(x, B.lo[0], B.hi[0]);
         (JD == 1 ? q.y : q.z) = wrapCoord(J, B.lo[JD], B.hi[JD]);
         (JD == 1 ? q.z : q.y) = cK;
         bad = stageDiffPart<JD, NC>(part, s, n, f, q, B) || bad;
      }
   }

   const bool jAct = inTile && (!DOX || i >= 1), xAct = DOX && inTile && j >= 1;
   RPt q = c;                                         // own cell, wrapped (it may be lo-1)
   q.x = wrapCoord(c.x, B.lo[0], B.hi[0]);
   (JD == 1 ? q.y : q.z) = wrapCoord(cJ, B.lo[JD], B.hi[JD]);
   if (JD == 2 && jAct) bad = bad || (ld<int>(f.nType_z, q) != Std_node);
   const bool anyBad = __syncthreads_or(bad);

   const int own = j*W + i;
   double FJ[nEq], Fx[nEq];
   HTR_UNROLL
   for (int e = 0; e < nEq; e++) { FJ[e] = 0.0; Fx[e] = 0.0; }
   const bool ok = !anyBad;
   if (ok && jAct) {
      if (JD == 1) diffFaceStd<1, JD, NC>(FJ, s, own, own + W, ld<double>(f.deta_s, q), 0, 1, 5, 6);
      else         diffFaceStd<2, JD, NC>(FJ, s, own, own + W, ld<double>(f.dzet_s, q), 0, 1, 2, 3);
   }
   if (DOX && ok && xAct) diffFaceStd<0, JD, NC>(Fx, s, own, own + 1, ld<double>(f.dcsi_s, q), 3, 2, 4, 6);
   double Fm[nEq];
   if (DOX) {
      HTR_UNROLL
      for (int e = 0; e < nEq; e++) Fm[e] = __shfl_up_sync(0xffffffffu, Fx[e], 1);
   }
   __syncthreads();
   double *sF = smem;
   HTR_UNROLL
   for (int e = 0; e < nEq; e++) sF[e*NT + tid] = FJ[e];
   __syncthreads();
   if (upd) {
      if (!HTR_D_HOIST) {
         HTR_UNROLL
         for (int e = 0; e < nEq; e++) ct[e] = Ct[e];
         mJ = ld<double>(m_dOf<JD>(f), c);
         if (DOX) mx = ld<double>(f.dcsi_d, c);
      }
      if (ok) {
         HTR_UNROLL
         for (int e = 0; e < nEq; e++) {
            double v = ct[e];
            v += mJ*(FJ[e] - sF[e*NT + tid - 32]);
            if (DOX) v += mx*(Fx[e] - Fm[e]);
            ct[e] = v;
         }
      }
      mask[maskIndex(c, tile)] = !ok;
      if (!ok) *flag = 1;
   }
   // rows of Conserved_t leave through shared memory as contiguous spans (a deferred tile rewrites its old values)
   if (f.Conserved_t.s[0] == 8*nEq) {
      __syncthreads();
      double *sOut = smem + j*(nEq*32);
      constexpr int X0 = DOX ? 1 : 0, NXC = 32 - X0;      // first updated thread column, updated cells per row
      if (upd) {
         HTR_UNROLL
         for (int e = 0; e < nEq; e++) sOut[(i - X0)*nEq + e] = ct[e];
      }
      __syncwarp();
      if (j >= 1 && cJ <= tile.hi[JD]) {
         RPt r0 = c; r0.x = cx0 + X0;
         const int left = tile.hi[0] - r0.x + 1;
         const int nval = nEq*(int)(left < NXC ? left : NXC);
         double *row = at<double>(f.Conserved_t, r0);
         HTR_UNROLL
         for (int e = 0; e < nEq; e++) { const int m = e*32 + i; if (m < nval) row[m] = sOut[m]; }
      }
   } else if (upd && ok) {
      HTR_UNROLL
      for (int e = 0; e < nEq; e++) Ct[e] = ct[e];
   }
   (void)sFlag;
}

} // namespace fast
} // namespace htr
