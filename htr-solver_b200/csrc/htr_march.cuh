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
// z-marching form of UpdateDerivatives for the smooth / standard-node part of the grid: TWO passes over
// Conserved_t instead of the four of htr_fast.cuh, in the reference's accumulation order
// (src/prometeo.rg:502-578:  0 -> +chem -> +Ez -> +Ey -> +Ex -> +Dz -> +Dy -> +Dx):
//
//   euler_z_march_kernel   Conserved_t  = 0 + chemistry source + Euler-z      (write only)
//   plane_march_kernel     Conserved_t += Euler-y; += Euler-x; += Dz; += Dy; += Dx   (one read-modify-write)
//
// euler_z_march_kernel: one thread owns an (x, y) column and marches in z with the six stencil planes of
// its column in REGISTERS (rho, p, H, u, Y_i per plane; the ring is rotated by unrolling the march six
// times), so a cell is read from HBM once, there is no shared-memory staging, no halo and no barrier;
// the face flux of the previous step is the minus face of the next cell.
//
// plane_march_kernel: one CTA owns a 32 x 8 (x, y) tile (31 x 7 cells updated: thread (i, j) owns the PLUS
// x- and y-faces of its cell, the minus faces come from lane i-1 by shuffle and from row j-1 through shared
// memory) and marches in z over a chunk of planes.  Per plane:
//   E phase  rho, p, H = (rhoE+p)/rho, Y_i of the cross-shaped 6-point stencil box (37 x 13 cells) staged
//            once; velocity comes from a ring of three velocity planes (p-1, p, p+1) that also feeds the
//            velocity derivatives of the diffusion fluxes, so every velocity plane is fetched once per CTA;
//   D phase  the 2-point operands of the diffusion fluxes (rho, mu, lam, T, X_i, D_i, rho Y_i) on the 33 x 9
//            box and the seven cell-centred velocity derivatives the in-plane faces interpolate, staged into
//            the SAME shared memory (the E data is dead by then); the z face p+1/2 takes its upper operands
//            (plane p+1, own cell) straight from global memory and its lower ones from the staged plane, and
//            is the minus face of the next step.
// Addressing: all fields share one cell pitch (checked on the host: makeUniform), so a staged cell is ONE
// precomputed 32-bit offset per thread plus a per-plane uniform offset; the periodic wrap is applied once,
// when those offsets are built, not per load.
//
// Arithmetic: Std_node rows of the stencil tables; expression for expression the per-face kernels of
// htr_euler.cuh / htr_diffusion.cuh (and of htr_fast.cuh), so the strict build stays bit-identical with the
// per-task sequence.  Cells the fast arithmetic does not cover (non-standard node types on the tile, a
// non-smooth Hybrid face) are flagged in `mask` and receive the reference's sequence of += from the masked
// generic kernels afterwards, exactly as in htr_fast.cuh.
#pragma once
#include "htr_fast.cuh"

// resident CTAs per SM the register allocation aims at (B200, 512^3 AirMix: Euler-z march 15.4 ms at 2, 14.4 ms at 3)
#ifndef HTR_MZ_MINB
#define HTR_MZ_MINB 3
#endif
// unroll factor of the Euler-z march: the plane ring is rotated by register moves (110 per step, ~9 % of the kernel's instructions);
// unrolling lets the compiler rename instead, but the copies of the step (flux + chemistry source, ~22 KB of SASS each) spill
// 0.5-0.8 KB per thread at 168 registers: measured on B200 at 512^3 (profiles/r2_v27_*) 14.0 ms at 1, 16.3 / 16.9 / 17.2 ms at 2 / 3 / 6
#ifndef HTR_MZ_UNROLL
#define HTR_MZ_UNROLL 1
#endif
#define HTR_PRAGMA(x) _Pragma(#x)
#define HTR_UNROLL_N(n) HTR_PRAGMA(unroll n)
#ifndef HTR_MP_MINB
#define HTR_MP_MINB 2
#endif
#ifndef HTR_ME_MINB
#define HTR_ME_MINB 3
#endif

namespace htr {
namespace march {

using fast::derivStd;
using fast::interpStd;

// Uniform-layout view of the fields (host: makeUniform).  Every pointer addresses cell tile.lo; a cell is
// n = x + y*SY + z*SZ (tile-relative coordinates, ghosts negative) and field F lives at F[n*ncomp + c].
struct UF {
   const double *rho, *p, *T, *mu, *lam;
   const double *Y, *X, *Di, *vel, *C;
   double *Ct;
   const double *me[3], *md[3], *ms[3];
   const int *nT[3];
   const uint8_t *sens[3];
   int SY, SZ;
   int n[3];                 // tile extents: tile = [0, n-1]
   int blo[3], bhi[3];       // Fluid_bounds relative to tile.lo (clamped to +-2^30)
};

// periodic wrap of a stencil coordinate that overshoots the box by less than its size (host-checked)
HTR_DEV int wrap1(int c, const int lo, const int hi) {
   if (c > hi) c -= hi - lo + 1;
   else if (c < lo) c += hi - lo + 1;
   return c;
}

// KG flux of the face between points 2 and 3 of a six-point stencil (Std_node).  g(k, v) returns variable v of stencil
// point k: v = 0 rho, 1 p, 2 H, 3..5 velocity, 6.. mass fractions.  Same operations per equation, in the same order, as
// kgFaceFlux / fast::kgStdFlux (src/prometeo_rhs.inl:42-135: S = (rho_a+rho_b)*(-(un_a+un_b))*(phi_a+phi_b) summed over the
// pairs (2,3) | (2,4) (1,3) | (2,5) (1,4) (0,3), weighted by KennedyCoeff), but with the equations in the OUTER loop: the
// six pair coefficients are formed once and one transported variable is live at a time (a fifth of the registers).
template<int dir, class Get>
HTR_DEV void kgFluxStream(double *Flux, const Get &g) {
   double rv[6];
   {
      double rho[6], un[6];
      HTR_UNROLL
      for (int k = 0; k < 6; k++) { rho[k] = g(k, 0); un[k] = g(k, 3 + dir); }
      rv[0] = (rho[2] + rho[3])*(-(un[2] + un[3]));
      rv[1] = (rho[2] + rho[4])*(-(un[2] + un[4]));
      rv[2] = (rho[1] + rho[3])*(-(un[1] + un[3]));
      rv[3] = (rho[2] + rho[5])*(-(un[2] + un[5]));
      rv[4] = (rho[1] + rho[4])*(-(un[1] + un[4]));
      rv[5] = (rho[0] + rho[3])*(-(un[0] + un[3]));
   }
   const double k0 = c_KennedyCoeff[Std_node][0], k1 = c_KennedyCoeff[Std_node][1], k2 = c_KennedyCoeff[Std_node][2];
   auto transported = [&](const int v) -> double {
      double ph[6];
      HTR_UNROLL
      for (int k = 0; k < 6; k++) ph[k] = g(k, v);
      double fl = 0.0;
      { double acc = 0.0; acc += rv[0]*(ph[2] + ph[3]); fl += k0*acc; }
      { double acc = 0.0; acc += rv[1]*(ph[2] + ph[4]); acc += rv[2]*(ph[1] + ph[3]); fl += k1*acc; }
      { double acc = 0.0; acc += rv[3]*(ph[2] + ph[5]); acc += rv[4]*(ph[1] + ph[4]); acc += rv[5]*(ph[0] + ph[3]); fl += k2*acc; }
      return fl;
   };
   HTR_UNROLL
   for (int i = 0; i < nSpec; i++) Flux[i] = 0.25*transported(6 + i);
   HTR_UNROLL
   for (int i = 0; i < 3; i++) Flux[irU + i] = 0.25*transported(3 + i);
   Flux[irE] = 0.25*transported(2);
   {
      double pr[6];
      HTR_UNROLL
      for (int k = 0; k < 6; k++) pr[k] = g(k, 1);
      double fl = 0.0;
      { double acc = 0.0; acc += (pr[2] + pr[3]); fl += k0*acc; }
      { double acc = 0.0; acc += (pr[2] + pr[4]); acc += (pr[1] + pr[3]); fl += k1*acc; }
      { double acc = 0.0; acc += (pr[2] + pr[5]); acc += (pr[1] + pr[4]); acc += (pr[0] + pr[3]); fl += k2*acc; }
      Flux[irU + dir] -= fl;
   }
}

// =============================================================================================
// Conserved_t = 0 (+ chemistry source) + Euler-z        block = 32 (x) x MZ_TY (y) columns
// =============================================================================================
constexpr int MZ_TY = 4, MZ_NT = 32*MZ_TY;

struct ZRaw { double rho, p, rhoE, u[3], Y[nSpec]; };
struct ZPt { double v[6 + nSpec]; };              // rho, p, H, u, Y: the variable order of kgFluxStream

HTR_DEV void loadZRaw(ZRaw &r, const UF &f, const int n) {
   r.rho = __ldg(f.rho + n);
   r.p = __ldg(f.p + n);
   r.rhoE = __ldg(f.C + (int64_t)n*nEq + irE);
   HTR_UNROLL
   for (int c = 0; c < 3; c++) r.u[c] = __ldg(f.vel + (int64_t)n*3 + c);
   HTR_UNROLL
   for (int s = 0; s < nSpec; s++) r.Y[s] = __ldg(f.Y + (int64_t)n*nSpec + s);
}
HTR_DEV void toPoint(ZPt &c, const ZRaw &r) {
   c.v[0] = r.rho; c.v[1] = r.p;
   c.v[2] = HTR_QDIV(r.rhoE + r.p, r.rho);       // production: reciprocal + multiply (1 ulp), htr_mix.cuh
   HTR_UNROLL
   for (int i = 0; i < 3; i++) c.v[3 + i] = r.u[i];
   HTR_UNROLL
   for (int s = 0; s < nSpec; s++) c.v[6 + s] = r.Y[s];
}

template<bool HYBRID, bool CHEM>
__global__ void __launch_bounds__(MZ_NT, HTR_MZ_MINB)
euler_z_march_kernel(const UF f, uint8_t *__restrict__ mask, int *__restrict__ flag, const int kch) {
   __shared__ double sOut[MZ_TY][32*nEq];
   const int i = threadIdx.x, j = threadIdx.y;
   const int x0 = (int)blockIdx.x*32;
   const int x = x0 + i, y = (int)blockIdx.y*MZ_TY + j;
   const int z0 = (int)blockIdx.z*kch;
   const int z1 = (z0 + kch - 1 < f.n[2] - 1) ? z0 + kch - 1 : f.n[2] - 1;
   const bool rowOK = y < f.n[1];                  // warp-uniform
   const bool act = rowOK && x < f.n[0];
   const int nxy = act ? x + y*f.SY : 0;           // inactive lanes read cell (0,0): always valid
   const int left = f.n[0] - x0;
   const int nval = nEq*(left < 32 ? left : 32);
   #define HTR_ZOFF(z) (nxy + wrap1((z), f.blo[2], f.bhi[2])*f.SZ)       /* 32-bit cell index (host-checked range) */

   // c[k] = plane z - 2 + k of the cell being updated; rotated by register moves every step (unrolling the march six
   // times instead removes the moves but makes the loop six times the instruction cache)
   ZPt c[6];
   HTR_UNROLL
   for (int k = 0; k < 6; k++) { ZRaw r; loadZRaw(r, f, HTR_ZOFF(z0 - 3 + k)); toPoint(c[k], r); }
   auto get = [&](const int k, const int v) -> double { return c[k].v[v]; };
   bool s0 = true, s1 = true;                       // rolling sensor values of the two planes below the face's upper cell
   if (HYBRID) { s0 = __ldg(f.sens[2] + HTR_ZOFF(z0 - 2)) != 0; s1 = __ldg(f.sens[2] + HTR_ZOFF(z0 - 1)) != 0; }
   double Fm[nEq];
   bool haveM;
   {  // face z0 - 1/2: stencil planes z0-3 .. z0+2
      bool sm = true;
      if (HYBRID) { const bool s2 = __ldg(f.sens[2] + HTR_ZOFF(z0)) != 0; sm = s0 && s1 && s2; s0 = s1; s1 = s2; }
      haveM = sm && (__ldg(f.nT[2] + HTR_ZOFF(z0 - 1)) == Std_node);
      kgFluxStream<2>(Fm, get);
   }
   ZRaw nx;                                          // plane z+3 of the next step, requested one step ahead
   loadZRaw(nx, f, HTR_ZOFF(z0 + 3));

   HTR_UNROLL_N(HTR_MZ_UNROLL)
   for (int z = z0; z <= z1; z++) {
      HTR_UNROLL
      for (int k = 0; k < 5; k++) c[k] = c[k + 1];
      toPoint(c[5], nx);
      if (z < z1) loadZRaw(nx, f, HTR_ZOFF(z + 4));
      const int n = HTR_ZOFF(z);
      double Tc = 0.0;
      if (CHEM) Tc = __ldg(f.T + n);
      const double mz = __ldg(f.me[2] + n);
      bool sm = true;
      if (HYBRID) { const bool s2 = __ldg(f.sens[2] + HTR_ZOFF(z + 1)) != 0; sm = s0 && s1 && s2; s0 = s1; s1 = s2; }
      const bool haveP = sm && (__ldg(f.nT[2] + n) == Std_node);
      double Fp[nEq];
      kgFluxStream<2>(Fp, get);
      double base[nEq];
      HTR_UNROLL
      for (int e = 0; e < nEq; e++) base[e] = 0.0;
      if (CHEM) {
         double w[nSpec];
         mix::GetProductionRates(w, c[2].v[0], c[2].v[1], Tc, &c[2].v[6]);
         HTR_UNROLL
         for (int e = 0; e < nSpec; e++) base[e] += w[e];
      }
      const bool deferred = !(haveP && haveM);
      if (!deferred) {
         HTR_UNROLL
         for (int e = 0; e < nEq; e++) base[e] += mz*(Fp[e] - Fm[e]);
      }
      HTR_UNROLL
      for (int e = 0; e < nEq; e++) Fm[e] = Fp[e];
      haveM = haveP;
      if (act) {
         mask[((int64_t)z*f.n[1] + y)*f.n[0] + x] = deferred;
         if (deferred) *flag = 1;
      }
      // the row of 32 cells leaves as one contiguous span of 32*nEq doubles
      double *so = sOut[j];
      HTR_UNROLL
      for (int e = 0; e < nEq; e++) so[i*nEq + e] = base[e];
      __syncwarp();
      if (rowOK) {
         double *row = f.Ct + ((int64_t)x0 + (int64_t)y*f.SY + (int64_t)z*f.SZ)*nEq;
         HTR_UNROLL
         for (int e = 0; e < nEq; e++) { const int m = e*32 + i; if (m < nval) row[m] = so[m]; }
      }
      __syncwarp();
   }
   #undef HTR_ZOFF
}

// =============================================================================================
// Conserved_t += Euler-y; += Euler-x; += diffusion z; += y; += x      block = 32 x 8 threads
// =============================================================================================
constexpr int PW = 37, PR = 13, NCE = PW*PR;          // stencil box: columns -2..34, rows -2..10 of the tile
constexpr int DW = 33, DR = 9, NCD = DW*DR;           // diffusion box: columns 0..32, rows 0..8
constexpr int PNT = 256;
constexpr int NG = 7;                                  // staged derivatives: dUdX dVdX dUdY dVdY dUdZ dVdZ dWdZ
// shared-memory map (doubles).  R1 is time-shared: the E data of a plane is dead when its D data is staged.
constexpr int E_RHO = 0, E_P = NCE, E_H = 2*NCE, E_Y = 3*NCE, E_END = NCE*(3 + nSpec);
constexpr int D_RHO = 0, D_MU = NCD, D_LAM = 2*NCD, D_T = 3*NCD, D_X = 4*NCD, D_DI = D_X + NCD*nSpec,
              D_RY = D_DI + NCD*nSpec, D_G = D_RY + NCD*nSpec, D_END = D_G + NG*NCD;
template<bool DIFF> struct PLay {
   static constexpr int NV = DIFF ? 3 : 1;                // velocity planes resident (ring p-1, p, p+1 | plane p)
   static constexpr int OFF_R1 = NV*NCE*3;
   // with diffusion: output rows beyond the E data, inside the (dead) D data; without: over the (dead) E data
   static constexpr int OFF_SO = DIFF ? E_END : 0;
   static constexpr int OFF_SB = DIFF ? E_END + 8*32*nEq : E_END;          // sensor bytes of the stencil box (E phase only)
   static constexpr int OFF_SF = DIFF ? D_END : E_END + 2*512/8;            // [nEq][PNT] y-face exchange, aliases nothing
   static constexpr int R1_DBL = OFF_SF + nEq*PNT;
   static constexpr size_t SMEM = sizeof(double)*(OFF_R1 + R1_DBL) + 2*PNT;
   static_assert(!DIFF || OFF_SB + 2*512/8 <= D_END, "sensor bytes must fit inside the time-shared region");
   static_assert(DIFF || 8*32*nEq <= E_END, "output rows must fit over the E data");
};

// Diffusion flux of the face between cells P and Q = P + 1 along dir (Std_node).  P(v), Q(v) return operand v of the
// cell: 0 rho, 1 mu, 2 lam, 3 T, 4..6 velocity, 7.. X_i, then D_i, then rho Y_i, then the four transverse derivatives
//   dir 0: dVdY dUdY dUdZ dWdZ    dir 1: dUdX dVdX dVdZ dWdZ    dir 2: dUdX dWdX dVdY dWdY
// Expression for expression fast::diffFaceStd / diffusionFaceFlux (src/prometeo_rhs.inl:1058-1215).
constexpr int DV_X = 7, DV_DI = 7 + nSpec, DV_RY = 7 + 2*nSpec, DV_G = 7 + 3*nSpec, DV_N = DV_G + 4;
template<int dir, class GP, class GQ>
HTR_DEV void diffFaceStream(double *Flux, const GP &P, const GQ &Q, const double m_s) {
   const double rho_s = interpStd(P(0), Q(0));
   double XP[nSpec], XQ[nSpec];
   HTR_UNROLL
   for (int i = 0; i < nSpec; i++) { XP[i] = P(DV_X + i); XQ[i] = Q(DV_X + i); }
   const double iMixW_s = 1.0/interpStd(mix::GetMolarWeightFromXi(XP), mix::GetMolarWeightFromXi(XQ));
   const double Tp = P(3), Tq = Q(3);
   const double T_s = interpStd(Tp, Tq);
   double heatFlux = interpStd(P(2), Q(2))*m_s*(Tp - Tq);

   double hi[nSpec];
   mix::GetSpeciesEnthalpies(hi, T_s);

   double ViCorr = 0.0;
   HTR_UNROLL
   for (int i = 0; i < nSpec; i++) {
      const double YiVi = interpStd(P(DV_DI + i), Q(DV_DI + i))*m_s*(XP[i] - XQ[i])*mix::GetSpeciesMolarWeight(i)*iMixW_s;
      Flux[i] = -rho_s*YiVi;
      heatFlux += rho_s*YiVi*hi[i];
      ViCorr += YiVi;
   }
   HTR_UNROLL
   for (int i = 0; i < nSpec; i++) {
      const double rhoYiViCorr = interpStd(P(DV_RY + i), Q(DV_RY + i))*ViCorr;
      Flux[i] += rhoYiViCorr;
      heatFlux -= rhoYiViCorr*hi[i];
   }

   double up[3], uq[3], dN[3];
   HTR_UNROLL
   for (int i = 0; i < 3; i++) { up[i] = P(4 + i); uq[i] = Q(4 + i); dN[i] = m_s*(uq[i] - up[i]); }
   const double mu_s = interpStd(P(1), Q(1));
   const double A = interpStd(P(DV_G + 0), Q(DV_G + 0));
   const double Bv = interpStd(P(DV_G + 1), Q(DV_G + 1));
   const double Cv = interpStd(P(DV_G + 2), Q(DV_G + 2));
   const double Dv = interpStd(P(DV_G + 3), Q(DV_G + 3));
   double sigma[3];
   if (dir == 0) {
      sigma[0] = mu_s*(4*dN[0] - 2*A - 2*Dv)/3;
      sigma[1] = mu_s*(dN[1] + Bv);
      sigma[2] = mu_s*(dN[2] + Cv);
   } else if (dir == 1) {
      sigma[0] = mu_s*(dN[0] + Bv);
      sigma[1] = mu_s*(4*dN[1] - 2*A - 2*Dv)/3;
      sigma[2] = mu_s*(dN[2] + Cv);
   } else {
      sigma[0] = mu_s*(dN[0] + Bv);
      sigma[1] = mu_s*(dN[1] + Dv);
      sigma[2] = mu_s*(4*dN[2] - 2*A - 2*Cv)/3;
   }
   HTR_UNROLL
   for (int i = 0; i < 3; i++) Flux[irU + i] = sigma[i];
   double uSigma = 0.0;
   HTR_UNROLL
   for (int i = 0; i < 3; i++) uSigma += interpStd(up[i], uq[i])*sigma[i];
   Flux[irE] = (uSigma - heatFlux);
}

// operands of the own cell of one plane in registers (the z face's upper cell; both cells in the prologue)
struct DOps { double v[DV_N]; };
HTR_DEV void loadDOps(DOps &o, const UF &f, const int n) {
   o.v[0] = __ldg(f.rho + n); o.v[1] = __ldg(f.mu + n); o.v[2] = __ldg(f.lam + n); o.v[3] = __ldg(f.T + n);
   HTR_UNROLL
   for (int s = 0; s < nSpec; s++) {
      o.v[DV_X + s] = __ldg(f.X + (int64_t)n*nSpec + s);
      o.v[DV_DI + s] = __ldg(f.Di + (int64_t)n*nSpec + s);
      o.v[DV_RY + s] = __ldg(f.C + (int64_t)n*nEq + s);
   }
}

// x/y derivatives of the velocity at stencil-box cell vi of one staged velocity plane, as the z face uses them:
// g = dUdX dWdX dVdY dWdY
HTR_DEV void zFaceDerivs(double *g, const double *__restrict__ v, const int vi, const double mx, const double my) {
   g[0] = derivStd(v[(vi - 1)*3 + 0], v[vi*3 + 0], v[(vi + 1)*3 + 0], mx);
   g[1] = derivStd(v[(vi - 1)*3 + 2], v[vi*3 + 2], v[(vi + 1)*3 + 2], mx);
   g[2] = derivStd(v[(vi - PW)*3 + 1], v[vi*3 + 1], v[(vi + PW)*3 + 1], my);
   g[3] = derivStd(v[(vi - PW)*3 + 2], v[vi*3 + 2], v[(vi + PW)*3 + 2], my);
}

template<bool HYBRID, bool DIFF>
__global__ void __launch_bounds__(PNT, DIFF ? HTR_MP_MINB : HTR_ME_MINB)
plane_march_kernel(const UF f, uint8_t *__restrict__ mask, int *__restrict__ flag, const int kch) {
   using L = PLay<DIFF>;
   constexpr int OFF_SO = L::OFF_SO;
   extern __shared__ double smem[];
   double *sV = smem;
   double *R1 = smem + L::OFF_R1;
   double *sF = R1 + L::OFF_SF;
   uint8_t *sSx = reinterpret_cast<uint8_t*>(R1 + L::OFF_SB);
   uint8_t *sSy = sSx + 512;
   uint8_t *sHave = reinterpret_cast<uint8_t*>(R1 + L::R1_DBL);
   uint8_t *sDef = sHave + PNT;

   const int i = threadIdx.x, j = threadIdx.y, tid = j*32 + i;
   const int cx0 = -1 + (int)blockIdx.x*31, cy0 = -1 + (int)blockIdx.y*7;
   const int x = cx0 + i, y = cy0 + j;
   const int z0 = (int)blockIdx.z*kch;
   const int z1 = (z0 + kch - 1 < f.n[2] - 1) ? z0 + kch - 1 : f.n[2] - 1;
   const bool faceAct = x <= f.n[0] - 1 && y <= f.n[1] - 1;
   const bool upd = faceAct && i >= 1 && j >= 1;
   const bool rowUpd = j >= 1 && y <= f.n[1] - 1;                  // warp-uniform: this row has updated cells
   const int ownE = (j + 2)*PW + i + 2, ownD = j*DW + i;
   // own cell (may be the ghost/wrapped column or row -1); inactive threads address cell (0,0)
   const int nOwn = faceAct ? wrap1(x, f.blo[0], f.bhi[0]) + wrap1(y, f.blo[1], f.bhi[1])*f.SY : 0;

   // ---- per-thread staging items, fixed for the whole march -------------------------------------------
   // stencil box: cells tid and tid + 256
   int eOfs[2]; bool eV[2], eE[2];
   HTR_UNROLL
   for (int k = 0; k < 2; k++) {
      const int be = tid + PNT*k;
      const int row = be/PW - 2, col = be - (row + 2)*PW - 2;
      const int gx = cx0 + col, gy = cy0 + row;
      eV[k] = be < NCE && gx <= f.n[0] + 2 && gy <= f.n[1] + 2;      // the cell exists (ghost layer of 3)
      eE[k] = eV[k] && ((row >= 1 && row <= 7) || (col >= 1 && col <= 31));
      eOfs[k] = eV[k] ? wrap1(gx, f.blo[0], f.bhi[0]) + wrap1(gy, f.blo[1], f.bhi[1])*f.SY : 0;
   }
   // diffusion box: cell tid in full (A); one of six parts of cell 256 + tid/6 (B)
   int dIdx[2], dOfs[2], dVi[2]; bool dV[2];
   int partB = -1;
   HTR_UNROLL
   for (int k = 0; k < 2; k++) {
      const int bd = (k == 0) ? tid : PNT + tid/6;
      const int row = bd/DW, col = bd - row*DW;
      const int gx = cx0 + col, gy = cy0 + row;
      dV[k] = bd < NCD && gx <= f.n[0] && gy <= f.n[1] && (k == 0 || tid < 6*(NCD - PNT));
      dIdx[k] = bd < NCD ? bd : 0;
      dVi[k] = (row + 2)*PW + col + 2;
      dOfs[k] = dV[k] ? wrap1(gx, f.blo[0], f.bhi[0]) + wrap1(gy, f.blo[1], f.bhi[1])*f.SY : 0;
   }
   partB = dV[1] ? tid%6 : -1;
   #define HTR_ZP(z) (wrap1((z), f.blo[2], f.bhi[2])*f.SZ)                /* 32-bit cell index of the plane (host-checked range) */

   // velocity plane z -> ring slot
   auto stageV = [&](double *v, const int zo) {
      HTR_UNROLL
      for (int k = 0; k < 2; k++)
         if (eV[k]) {
            const int64_t n = (int64_t)(eOfs[k] + zo)*3;
            const int be = tid + PNT*k;
            HTR_UNROLL
            for (int c = 0; c < 3; c++) v[be*3 + c] = __ldg(f.vel + n + c);
         }
   };

   // ---- prologue ------------------------------------------------------------------------------------------
   // node types of the whole chunk (planes z0-1 .. z1+1 of the diffusion box): all loads are independent, one
   // barrier.  A chunk with any node the Std_node arithmetic does not cover is left to the generic kernels.
   {
      bool bad = false;
      for (int z = z0 - 1; z <= z1 + 1; z++) {
         const int zo = HTR_ZP(z);
         if (dV[0]) {
            const int n = dOfs[0] + zo;
            bad = bad || (__ldg(f.nT[0] + n) != Std_node) || (__ldg(f.nT[1] + n) != Std_node) || (__ldg(f.nT[2] + n) != Std_node);
         }
         if (partB == 0) {
            const int n = dOfs[1] + zo;
            bad = bad || (__ldg(f.nT[0] + n) != Std_node) || (__ldg(f.nT[1] + n) != Std_node) || (__ldg(f.nT[2] + n) != Std_node);
         }
      }
      if (DIFF) { stageV(sV, HTR_ZP(z0 - 1)); stageV(sV + NCE*3, HTR_ZP(z0)); }
      if (__syncthreads_or(bad)) {
         if (upd) {
            for (int p = z0; p <= z1; p++) mask[((int64_t)p*f.n[1] + y)*f.n[0] + x] = 1;
            *flag = 1;
         }
         return;
      }
   }
   // velocity planes z0-1, z0 are staged (ring slot of plane z: (z - z0 + 1) mod 3); diffusion z face z0-1/2
   double FzM[nEq], gC[4];                               // diffusion z flux of face p-1/2; z-face derivatives of plane p
   if (DIFF) {
      DOps P, Q;
      const int nm = nOwn + HTR_ZP(z0 - 1), nc = nOwn + HTR_ZP(z0);
      loadDOps(P, f, nm); loadDOps(Q, f, nc);
      HTR_UNROLL
      for (int c = 0; c < 3; c++) { P.v[4 + c] = sV[ownE*3 + c]; Q.v[4 + c] = sV[NCE*3 + ownE*3 + c]; }
      zFaceDerivs(&P.v[DV_G], sV, ownE, __ldg(f.md[0] + nm), __ldg(f.md[1] + nm));
      zFaceDerivs(gC, sV + NCE*3, ownE, __ldg(f.md[0] + nc), __ldg(f.md[1] + nc));
      HTR_UNROLL
      for (int k = 0; k < 4; k++) Q.v[DV_G + k] = gC[k];
      diffFaceStream<2>(FzM, [&](const int v) { return P.v[v]; }, [&](const int v) { return Q.v[v]; }, __ldg(f.ms[2] + nm));
   }

   #pragma unroll 1
   for (int p = z0; p <= z1; p++) {
      const int k3 = DIFF ? (p - z0)%3 : 0;
      const double *vm = sV + k3*(NCE*3);                           // plane p-1
      const double *vc = DIFF ? sV + ((k3 + 1)%3)*(NCE*3) : sV;     // plane p
      double *vn = DIFF ? sV + ((k3 + 2)%3)*(NCE*3) : sV;           // plane p+1
      const int zo = HTR_ZP(p), zn = HTR_ZP(p + 1);
      const int nO = nOwn + zo;

      // ---- (1) stage the E data of plane p, the sensors, the velocity of plane p+1 (DIFF) or p --------------
      HTR_UNROLL
      for (int k = 0; k < 2; k++) {
         const int be = tid + PNT*k;
         if (eE[k]) {
            const int n = eOfs[k] + zo;
            const double rho = __ldg(f.rho + n), pr = __ldg(f.p + n);
            R1[E_RHO + be] = rho;
            R1[E_P + be] = pr;
            R1[E_H + be] = (__ldg(f.C + (int64_t)n*nEq + irE) + pr)/rho;
            HTR_UNROLL
            for (int s = 0; s < nSpec; s++) R1[E_Y + be*nSpec + s] = __ldg(f.Y + (int64_t)n*nSpec + s);
            if (HYBRID) { sSx[be] = __ldg(f.sens[0] + n); sSy[be] = __ldg(f.sens[1] + n); }
         }
      }
      if (DIFF) stageV(vn, zn); else stageV(sV, zo);
      __syncthreads();                                                             // B1

      // ---- (2) Euler y face (published through sF), Euler x face (registers) ------------------------------------
      bool haveY = false, haveX = false;
      if (faceAct) {
         haveY = !HYBRID || (sSy[ownE - PW] && sSy[ownE] && sSy[ownE + PW]);
         haveX = !HYBRID || (sSx[ownE - 1] && sSx[ownE] && sSx[ownE + 1]);
      }
      {
         double Fy[nEq];
         if (haveY)
            kgFluxStream<1>(Fy, [&](const int k, const int v) -> double {
               const int idx = ownE + (k - 2)*PW;
               return v < 3 ? R1[v*NCE + idx] : (v < 6 ? vc[idx*3 + (v - 3)] : R1[E_Y + idx*nSpec + (v - 6)]); });
         else { HTR_UNROLL for (int e = 0; e < nEq; e++) Fy[e] = 0.0; }
         HTR_UNROLL
         for (int e = 0; e < nEq; e++) sF[e*PNT + tid] = Fy[e];
         sHave[tid] = haveY;
      }
      double Fx[nEq];
      if (haveX)
         kgFluxStream<0>(Fx, [&](const int k, const int v) -> double {
            const int idx = ownE + (k - 2);
            return v < 3 ? R1[v*NCE + idx] : (v < 6 ? vc[idx*3 + (v - 3)] : R1[E_Y + idx*nSpec + (v - 6)]); });
      else { HTR_UNROLL for (int e = 0; e < nEq; e++) Fx[e] = 0.0; }
      double ct[nEq];
      if (upd) {
         HTR_UNROLL
         for (int e = 0; e < nEq; e++) ct[e] = f.Ct[(int64_t)nO*nEq + e];
      }
      __syncthreads();                                                             // B2
      bool okE = false;
      {
         const bool haveXm = __shfl_up_sync(0xffffffffu, (int)haveX, 1) != 0;
         const double my = upd ? __ldg(f.me[1] + nO) : 0.0, mx = upd ? __ldg(f.me[0] + nO) : 0.0;
         if (upd) okE = haveX && haveXm && haveY && sHave[tid - 32];
         HTR_UNROLL
         for (int e = 0; e < nEq; e++) {
            const double FxM = __shfl_up_sync(0xffffffffu, Fx[e], 1);
            if (upd) {
               double v = ct[e];
               v += my*(sF[e*PNT + tid] - sF[e*PNT + tid - 32]);
               v += mx*(Fx[e] - FxM);
               ct[e] = v;
            }
         }
      }

      if (DIFF) {
         const int nN = nOwn + zn;
         // ---- (3) stage the diffusion operands of plane p and the velocity derivatives (the E data is dead) ----
         if (dV[0]) {
            const int idx = dIdx[0], vi = dVi[0];
            const int n = dOfs[0] + zo;
            R1[D_RHO + idx] = __ldg(f.rho + n); R1[D_MU + idx] = __ldg(f.mu + n);
            R1[D_LAM + idx] = __ldg(f.lam + n); R1[D_T + idx] = __ldg(f.T + n);
            HTR_UNROLL
            for (int s = 0; s < nSpec; s++) {
               R1[D_X + idx*nSpec + s] = __ldg(f.X + (int64_t)n*nSpec + s);
               R1[D_DI + idx*nSpec + s] = __ldg(f.Di + (int64_t)n*nSpec + s);
               R1[D_RY + idx*nSpec + s] = __ldg(f.C + (int64_t)n*nEq + s);
            }
            const double mx = __ldg(f.md[0] + n), my = __ldg(f.md[1] + n), mz = __ldg(f.md[2] + n);
            R1[D_G + 0*NCD + idx] = derivStd(vc[(vi - 1)*3 + 0], vc[vi*3 + 0], vc[(vi + 1)*3 + 0], mx);
            R1[D_G + 1*NCD + idx] = derivStd(vc[(vi - 1)*3 + 1], vc[vi*3 + 1], vc[(vi + 1)*3 + 1], mx);
            R1[D_G + 2*NCD + idx] = derivStd(vc[(vi - PW)*3 + 0], vc[vi*3 + 0], vc[(vi + PW)*3 + 0], my);
            R1[D_G + 3*NCD + idx] = derivStd(vc[(vi - PW)*3 + 1], vc[vi*3 + 1], vc[(vi + PW)*3 + 1], my);
            HTR_UNROLL
            for (int c = 0; c < 3; c++)
               R1[D_G + (4 + c)*NCD + idx] = derivStd(vm[vi*3 + c], vc[vi*3 + c], vn[vi*3 + c], mz);
         }
         if (partB >= 0) {
            const int idx = dIdx[1], vi = dVi[1];
            const int n = dOfs[1] + zo;
            if (partB == 0) {
               R1[D_RHO + idx] = __ldg(f.rho + n); R1[D_MU + idx] = __ldg(f.mu + n);
               R1[D_LAM + idx] = __ldg(f.lam + n); R1[D_T + idx] = __ldg(f.T + n);
            } else if (partB == 1) {
               HTR_UNROLL
               for (int s = 0; s < nSpec; s++) R1[D_X + idx*nSpec + s] = __ldg(f.X + (int64_t)n*nSpec + s);
            } else if (partB == 2) {
               HTR_UNROLL
               for (int s = 0; s < nSpec; s++) R1[D_DI + idx*nSpec + s] = __ldg(f.Di + (int64_t)n*nSpec + s);
            } else if (partB == 3) {
               HTR_UNROLL
               for (int s = 0; s < nSpec; s++) R1[D_RY + idx*nSpec + s] = __ldg(f.C + (int64_t)n*nEq + s);
            } else if (partB == 4) {
               const double mx = __ldg(f.md[0] + n), my = __ldg(f.md[1] + n);
               R1[D_G + 0*NCD + idx] = derivStd(vc[(vi - 1)*3 + 0], vc[vi*3 + 0], vc[(vi + 1)*3 + 0], mx);
               R1[D_G + 1*NCD + idx] = derivStd(vc[(vi - 1)*3 + 1], vc[vi*3 + 1], vc[(vi + 1)*3 + 1], mx);
               R1[D_G + 2*NCD + idx] = derivStd(vc[(vi - PW)*3 + 0], vc[vi*3 + 0], vc[(vi + PW)*3 + 0], my);
               R1[D_G + 3*NCD + idx] = derivStd(vc[(vi - PW)*3 + 1], vc[vi*3 + 1], vc[(vi + PW)*3 + 1], my);
            } else {
               const double mz = __ldg(f.md[2] + n);
               HTR_UNROLL
               for (int c = 0; c < 3; c++)
                  R1[D_G + (4 + c)*NCD + idx] = derivStd(vm[vi*3 + c], vc[vi*3 + c], vn[vi*3 + c], mz);
            }
         }
         // upper operands of the z face p+1/2 (own cell of plane p+1): requested before the barrier
         DOps Qz;
         double gN[4];
         if (faceAct) {
            loadDOps(Qz, f, nN);
            HTR_UNROLL
            for (int c = 0; c < 3; c++) Qz.v[4 + c] = vn[ownE*3 + c];
            zFaceDerivs(gN, vn, ownE, __ldg(f.md[0] + nN), __ldg(f.md[1] + nN));
            HTR_UNROLL
            for (int k = 0; k < 4; k++) Qz.v[DV_G + k] = gN[k];
         } else {
            HTR_UNROLL
            for (int k = 0; k < 4; k++) gN[k] = 0.0;
         }
         __syncthreads();                                                          // B4
         // ---- (4) diffusion faces.  z (p+1/2) is applied at once and becomes the next step's minus face; y is
         //          published through sF; x stays in registers until the barrier ----------------------------------
         // staged operand v of diffusion-box cell idx (velocity from the ring); the derivative slots depend on the face
         #define HTR_DGET(idx, vidx, gA, gB, gC_, gD) [&](const int v) -> double { \
            return v < 4 ? R1[v*NCD + (idx)] : (v < 7 ? vc[(vidx)*3 + (v - 4)] : (v < DV_G ? R1[D_X + (v - 7)/nSpec*(NCD*nSpec) + (idx)*nSpec + (v - 7)%nSpec] \
                   : R1[D_G + (v == DV_G ? (gA) : (v == DV_G + 1 ? (gB) : (v == DV_G + 2 ? (gC_) : (gD))))*NCD + (idx)])); }
         if (faceAct) {
            double Dz[nEq];
            diffFaceStream<2>(Dz, [&](const int v) -> double {
                                  return v < 4 ? R1[v*NCD + ownD] : (v < 7 ? vc[ownE*3 + (v - 4)]
                                       : (v < DV_G ? R1[D_X + (v - 7)/nSpec*(NCD*nSpec) + ownD*nSpec + (v - 7)%nSpec] : gC[v - DV_G])); },
                               [&](const int v) -> double { return Qz.v[v]; }, __ldg(f.ms[2] + nO));
            if (upd) {
               const double mz = __ldg(f.md[2] + nO);
               HTR_UNROLL
               for (int e = 0; e < nEq; e++) ct[e] += mz*(Dz[e] - FzM[e]);
            }
            HTR_UNROLL
            for (int e = 0; e < nEq; e++) FzM[e] = Dz[e];
         }
         HTR_UNROLL
         for (int k = 0; k < 4; k++) gC[k] = gN[k];
         {
            double Dy[nEq];
            if (faceAct) diffFaceStream<1>(Dy, HTR_DGET(ownD, ownE, 0, 1, 5, 6), HTR_DGET(ownD + DW, ownE + PW, 0, 1, 5, 6), __ldg(f.ms[1] + nO));
            else { HTR_UNROLL for (int e = 0; e < nEq; e++) Dy[e] = 0.0; }
            HTR_UNROLL
            for (int e = 0; e < nEq; e++) sF[e*PNT + tid] = Dy[e];
         }
         double Dx[nEq];
         if (faceAct) diffFaceStream<0>(Dx, HTR_DGET(ownD, ownE, 3, 2, 4, 6), HTR_DGET(ownD + 1, ownE + 1, 3, 2, 4, 6), __ldg(f.ms[0] + nO));
         else { HTR_UNROLL for (int e = 0; e < nEq; e++) Dx[e] = 0.0; }
         #undef HTR_DGET
         __syncthreads();                                                          // B5: sF visible, the D data has been read
         {
            const double my = upd ? __ldg(f.md[1] + nO) : 0.0, mx = upd ? __ldg(f.md[0] + nO) : 0.0;
            HTR_UNROLL
            for (int e = 0; e < nEq; e++) {
               const double DxM = __shfl_up_sync(0xffffffffu, Dx[e], 1);
               if (upd) {
                  double v = ct[e];
                  v += my*(sF[e*PNT + tid] - sF[e*PNT + tid - 32]);
                  v += mx*(Dx[e] - DxM);
                  ct[e] = v;
               }
            }
         }
      }

      // ---- (5) Conserved_t rows leave as contiguous spans; deferred cells keep their value -----------------
      const bool deferred = !okE;
      if (upd) {
         mask[((int64_t)p*f.n[1] + y)*f.n[0] + x] = deferred;
         if (deferred) *flag = 1;
      }
      {
         double *so = R1 + OFF_SO + j*(32*nEq);
         sDef[tid] = deferred;
         if (upd && !deferred) {
            HTR_UNROLL
            for (int e = 0; e < nEq; e++) so[(i - 1)*nEq + e] = ct[e];
         }
         const bool anyDef = __any_sync(0xffffffffu, upd && deferred);
         __syncwarp();
         if (rowUpd) {
            const int left = f.n[0] - (cx0 + 1);
            const int nval = nEq*(left < 31 ? left : 31);
            double *row = f.Ct + ((int64_t)(cx0 + 1) + (int64_t)y*f.SY + (int64_t)p*f.SZ)*nEq;
            if (!anyDef) {
               HTR_UNROLL
               for (int e = 0; e < nEq; e++) { const int m = e*32 + i; if (m < nval) row[m] = so[m]; }
            } else {
               HTR_UNROLL
               for (int e = 0; e < nEq; e++) { const int m = e*32 + i; if (m < nval && !sDef[j*32 + 1 + m/nEq]) row[m] = so[m]; }
            }
         }
      }
      if (!DIFF) __syncthreads();      // without the D phase: the E data / sF of this plane have been read before the next plane is staged
   }
   #undef HTR_ZP
}


// =============================================================================================
// Conserved_t += diffusion z; += y; += x in ONE pass, without marching: the (x, y) tile of one plane is staged exactly
// like fast::diffusion_staged_kernel<1> (same shared-memory layout, 3 CTAs per SM); the two z faces of a cell take the
// operands of the planes p-1 and p+1 (own column only: 2 x 27 values + the velocity neighbours of the derivatives)
// straight from global memory -- the CTAs of the neighbouring planes stage the same cells, so these are L2 hits, and
// diffusion z no longer needs a pass of its own over the state and over Conserved_t (881 -> ~520 B/cell of DRAM traffic).
// The price is one redundant face per cell (each z face is evaluated by both cells it separates).
// Accumulation order per cell: += Dz, += Dy, += Dx (src/prometeo.rg:560-578).
// =============================================================================================
// x/y derivatives of the velocity at cell q (wrapped) read from global memory, as the z face wants them: dUdX dWdX dVdY dWdY
HTR_DEV void zDerivsGlobal(double *g, const rfluid &f, const RPt &q, const rbox &B) {
   RPt a = q, b = q;
   a.x = fast::wrapCoord(q.x - 1, B.lo[0], B.hi[0]); b.x = fast::wrapCoord(q.x + 1, B.lo[0], B.hi[0]);
   const double mx = ld<double>(f.dcsi_d, q);
   g[0] = derivStd(ldc<double>(f.velocity, a, 0), ldc<double>(f.velocity, q, 0), ldc<double>(f.velocity, b, 0), mx);
   g[1] = derivStd(ldc<double>(f.velocity, a, 2), ldc<double>(f.velocity, q, 2), ldc<double>(f.velocity, b, 2), mx);
   a = q; b = q;
   a.y = fast::wrapCoord(q.y - 1, B.lo[1], B.hi[1]); b.y = fast::wrapCoord(q.y + 1, B.lo[1], B.hi[1]);
   const double my = ld<double>(f.deta_d, q);
   g[2] = derivStd(ldc<double>(f.velocity, a, 1), ldc<double>(f.velocity, q, 1), ldc<double>(f.velocity, b, 1), my);
   g[3] = derivStd(ldc<double>(f.velocity, a, 2), ldc<double>(f.velocity, q, 2), ldc<double>(f.velocity, b, 2), my);
}
// operand v (order of diffFaceStream) of cell q straight from global memory, loaded where it is used
#define HTR_GOPS(q, g) [&](const int v) -> double { \
   return v == 0 ? ld<double>(f.rho, q) : (v == 1 ? ld<double>(f.mu, q) : (v == 2 ? ld<double>(f.lam, q) : (v == 3 ? ld<double>(f.temperature, q) \
        : (v < 7 ? ldc<double>(f.velocity, q, v - 4) : (v < DV_DI ? ldc<double>(f.MolarFracs, q, v - DV_X) : (v < DV_RY ? ldc<double>(f.Di, q, v - DV_DI) \
        : (v < DV_G ? ldc<double>(f.Conserved, q, v - DV_RY) : (g)[v - DV_G]))))))); }

template<int TJ>
__global__ void __launch_bounds__(32*TJ, HTR_D_MINB)
diffusion3_staged_kernel(const rfluid f, const rbox tile, const rbox B, uint8_t *__restrict__ mask, int *__restrict__ flag) {
   using T = fast::DTile<1, TJ>;
   using L = fast::DLay<1>;
   constexpr int W = T::W, NC = T::NC, NT = T::NT;
   extern __shared__ double smem[];
   double *s = smem;
   static_assert(L::NVD*NC >= nEq*NT, "flux exchange reuses the staging buffer");

   const int i = threadIdx.x, j = threadIdx.y;
   const int tid = j*32 + i;
   const int cx0 = tile.lo[0] - 1 + (int)blockIdx.x*31;
   const int cy0 = tile.lo[1] - 1 + (int)blockIdx.y*(TJ - 1);
   RPt c;
   c.x = cx0 + i; c.y = cy0 + j; c.z = tile.lo[2] + blockIdx.z;
   const bool inTile = c.x <= tile.hi[0] && c.y <= tile.hi[1];
   const bool upd = inTile && j >= 1 && i >= 1;
   bool bad = false;
   for (int item = tid; item < 4*NC; item += NT) {
      const int part = item/NC, n = item - part*NC;
      const int rr = n/W, cc = n - rr*W;
      const int x = cx0 + cc, y = cy0 + rr;
      if (x <= tile.hi[0] + 1 && y <= tile.hi[1] + 1) {
         const RPt q = {fast::wrapCoord(x, B.lo[0], B.hi[0]), fast::wrapCoord(y, B.lo[1], B.hi[1]), c.z};
         bad = fast::stageDiffPart<1, NC>(part, s, n, f, q, B) || bad;
      }
   }
   // own cell in the planes below and above (updated cells lie inside the tile: no wrap in x, y)
   const RPt qm = {c.x, c.y, fast::wrapCoord(c.z - 1, B.lo[2], B.hi[2])}, qp = {c.x, c.y, fast::wrapCoord(c.z + 1, B.lo[2], B.hi[2])};
   if (upd) {
      // the z faces p-1/2 and p+1/2 are owned by the cells p-1 and p; the derivatives of p-1 and p+1 use their x / y node types
      bad = bad || (ld<int>(f.nType_z, qm) != Std_node) || (ld<int>(f.nType_z, c) != Std_node) ||
            (ld<int>(f.nType_x, qm) != Std_node) || (ld<int>(f.nType_y, qm) != Std_node) ||
            (ld<int>(f.nType_x, qp) != Std_node) || (ld<int>(f.nType_y, qp) != Std_node);
   }
   const bool jAct = inTile && i >= 1, xAct = inTile && j >= 1;
   RPt q = c;                                         // own cell, wrapped (it may be lo-1)
   q.x = fast::wrapCoord(c.x, B.lo[0], B.hi[0]);
   q.y = fast::wrapCoord(c.y, B.lo[1], B.hi[1]);
   const bool anyBad = __syncthreads_or(bad);
   const bool ok = !anyBad;
   const int own = j*W + i;
   // staged operand v of tile cell idx; the four derivative slots of a face come from `gexpr`
   #define HTR_SOPS(idx, gexpr) [&](const int v) -> double { \
      return v < 4 ? s[v*NC + (idx)] : (v < 7 ? s[(L::O_U + v - 4)*NC + (idx)] : (v < DV_DI ? s[(L::O_X + v - DV_X)*NC + (idx)] \
           : (v < DV_RY ? s[(L::O_DI + v - DV_DI)*NC + (idx)] : (v < DV_G ? s[(L::O_RY + v - DV_RY)*NC + (idx)] : (gexpr))))); }
   #define HTR_SG(idx, a, b, cc, d) s[(L::O_G + (v == DV_G ? (a) : (v == DV_G + 1 ? (b) : (v == DV_G + 2 ? (cc) : (d)))))*NC + (idx)]

   double *Ct = at<double>(f.Conserved_t, c);
   double ct[nEq], mz = 0.0, my = 0.0, mx = 0.0;
   if (upd) {
      HTR_UNROLL
      for (int e = 0; e < nEq; e++) ct[e] = Ct[e];
      mz = ld<double>(f.dzet_d, c); my = ld<double>(f.deta_d, c); mx = ld<double>(f.dcsi_d, c);
   }
   // ---- z faces (both, from the planes below and above), applied first: += Dz ---------------------------------------
   if (ok && upd) {
      // z-face derivatives of the own cell in plane p: dUdX and dVdY are staged, dWdX and dWdY come from the staged velocity
      double gC[4];
      gC[0] = s[(L::O_G + 0)*NC + own];
      gC[1] = derivStd(s[(L::O_U + 2)*NC + own - 1], s[(L::O_U + 2)*NC + own], s[(L::O_U + 2)*NC + own + 1], mx);
      gC[2] = s[(L::O_G + 3)*NC + own];
      gC[3] = derivStd(s[(L::O_U + 2)*NC + own - W], s[(L::O_U + 2)*NC + own], s[(L::O_U + 2)*NC + own + W], my);
      double FzM[nEq], FzP[nEq], g[4];
      zDerivsGlobal(g, f, qm, B);
      diffFaceStream<2>(FzM, HTR_GOPS(qm, g), HTR_SOPS(own, gC[v - DV_G]), ld<double>(f.dzet_s, qm));
      zDerivsGlobal(g, f, qp, B);
      diffFaceStream<2>(FzP, HTR_SOPS(own, gC[v - DV_G]), HTR_GOPS(qp, g), ld<double>(f.dzet_s, c));
      HTR_UNROLL
      for (int e = 0; e < nEq; e++) ct[e] += mz*(FzP[e] - FzM[e]);
   }
   // ---- y and x faces of the plane ----------------------------------------------------------------------------------------
   double FJ[nEq], Fx[nEq];
   HTR_UNROLL
   for (int e = 0; e < nEq; e++) { FJ[e] = 0.0; Fx[e] = 0.0; }
   if (ok && jAct) diffFaceStream<1>(FJ, HTR_SOPS(own, HTR_SG(own, 0, 1, 5, 6)), HTR_SOPS(own + W, HTR_SG(own + W, 0, 1, 5, 6)), ld<double>(f.deta_s, q));
   if (ok && xAct) diffFaceStream<0>(Fx, HTR_SOPS(own, HTR_SG(own, 3, 2, 4, 6)), HTR_SOPS(own + 1, HTR_SG(own + 1, 3, 2, 4, 6)), ld<double>(f.dcsi_s, q));
   #undef HTR_SOPS
   #undef HTR_SG
   double Fm[nEq];
   HTR_UNROLL
   for (int e = 0; e < nEq; e++) Fm[e] = __shfl_up_sync(0xffffffffu, Fx[e], 1);
   __syncthreads();
   double *sF = smem;
   HTR_UNROLL
   for (int e = 0; e < nEq; e++) sF[e*NT + tid] = FJ[e];
   __syncthreads();
   if (upd) {
      if (ok) {
         HTR_UNROLL
         for (int e = 0; e < nEq; e++) {
            double v = ct[e];
            v += my*(FJ[e] - sF[e*NT + tid - 32]);
            v += mx*(Fx[e] - Fm[e]);
            ct[e] = v;
         }
      }
      mask[fast::maskIndex(c, tile)] = !ok;
      if (!ok) *flag = 1;
   }
   // rows of Conserved_t leave through shared memory as contiguous spans (a deferred tile rewrites its old values)
   if (f.Conserved_t.s[0] == 8*nEq) {
      __syncthreads();
      double *sOut = smem + j*(nEq*32);
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
   } else if (upd && ok) {
      HTR_UNROLL
      for (int e = 0; e < nEq; e++) Ct[e] = ct[e];
   }
}
#undef HTR_GOPS

} // namespace march
} // namespace htr
