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
// Diffusion face flux at i+1/2 along `dir` (viscous stress, Fourier + species-enthalpy heat
// flux, Fickian species flux with correction velocity), 2nd-order staggered.
// Arithmetic of src/prometeo_rhs.inl:1058-1215 (GetSigma, GetDiffusionFlux); each face is
// evaluated independently: the two transverse velocity gradients the reference carries
// from face to face (vGrad[0] <- vGrad[1]) are simply recomputed at p and p+1.
#pragma once
#include "htr_access.cuh"
#include "htr_mix.cuh"
#include "htr_stencil.cuh"

namespace htr {

// d(velocity)/d(xi_D) at cell q, all three components
template<int D>
HTR_DEV void velocityDeriv(double *g, const htr_fluid &f, const Pt &q, const htr_rect &B) {
   const int t = ld<int>(nTypeOf<D>(f), q);
   const double m = ld<double>(m_dOf<D>(f), q);
   const Pt qm = shift<D>(q, offM1(t), B);
   const Pt qp = shift<D>(q, offP1(t), B);
   HTR_UNROLL
   for (int i = 0; i < 3; i++)
      g[i] = getDeriv(t, ldc<double>(f.velocity, qm, i), ldc<double>(f.velocity, q, i), ldc<double>(f.velocity, qp, i), m);
}

// spanFirst: this is the plus face of the first cell of the task's span.  The reference carries the transverse
// gradients from face to face (vGrad[0] <- vGrad[1], src/prometeo_rhs.inl:1211-1212) assuming the span's first
// minus face lies one cell below; on a left-boundary node (offM1 = 0: L_S, L_C) that face is the cell's own plus
// face, so the first plus face of the span is evaluated with the gradients of p+1 on both sides
// (src/prometeo_rhs.inl:1249-1268).  Reproduced here so that boundary cells match the reference.
template<int dir>
HTR_DEV void diffusionFaceFlux(double *Flux, const htr_fluid &f, const Pt &p, const htr_rect &B, const bool spanFirst = false) {
   constexpr int D1 = T1(dir), D2 = T2(dir);
   const int t = ld<int>(nTypeOf<dir>(f), p);
   const double m_s = ld<double>(m_sOf<dir>(f), p);
   const Pt pp1 = shift<dir>(p, 1, B);

   double g1[2][3], g2[2][3];     // transverse gradients at p and p+1
   velocityDeriv<D1>(g1[0], f, p, B);   velocityDeriv<D2>(g2[0], f, p, B);
   velocityDeriv<D1>(g1[1], f, pp1, B); velocityDeriv<D2>(g2[1], f, pp1, B);
   if (spanFirst && offM1(t) == 0) {
      HTR_UNROLL
      for (int i = 0; i < 3; i++) { g1[0][i] = g1[1][i]; g2[0][i] = g2[1][i]; }
   }

   double Xp[nSpec], Xq[nSpec];
   HTR_UNROLL
   for (int i = 0; i < nSpec; i++) { Xp[i] = ldc<double>(f.MolarFracs, p, i); Xq[i] = ldc<double>(f.MolarFracs, pp1, i); }

   const double rho_s = Interp2Staggered(t, ld<double>(f.rho, p), ld<double>(f.rho, pp1));
   const double iMixW_s = 1.0/Interp2Staggered(t, mix::GetMolarWeightFromXi(Xp), mix::GetMolarWeightFromXi(Xq));
   const double Tp = ld<double>(f.temperature, p), Tq = ld<double>(f.temperature, pp1);
   const double T_s = Interp2Staggered(t, Tp, Tq);
   double heatFlux = Interp2Staggered(t, ld<double>(f.lam, p), ld<double>(f.lam, pp1))*m_s*(Tp - Tq);

   double hi[nSpec];
   mix::GetSpeciesEnthalpies(hi, T_s);

   double ViCorr = 0.0;
   HTR_UNROLL
   for (int i = 0; i < nSpec; i++) {
      const double YiVi = Interp2Staggered(t, ldc<double>(f.Di, p, i), ldc<double>(f.Di, pp1, i))*
                          m_s*(Xp[i] - Xq[i])*mix::GetSpeciesMolarWeight(i)*iMixW_s;
      Flux[i] = -rho_s*YiVi;
      heatFlux += rho_s*YiVi*hi[i];
      ViCorr += YiVi;
   }
   HTR_UNROLL
   for (int i = 0; i < nSpec; i++) {
      const double rhoYiViCorr = Interp2Staggered(t, ldc<double>(f.Conserved, p, i), ldc<double>(f.Conserved, pp1, i))*ViCorr;
      Flux[i] += rhoYiViCorr;
      heatFlux -= rhoYiViCorr*hi[i];
   }

   // viscous stress on the face
   double up[3], uq[3];
   HTR_UNROLL
   for (int i = 0; i < 3; i++) { up[i] = ldc<double>(f.velocity, p, i); uq[i] = ldc<double>(f.velocity, pp1, i); }
   const double mu_s = Interp2Staggered(t, ld<double>(f.mu, p), ld<double>(f.mu, pp1));
   double dN[3];                       // d(u,v,w)/d(normal) on the face
   HTR_UNROLL
   for (int i = 0; i < 3; i++) dN[i] = m_s*(uq[i] - up[i]);
   // transverse derivatives interpolated to the face: d u_a / d x_b
   #define HTR_TG(g, a) Interp2Staggered(t, g[0][a], g[1][a])
   double sigma[3];
   if (dir == 0) {
      const double dVdY = HTR_TG(g1, 1), dUdY = HTR_TG(g1, 0), dUdZ = HTR_TG(g2, 0), dWdZ = HTR_TG(g2, 2);
      sigma[0] = mu_s*(4*dN[0] - 2*dVdY - 2*dWdZ)/3;
      sigma[1] = mu_s*(dN[1] + dUdY);
      sigma[2] = mu_s*(dN[2] + dUdZ);
   } else if (dir == 1) {
      const double dUdX = HTR_TG(g1, 0), dVdX = HTR_TG(g1, 1), dVdZ = HTR_TG(g2, 1), dWdZ = HTR_TG(g2, 2);
      sigma[0] = mu_s*(dN[0] + dVdX);
      sigma[1] = mu_s*(4*dN[1] - 2*dUdX - 2*dWdZ)/3;
      sigma[2] = mu_s*(dN[2] + dVdZ);
   } else {
      const double dUdX = HTR_TG(g1, 0), dWdX = HTR_TG(g1, 2), dVdY = HTR_TG(g2, 1), dWdY = HTR_TG(g2, 2);
      sigma[0] = mu_s*(dN[0] + dWdX);
      sigma[1] = mu_s*(dN[1] + dWdY);
      sigma[2] = mu_s*(4*dN[2] - 2*dUdX - 2*dVdY)/3;
   }
   #undef HTR_TG
   HTR_UNROLL
   for (int i = 0; i < 3; i++) Flux[irU + i] = sigma[i];
   double uSigma = 0.0;
   HTR_UNROLL
   for (int i = 0; i < 3; i++) uSigma += Interp2Staggered(t, up[i], uq[i])*sigma[i];
   Flux[irE] = (uSigma - heatFlux);
}

} // namespace htr
