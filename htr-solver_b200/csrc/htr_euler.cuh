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
// Euler face fluxes at i+1/2 along `dir`:
//   kgFaceFlux    -- 6th-order Kennedy-Gruber / Pirozzoli skew-symmetric split form
//   tenoFaceFlux  -- Roe-linearised characteristic-wise Lax-Friedrichs split TENO6(-A/-LAD)
//                    with the multispecies first-order positivity blend
// One thread evaluates one face directly from the 6-point stencil (no rolling buffer), so every
// face of the grid is independent work.  The arithmetic of each face is the reference's:
//   KG sums/weights   src/prometeo_rhs.inl:42-135   (pair (p-m, p-m+l+1) <-> slot (l, m))
//   Roe state         src/prometeo_rhs.inl:200-279
//   projections       src/prometeo_rhs.inl:436-572
//   split + recon.    src/prometeo_rhs.inl:586-759
// Fluxes carry the reference's sign convention (already negated: they sit on the RHS).
#pragma once
#include "htr_access.cuh"
#include "htr_mix.cuh"
#include "htr_stencil.cuh"

namespace htr {

// ---------------------------------------------------------------------------------------------
// Kennedy-Gruber
// ---------------------------------------------------------------------------------------------
struct KGPoint {            // what the split form needs from one cell
   double rho, un, p, H;
   double Y[nSpec];
   double u[3];
};

template<int dir>
HTR_DEV void loadKG(KGPoint &c, const htr_fluid &f, const Pt &p) {
   c.rho = ld<double>(f.rho, p);
   c.p = ld<double>(f.pressure, p);
   HTR_UNROLL
   for (int i = 0; i < 3; i++) c.u[i] = ldc<double>(f.velocity, p, i);
   c.un = c.u[dir];
   HTR_UNROLL
   for (int i = 0; i < nSpec; i++) c.Y[i] = ldc<double>(f.MassFracs, p, i);
   c.H = (ldc<double>(f.Conserved, p, irE) + c.p)/c.rho;
}

// product sums of the pair (a, b); S[nEq] is the pressure sum
HTR_DEV void kgPair(double *S, const KGPoint &a, const KGPoint &b) {
   const double rhom = a.rho + b.rho;
   const double vm = -(a.un + b.un);
   HTR_UNROLL
   for (int i = 0; i < nSpec; i++) S[i] = rhom*vm*(a.Y[i] + b.Y[i]);
   HTR_UNROLL
   for (int i = 0; i < 3; i++) S[irU + i] = rhom*vm*(a.u[i] + b.u[i]);
   S[irE] = rhom*vm*(a.H + b.H);
   S[nEq] = (a.p + b.p);
}

template<int dir>
HTR_DEV void kgFaceFlux(double *Flux, const htr_fluid &f, const Pt &p, int t, const htr_rect &B) {
   if (t == L_S_node || t == Rm1_S_node) {
      // staggered boundary face: plain flux of the wall-adjacent state
      const Pt q = (t == L_S_node) ? p : shift<dir>(p, offP1(t), B);
      const double un = ldc<double>(f.velocity, q, dir);
      const double pr = ld<double>(f.pressure, q);
      HTR_UNROLL
      for (int i = 0; i < nEq; i++) Flux[i] = -(ldc<double>(f.Conserved, q, i)*un);
      Flux[irU + dir] -= pr;
      Flux[irE] -= pr*un;
      return;
   }
   const int order = c_KennedyOrder[t];
   // cells p-2 .. p+3 (plain offsets, as the reference's sums use)
   KGPoint c[6];
   HTR_UNROLL
   for (int k = 0; k < 6; k++)
      if (k >= 3 - order && k <= 2 + order) loadKG<dir>(c[k], f, shift<dir>(p, k - 2, B));
   double fl[nEq + 1];
   HTR_UNROLL
   for (int i = 0; i < nEq + 1; i++) fl[i] = 0.0;
   HTR_UNROLL
   for (int l = 0; l < 3; l++) {
      if (l < order) {
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
         const double cl = c_KennedyCoeff[t][l];
         HTR_UNROLL
         for (int i = 0; i < nEq + 1; i++) fl[i] += cl*acc[i];
      }
   }
   HTR_UNROLL
   for (int i = 0; i < nEq; i++) Flux[i] = 0.25*fl[i];
   Flux[irU + dir] -= fl[nEq];
}

// ---------------------------------------------------------------------------------------------
// Characteristic TENO
// ---------------------------------------------------------------------------------------------
struct RoeState {
   double rho, H, e, dpde, a2, a;
   double Yi[nSpec], dpdrhoi[nSpec];
   double u[3];
};

struct CellState {          // primitive + conserved data of one stencil cell
   double U[nEq];
   double un, p;
};

template<int dir>
HTR_DEV void loadCell(CellState &c, const htr_fluid &f, const Pt &p) {
   HTR_UNROLL
   for (int i = 0; i < nEq; i++) c.U[i] = ldc<double>(f.Conserved, p, i);
   c.un = ldc<double>(f.velocity, p, dir);
   c.p = ld<double>(f.pressure, p);
}

// Roe average of the L|R states with the Shuen-Liou-van Leer pressure-jump consistent derivatives
HTR_DEV void roeAverage(RoeState &R, const double *UL, const double *UR, const double *YL, const double *YR,
                        double pL, double pR, const double *uL, const double *uR, double rhoL, double rhoR) {
   const double eL = UL[irE]/rhoL - 0.5*mod2(uL);
   const double eR = UR[irE]/rhoR - 0.5*mod2(uR);
   const double HL = (UL[irE] + pL)/rhoL;
   const double HR = (UR[irE] + pR)/rhoR;
   const double fL = sqrt(rhoL)/(sqrt(rhoL) + sqrt(rhoR));
   const double fR = sqrt(rhoR)/(sqrt(rhoL) + sqrt(rhoR));
   R.rho = sqrt(rhoL*rhoR);
   HTR_UNROLL
   for (int i = 0; i < nSpec; i++) R.Yi[i] = YL[i]*fL + YR[i]*fR;
   HTR_UNROLL
   for (int i = 0; i < 3; i++) R.u[i] = uL[i]*fL + uR[i]*fR;
   R.H = HL*fL + HR*fR;
   R.e = eL*fL + eR*fR;

   const double PovRho = R.H - R.e - 0.5*mod2(R.u);
   const double MixW = mix::GetMolarWeightFromYi(R.Yi);
   const double Teos = mix::GetTFromRhoAndP(R.rho, MixW, PovRho*R.rho);
   const double T = mix::GetTFromInternalEnergy(R.e, Teos, R.Yi);
   const double gamma = mix::GetGamma(T, MixW, R.Yi);
   R.dpde = mix::Getdpde(R.rho, gamma);
   mix::Getdpdrhoi(R.dpdrhoi, gamma, T, R.Yi);

   const double dp = pR - pL;
   const double de = eR - eL;
   double drhoi[nSpec];
   HTR_UNROLL
   for (int i = 0; i < nSpec; i++) drhoi[i] = rhoR*YR[i] - rhoL*YL[i];
   double dpError = dp - de*R.dpde;
   double fact = de*R.dpde; fact *= fact;
   HTR_UNROLL
   for (int i = 0; i < nSpec; i++) {
      const double dpi = drhoi[i]*R.dpdrhoi[i];
      dpError -= dpi;
      fact += dpi*dpi;
   }
   fact = dpError/fmax(fact, 1e-6);
   HTR_UNROLL
   for (int i = 0; i < nSpec; i++) R.dpdrhoi[i] = R.dpdrhoi[i]*(1.0 + R.dpdrhoi[i]*drhoi[i]*fact);
   R.dpde = R.dpde*(1.0 + R.dpde*de*fact);

   R.a2 = PovRho/R.rho*R.dpde;
   HTR_UNROLL
   for (int i = 0; i < nSpec; i++) R.a2 += R.Yi[i]*R.dpdrhoi[i];
   R.a = sqrt(R.a2);
}

// r = L q : physical -> characteristic ([u-a, species 1..N, shear 1, shear 2, u+a])
template<int dir>
HTR_DEV void toCharacteristic(double *r, const double *q, const RoeState &R) {
   constexpr int iN = dir, iT1 = T1(dir), iT2 = T2(dir);
   const double iaRoe = 1.0/R.a;
   const double iaRoe2 = 1.0/R.a2;
   const double Coeff = (R.e - 0.5*mod2(R.u))*R.dpde/R.rho;
   double sum1 = 0.0;
   HTR_UNROLL
   for (int i = 0; i < nSpec; i++) sum1 += (Coeff - R.dpdrhoi[i])*iaRoe2*q[i];
   double sumQsp = 0.0;
   HTR_UNROLL
   for (int i = 0; i < nSpec; i++) sumQsp += q[i];
   const double d = R.dpde/(R.rho*R.a2);
   const double c[3] = {R.u[0]*d, R.u[1]*d, R.u[2]*d};

   r[0] = sumQsp*R.u[iN]*iaRoe - sum1;
   r[0] -= (c[iN] + iaRoe)*q[nSpec + iN];
   r[0] -= c[iT1]*q[nSpec + iT1];
   r[0] -= c[iT2]*q[nSpec + iT2];
   r[0] += d*q[nSpec + 3];
   r[0] *= 0.5;
   HTR_UNROLL
   for (int i = 1; i < nSpec + 1; i++) {
      r[i] = R.Yi[i - 1]*sum1;
      r[i] += q[i - 1];
      HTR_UNROLL
      for (int j = 0; j < 3; j++) r[i] += R.Yi[i - 1]*c[j]*q[nSpec + j];
      r[i] += -R.Yi[i - 1]*d*q[nSpec + 3];
   }
   r[nSpec + 1] = -R.u[iT1]*sumQsp;
   r[nSpec + 1] += q[nSpec + iT1];
   r[nSpec + 2] = -R.u[iT2]*sumQsp;
   r[nSpec + 2] += q[nSpec + iT2];
   r[nSpec + 3] = -sumQsp*R.u[iN]*iaRoe - sum1;
   r[nSpec + 3] -= (c[iN] - iaRoe)*q[nSpec + iN];
   r[nSpec + 3] -= c[iT1]*q[nSpec + iT1];
   r[nSpec + 3] -= c[iT2]*q[nSpec + iT2];
   r[nSpec + 3] += d*q[nSpec + 3];
   r[nSpec + 3] *= 0.5;
}

// r = K q : characteristic -> physical
template<int dir>
HTR_DEV void fromCharacteristic(double *r, const double *q, const RoeState &R) {
   constexpr int iN = dir, iT1 = T1(dir), iT2 = T2(dir);
   HTR_UNROLL
   for (int i = 0; i < nEq; i++) r[i] = 0.0;
   double sumQsp = 0.0;
   HTR_UNROLL
   for (int i = 0; i < nSpec; i++) sumQsp += q[i + 1];
   HTR_UNROLL
   for (int i = 0; i < nSpec; i++) {
      r[i] += R.Yi[i]*q[0];
      r[i] += q[i + 1];
      r[i] += R.Yi[i]*q[nSpec + 3];
   }
   r[nSpec + iN] += (R.u[iN] - R.a)*q[0];
   r[nSpec + iN] += R.u[iN]*sumQsp;
   r[nSpec + iN] += (R.u[iN] + R.a)*q[nSpec + 3];

   r[nSpec + iT1] += R.u[iT1]*q[0];
   r[nSpec + iT1] += R.u[iT1]*sumQsp;
   r[nSpec + iT1] += q[nSpec + 1];
   r[nSpec + iT1] += R.u[iT1]*q[nSpec + 3];

   r[nSpec + iT2] += R.u[iT2]*q[0];
   r[nSpec + iT2] += R.u[iT2]*sumQsp;
   r[nSpec + iT2] += q[nSpec + 2];
   r[nSpec + iT2] += R.u[iT2]*q[nSpec + 3];

   r[nSpec + 3] += (R.H - R.u[iN]*R.a)*q[0];
   const double dedp = 1.0/R.dpde;
   const double ERoe = R.e + 0.5*mod2(R.u);
   HTR_UNROLL
   for (int i = 0; i < nSpec; i++) r[nSpec + 3] += (ERoe - R.rho*R.dpdrhoi[i]*dedp)*q[i + 1];
   r[nSpec + 3] += R.u[iT1]*q[nSpec + 1];
   r[nSpec + 3] += R.u[iT2]*q[nSpec + 2];
   r[nSpec + 3] += (R.H + R.u[iN]*R.a)*q[nSpec + 3];
}

// Lax-Friedrichs split characteristic fluxes of one stencil cell
template<int dir>
HTR_DEV void splitFlux(double *Fp, double *Fm, const RoeState &R, const CellState &c,
                       double Lam1, double Lam, double LamN) {
   double Fl[nEq];
   HTR_UNROLL
   for (int i = 0; i < nEq; i++) Fl[i] = c.U[i]*c.un;
   Fl[irU + dir] += c.p;
   Fl[irE] += c.p*c.un;
   double Q[nEq], F[nEq];
   toCharacteristic<dir>(Q, c.U, R);
   toCharacteristic<dir>(F, Fl, R);
   Fp[0] = F[0] + Lam1*Q[0];
   Fm[0] = F[0] - Lam1*Q[0];
   HTR_UNROLL
   for (int i = 1; i < nEq - 1; i++) { Fp[i] = F[i] + Lam*Q[i]; Fm[i] = F[i] - Lam*Q[i]; }
   Fp[nEq - 1] = F[nEq - 1] + LamN*Q[nEq - 1];
   Fm[nEq - 1] = F[nEq - 1] - LamN*Q[nEq - 1];
}

template<int dir, int OP>
HTR_DEV void tenoFaceFlux(double *Flux, const htr_fluid &f, const Pt &p, int t, const htr_rect &B,
                          double RK0, double RK1, double lim_f) {
   Pt s[6];
   s[0] = shift<dir>(p, offM2(t), B);
   s[1] = shift<dir>(p, offM1(t), B);
   s[2] = p;
   s[3] = shift<dir>(p, offP1(t), B);
   s[4] = shift<dir>(p, offP2(t), B);
   s[5] = shift<dir>(p, offP3(t), B);

   CellState c[6];
   double Lam1 = 0.0, Lam = 0.0, LamN = 0.0;   // the reference folds max() over |.| values, all >= 0
   HTR_UNROLL
   for (int k = 0; k < 6; k++) {
      loadCell<dir>(c[k], f, s[k]);
      const double a = ld<double>(f.SoS, s[k]);
      Lam1 = (k == 0) ? fabs(c[k].un - a) : fmax(Lam1, fabs(c[k].un - a));
      Lam  = (k == 0) ? fabs(c[k].un)     : fmax(Lam,  fabs(c[k].un));
      LamN = (k == 0) ? fabs(c[k].un + a) : fmax(LamN, fabs(c[k].un + a));
   }

   RoeState R;
   {
      double YL[nSpec], YR[nSpec], uL[3], uR[3];
      HTR_UNROLL
      for (int i = 0; i < nSpec; i++) { YL[i] = ldc<double>(f.MassFracs, s[2], i); YR[i] = ldc<double>(f.MassFracs, s[3], i); }
      HTR_UNROLL
      for (int i = 0; i < 3; i++) { uL[i] = ldc<double>(f.velocity, s[2], i); uR[i] = ldc<double>(f.velocity, s[3], i); }
      roeAverage(R, c[2].U, c[3].U, YL, YR, c[2].p, c[3].p, uL, uR, ld<double>(f.rho, s[2]), ld<double>(f.rho, s[3]));
   }

   double Fp[6][nEq], Fm[6][nEq];
   HTR_UNROLL
   for (int k = 0; k < 6; k++) splitFlux<dir>(Fp[k], Fm[k], R, c[k], Lam1, Lam, LamN);

   double F[nEq];
   HTR_UNROLL
   for (int i = 0; i < nEq; i++) {
      const double FPlus  = tenoReconstruct<OP, true >(Fp[0][i], Fp[1][i], Fp[2][i], Fp[3][i], Fp[4][i], Fp[5][i], t);
      const double FMinus = tenoReconstruct<OP, false>(Fm[0][i], Fm[1][i], Fm[2][i], Fm[3][i], Fm[4][i], Fm[5][i], t);
      F[i] = -0.5*(FPlus + FMinus);
   }
   fromCharacteristic<dir>(Flux, F, R);

   if (nSpec > 1) {
      // positivity of the partial densities after this stage: blend with the first-order flux
      constexpr double eps_l = -1e-8;
      bool limit = false;
      HTR_UNROLL
      for (int i = 0; i < nSpec; i++) {
         const double dRhoi = lim_f*Flux[i];
         if (((RK0*ldc<double>(f.Conserved_old, s[2], i) + RK1*c[2].U[i] + dRhoi) < eps_l) ||
             ((RK0*ldc<double>(f.Conserved_old, s[3], i) + RK1*c[3].U[i] - dRhoi) < eps_l))
            limit = true;
      }
      if (limit) {
         constexpr double eps = 1e-60;
         double theta = 1.0;
         HTR_UNROLL
         for (int i = 0; i < nEq; i++) F[i] = -0.5*(Fp[2][i] + Fm[3][i]);
         double FLO[nEq];
         fromCharacteristic<dir>(FLO, F, R);
         for (int i = 0; i < nSpec; i++) {
            const double dRhoi = lim_f*Flux[i];
            const double dRhoi_lo = lim_f*FLO[i];
            double theta_p = 1.0;
            const double rhoi_p = RK0*ldc<double>(f.Conserved_old, s[2], i) + RK1*c[2].U[i];
            const double rhoHO_p = rhoi_p + dRhoi;
            if (rhoHO_p < eps) {
               const double rhoLO = rhoi_p + dRhoi_lo;
               theta_p = (fabs(rhoHO_p - rhoLO) > eps) ? fmax(0.0, (eps - rhoLO)/(rhoHO_p - rhoLO)) : 1.0;
            }
            double theta_pp1 = 1.0;
            const double rhoi_pp1 = RK0*ldc<double>(f.Conserved_old, s[3], i) + RK1*c[3].U[i];
            const double rhoHO_pp1 = rhoi_pp1 - dRhoi;
            if (rhoHO_pp1 < eps) {
               const double rhoLO = rhoi_pp1 - dRhoi_lo;
               theta_pp1 = (fabs(rhoHO_pp1 - rhoLO) > eps) ? fmax(0.0, (eps - rhoLO)/(rhoHO_pp1 - rhoLO)) : 1.0;
            }
            theta = fmin(fmin(theta_p, theta_pp1), theta);
         }
         HTR_UNROLL
         for (int i = 0; i < nEq; i++) Flux[i] = Flux[i]*theta + (1 - theta)*FLO[i];
      }
   }
}

} // namespace htr
