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
// Stencil operators: node-type tables in __constant__ memory, staggered interpolation,
// centred derivative, TENO6 / TENO6-A / TENO6-LAD reconstruction and the TENO-A shock sensor.
// Numerics follow src/prometeo_metric.inl:70-149 (offsets, Interp2Staggered, getDeriv),
// :297-426 (TENOsensor), :431-1037 (reconstruction operators); table values are the
// reference's (generated/stencil_tables.inc).
#pragma once
#include "htr_config.h"

namespace htr {

#define HTR_TABLE(name, type, dims, ...) __constant__ type c_##name dims = __VA_ARGS__;
#include "generated/stencil_tables.inc"
#undef HTR_TABLE

#ifndef HTR_DEV
#define HTR_DEV __device__ __forceinline__
#define HTR_UNROLL _Pragma("unroll")
#endif

enum : int { Std_node = 0, L_S_node = 1, Lp1_S_node = 2, Lp2_S_node = 3, Rm3_S_node = 4, Rm2_S_node = 5,
             Rm1_S_node = 6, R_S_node = 7, L_C_node = 8, Lp1_C_node = 9, Rm2_C_node = 10, Rm1_C_node = 11,
             R_C_node = 12 };

HTR_DEV int offM2(int t) { return c_Cp[t][0]; }
HTR_DEV int offM1(int t) { return c_Cp[t][1]; }
HTR_DEV int offP1(int t) { return c_Cp[t][2]; }
HTR_DEV int offP2(int t) { return c_Cp[t][3]; }
HTR_DEV int offP3(int t) { return c_Cp[t][4]; }

// value at the i+1/2 staggered point (float weights, promoted)
HTR_DEV double Interp2Staggered(int t, double x1, double x2) { return c_Interp[t][0]*x1 + c_Interp[t][1]*x2; }
// cell-centred first derivative; m is the metric 1/dx
HTR_DEV double getDeriv(int t, double xm1, double x, double xp1, double m) {
   return (c_Grad[t][0]*(x - xm1) + c_Grad[t][1]*(xp1 - x))*m;
}

// x^6.  HTR_EXACT_POW selects the libdevice pow (the reference calls pow(x, 6)); the default is three
// multiplications, <= 3 ulp from the correctly rounded value, used only inside cut-off comparisons.
HTR_DEV double pow6(double x) {
#ifdef HTR_EXACT_POW
   return pow(x, 6.0);
#else
   const double x2 = x*x; const double x3 = x2*x; return x3*x3;
#endif
}
HTR_DEV double pow4(double x) {
#ifdef HTR_EXACT_POW
   return pow(x, 4.0);
#else
   const double x2 = x*x; return x2*x2;
#endif
}
HTR_DEV double pow12(double x) {
#ifdef HTR_EXACT_POW
   return pow(x, 12.0);
#else
   const double x2 = x*x; const double x4 = x2*x2; return x4*x4*x4;
#endif
}

// a/b for quotients that only feed cut-off comparisons (the chi_k = tau6/(beta_k + eps) of the TENO weights, the Ren
// indicator).  In smooth flow tau6 is tiny (1e-20 ... 1e-40), which sends the IEEE division through its slow path
// (__cuda_sm20_div_rn_f64_full: ncu attributed 30 % of the executed instructions of the TENO kernels to it,
// profiles/r2_v22_*); the reciprocal of the well-scaled denominator never leaves the fast path, and the product is
// within 1 ulp of the quotient.  The strict build keeps the division (same bits as the reference's CPU code).
HTR_DEV double cutDiv(double a, double b) {
#ifdef HTR_EXACT_POW
   return a/b;
#else
   return a*__drcp_rn(b);
#endif
}

constexpr double tenPow(int n) {   // 10^n by repeated multiplication, like the reference's my_10_pow
   double v = 1;
   for (int i = 0; i < (n < 0 ? -n : n); i++) v = (n < 0 ? 1.0e-1 : 1.0e1)*v;
   return v;
}

// Jiang-Shu smoothness constants
constexpr double C13 = 13.0/12.0, C23 = 3.0/12.0, C14 = 1.0/36.0, C24 = 13.0/12.0, C34 = 781.0/720.0;

struct Smoothness { double s1, s2, s3, s4, tau6; };

// smoothness of the 3 three-point and the 1 four-point candidate stencils seen from the upwind side.
// Plus: (ym2..yp3) as given;  Minus: mirrored about the face.
template<bool plus>
HTR_DEV Smoothness smoothness(double ym2, double ym1, double y, double yp1, double yp2, double yp3) {
   Smoothness r;
   double a1, a2, a3;
   if (plus) {
      a1 = (ym1 - 2*y + yp1);   a2 = (ym1 - yp1);             r.s1 = C13*a1*a1 + C23*a2*a2;
      a1 = (y - 2*yp1 + yp2);   a2 = (3*y - 4*yp1 + yp2);     r.s2 = C13*a1*a1 + C23*a2*a2;
      a1 = (ym2 - 2*ym1 + y);   a2 = (3*y - 4*ym1 + ym2);     r.s3 = C13*a1*a1 + C23*a2*a2;
      a1 = (-11*y + 18*yp1 - 9*yp2 + 2*yp3);
      a2 = (2*y - 5*yp1 + 4*yp2 - yp3);
      a3 = (-y + 3*yp1 - 3*yp2 + yp3);
   } else {
      a1 = (yp2 - 2*yp1 + y);   a2 = (yp2 - y);               r.s1 = C13*a1*a1 + C23*a2*a2;
      a1 = (yp1 - 2*y + ym1);   a2 = (3*yp1 - 4*y + ym1);     r.s2 = C13*a1*a1 + C23*a2*a2;
      a1 = (yp3 - 2*yp2 + yp1); a2 = (3*yp1 - 4*yp2 + yp3);   r.s3 = C13*a1*a1 + C23*a2*a2;
      a1 = (-11*yp1 + 18*y - 9*ym1 + 2*ym2);
      a2 = (2*yp1 - 5*y + 4*ym1 - ym2);
      a3 = (-yp1 + 3*y - 3*ym1 + ym2);
   }
   r.s4 = C14*a1*a1 + C24*a2*a2 + C34*a3*a3;
   r.tau6 = fabs(r.s4 - (r.s3 + r.s2 + 4*r.s1)/6);
   return r;
}

// Blend the candidate reconstructions with the surviving optimal weights (b_k in {0,1}).
template<bool plus>
HTR_DEV double assemble(double ym2, double ym1, double y, double yp1, double yp2, double yp3, int t,
                        int b1, int b2, int b3, int b4) {
   const double *Co = plus ? c_Coeffs_Plus[t] : c_Coeffs_Minus[t];
   const double *Re = plus ? c_Recon_Plus[t] : c_Recon_Minus[t];
   const double y0 = plus ? y : yp1;
   double var[4];
   HTR_UNROLL
   for (int k = 0; k < 4; k++)
      var[k] = ym2*Re[6*k + 0] + ym1*Re[6*k + 1] + y*Re[6*k + 2] + yp1*Re[6*k + 3] + yp2*Re[6*k + 4] + yp3*Re[6*k + 5] - y0;
   const double a1 = Co[0]*b1, a2 = Co[1]*b2, a3 = Co[2]*b3, a4 = Co[3]*b4;
   const double a = 1.0/(a1 + a2 + a3 + a4);
   return y0 + (a1*a)*var[0] + (a2*a)*var[1] + (a3*a)*var[2] + (a4*a)*var[3];
}

template<bool plus>
HTR_DEV void dropUnusedStencils(int t, double &a1, double &a2, double &a3, double &a4) {
   const double *Co = plus ? c_Coeffs_Plus[t] : c_Coeffs_Minus[t];
   if (Co[0] < 1e-10) a1 = 0.0;
   if (Co[1] < 1e-10) a2 = 0.0;
   if (Co[2] < 1e-10) a3 = 0.0;
   if (Co[3] < 1e-10) a4 = 0.0;
}

enum { OP_TENO = 0, OP_TENOA = 1, OP_TENOLAD = 2 };

// One-sided sixth-order targeted-ENO reconstruction at the i+1/2 face.
//  OP_TENO    : fixed cut-off 1e-6, eps 1e-10
//  OP_TENOA   : cut-off adapted by a Ren smoothness indicator on standard nodes (eps 1e-8)
//  OP_TENOLAD : cut-off adapted by the largest chi_k = tau6/(beta_k+eps) (eps 1e-16)
template<int OP, bool plus>
HTR_DEV double tenoReconstruct(double ym2, double ym1, double y, double yp1, double yp2, double yp3, int t) {
   constexpr double eps = (OP == OP_TENO) ? tenPow(-10) : (OP == OP_TENOA) ? tenPow(-8) : tenPow(-16);
   const Smoothness S = smoothness<plus>(ym2, ym1, y, yp1, yp2, yp3);
   const double chi1 = cutDiv(S.tau6, S.s1 + eps);
   const double chi2 = cutDiv(S.tau6, S.s2 + eps);
   const double chi3 = cutDiv(S.tau6, S.s3 + eps);
   const double chi4 = cutDiv(S.tau6, S.s4 + eps);
   double a1 = pow6(1 + chi1), a2 = pow6(1 + chi2), a3 = pow6(1 + chi3), a4 = pow6(1 + chi4);
   double cut_off;
   if (OP == OP_TENO) {
      dropUnusedStencils<plus>(t, a1, a2, a3, a4);
      cut_off = 1e-6;
   } else if (OP == OP_TENOA) {
      if (t == Std_node) {
         constexpr double Ren_r_c = 0.2;
         constexpr double Ren_eps = 0.9*Ren_r_c*1e-16/(1.0 - 0.9*Ren_r_c);
         constexpr double Ren_irc = 1.0/Ren_r_c;
         constexpr double Smooth_pow = 9.5, Shock_pow = 6.0, Diff_pow = Smooth_pow - Shock_pow;
         const double v2 = fabs(ym1 - ym2), v3 = fabs(y - ym1), v4 = fabs(yp1 - y), v5 = fabs(yp2 - yp1), v6 = fabs(yp3 - yp2);
         double eta = fmin(cutDiv(2*v2*v3 + Ren_eps, v2*v2 + v3*v3 + Ren_eps),
                           cutDiv(2*v3*v4 + Ren_eps, v3*v3 + v4*v4 + Ren_eps));
         eta = fmin(eta, cutDiv(2*v4*v5 + Ren_eps, v4*v4 + v5*v5 + Ren_eps));
         eta = fmin(eta, cutDiv(2*v5*v6 + Ren_eps, v5*v5 + v6*v6 + Ren_eps));
         const double delta = 1 - fmin(eta*Ren_irc, 1.0);
         const double decay = pow4(1 - delta)*(1 + 4*delta);
         const int power = min(max(int(floor(Smooth_pow - Diff_pow*(1 - decay))), 0), 13);
         cut_off = c_p10[power];
      } else {
         dropUnusedStencils<plus>(t, a1, a2, a3, a4);
         cut_off = c_p10[4];
      }
   } else {
      dropUnusedStencils<plus>(t, a1, a2, a3, a4);
      constexpr double Smooth_pow = 12.5, Shock_pow = 3.0, iH = 1e-1, Diff_pow = Smooth_pow - Shock_pow;
      const double chiMax = fmax(chi1, fmax(chi2, fmax(chi3, chi4)));
      const double theta = 1.0/(1 + chiMax*iH);
      const int power = min(max(int(floor(Shock_pow + Diff_pow*theta)), 0), 13);
      cut_off = c_p10[power];
   }
   const double a = 1.0/(a1 + a2 + a3 + a4);
   const int b1 = (a1*a < cut_off) ? 0 : 1;
   const int b2 = (a2*a < cut_off) ? 0 : 1;
   const int b3 = (a3*a < cut_off) ? 0 : 1;
   const int b4 = (a4*a < cut_off) ? 0 : 1;
   return assemble<plus>(ym2, ym1, y, yp1, yp2, yp3, t, b1, b2, b3, b4);
}

// The same reconstruction with the upwind side as a RUN-TIME flag, for the lane-cooperative TENO kernels
// (htr_teno_coop.cuh), where the lanes of a warp reconstruct plus and minus fields side by side and must not diverge.
// Bit-identical with tenoReconstruct<OP, plus>: smoothness<false>(a..f) is smoothness<true>(f..a) expression for expression
// (same operand order in every sum), the Ren indicator and `assemble` use the inputs in the given order with the tables
// of the chosen side.
template<int OP>
HTR_DEV double tenoReconstructRT(const bool plus, double ym2, double ym1, double y, double yp1, double yp2, double yp3, int t) {
   constexpr double eps = (OP == OP_TENO) ? tenPow(-10) : (OP == OP_TENOA) ? tenPow(-8) : tenPow(-16);
   const Smoothness S = smoothness<true>(plus ? ym2 : yp3, plus ? ym1 : yp2, plus ? y : yp1, plus ? yp1 : y, plus ? yp2 : ym1, plus ? yp3 : ym2);
   double Co[4];
   HTR_UNROLL
   for (int k = 0; k < 4; k++) Co[k] = plus ? c_Coeffs_Plus[t][k] : c_Coeffs_Minus[t][k];
   const double chi1 = cutDiv(S.tau6, S.s1 + eps);
   const double chi2 = cutDiv(S.tau6, S.s2 + eps);
   const double chi3 = cutDiv(S.tau6, S.s3 + eps);
   const double chi4 = cutDiv(S.tau6, S.s4 + eps);
   double a1 = pow6(1 + chi1), a2 = pow6(1 + chi2), a3 = pow6(1 + chi3), a4 = pow6(1 + chi4);
   auto drop = [&]() {
      if (Co[0] < 1e-10) a1 = 0.0;
      if (Co[1] < 1e-10) a2 = 0.0;
      if (Co[2] < 1e-10) a3 = 0.0;
      if (Co[3] < 1e-10) a4 = 0.0;
   };
   double cut_off;
   if (OP == OP_TENO) {
      drop();
      cut_off = 1e-6;
   } else if (OP == OP_TENOA) {
      if (t == Std_node) {
         constexpr double Ren_r_c = 0.2;
         constexpr double Ren_eps = 0.9*Ren_r_c*1e-16/(1.0 - 0.9*Ren_r_c);
         constexpr double Ren_irc = 1.0/Ren_r_c;
         constexpr double Smooth_pow = 9.5, Shock_pow = 6.0, Diff_pow = Smooth_pow - Shock_pow;
         const double v2 = fabs(ym1 - ym2), v3 = fabs(y - ym1), v4 = fabs(yp1 - y), v5 = fabs(yp2 - yp1), v6 = fabs(yp3 - yp2);
         double eta = fmin(cutDiv(2*v2*v3 + Ren_eps, v2*v2 + v3*v3 + Ren_eps),
                           cutDiv(2*v3*v4 + Ren_eps, v3*v3 + v4*v4 + Ren_eps));
         eta = fmin(eta, cutDiv(2*v4*v5 + Ren_eps, v4*v4 + v5*v5 + Ren_eps));
         eta = fmin(eta, cutDiv(2*v5*v6 + Ren_eps, v5*v5 + v6*v6 + Ren_eps));
         const double delta = 1 - fmin(eta*Ren_irc, 1.0);
         const double decay = pow4(1 - delta)*(1 + 4*delta);
         const int power = min(max(int(floor(Smooth_pow - Diff_pow*(1 - decay))), 0), 13);
         cut_off = c_p10[power];
      } else {
         drop();
         cut_off = c_p10[4];
      }
   } else {
      drop();
      constexpr double Smooth_pow = 12.5, Shock_pow = 3.0, iH = 1e-1, Diff_pow = Smooth_pow - Shock_pow;
      const double chiMax = fmax(chi1, fmax(chi2, fmax(chi3, chi4)));
      const double theta = 1.0/(1 + chiMax*iH);
      const int power = min(max(int(floor(Shock_pow + Diff_pow*theta)), 0), 13);
      cut_off = c_p10[power];
   }
   const double a = 1.0/(a1 + a2 + a3 + a4);
   const int b1 = (a1*a < cut_off) ? 0 : 1;
   const int b2 = (a2*a < cut_off) ? 0 : 1;
   const int b3 = (a3*a < cut_off) ? 0 : 1;
   const int b4 = (a4*a < cut_off) ? 0 : 1;
   // assemble<plus> with the side chosen at run time
   const double y0 = plus ? y : yp1;
   double var[4];
   HTR_UNROLL
   for (int k = 0; k < 4; k++) {
      double Re[6];
      HTR_UNROLL
      for (int j = 0; j < 6; j++) Re[j] = plus ? c_Recon_Plus[t][6*k + j] : c_Recon_Minus[t][6*k + j];
      var[k] = ym2*Re[0] + ym1*Re[1] + y*Re[2] + yp1*Re[3] + yp2*Re[4] + yp3*Re[5] - y0;
   }
   const double w1 = Co[0]*b1, w2 = Co[1]*b2, w3 = Co[2]*b3, w4 = Co[3]*b4;
   const double w = 1.0/(w1 + w2 + w3 + w4);
   return y0 + (w1*w)*var[0] + (w2*w)*var[1] + (w3*w)*var[2] + (w4*w)*var[3];
}

// TENO-A smoothness test used by the shock sensor: true = smooth (central flux allowed).
// Phi is the largest Ducros value over the stencil.
HTR_DEV bool tenoASensor(double ym2, double ym1, double y, double yp1, double yp2, double yp3, int t, double Phi) {
   constexpr double Ren_r_c = 0.2;
   constexpr double Ren_eps = 0.9*Ren_r_c*1e-6/(1.0 - 0.9*Ren_r_c);
   constexpr double Ren_irc = 1.0/Ren_r_c;
   if ((t == L_S_node) || (t == Rm1_S_node)) return true;
   if (t == L_C_node) {
      const double v1 = yp1 - y, v2 = yp2 - yp1;
      const double eta = (fabs(2.0*v1*v2) + Ren_eps)/(v1*v1 + v2*v2 + Ren_eps);
      return ((1.0 - fmin(1.0, Ren_irc*eta)) < 0.5);
   }
   if (t == Rm1_C_node) {
      const double v1 = y - ym1, v2 = yp1 - y;
      const double eta = (fabs(2.0*v1*v2) + Ren_eps)/(v1*v1 + v2*v2 + Ren_eps);
      return ((1.0 - fmin(1.0, Ren_irc*eta)) < 0.5);
   }
   constexpr double eps = 1e-8, Smooth_pow = 12.5, Shock_pow = 1.0, Diff_pow = Smooth_pow - Shock_pow;
   const Smoothness S = smoothness<true>(ym2, ym1, y, yp1, yp2, yp3);
   const double a1 = pow6(1 + cutDiv(S.tau6, S.s1 + eps));
   const double a2 = pow6(1 + cutDiv(S.tau6, S.s2 + eps));
   const double a3 = pow6(1 + cutDiv(S.tau6, S.s3 + eps));
   const double a4 = pow6(1 + cutDiv(S.tau6, S.s4 + eps));
   const double a = 1.0/(a1 + a2 + a3 + a4);
   const double decay = pow12(1 - Phi)*(1 + 12*Phi);
   const int power = min(max(int(floor(Smooth_pow - Diff_pow*(1 - decay))), 0), 13);
   const double cut_off = c_p10[power];
   return ((a1*a > cut_off) && (a2*a > cut_off) && (a3*a > cut_off) && (a4*a > cut_off));
}

} // namespace htr
