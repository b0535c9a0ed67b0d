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
// Mixture thermodynamics, transport and kinetics evaluated in registers from the parameter
// block in __constant__ memory.  Free functions over plain arrays; one set per mixture family.
//
// Algorithms follow the reference's `Mix` methods, operation order kept so results agree to
// round-off:  ConstPropMix  src/Mixtures/ConstPropMix.inl:83-167
//             multi-comp.   src/Mixtures/MultiComponent.inl:59-435, Species.inl:33-357,
//                           Reaction.inl:34-212
#pragma once
#include "htr_mix_layout.h"

namespace htr {

__constant__ MixData c_mix;
#if HTR_MULTICOMP
// Quotients of table constants, evaluated once on the host with the reference's own expressions
// (src/Mixtures/Species.inl: RGAS/W, cp[3]*0.50, cp[4]/3, cp[5]*0.25, cp[6]/5, cp[5]/6) for the ranges H, M, L.
struct SpecDerived { double rOvW; double q[3][5]; };
__constant__ SpecDerived c_der[HTR_NSPEC];
// Temperature-independent parts of the transport models (src/Mixtures/Species.inl:205-357, MultiComponent.inl:150-260),
// evaluated once on the host with the reference's expressions and libm (htr_set_mixture): powers and roots of
// collision diameters, molar-weight ratios and well depths that the reference recomputes per cell.
struct TransDerived {
   double sigma2[HTR_NSPEC];                 // pow(sigma, 2)
   double fz298[HTR_NSPEC];                  // GetFZrot(298)
   double w25[HTR_NSPEC][HTR_NSPEC];         // pow(W[j]/W[i], 0.25)
   double s8[HTR_NSPEC][HTR_NSPEC];          // sqrt(8*(1 + W[i]/W[j]))
   double sigmaij[HTR_NSPEC][HTR_NSPEC];     // Stockmayer-corrected pair diameter, i > j
   double kboEpsij[HTR_NSPEC][HTR_NSPEC];
   double invWij[HTR_NSPEC][HTR_NSPEC];
   double pi15, kb3;                         // pow(PI, 1.5), pow(kb, 3)
   // production build only (GetTransport): the temperature-independent factors of the kinetic-theory formulas
   double cmu[HTR_NSPEC];                    // mu_i = cmu_i sqrt(T)/Omega_mu(T*_i)
   double is8[HTR_NSPEC][HTR_NSPEC];         // 1/s8
   double cdif[HTR_NSPEC][HTR_NSPEC];        // 1/D_ij = cdif_ij P Omega_D(T*_ij)/T^1.5, i > j
   double zf[HTR_NSPEC];                     // Z298 * fz298
};
__constant__ TransDerived c_tr;
#endif

#define HTR_DEV __device__ __forceinline__
#define HTR_UNROLL _Pragma("unroll")

// a/b where 1 ulp is inside the tolerance of the production build (post-stage recovery, transport): reciprocal + multiply,
// about half the instructions of the IEEE division sequence (the fused RK stage executed ~80 divisions per cell, ~40 % of its
// instructions).  The strict build keeps the division: same bits as the reference's CPU code.
#ifdef HTR_EXACT_POW
#define HTR_QDIV(a, b) ((a)/(b))
#else
#define HTR_QDIV(a, b) ((a)*__drcp_rn(b))
#endif

#if !HTR_MULTICOMP
// ------------------------------------------------------------------------------------------
// Calorically perfect gas, one species
// ------------------------------------------------------------------------------------------
namespace mix {
HTR_DEV void   ClipYi(double *Yi) {}
HTR_DEV double GetMolarWeightFromYi(const double *Yi) { return RGAS/c_mix.R; }
HTR_DEV double GetMolarWeightFromXi(const double *Xi) { return RGAS/c_mix.R; }
HTR_DEV void   GetMolarFractions(double *Xi, double MixW, const double *Yi) { Xi[0] = Yi[0]; }
HTR_DEV void   GetMassFractions(double *Yi, double MixW, const double *Xi) { Yi[0] = Xi[0]; }
HTR_DEV double GetRhoFromRhoYi(const double *rhoYi) { return rhoYi[0]; }
HTR_DEV void   GetYi(double *Yi, double rho, const double *rhoYi) { Yi[0] = rhoYi[0]/rho; }
HTR_DEV double GetRho(double P, double T, double MixW) { return P/(c_mix.R*T); }
HTR_DEV double GetHeatCapacity(double T, const double *Yi) { return c_mix.gamma/(c_mix.gamma - 1)*c_mix.R; }
HTR_DEV double GetSpeciesEnthalpy(int i, double T) { return c_mix.gamma/(c_mix.gamma - 1)*c_mix.R*T; }
HTR_DEV void   GetSpeciesEnthalpies(double *hi, double T) { hi[0] = c_mix.gamma/(c_mix.gamma - 1)*c_mix.R*T; }
HTR_DEV double GetSpeciesMolarWeight(int i) { return RGAS/c_mix.R; }
HTR_DEV double GetInternalEnergy(double T, const double *Yi) { return T*c_mix.R/(c_mix.gamma - 1.0); }
HTR_DEV double GetTFromInternalEnergy(double e0, double T, const double *Yi) { return e0*(c_mix.gamma - 1.0)/c_mix.R; }
HTR_DEV double GetTFromRhoAndP(double rho, double MixW, double P) { return P*MixW/(rho*RGAS); }
HTR_DEV double GetPFromRhoAndT(double rho, double MixW, double T) { return rho*RGAS*T/MixW; }
HTR_DEV double GetViscosity(double T, const double *Xi) {
   if (c_mix.viscosityModel == 0) return c_mix.constantVisc;
   if (c_mix.viscosityModel == 1) return c_mix.powerlawViscRef*pow(T/c_mix.powerlawTempRef, 0.7);
   return (c_mix.sutherlandViscRef*pow(T/c_mix.sutherlandTempRef, 1.5))*
          ((c_mix.sutherlandTempRef + c_mix.sutherlandSRef)/(T + c_mix.sutherlandSRef));
}
HTR_DEV double GetHeatConductivity(double T, const double *Xi) {
   const double cp = c_mix.gamma/(c_mix.gamma - 1)*c_mix.R;
   return cp/c_mix.Prandtl*GetViscosity(T, Xi);
}
HTR_DEV double GetGamma(double T, double MixW, const double *Yi) { return c_mix.gamma; }
HTR_DEV double GetSpeedOfSound(double T, double gamma, double MixW) { return sqrt(gamma*c_mix.R*T); }
HTR_DEV void   GetDiffusivity(double *Di, double P, double T, double MixW, const double *Xi) { Di[0] = 0.0; }
HTR_DEV void   GetProductionRates(double *w, double rho, double P, double T, const double *Yi) { w[0] = 0.0; }
HTR_DEV double Getdpde(double rho, double gamma) { return rho*(gamma - 1); }
HTR_DEV void   Getdpdrhoi(double *dpdrhoi, double gamma, double T, const double *Yi) { dpdrhoi[0] = c_mix.R*T; }
} // namespace mix

#else
// ------------------------------------------------------------------------------------------
// Multi-component ideal-gas mixture with NASA-9 thermodynamics
// ------------------------------------------------------------------------------------------
namespace spec {
// polynomial range for temperature T (dimensional); switch temperatures are floats promoted to double:
// 0 = high (cpH), 1 = middle (cpM), 2 = low (cpL), which are contiguous in NasaPoly
HTR_DEV int rangeIndex(const SpeciesData &s, double T) {
   return (T > s.cp.TSwitch2) ? 0 : ((T > s.cp.TSwitch1) ? 1 : 2);
}
HTR_DEV const double *range(const SpeciesData &s, double T) { return s.cp.cpH + 9*rangeIndex(s, T); }
// The quotients of table constants that the reference evaluates per call (RGAS/W, c4/3, c6/5, c5/6) come from
// c_der: the same IEEE divisions done once on the host (htr_set_mixture), hence the same bits.  log T and 1/T
// are taken once per temperature by the callers (same bits as taking them per species).
HTR_DEV double GetCp(const SpeciesData &s, const SpecDerived &d, double T, double Tinv) {
   const double *c = range(s, T);
   return d.rOvW*(c[0]*Tinv*Tinv + c[1]*Tinv + c[2] + T*(c[3] + T*(c[4] + T*(c[5] + T*c[6]))));
}
HTR_DEV double GetEnthalpy(const SpeciesData &s, const SpecDerived &d, double T, double Tinv, double logT) {
   const int r = rangeIndex(s, T);
   const double *c = s.cp.cpH + 9*r;
   const double *q = d.q[r];
   const double E = -c[0]*Tinv + c[1]*logT + c[7] + T*(c[2] + T*(q[0] + T*(q[1] + T*(q[2] + q[3]*T))));
   return E*d.rOvW;
}
HTR_DEV double GetFreeEnthalpy(const SpeciesData &s, const SpecDerived &d, double T, double Tinv, double logT) {   // H/(RT) - S/R
   const int r = rangeIndex(s, T);
   const double *c = s.cp.cpH + 9*r;
   const double *q = d.q[r];
   const double G = -0.5*c[0]*Tinv*Tinv + c[1]*Tinv*(1.0 + logT) + c[2]*(1.0 - logT) + c[7]*Tinv - c[8];
   return G - 0.5*T*(c[3] + T*(q[1] + T*(q[4] + 0.1*T*c[6])));
}
HTR_DEV double GetCp(const SpeciesData &s, const SpecDerived &d, double T) { return GetCp(s, d, T, 1.0/T); }
// reduced collision integrals, rational fits in T* = T kb/eps
HTR_DEV double omega_mu(double T) {
   const double num = 3.3530622607 + T*(2.53272006 + T*(2.9024238575 + T*0.11186138893));
   const double den = 0.8662326188 + T*(1.3913958626 + T*(3.158490576 + T*(0.18973411754 + T*0.00018682962894)));
   return num/den;
}
HTR_DEV double omega_D(double T) {
   const double num = 6.8728271691 + T*(9.4122316321 + T*(7.7442359037 + T*0.23424661229));
   const double den = 1.45337701568 + T*(5.2269794238 + T*(9.7108519575 + T*(0.46539437353 + T*0.00041908394781)));
   return num/den;
}
// T^3 and x^1.5: libdevice pow in the strict build (the reference calls pow), products otherwise
HTR_DEV double cube(double x) {
#ifdef HTR_EXACT_POW
   return pow(x, 3);
#else
   return x*x*x;
#endif
}
HTR_DEV double pow15(double x) {
#ifdef HTR_EXACT_POW
   return pow(x, 1.5);
#else
   return x*sqrt(x);
#endif
}
HTR_DEV double GetMu(const SpeciesData &s, int i, double T) {
   const double num = 5*sqrt(PI*s.W/Na*kb*T);
   const double den = 16*PI*c_tr.sigma2[i]*omega_mu(T*s.diff.kbOveps);
   return num/den;
}
// Stockmayer-corrected binary diffusion coefficient of the pair (i, j), i > j; T3 = T^3
HTR_DEV double GetDif(int i, int j, double P, double T3, double T) {
   const double sigmaij = c_tr.sigmaij[i][j];
   const double omega11 = omega_D(T*c_tr.kboEpsij[i][j]);
   const double num = 3*sqrt(2*PI*c_tr.kb3*T3*Na*c_tr.invWij[i][j]);
   const double den = 16*PI*P*sigmaij*sigmaij*omega11;
   return num/den;
}
HTR_DEV double GetSelfDiffusion(const SpeciesData &s, int i, double T) {
   const double num = 3*sqrt(PI*kb*T*s.W/Na);
   const double den = 8*PI*c_tr.sigma2[i]*omega_D(T*s.diff.kbOveps);
   return num/den;
}
HTR_DEV double GetFZrot(const SpeciesData &s, double T) {
   const double tmp = 1.0/(s.diff.kbOveps*T);
   return 1 + 0.5*c_tr.pi15*sqrt(tmp) + (2 + 0.25*PI*PI)*tmp + c_tr.pi15*pow15(tmp);
}
// Species conductivity: monatomic (Eucken 15/4) or modified-Eucken with rotational relaxation
HTR_DEV double GetLam(const SpeciesData &s, const SpecDerived &d, int i, double T) {
   const double mu = GetMu(s, i, T);
   if (s.diff.Geom == 0) return 15.0/4*mu*RGAS/s.W;
   const bool lin = (s.diff.Geom == 1);
   const double CvROvR = lin ? 1.0 : 1.5;
   const double CvT = 1.5*RGAS;
   const double CvR = CvROvR*RGAS;
   const double CvV = GetCp(s, d, T)*s.W - (lin ? 3.5 : 4.0)*RGAS;
   const double Dkk = GetSelfDiffusion(s, i, T);
   const double fV = Dkk/mu;
   const double Zrot = s.diff.Z298*c_tr.fz298[i]/GetFZrot(s, T);
   const double A = 2.5 - fV;
   const double B = Zrot + 2/PI*(5./3*CvROvR + fV);
   const double fT = 2.5*(1. - 2*CvR*A/(PI*CvT*B));
   const double fR = fV*(1. + 2*A/(PI*B));
   return mu/s.W*(fT*CvT + fR*CvR + fV*CvV);
}
} // namespace spec

namespace mix {
HTR_DEV void ClipYi(double *Yi) {
   HTR_UNROLL
   for (int i = 0; i < nSpec; i++) { Yi[i] = fmax(Yi[i], 1e-60); Yi[i] = fmin(Yi[i], 1.0); }
}
HTR_DEV double GetMolarWeightFromYi(const double *Yi) {
   double s = 0.0;
   HTR_UNROLL
   for (int i = 0; i < nSpec; i++) s += HTR_QDIV(Yi[i], c_mix.species[i].W);
   return 1.0/s;
}
HTR_DEV double GetMolarWeightFromXi(const double *Xi) {
   double s = 0.0;
   HTR_UNROLL
   for (int i = 0; i < nSpec; i++) s += Xi[i]*c_mix.species[i].W;
   return s;
}
HTR_DEV void GetMolarFractions(double *Xi, double MixW, const double *Yi) {
   HTR_UNROLL
   for (int i = 0; i < nSpec; i++) Xi[i] = HTR_QDIV(Yi[i]*MixW, c_mix.species[i].W);
}
HTR_DEV void GetMassFractions(double *Yi, double MixW, const double *Xi) {
   HTR_UNROLL
   for (int i = 0; i < nSpec; i++) Yi[i] = HTR_QDIV(Xi[i]*c_mix.species[i].W, MixW);
}
HTR_DEV double GetRhoFromRhoYi(const double *rhoYi) {
   double rho = 0.0;
   HTR_UNROLL
   for (int i = 0; i < nSpec; i++) rho += rhoYi[i];
   return rho;
}
HTR_DEV void GetYi(double *Yi, double rho, const double *rhoYi) {
   const double irho = 1.0/rho;
   HTR_UNROLL
   for (int i = 0; i < nSpec; i++) Yi[i] = irho*rhoYi[i];
}
HTR_DEV double GetRho(double P, double T, double MixW) { return P*MixW*c_mix.iMixWRef/T; }
HTR_DEV double GetHeatCapacity(double Tn, const double *Yi) {
   const double T = Tn*c_mix.TRef;
   const double Tinv = 1.0/T;
   double cp = 0.0;
   HTR_UNROLL
   for (int i = 0; i < nSpec; i++) cp += Yi[i]*spec::GetCp(c_mix.species[i], c_der[i], T, Tinv);
   return cp*c_mix.iCpRef;
}
HTR_DEV double GetSpeciesEnthalpy(int i, double Tn) {
   const double T = Tn*c_mix.TRef;
   return spec::GetEnthalpy(c_mix.species[i], c_der[i], T, 1.0/T, log(T))*c_mix.ieRef;
}
// all species at one temperature: one log, one reciprocal
HTR_DEV void GetSpeciesEnthalpies(double *hi, double Tn) {
   const double T = Tn*c_mix.TRef;
   const double Tinv = 1.0/T, logT = log(T);
   HTR_UNROLL
   for (int i = 0; i < nSpec; i++) hi[i] = spec::GetEnthalpy(c_mix.species[i], c_der[i], T, Tinv, logT)*c_mix.ieRef;
}
HTR_DEV double GetSpeciesMolarWeight(int i) { return c_mix.species[i].W; }
HTR_DEV double GetInternalEnergy(double Tn, const double *Yi) {
   const double T = Tn*c_mix.TRef;
   const double Tinv = 1.0/T, logT = log(T);
   double e = 0.0;
   HTR_UNROLL
   for (int i = 0; i < nSpec; i++) e += Yi[i]*(spec::GetEnthalpy(c_mix.species[i], c_der[i], T, Tinv, logT) - HTR_QDIV(RGAS*T, c_mix.species[i].W));
   return e*c_mix.ieRef;
}
// Newton iteration on e(T) = e0 from the caller's guess (step tolerance 1e-8 on non-dimensional T)
HTR_DEV double GetTFromInternalEnergy(double e0, double T, const double *Yi) {
   double dfdT = 1.0;
   for (int j = 0; j < 1000; j++) {
      const double f = e0 - GetInternalEnergy(T, Yi);
      if (fabs(HTR_QDIV(f, dfdT)) < 1e-8) break;
      dfdT = 0.0;
      const double Td = T*c_mix.TRef, Tdinv = 1.0/Td;
      HTR_UNROLL
      for (int i = 0; i < nSpec; i++)
         dfdT += Yi[i]*(spec::GetCp(c_mix.species[i], c_der[i], Td, Tdinv) - c_der[i].rOvW)*c_mix.iCpRef;
      T += HTR_QDIV(f, dfdT);
   }
   return T;
}
HTR_DEV double GetTFromRhoAndP(double rho, double MixW, double P) { return P*MixW*c_mix.iMixWRef/rho; }
HTR_DEV double GetPFromRhoAndT(double rho, double MixW, double T) { return rho*T/(MixW*c_mix.iMixWRef); }
// Wilke mixing rule
HTR_DEV double GetViscosity(double Tn, const double *Xi) {
   const double T = Tn*c_mix.TRef;
   double muk[nSpec];
   HTR_UNROLL
   for (int i = 0; i < nSpec; i++) muk[i] = spec::GetMu(c_mix.species[i], i, T);
   double mu = 0.0;
   HTR_UNROLL
   for (int i = 0; i < nSpec; i++) {
      double den = 0.0;
      HTR_UNROLL
      for (int j = 0; j < nSpec; j++) {
         double Phi = 1 + sqrt(muk[i]/muk[j])*c_tr.w25[i][j];
         Phi *= Phi/c_tr.s8[i][j];
         den += Xi[j]*Phi;
      }
      mu += Xi[i]*muk[i]/den;
   }
   return mu*c_mix.imuRef;
}
HTR_DEV double GetHeatConductivity(double Tn, const double *Xi) {
   const double T = Tn*c_mix.TRef;
   double a = 0.0, b = 0.0;
   HTR_UNROLL
   for (int i = 0; i < nSpec; i++) {
      const double lami = spec::GetLam(c_mix.species[i], c_der[i], i, T);
      a += Xi[i]*lami;
      b += Xi[i]/lami;
   }
   return 0.5*(a + 1.0/b)*c_mix.ilamRef;
}
HTR_DEV double GetGamma(double T, double MixW, const double *Yi) {
   const double cp = GetHeatCapacity(T, Yi);
   const double Wr = MixW*c_mix.iMixWRef;
   return cp/(cp - 1.0/Wr);
}
HTR_DEV double GetSpeedOfSound(double T, double gamma, double MixW) { return sqrt(gamma*T/(MixW*c_mix.iMixWRef)); }
HTR_DEV void GetDiffusivity(double *Di, double Pn, double Tn, double MixW, const double *Xi) {
   const double T = Tn*c_mix.TRef;
   const double P = Pn*c_mix.PRef;
   const double T3 = spec::cube(T);
   HTR_UNROLL
   for (int i = 0; i < nSpec; i++) Di[i] = 0.0;
   HTR_UNROLL
   for (int i = 0; i < nSpec; i++) {
      HTR_UNROLL
      for (int j = 0; j < i; j++) {
         const double invDij = 1.0/spec::GetDif(i, j, P, T3, T);
         Di[i] += Xi[j]*invDij;
         Di[j] += Xi[i]*invDij;
      }
   }
   HTR_UNROLL
   for (int i = 0; i < nSpec; i++) {
      double num = 0.0;
      HTR_UNROLL
      for (int j = 0; j < nSpec; j++)
         if (j != i) num += Xi[j]*c_mix.species[j].W;
      Di[i] = num/(MixW*Di[i])*c_mix.iDiRef;
   }
}
#ifndef HTR_EXACT_POW
// mu, lam and Di of one state in one pass (production build).  Same kinetic-theory formulas as GetViscosity /
// GetHeatConductivity / GetDiffusivity above, regrouped so that every cell evaluates one sqrt(T), one division per
// collision integral and reciprocal square roots instead of quotient-then-root chains:
//    mu_i  = cmu_i sqrt(T)/Omega_mu_i                         D_kk/mu_k = 1.2 Omega_mu_k/Omega_D_k
//    sqrt(mu_i/mu_j) = sqrt(mu_i) rsqrt(mu_j)                 1/D_ij = cdif_ij P Omega_D_ij/T^1.5
// The regrouping changes rounding only (a few ulp per property; parity tests hold it to 1e-10 against the reference).
HTR_DEV void GetTransport(double &muOut, double &lamOut, double *Di, double Pn, double Tn, double MixW, const double *Xi) {
   const double T = Tn*c_mix.TRef;
   const double P = Pn*c_mix.PRef;
   const double sqrtT = sqrt(T);
   const double Tinv = 1.0/T;
   double muk[nSpec], a = 0.0, b = 0.0;
   HTR_UNROLL
   for (int i = 0; i < nSpec; i++) {
      const SpeciesData &s = c_mix.species[i];
      const double Ts = T*s.diff.kbOveps;
      const double nm = 3.3530622607 + Ts*(2.53272006 + Ts*(2.9024238575 + Ts*0.11186138893));
      const double dm = 0.8662326188 + Ts*(1.3913958626 + Ts*(3.158490576 + Ts*(0.18973411754 + Ts*0.00018682962894)));
      double lami;
      if (s.diff.Geom == 0) {
         muk[i] = HTR_QDIV(c_tr.cmu[i]*sqrtT*dm, nm);
         lami = 15.0/4*muk[i]*c_der[i].rOvW;
      } else {
         const double nd = 6.8728271691 + Ts*(9.4122316321 + Ts*(7.7442359037 + Ts*0.23424661229));
         const double dd = 1.45337701568 + Ts*(5.2269794238 + Ts*(9.7108519575 + Ts*(0.46539437353 + Ts*0.00041908394781)));
         const double inv = 1.0/(nm*dm*nd);                    // one division for both quotients
         muk[i] = c_tr.cmu[i]*sqrtT*(dm*dm*nd*inv);
         const double fV = 1.2*(nm*nm*dd*inv);
         const bool lin = (s.diff.Geom == 1);
         const double CvROvR = lin ? 1.0 : 1.5;
         const double CvT = 1.5*RGAS;
         const double CvR = CvROvR*RGAS;
         const double CvV = spec::GetCp(s, c_der[i], T, Tinv)*s.W - (lin ? 3.5 : 4.0)*RGAS;
         const double rs = rsqrt(Ts), tmp = rs*rs;
         const double F = 1 + 0.5*c_tr.pi15*rs + (2 + 0.25*PI*PI)*tmp + c_tr.pi15*tmp*rs;
         const double A = 2.5 - fV;
         const double c = 2/PI*(5./3*CvROvR + fV);
         const double iB = HTR_QDIV(F, c_tr.zf[i] + c*F);              // 1/(Zrot + c), Zrot = zf/F
         const double fT = 2.5*(1. - HTR_QDIV(2*CvR*A*iB, PI*CvT));
         const double fR = fV*(1. + HTR_QDIV(2*A*iB, PI));
         lami = HTR_QDIV(muk[i]*(fT*CvT + fR*CvR + fV*CvV), s.W);
      }
      a += Xi[i]*lami;
      b += HTR_QDIV(Xi[i], lami);
   }
   lamOut = 0.5*(a + 1.0/b)*c_mix.ilamRef;
   double sm[nSpec], ism[nSpec];
   HTR_UNROLL
   for (int i = 0; i < nSpec; i++) { ism[i] = rsqrt(muk[i]); sm[i] = muk[i]*ism[i]; }
   double mu = 0.0;
   HTR_UNROLL
   for (int i = 0; i < nSpec; i++) {
      double den = 0.0;
      HTR_UNROLL
      for (int j = 0; j < nSpec; j++) {
         const double Phi = 1 + sm[i]*ism[j]*c_tr.w25[i][j];
         den += Xi[j]*(Phi*Phi*c_tr.is8[i][j]);
      }
      mu += HTR_QDIV(Xi[i]*muk[i], den);
   }
   muOut = mu*c_mix.imuRef;
   const double pt = P*Tinv/sqrtT;
   HTR_UNROLL
   for (int i = 0; i < nSpec; i++) Di[i] = 0.0;
   HTR_UNROLL
   for (int i = 0; i < nSpec; i++) {
      HTR_UNROLL
      for (int j = 0; j < i; j++) {
         const double Ts = T*c_tr.kboEpsij[i][j];
         const double nd = 6.8728271691 + Ts*(9.4122316321 + Ts*(7.7442359037 + Ts*0.23424661229));
         const double dd = 1.45337701568 + Ts*(5.2269794238 + Ts*(9.7108519575 + Ts*(0.46539437353 + Ts*0.00041908394781)));
         const double invDij = HTR_QDIV(c_tr.cdif[i][j]*pt*nd, dd);
         Di[i] += Xi[j]*invDij;
         Di[j] += Xi[i]*invDij;
      }
   }
   const double iW = c_mix.iDiRef/MixW;
   HTR_UNROLL
   for (int i = 0; i < nSpec; i++) {
      double num = 0.0;
      HTR_UNROLL
      for (int j = 0; j < nSpec; j++)
         if (j != i) num += Xi[j]*c_mix.species[j].W;
      Di[i] = HTR_QDIV(num*iW, Di[i]);
   }
}
#endif
HTR_DEV double Getdpde(double rho, double gamma) { return rho*(gamma - 1); }
HTR_DEV void Getdpdrhoi(double *dpdrhoi, double gamma, double Tn, const double *Yi) {
   const double T = Tn*c_mix.TRef;
   const double Tinv = 1.0/T, logT = log(T);
   double e = 0.0;
   HTR_UNROLL
   for (int i = 0; i < nSpec; i++) {
      dpdrhoi[i] = spec::GetEnthalpy(c_mix.species[i], c_der[i], T, Tinv, logT) - RGAS*T/c_mix.species[i].W;
      e += Yi[i]*dpdrhoi[i];
   }
   HTR_UNROLL
   for (int i = 0; i < nSpec; i++)
      dpdrhoi[i] = (RGAS*T/c_mix.species[i].W + (gamma - 1)*(e - dpdrhoi[i]))*c_mix.ieRef;
}

// ---- finite-rate kinetics ----
HTR_DEV double rateCoeff(const Arrhenius &k, double T) {
   double Kf = k.A;
   if (k.n != 0.0) Kf *= pow(T, (double)k.n);
   if (fabs(k.EovR) > 1e-5) Kf *= exp(-k.EovR/T);
   return Kf;
}
HTR_DEV double massAction(const StoichTerm *t, int n, const double *C) {
   double a = 1.0;
   for (int i = 0; i < n; i++) {
      const int ind = t[i].ind;
      if (t[i].nu == 1) a *= C[ind];
      else              a *= pow(C[ind], (double)t[i].nu);
   }
   return a;
}
#ifdef HTR_FWD_ORDERS
// forward rate with specified reaction orders (src/Mixtures/Reaction.inl:110-124 under FWD_ORDERS)
HTR_DEV double massAction(const ReactTerm *t, int n, const double *C) {
   double a = 1.0;
   for (int i = 0; i < n; i++) {
      const int ind = t[i].ind;
      if (t[i].ord == 1) a *= C[ind];
      else               a *= pow(C[ind], (double)t[i].ord);
   }
   return a;
}
#endif
template<class R>
HTR_DEV double backwardRateCoeff(const R &r, double Kf, double P, double T, const double *G) {
   double sumNu = 0.0, sumNuG = 0.0;
   for (int i = 0; i < r.Neducts; i++) { sumNu -= r.educts[i].nu; sumNuG -= r.educts[i].nu*G[r.educts[i].ind]; }
   for (int i = 0; i < r.Npducts; i++) { sumNu += r.pducts[i].nu; sumNuG += r.pducts[i].nu*G[r.pducts[i].ind]; }
   const double lnKc = -sumNuG - sumNu*(log(T) + log(RGAS/P));
   return Kf*exp(-lnKc);
}
template<class R>
HTR_DEV double netRate(const R &r, double P, double T, const double *C, const double *G) {
   const double Kf = rateCoeff(r.k, T);
   const double a = massAction(r.educts, r.Neducts, C);
   double Kb = 0.0, b = 1.0;
   if (r.has_backward) {
      Kb = backwardRateCoeff(r, Kf, P, T, G);
      b = massAction(r.pducts, r.Npducts, C);
   }
   return Kf*a - Kb*b;
}
template<class R>
HTR_DEV void scatter(const R &r, double rate, double *w) {
   for (int i = 0; i < r.Neducts; i++) w[r.educts[i].ind] -= r.educts[i].nu*rate;
   for (int i = 0; i < r.Npducts; i++) w[r.pducts[i].ind] += r.pducts[i].nu*rate;
}
#ifndef HTR_EXACT_POW
// ---- production kinetics: same rate laws, register-only evaluation -------------------------------
// Differences from the operation-for-operation form below (kept for the strict build):
//   * T^n e^{-E/T} (and the equilibrium factor of the backward rate) is ONE exp of the summed
//     exponent, with log T evaluated once per cell -- no pow(); the result differs from
//     pow()*exp()*exp() by the rounding of the exponent sum (<= ~2e-14 relative, tolerance 1e-10);
//   * the structure of the mechanism (species indices, stoichiometric coefficients, which Arrhenius terms
//     exist) is compile-time (generated/<Mix>_mechanism.inc), so every array index is a literal after
//     unrolling: no local memory, no select chains; nu = 2 is a multiplication.
HTR_DEV double powNu(double c, float nu) { return (nu == 1.0f) ? c : ((nu == 2.0f) ? c*c : ((nu == 0.5f) ? sqrt(c) : pow(c, (double)nu))); }
HTR_DEV void GetProductionRates(double *w, double rhon, double Pn, double Tn, const double *Yi) {
   // structure of the mechanism as compile-time tables (generated/<Mix>_mechanism.inc); htr_set_mixture
   // refuses a mixture block whose reaction tables differ from the ones these were generated from
#if defined(HTR_MIX_AirMix)
#include "generated/AirMix_mechanism.inc"
#elif defined(HTR_MIX_CH41StMix)
#include "generated/CH41StMix_mechanism.inc"
#endif
   static_assert(MECH_NR == HTR_NREAC + HTR_NTBREAC, "mechanism tables out of date");
   const double T = Tn*c_mix.TRef;
   const double P = Pn*c_mix.PRef;
   const double rho = rhon*c_mix.rhoRef;
   const double logT = log(T), iT = 1.0/T;
   const double logTRP = logT + log(HTR_QDIV(RGAS, P));
   double G[nSpec], C[nSpec];
   HTR_UNROLL
   for (int i = 0; i < nSpec; i++) {
      w[i] = 0.0;
      C[i] = HTR_QDIV(Yi[i]*rho, c_mix.species[i].W);
      G[i] = spec::GetFreeEnthalpy(c_mix.species[i], c_der[i], T, iT, logT);
   }
   const double cM = HTR_QDIV(P, RGAS*T);
   HTR_UNROLL
   for (int r = 0; r < MECH_NR; r++) {
#if HTR_NTBREAC > 0
      const Arrhenius k = (r < HTR_NREAC) ? c_mix.reactions[r < HTR_NREAC ? r : 0].k
                                          : c_mix.tbReactions[r >= HTR_NREAC ? r - HTR_NREAC : 0].k;
#else
      const Arrhenius k = c_mix.reactions[r].k;
#endif
      double ex = 0.0;
      if (MECH_HASN[r]) ex += k.n*logT;
      if (MECH_HASE[r]) ex -= k.EovR*iT;
      double a = 1.0;
      HTR_UNROLL
      for (int i = 0; i < HTR_MAX_REACTANTS; i++) if (i < MECH_NE[r]) a *= powNu(C[MECH_EI[r][i]], MECH_EO[r][i]);   // EO = EN unless FWD_ORDERS
      double rate = k.A*exp(ex)*a;
      if (MECH_BW[r]) {
         double sumNu = 0.0, sumNuG = 0.0, b = 1.0;
         HTR_UNROLL
         for (int i = 0; i < HTR_MAX_REACTANTS; i++) if (i < MECH_NE[r]) { sumNu -= MECH_EN[r][i]; sumNuG -= MECH_EN[r][i]*G[MECH_EI[r][i]]; }
         HTR_UNROLL
         for (int i = 0; i < HTR_MAX_PRODUCTS; i++) if (i < MECH_NP[r]) {
            sumNu += MECH_PN[r][i]; sumNuG += MECH_PN[r][i]*G[MECH_PI[r][i]];
            b *= powNu(C[MECH_PI[r][i]], MECH_PN[r][i]);
         }
         rate -= k.A*exp(ex + sumNuG + sumNu*logTRP)*b;
      }
      if (r >= HTR_NREAC && MECH_NT[r] > 0) {
         double c = cM;
         HTR_UNROLL
         for (int i = 0; i < (HTR_MAX_TB > 0 ? HTR_MAX_TB : 1); i++) if (i < MECH_NT[r]) c += C[MECH_TI[r][i]]*(MECH_TN[r][i] - 1.0);
         rate *= c;
      }
      HTR_UNROLL
      for (int i = 0; i < HTR_MAX_REACTANTS; i++) if (i < MECH_NE[r]) w[MECH_EI[r][i]] -= MECH_EN[r][i]*rate;
      HTR_UNROLL
      for (int i = 0; i < HTR_MAX_PRODUCTS; i++) if (i < MECH_NP[r]) w[MECH_PI[r][i]] += MECH_PN[r][i]*rate;
   }
   HTR_UNROLL
   for (int i = 0; i < nSpec; i++) w[i] *= c_mix.species[i].W*c_mix.iwiRef;
}
#else
// w_i in computational units from non-dimensional rho, P, T and mass fractions
HTR_DEV void GetProductionRates(double *w, double rhon, double Pn, double Tn, const double *Yi) {
   const double T = Tn*c_mix.TRef;
   const double P = Pn*c_mix.PRef;
   const double rho = rhon*c_mix.rhoRef;
   double G[nSpec], C[nSpec];
   HTR_UNROLL
   for (int i = 0; i < nSpec; i++) {
      w[i] = 0.0;
      C[i] = Yi[i]*rho/c_mix.species[i].W;
      G[i] = spec::GetFreeEnthalpy(c_mix.species[i], c_der[i], T, 1.0/T, log(T));
   }
#if HTR_NREAC > 0
   HTR_UNROLL
   for (int r = 0; r < HTR_NREAC; r++)
      scatter(c_mix.reactions[r], netRate(c_mix.reactions[r], P, T, C, G), w);
#endif
#if HTR_NTBREAC > 0
   HTR_UNROLL
   for (int r = 0; r < HTR_NTBREAC; r++) {
      const CollisionReaction &R = c_mix.tbReactions[r];
      double c = 1.0;
      if (R.Nthirdb > 0) {
         c = P/(RGAS*T);
         for (int i = 0; i < R.Nthirdb; i++) c += C[R.thirdb[i].ind]*(R.thirdb[i].nu - 1.0);
      }
      scatter(R, c*netRate(R, P, T, C, G), w);
   }
#endif
   HTR_UNROLL
   for (int i = 0; i < nSpec; i++) w[i] *= c_mix.species[i].W*c_mix.iwiRef;
}
#endif // HTR_EXACT_POW
} // namespace mix
#endif // HTR_MULTICOMP

} // namespace htr
