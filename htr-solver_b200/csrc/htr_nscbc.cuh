// Copyright (c) "2019, by Stanford University
//               Developer: Mario Di Renzo
//               Affiliation: Center for Turbulence Research, Stanford University
//               URL: https://ctr.stanford.edu
//               Citation: Di Renzo, M., Lin, F., and Urzay, J. (2020).
//                         HTR solver: An open-source exascale-oriented task-based
//                         multi-GPU high-order code for hypersonic aerothermodynamics.
//                         Computer Physics Communications 255, 107262"
// All rights reserved.  (BSD 3-clause; the LODI expressions below restate src/prometeo_rhs.inl:1285-1930 of the HTR solver,
// whose notice this is; redistribution conditions and disclaimer as in that file.)
//
// ---------------------------------------------------------------------------------------------------------------
// Non-periodic boundaries: the NSCBC right-hand-side tasks (LODI corrections of Conserved_t on the boundary cells) and the
// boundary-cell primitive updates of src/prometeo_bc.{hpp,cc,cu}.  Boundary planes are thin (one cell along `dir`), so
// these are plain one-thread-per-cell kernels: they are launch-latency bound at any realistic size and keep the
// reference's expression order term for term (the strict build is bit-identical with the reference's CPU tasks).
// Included by htr_kernels.cu after the point kernels (uses cellOf, primitivesFromConserved, velocityDeriv).
#pragma once

namespace htr {

// one-sided first derivatives at a boundary node (src/prometeo_metric.inl:93-102)
HTR_DEV double derivLeftBC(int t, double x, double xp1, double m) { return (c_Grad[t][1]*(xp1 - x))*m; }
HTR_DEV double derivRightBC(int t, double xm1, double x, double m) { return (c_Grad[t][0]*(x - xm1))*m; }
// one-sided derivative towards the interior: SIDE 0 = minus boundary (interior at +1), 1 = plus boundary (interior at -1)
template<int SIDE>
HTR_DEV double derivBC(int t, double bnd, double inn, double m) {
   return SIDE == 0 ? derivLeftBC(t, bnd, inn, m) : derivRightBC(t, inn, bnd, m);
}
template<int dir, int SIDE>
HTR_DEV Pt interiorOf(const Pt &p) {     // getPIntBC (src/Utils/PointDomain_helper.hpp)
   Pt q = p;
   ((dir == 0) ? q.x : ((dir == 1) ? q.y : q.z)) += (SIDE == 0 ? 1 : -1);
   return q;
}
// d(velocity component c)/d(xi_D) at q
template<int D>
HTR_DEV double velocityDerivComp(const htr_fluid &f, const Pt &q, int c, const htr_rect &B) {
   const int t = ld<int>(nTypeOf<D>(f), q);
   const Pt qm = shift<D>(q, offM1(t), B), qp = shift<D>(q, offP1(t), B);
   return getDeriv(t, ldc<double>(f.velocity, qm, c), ldc<double>(f.velocity, q, c), ldc<double>(f.velocity, qp, c), ld<double>(m_dOf<D>(f), q));
}

// ---- inflow (src/prometeo_rhs.inl:1288-1412) ---------------------------------------------------------------
template<int dir, int SIDE>
HTR_DEV void nscbcInflow(double *RHS, const htr_fluid &f, const htr_bc_fields &b, const Pt &p) {
   const double SoS = ld<double>(f.SoS, p), rho = ld<double>(f.rho, p), T = ld<double>(f.temperature, p);
   const double un = ldc<double>(f.velocity, p, dir);
   if (SIDE == 0 ? (un >= SoS) : (un <= -SoS)) {       // supersonic inlet
      HTR_UNROLL
      for (int l = 0; l < nEq; l++) RHS[l] = 0.0;
      return;
   }
   const Pt pi = interiorOf<dir, SIDE>(p);
   const int t = ld<int>(nTypeOf<dir>(f), p);
   const double m = ld<double>(m_dOf<dir>(f), p);
   double Yi[nSpec];
   HTR_UNROLL
   for (int s = 0; s < nSpec; s++) Yi[s] = ldc<double>(f.MassFracs, p, s);
   const double MixW_bnd = mix::GetMolarWeightFromYi(Yi);
   const double Cp_bnd = mix::GetHeatCapacity(T, Yi);
   const double lambda = un;                            // characteristic velocity entering the domain
   const double dp_dn = derivBC<SIDE>(t, ld<double>(f.pressure, p), ld<double>(f.pressure, pi), m);
   const double du_dn = derivBC<SIDE>(t, un, ldc<double>(f.velocity, pi, dir), m);
   const double dudtN = ldc<double>(b.dudtBoundary, p, dir);
   double L1, LN;
   if (SIDE == 0) {
      const double lambda_1 = un - SoS;                 // leaving the domain
      L1 = lambda_1*(dp_dn - rho*SoS*du_dn);
      LN = L1 - 2*rho*SoS*dudtN;
   } else {
      const double lambda_N = un + SoS;
      LN = lambda_N*(dp_dn + rho*SoS*du_dn);
      L1 = LN + 2*rho*SoS*dudtN;
   }
   double LS[nSpec];
   HTR_UNROLL
   for (int s = 0; s < nSpec; s++) LS[s] = lambda*derivBC<SIDE>(t, Yi[s], ldc<double>(f.MassFracs, pi, s), m);
   double L2 = ld<double>(b.dTdtBoundary, p)/T + (LN + L1)/(2*rho*Cp_bnd*T);
   HTR_UNROLL
   for (int s = 0; s < nSpec; s++) L2 -= MixW_bnd/mix::GetSpeciesMolarWeight(s)*LS[s];
   L2 *= -rho*SoS*SoS;
   const double d1 = (0.5*(L1 + LN) - L2)/(SoS*SoS);
   HTR_UNROLL
   for (int s = 0; s < nSpec; s++) RHS[s] -= (d1*Yi[s] + rho*LS[s]);
}

// Common tail of the outflow / far-field forms: the LODI source of all equations from d1, dM, dN, LS
// (src/prometeo_rhs.inl:1519-1537 and the identical blocks of the other tasks); `visc` = (dtau_dn, viscous_heating) or zeros.
template<int dir>
HTR_DEV void lodiUpdate(double *RHS, const double *Yi, const double *u, const double *U, double rho, double P, double T,
                        double cpT, double MixW, double d1, const double *dM, double dN, const double *LS,
                        bool viscous, double dtau_dn, double viscous_heating) {
   constexpr int iN = dir, iT1 = T1(dir), iT2 = T2(dir);
   HTR_UNROLL
   for (int s = 0; s < nSpec; s++) RHS[s] -= (d1*Yi[s] + rho*LS[s]);
   if (viscous) RHS[irU + iN] -= (u[iN]*d1 + rho*dM[iN] - dtau_dn);
   else         RHS[irU + iN] -= (u[iN]*d1 + rho*dM[iN]);
   RHS[irU + iT1] -= (u[iT1]*d1 + rho*dM[iT1]);
   RHS[irU + iT2] -= (u[iT2]*d1 + rho*dM[iT2]);
   if (viscous)
      RHS[irE] -= ((U[irE] + P)*d1/rho + U[irU + iN]*dM[iN] + U[irU + iT1]*dM[iT1] + U[irU + iT2]*dM[iT2] + cpT*dN - viscous_heating);
   else
      RHS[irE] -= ((U[irE] + P)*d1/rho + U[irU + iN]*dM[iN] + U[irU + iT1]*dM[iT1] + U[irU + iT2]*dM[iT2] + cpT*dN);
   const double cpT_Wmix = cpT*MixW;
   HTR_UNROLL
   for (int s = 0; s < nSpec; s++) {
      const double dEdYi = mix::GetSpeciesEnthalpy(s, T) - cpT_Wmix/mix::GetSpeciesMolarWeight(s);
      RHS[irE] -= (rho*LS[s]*dEdYi);
   }
}

// relaxation coefficient of the pressure wave entering through a subsonic outlet
HTR_DEV double outflowK(double MaxMach, double SoS, double un, double LengthScale) {
   const double sigma = 0.25;
   return (MaxMach < 0.99) ? sigma*(1.0 - MaxMach*MaxMach)*SoS/LengthScale : sigma*(SoS - (un*un)/SoS)/LengthScale;
}

// ---- outflow (src/prometeo_rhs.inl:1418-1638) ---------------------------------------------------------------
template<int dir, int SIDE>
HTR_DEV void nscbcOutflow(double *RHS, const htr_fluid &f, const Pt &p, const htr_rect &B,
                          double MaxMach, double LengthScale, double PInf) {
   constexpr int iN = dir, iT1 = T1(dir), iT2 = T2(dir);
   const Pt pi = interiorOf<dir, SIDE>(p);
   const int t = ld<int>(nTypeOf<dir>(f), p);
   const double m = ld<double>(m_dOf<dir>(f), p);
   const double SoS = ld<double>(f.SoS, p), rho = ld<double>(f.rho, p), T = ld<double>(f.temperature, p), P = ld<double>(f.pressure, p);
   double u[3], ui[3], Yi[nSpec], U[nEq];
   HTR_UNROLL
   for (int i = 0; i < 3; i++) { u[i] = ldc<double>(f.velocity, p, i); ui[i] = ldc<double>(f.velocity, pi, i); }
   HTR_UNROLL
   for (int s = 0; s < nSpec; s++) Yi[s] = ldc<double>(f.MassFracs, p, s);
   HTR_UNROLL
   for (int i = 0; i < nEq; i++) U[i] = ldc<double>(f.Conserved, p, i);

   const double dp_dn  = derivBC<SIDE>(t, P, ld<double>(f.pressure, pi), m);
   const double dun_dn = derivBC<SIDE>(t, u[iN], ui[iN], m);
   const double duT1dn = derivBC<SIDE>(t, u[iT1], ui[iT1], m);
   const double duT2dn = derivBC<SIDE>(t, u[iT2], ui[iT2], m);

   const double lambda_1 = u[iN] - SoS;
   const double lambda   = u[iN];
   const double lambda_s = (SIDE == 0) ? fmin(u[iN], 0.0) : fmax(u[iN], 0.0);     // species: no extrapolation on backflow
   const double lambda_N = u[iN] + SoS;

   double L1, LN;
   if (SIDE == 0) {
      L1 = lambda_1*(dp_dn - rho*SoS*dun_dn);
      if (lambda_N < 0) LN = lambda_N*(dp_dn + rho*SoS*dun_dn);                     // supersonic
      else              LN = outflowK(MaxMach, SoS, u[iN], LengthScale)*(P - PInf);
   } else {
      if (lambda_1 > 0) L1 = lambda_1*(dp_dn - rho*SoS*dun_dn);
      else              L1 = outflowK(MaxMach, SoS, u[iN], LengthScale)*(P - PInf);
      LN = lambda_N*(dp_dn + rho*SoS*dun_dn);
   }
   const double LM = lambda*(dp_dn - SoS*SoS*derivBC<SIDE>(t, rho, ld<double>(f.rho, pi), m));
   double LS[nSpec];
   HTR_UNROLL
   for (int s = 0; s < nSpec; s++) LS[s] = lambda_s*derivBC<SIDE>(t, Yi[s], ldc<double>(f.MassFracs, pi, s), m);

   const double d1 = (0.5*(L1 + LN) - LM)/(SoS*SoS);
   double dM[3];
   dM[iN ] = (LN - L1)/(2*rho*SoS);
   dM[iT1] = lambda*duT1dn;
   dM[iT2] = lambda*duT2dn;
   const double dN = LM/(SoS*SoS);

   // viscous terms
   double vGradT1[3], vGradT2[3];
   velocityDeriv<iT1>(vGradT1, f, p, B);
   velocityDeriv<iT2>(vGradT2, f, p, B);
   const double mu = ld<double>(f.mu, p);
   const double tauNN_bnd = mu*(4*dun_dn - 2*vGradT1[iT1] - 2*vGradT2[iT2])/3;
   const double tauNN_int = ld<double>(f.mu, pi)*(4*velocityDerivComp<dir>(f, pi, iN, B)
                                                - 2*velocityDerivComp<iT1>(f, pi, iT1, B)
                                                - 2*velocityDerivComp<iT2>(f, pi, iT2, B))/3;
   const double dtau_dn = derivBC<SIDE>(t, tauNN_bnd, tauNN_int, m);
   const double viscous_heating = derivBC<SIDE>(t, u[iN]*tauNN_bnd, ui[iN]*tauNN_int, m) +
                                  duT1dn*mu*(duT1dn + vGradT1[iN]) +
                                  duT2dn*mu*(duT2dn + vGradT2[iN]);

   const double cpT = mix::GetHeatCapacity(T, Yi)*T;
   lodiUpdate<dir>(RHS, Yi, u, U, rho, P, T, cpT, mix::GetMolarWeightFromYi(Yi), d1, dM, dN, LS, true, dtau_dn, viscous_heating);
}

// ---- far field (src/prometeo_rhs.inl:1644-1925) -------------------------------------------------------------
// TInf, vInf, XiInf: the far-field state (the *_profile fields, or the constant states of the incoming-shock boundary)
template<int dir, int SIDE>
HTR_DEV void nscbcFarField(double *RHS, const htr_fluid &f, const Pt &p, double TInf, const double *vInf, const double *XiInf,
                           double PInf, double MaxMach, double LengthScale) {
   constexpr int iN = dir, iT1 = T1(dir), iT2 = T2(dir);
   const Pt pi = interiorOf<dir, SIDE>(p);
   const int t = ld<int>(nTypeOf<dir>(f), p);
   const double m = ld<double>(m_dOf<dir>(f), p);
   const double SoS = ld<double>(f.SoS, p), rho = ld<double>(f.rho, p), T = ld<double>(f.temperature, p), P = ld<double>(f.pressure, p);
   double u[3], Yi[nSpec], U[nEq];
   HTR_UNROLL
   for (int i = 0; i < 3; i++) u[i] = ldc<double>(f.velocity, p, i);
   HTR_UNROLL
   for (int s = 0; s < nSpec; s++) Yi[s] = ldc<double>(f.MassFracs, p, s);
   HTR_UNROLL
   for (int i = 0; i < nEq; i++) U[i] = ldc<double>(f.Conserved, p, i);

   const double lambda_1 = u[iN] - SoS;
   const double lambda   = u[iN];
   const double lambda_N = u[iN] + SoS;
   const double dp_dn  = derivBC<SIDE>(t, P, ld<double>(f.pressure, pi), m);
   const double dun_dn = derivBC<SIDE>(t, u[iN], ldc<double>(f.velocity, pi, iN), m);
   const double MixW = mix::GetMolarWeightFromYi(Yi);
   const double cpT = mix::GetHeatCapacity(T, Yi)*T;
   const double MaInflow = (SIDE == 0) ? fmax(lambda/SoS, 0.0) : fmax(-lambda/SoS, 0.0);
   constexpr double sigma = 0.25;
   double LS[nSpec], dM[3];
   if (MaInflow > 0) {
      constexpr double coeff = 39;
      const double Kp = sigma*(1 + coeff*MaInflow*MaInflow)*SoS/LengthScale;
      if (MaInflow > 1) {
         // supersonic inflow: everything weakly imposed
         const double rhoTarget = mix::GetRho(PInf, TInf, MixW);
         HTR_UNROLL
         for (int s = 0; s < nSpec; s++) RHS[s] += Kp*(rhoTarget*XiInf[s]*mix::GetSpeciesMolarWeight(s)/MixW - U[s]);
         HTR_UNROLL
         for (int i = 0; i < 3; i++) RHS[irU + i] += Kp*(rhoTarget*vInf[i] - U[irU + i]);
         RHS[irE] += Kp*(rhoTarget*(0.5*mod2(vInf) + mix::GetInternalEnergy(TInf, Yi)) - U[irE]);
         return;
      }
      // subsonic inflow
      const double K = sigma*(coeff + 1)*MaInflow*MaInflow*SoS/LengthScale;
      double L1, LN;
      if (SIDE == 0) { L1 = lambda_1*(dp_dn - rho*SoS*dun_dn); LN = Kp*(P - PInf); }
      else           { L1 = Kp*(P - PInf); LN = lambda_N*(dp_dn + rho*SoS*dun_dn); }
      HTR_UNROLL
      for (int s = 0; s < nSpec; s++) LS[s] = K*(Yi[s] - XiInf[s]*mix::GetSpeciesMolarWeight(s)/MixW);
      double L2 = K*(TInf/T - 1) + (LN + L1)/(2*rho*cpT);
      HTR_UNROLL
      for (int s = 0; s < nSpec; s++) L2 -= MixW/mix::GetSpeciesMolarWeight(s)*LS[s];
      L2 *= -rho*SoS*SoS;
      const double d1 = (0.5*(L1 + LN) - L2)/(SoS*SoS);
      dM[iN ] = (LN - L1)/(2*rho*SoS);
      dM[iT1] = K*(u[iT1] - vInf[iT1]);
      dM[iT2] = K*(u[iT2] - vInf[iT2]);
      const double dN = L2/(SoS*SoS);
      lodiUpdate<dir>(RHS, Yi, u, U, rho, P, T, cpT, MixW, d1, dM, dN, LS, false, 0.0, 0.0);
      return;
   }
   // outflow
   double L1, LN;
   if (SIDE == 0) {
      L1 = lambda_1*(dp_dn - rho*SoS*dun_dn);
      if (lambda_N < 0) LN = lambda_N*(dp_dn + rho*SoS*dun_dn);
      else              LN = outflowK(MaxMach, SoS, u[iN], LengthScale)*(P - PInf);
   } else {
      if (lambda_1 > 0) L1 = lambda_1*(dp_dn - rho*SoS*dun_dn);
      else              L1 = outflowK(MaxMach, SoS, u[iN], LengthScale)*(P - PInf);
      LN = lambda_N*(dp_dn + rho*SoS*dun_dn);
   }
   const double L2 = lambda*(dp_dn - SoS*SoS*derivBC<SIDE>(t, rho, ld<double>(f.rho, pi), m));
   HTR_UNROLL
   for (int s = 0; s < nSpec; s++) LS[s] = lambda*derivBC<SIDE>(t, Yi[s], ldc<double>(f.MassFracs, pi, s), m);
   const double d1 = (0.5*(L1 + LN) - L2)/(SoS*SoS);
   dM[iN ] = (LN - L1)/(2*rho*SoS);
   dM[iT1] = lambda*derivBC<SIDE>(t, u[iT1], ldc<double>(f.velocity, pi, iT1), m);
   dM[iT2] = lambda*derivBC<SIDE>(t, u[iT2], ldc<double>(f.velocity, pi, iT2), m);
   const double dN = L2/(SoS*SoS);
   lodiUpdate<dir>(RHS, Yi, u, U, rho, P, T, cpT, MixW, d1, dM, dN, LS, false, 0.0, 0.0);
}

// ---- kernels: one thread per boundary cell ----------------------------------------------------------------------
struct ShockParams {       // IncomingShockParams (src/prometeo_bc_types.h:46-59)
   int iShock;
   double Xi[nSpec];
   double P0, T0, v0[3], P1, T1, v1[3];
};

template<int dir, int SIDE>
__global__ void __launch_bounds__(128) nscbc_inflow_kernel(const htr_fluid f, const htr_bc_fields b, const htr_rect bc) {
   Pt p; if (!cellOf(p, bc)) return;
   double *Ct = at<double>(f.Conserved_t, p);
   double RHS[nEq];
   HTR_UNROLL
   for (int i = 0; i < nEq; i++) RHS[i] = Ct[i];
   nscbcInflow<dir, SIDE>(RHS, f, b, p);
   HTR_UNROLL
   for (int i = 0; i < nEq; i++) Ct[i] = RHS[i];
}

template<int dir, int SIDE>
__global__ void __launch_bounds__(128) nscbc_outflow_kernel(const htr_fluid f, const htr_rect bc, const htr_rect B,
                                                            const double MaxMach, const double LengthScale, const double PInf) {
   Pt p; if (!cellOf(p, bc)) return;
   double *Ct = at<double>(f.Conserved_t, p);
   double RHS[nEq];
   HTR_UNROLL
   for (int i = 0; i < nEq; i++) RHS[i] = Ct[i];
   nscbcOutflow<dir, SIDE>(RHS, f, p, B, MaxMach, LengthScale, PInf);
   HTR_UNROLL
   for (int i = 0; i < nEq; i++) Ct[i] = RHS[i];
}

template<int dir, int SIDE>
__global__ void __launch_bounds__(128) nscbc_farfield_kernel(const htr_fluid f, const htr_bc_fields b, const htr_rect bc,
                                                             const double MaxMach, const double LengthScale, const double PInf) {
   Pt p; if (!cellOf(p, bc)) return;
   double *Ct = at<double>(f.Conserved_t, p);
   double RHS[nEq], vInf[3], XiInf[nSpec];
   HTR_UNROLL
   for (int i = 0; i < nEq; i++) RHS[i] = Ct[i];
   HTR_UNROLL
   for (int i = 0; i < 3; i++) vInf[i] = ldc<double>(b.velocity_profile, p, i);
   HTR_UNROLL
   for (int s = 0; s < nSpec; s++) XiInf[s] = ldc<double>(b.MolarFracs_profile, p, s);
   nscbcFarField<dir, SIDE>(RHS, f, p, ld<double>(b.temperature_profile, p), vInf, XiInf, PInf, MaxMach, LengthScale);
   HTR_UNROLL
   for (int i = 0; i < nEq; i++) Ct[i] = RHS[i];
}

// UpdateUsingFluxIncomingShockTask<Ydir> (src/prometeo_rhs.cc:1076-1138): far-field form of the y-plus side with constant states
__global__ void __launch_bounds__(128) nscbc_incoming_shock_kernel(const htr_fluid f, const htr_rect bc, const ShockParams P,
                                                                   const double MaxMach, const double LengthScale) {
   Pt p; if (!cellOf(p, bc)) return;
   double *Ct = at<double>(f.Conserved_t, p);
   double RHS[nEq];
   HTR_UNROLL
   for (int i = 0; i < nEq; i++) RHS[i] = Ct[i];
   if (p.x < P.iShock) nscbcFarField<1, 1>(RHS, f, p, P.T0, P.v0, P.Xi, P.P0, MaxMach, LengthScale);
   else                nscbcFarField<1, 1>(RHS, f, p, P.T1, P.v1, P.Xi, P.P1, MaxMach, LengthScale);
   HTR_UNROLL
   for (int i = 0; i < nEq; i++) Ct[i] = RHS[i];
}

// SetNSCBC_InflowBCTask<dir> (src/prometeo_bc.cc:80-135)
template<int dir>
__global__ void __launch_bounds__(128) set_nscbc_inflow_bc_kernel(const htr_fluid f, const htr_bc_fields b, const htr_rect bc, const double Pbc) {
   Pt p; if (!cellOf(p, bc)) return;
   double Xi[nSpec], v[3];
   HTR_UNROLL
   for (int s = 0; s < nSpec; s++) Xi[s] = ldc<double>(b.MolarFracs_profile, p, s);
   HTR_UNROLL
   for (int i = 0; i < 3; i++) v[i] = ldc<double>(b.velocity_profile, p, i);
   const double T = ld<double>(b.temperature_profile, p);
   double *oX = at<double>(f.MolarFracs, p), *ov = at<double>(f.velocity, p);
   HTR_UNROLL
   for (int s = 0; s < nSpec; s++) oX[s] = Xi[s];
   HTR_UNROLL
   for (int i = 0; i < 3; i++) ov[i] = v[i];
   *at<double>(f.temperature, p) = T;
   double P = Pbc;                                            // supersonic: everything is imposed
   if (!(fabs(v[dir]) >= ld<double>(f.SoS, p))) {
      // setInflowPressure (src/prometeo_bc.hpp:124-136): from the transported density
      double rhoYi[nSpec];
      HTR_UNROLL
      for (int s = 0; s < nSpec; s++) rhoYi[s] = ldc<double>(f.Conserved, p, s);
      const double rho = mix::GetRhoFromRhoYi(rhoYi);
      const double MixW = mix::GetMolarWeightFromXi(Xi);
      P = mix::GetPFromRhoAndT(rho, MixW, T);
   }
   *at<double>(f.pressure, p) = P;
}

// SetIncomingShockBCTask (src/prometeo_bc.cc:203-264)
__global__ void __launch_bounds__(128) set_incoming_shock_bc_kernel(const htr_fluid f, const htr_rect bc, const ShockParams S) {
   Pt p; if (!cellOf(p, bc)) return;
   double Xi[nSpec], u[3], T, P;
   if (p.x < S.iShock - 1 || p.x > S.iShock + 1) {            // treated as an outflow
      double U[nEq], rho, Yi[nSpec], MixW;
      HTR_UNROLL
      for (int i = 0; i < nEq; i++) U[i] = ldc<double>(f.Conserved, p, i);
      T = *at<double>(f.temperature, p);
      primitivesFromConserved<true>(U, rho, Yi, MixW, u, T, P);
      mix::GetMolarFractions(Xi, MixW, Yi);
   } else {                                                   // the shock is injected over three cells
      const double w0 = (p.x == S.iShock - 1) ? 0.75 : ((p.x == S.iShock) ? 0.25 : 0.0);
      const double w1 = (p.x == S.iShock - 1) ? 0.25 : ((p.x == S.iShock) ? 0.75 : 1.0);
      HTR_UNROLL
      for (int s = 0; s < nSpec; s++) Xi[s] = S.Xi[s];
      if (p.x == S.iShock + 1) {
         HTR_UNROLL
         for (int i = 0; i < 3; i++) u[i] = S.v1[i];
         T = S.T1; P = S.P1;
      } else {
         HTR_UNROLL
         for (int i = 0; i < 3; i++) u[i] = w0*S.v0[i] + w1*S.v1[i];
         T = w0*S.T0 + w1*S.T1;
         P = w0*S.P0 + w1*S.P1;
      }
   }
   double *oX = at<double>(f.MolarFracs, p), *ov = at<double>(f.velocity, p);
   HTR_UNROLL
   for (int s = 0; s < nSpec; s++) oX[s] = Xi[s];
   HTR_UNROLL
   for (int i = 0; i < 3; i++) ov[i] = u[i];
   *at<double>(f.temperature, p) = T;
   *at<double>(f.pressure, p) = P;
}

} // namespace htr

extern "C" {

#define HTR_NSCBC_DISPATCH(KERNEL, ...) \
   switch (2*dir + side) { \
   case 0: KERNEL<0, 0><<<blocksFor(*bc, 128), 128, 0, st>>>(__VA_ARGS__); break; \
   case 1: KERNEL<0, 1><<<blocksFor(*bc, 128), 128, 0, st>>>(__VA_ARGS__); break; \
   case 2: KERNEL<1, 0><<<blocksFor(*bc, 128), 128, 0, st>>>(__VA_ARGS__); break; \
   case 3: KERNEL<1, 1><<<blocksFor(*bc, 128), 128, 0, st>>>(__VA_ARGS__); break; \
   case 4: KERNEL<2, 0><<<blocksFor(*bc, 128), 128, 0, st>>>(__VA_ARGS__); break; \
   case 5: KERNEL<2, 1><<<blocksFor(*bc, 128), 128, 0, st>>>(__VA_ARGS__); break; \
   }

int htr_nscbc_inflow(int dir, int side, const htr_fluid *f, const htr_bc_fields *b, const htr_rect *bc, void *stream) {
   HTR_CHECK_COMMON(f, bc)
   if (!b || dir < 0 || dir > 2 || side < 0 || side > 1 || !b->dudtBoundary.base || !b->dTdtBoundary.base) return HTR_EINVAL;
   cudaStream_t st = (cudaStream_t)stream;
   HTR_NSCBC_DISPATCH(nscbc_inflow_kernel, *f, *b, *bc)
   return cuerr(cudaGetLastError());
}

int htr_nscbc_outflow(int dir, int side, const htr_fluid *f, const htr_rect *bc, const htr_rect *bounds,
                      double MaxMach, double LengthScale, double PInf, void *stream) {
   HTR_CHECK_COMMON(f, bc)
   if (!bounds || dir < 0 || dir > 2 || side < 0 || side > 1) return HTR_EINVAL;
   cudaStream_t st = (cudaStream_t)stream;
   HTR_NSCBC_DISPATCH(nscbc_outflow_kernel, *f, *bc, *bounds, MaxMach, LengthScale, PInf)
   return cuerr(cudaGetLastError());
}

int htr_nscbc_farfield(int dir, int side, const htr_fluid *f, const htr_bc_fields *b, const htr_rect *bc,
                       double MaxMach, double LengthScale, double PInf, void *stream) {
   HTR_CHECK_COMMON(f, bc)
   if (!b || dir < 0 || dir > 2 || side < 0 || side > 1 ||
       !b->velocity_profile.base || !b->temperature_profile.base || !b->MolarFracs_profile.base) return HTR_EINVAL;
   cudaStream_t st = (cudaStream_t)stream;
   HTR_NSCBC_DISPATCH(nscbc_farfield_kernel, *f, *b, *bc, MaxMach, LengthScale, PInf)
   return cuerr(cudaGetLastError());
}

static ShockParams shockParamsOf(const double *q) {
   ShockParams P;
   P.iShock = (int)q[0];
   for (int s = 0; s < nSpec; s++) P.Xi[s] = q[1 + s];
   const double *r = q + 1 + nSpec;
   P.P0 = r[0]; P.T0 = r[1]; for (int i = 0; i < 3; i++) P.v0[i] = r[2 + i];
   P.P1 = r[5]; P.T1 = r[6]; for (int i = 0; i < 3; i++) P.v1[i] = r[7 + i];
   return P;
}

int htr_nscbc_incoming_shock(const htr_fluid *f, const htr_rect *bc, double MaxMach, double LengthScale,
                             const double *params, void *stream) {
   HTR_CHECK_COMMON(f, bc)
   if (!params) return HTR_EINVAL;
   nscbc_incoming_shock_kernel<<<blocksFor(*bc, 128), 128, 0, (cudaStream_t)stream>>>(*f, *bc, shockParamsOf(params), MaxMach, LengthScale);
   return cuerr(cudaGetLastError());
}

int htr_set_nscbc_inflow_bc(int dir, const htr_fluid *f, const htr_bc_fields *b, const htr_rect *bc, double Pbc, void *stream) {
   HTR_CHECK_COMMON(f, bc)
   if (!b || dir < 0 || dir > 2 || !b->velocity_profile.base || !b->temperature_profile.base || !b->MolarFracs_profile.base) return HTR_EINVAL;
   cudaStream_t st = (cudaStream_t)stream;
   switch (dir) {
   case 0: set_nscbc_inflow_bc_kernel<0><<<blocksFor(*bc, 128), 128, 0, st>>>(*f, *b, *bc, Pbc); break;
   case 1: set_nscbc_inflow_bc_kernel<1><<<blocksFor(*bc, 128), 128, 0, st>>>(*f, *b, *bc, Pbc); break;
   case 2: set_nscbc_inflow_bc_kernel<2><<<blocksFor(*bc, 128), 128, 0, st>>>(*f, *b, *bc, Pbc); break;
   }
   return cuerr(cudaGetLastError());
}

// the task is UpdatePrimitive over the boundary cells: the kernel of htr_update_primitive_from_conserved
int htr_set_nscbc_outflow_bc(const htr_fluid *f, const htr_rect *bc, void *stream) {
   return htr_update_primitive_from_conserved(f, bc, stream);
}

int htr_set_incoming_shock_bc(const htr_fluid *f, const htr_rect *bc, const double *params, void *stream) {
   HTR_CHECK_COMMON(f, bc)
   if (!params) return HTR_EINVAL;
   set_incoming_shock_bc_kernel<<<blocksFor(*bc, 128), 128, 0, (cudaStream_t)stream>>>(*f, *bc, shockParamsOf(params));
   return cuerr(cudaGetLastError());
}

} // extern "C"
