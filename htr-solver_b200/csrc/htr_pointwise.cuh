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
// Cell-local operators: Ducros and shock sensors, chemistry source and implicit chemistry
// integration, conserved <-> primitive <-> properties, spectral radius.
//   sensors     src/prometeo_sensor.inl:52-77, src/prometeo_sensor.cc:83-138
//   chemistry   src/prometeo_chem.hpp:53-92 (ImplicitSolver::rhs), src/prometeo_chem.cc:36-112,
//               src/Utils/math_utils.hpp:88-189 (LU), :196-441 (Rosenbrock)
//   variables   src/prometeo_variables.hpp:62-89, :123-135, :170-192
//   cfl         src/prometeo_cfl.hpp:70-177
#pragma once
#include "htr_access.cuh"
#include "htr_mix.cuh"
#include "htr_stencil.cuh"
#include "htr_diffusion.cuh"

namespace htr {

// ---- sensors -------------------------------------------------------------------------------
HTR_DEV double ducrosSensor(const htr_fluid &f, const Pt &p, const htr_rect &B, double eps) {
   double gx[3], gy[3], gz[3];
   velocityDeriv<0>(gx, f, p, B);
   velocityDeriv<1>(gy, f, p, B);
   velocityDeriv<2>(gz, f, p, B);
   const double div = gx[0] + gy[1] + gz[2];
   const double div2 = div*div;
   const double omz = gx[1] - gy[0];
   const double omy = gy[2] - gz[1];
   const double omx = gz[0] - gx[2];
   const double om2 = omx*omx + omy*omy + omz*omz;
   return div2/(div2 + om2 + eps);
}

template<int dir>
HTR_DEV bool shockSensor(const htr_fluid &f, const Pt &p, const htr_rect &B) {
   const int t = ld<int>(nTypeOf<dir>(f), p);
   Pt s[6];
   s[0] = shift<dir>(p, offM2(t), B); s[1] = shift<dir>(p, offM1(t), B); s[2] = p;
   s[3] = shift<dir>(p, offP1(t), B); s[4] = shift<dir>(p, offP2(t), B); s[5] = shift<dir>(p, offP3(t), B);
   double Phi = ld<double>(f.DucrosSensor, s[0]);
   HTR_UNROLL
   for (int k = 1; k < 6; k++) Phi = fmax(Phi, ld<double>(f.DucrosSensor, s[k]));
   bool smooth = true;
   HTR_UNROLL
   for (int h = 0; h < nSpec; h++)
      smooth = smooth && tenoASensor(ldc<double>(f.Conserved, s[0], h), ldc<double>(f.Conserved, s[1], h),
                                     ldc<double>(f.Conserved, s[2], h), ldc<double>(f.Conserved, s[3], h),
                                     ldc<double>(f.Conserved, s[4], h), ldc<double>(f.Conserved, s[5], h), t, Phi);
   return smooth;
}

// ---- variables -----------------------------------------------------------------------------
// rho, Y, u, T, p from a conserved state; T enters as the Newton guess and leaves converged
template<bool NEWTON = true>
HTR_DEV void primitivesFromConserved(const double *U, double &rho, double *Yi, double &MixW, double *u,
                                     double &T, double &P) {
   double rhoYi[nSpec];
   HTR_UNROLL
   for (int i = 0; i < nSpec; i++) rhoYi[i] = U[i];
   rho = mix::GetRhoFromRhoYi(rhoYi);
   mix::GetYi(Yi, rho, rhoYi);
   mix::ClipYi(Yi);
   MixW = mix::GetMolarWeightFromYi(Yi);
   const double rhoInv = 1.0/rho;
   HTR_UNROLL
   for (int i = 0; i < 3; i++) u[i] = U[i + irU]*rhoInv;
   const double kineticEnergy = 0.5*mod2(u);
   const double e = U[irE]*rhoInv - kineticEnergy;
   if (NEWTON) T = mix::GetTFromInternalEnergy(e, T, Yi);      // else: T is the converged value already (ghost copies)
   P = mix::GetPFromRhoAndT(rho, MixW, T);
}

// ---- chemistry -----------------------------------------------------------------------------
// One evaluation of dU/dt = w(U) + S (S frozen: the explicit RHS of the other terms).  Deliberately NOT inlined: the
// integrator calls it from 2 nEq + 4 places per step; inlined and unrolled that was 30 000 SASS instructions (480 KB,
// far beyond the instruction cache: `no_instruction` was the top stall of update_chemistry_kernel, profiles/r2_*).
__device__ __noinline__ void chemRhs(double *r, const double *x, const double *S, double *T) {
   double rho, Yi[nSpec], MixW, u[3], P;
   double Tl = *T;
   primitivesFromConserved(x, rho, Yi, MixW, u, Tl, P);
   *T = Tl;
   double w[nSpec];
   mix::GetProductionRates(w, rho, P, Tl, Yi);
   HTR_UNROLL
   for (int i = 0; i < nSpec; i++) r[i] = w[i] + S[i];
   HTR_UNROLL
   for (int i = nSpec; i < nEq; i++) r[i] = S[i];
}
struct ChemProblem {
   const double *S;
   double T;                    // running Newton guess, updated by every evaluation
   int calls;                   // rhs() evaluations: (2 nEq + 1) per step + 3 per attempt, a signature of the step sequence
   __device__ void rhs(double *r, const double *x) { calls++; chemRhs(r, x, S, &T); }
};

// In-place LU factorisation with implicit-scaling partial pivoting (Crout) and solve.
// A is column-major n x n like the reference's MyMatrix.
template<int n>
struct LU {
   double A[n*n];
   int ind[n];
   __device__ double &a(int i, int j) { return A[i + j*n]; }
   __device__ __noinline__ void factor() {
      int imax = 0;
      double vv[n];
      for (int i = 0; i < n; i++) {
         double big = 0.0;
         for (int j = 0; j < n; j++) big = fmax(fabs(a(i, j)), big);
         vv[i] = 1.0/big;
      }
      for (int j = 0; j < n; j++) {
         for (int i = 0; i < j; i++) {
            double sum = a(i, j);
            for (int k = 0; k < i; k++) sum -= a(i, k)*a(k, j);
            a(i, j) = sum;
         }
         double big = 0.0;
         for (int i = j; i < n; i++) {
            double sum = a(i, j);
            for (int k = 0; k < j; k++) sum -= a(i, k)*a(k, j);
            a(i, j) = sum;
            const double dum = vv[i]*fabs(sum);
            if (dum >= big) { big = dum; imax = i; }
         }
         if (j != imax) {
            for (int k = 0; k < n; k++) { const double dum = a(imax, k); a(imax, k) = a(j, k); a(j, k) = dum; }
            vv[imax] = vv[j];
         }
         ind[j] = imax;
         if (a(j, j) == 0.0) a(j, j) = 1.0e-20;
         if (j != n - 1) {
            const double dum = 1.0/a(j, j);
            for (int i = j + 1; i < n; i++) a(i, j) *= dum;
         }
      }
   }
   __device__ __noinline__ void solve(double *b) {
      int ii = 0;
      for (int i = 0; i < n; i++) {
         const int ip = ind[i];
         double sum = b[ip];
         b[ip] = b[i];
         if (ii != 0) { for (int j = ii - 1; j < i; j++) sum -= a(i, j)*b[j]; }
         else if (sum != 0.0) ii = i + 1;
         b[i] = sum;
      }
      for (int i = n - 1; i > -1; i--) {
         double sum = b[i];
         for (int j = i + 1; j < n; j++) sum -= a(i, j)*b[j];
         b[i] = sum/a(i, i);
      }
   }
};

// Adaptive 4-stage Kaps-Rentrop Rosenbrock integration of `prob` over DelT, central-difference Jacobian.
// Returns false if the step or iteration caps were hit (the reference asserts).
template<int n, class Problem>
__device__ bool rosenbrockSolve(Problem &prob, double *x, double dtTry, double DelT) {
   constexpr int MAXSTP = 100000, MAXITS = 100;
   constexpr double TOL = 1e-10, iTOL = 1.0/TOL, SAFETY = 0.9, GROW = 1.5, PGROW = -0.25, SHRNK = 0.5,
      PSHRNK = -1.0/3.0, ERRCON = 0.1296, GAM = 1.0/2.0, A21 = 2.0, A31 = 48.0/25.0, A32 = 6.0/25.0,
      C21 = -8.0, C31 = 372.0/25.0, C32 = 12.0/5.0, C41 = -112.0/125.0, C42 = -54.0/125.0, C43 = -2.0/5.0,
      B1 = 19.0/9.0, B2 = 1.0/2.0, B3 = 25.0/108.0, B4 = 125.0/108.0, E1 = 17.0/54.0, E2 = 7.0/36.0, E3 = 0.0,
      E4 = 125.0/108.0, C1X = 1.0/2.0, C2X = -3.0/2.0, C3X = 121.0/50.0, C4X = 29.0/250.0;
   double Jac[n*n];
   LU<n> lu;
   double dx[n], xsav[n], dxsav[n], err[n];
   bool finish = false;
   double time = 0.0;
   double dt = dtTry;
   #pragma unroll 1
   for (int s = 0; s < MAXSTP; s++) {
      const double t0 = time;
      double dtNext = 0.0;
      {  // Jacobian by central differences, step x_j*1e-6 + 1e-14
         double tmp[n], fp[n], fm[n];
         for (int i = 0; i < n; i++) tmp[i] = x[i];
         #pragma unroll 1
         for (int j = 0; j < n; j++) {
            const double h = x[j]*1.0e-6 + 1.0e-14;
            x[j] = tmp[j] + h;
            const double hp = x[j] - tmp[j];
            prob.rhs(fp, x);
            x[j] = tmp[j] - h;
            const double hm = tmp[j] - x[j];
            prob.rhs(fm, x);
            x[j] = tmp[j];
            for (int i = 0; i < n; i++) Jac[i + j*n] = (fp[i] - fm[i])/(hp + hm);
         }
      }
      prob.rhs(dx, x);
      for (int i = 0; i < n; i++) { xsav[i] = x[i]; dxsav[i] = dx[i]; }
      bool converged = false;
      #pragma unroll 1
      for (int jtry = 0; jtry < MAXITS; jtry++) {
         double g1[n], g2[n], g3[n], g4[n];
         for (int i = 0; i < n; i++) {
            for (int j = 0; j < n; j++) lu.a(i, j) = -Jac[i + j*n];
            lu.a(i, i) += 1.0/(GAM*dt);
         }
         lu.factor();
         for (int i = 0; i < n; i++) g1[i] = dxsav[i] + dt*C1X*dx[i];
         lu.solve(g1);
         for (int i = 0; i < n; i++) x[i] = xsav[i] + A21*g1[i];
         prob.rhs(dx, x);
         for (int i = 0; i < n; i++) g2[i] = dx[i] + dt*C2X*dx[i] + C21*g1[i]/dt;
         lu.solve(g2);
         for (int i = 0; i < n; i++) x[i] = xsav[i] + A31*g1[i] + A32*g2[i];
         prob.rhs(dx, x);
         for (int i = 0; i < n; i++) g3[i] = dx[i] + dt*C3X*dx[i] + (C31*g1[i] + C32*g2[i])/dt;
         lu.solve(g3);
         for (int i = 0; i < n; i++) g4[i] = dx[i] + dt*C4X*dx[i] + (C41*g1[i] + C42*g2[i] + C43*g3[i])/dt;
         lu.solve(g4);
         for (int i = 0; i < n; i++) x[i] = xsav[i] + B1*g1[i] + B2*g2[i] + B3*g3[i] + B4*g4[i];
         for (int i = 0; i < n; i++) err[i] = E1*g1[i] + E2*g2[i] + E3*g3[i] + E4*g4[i];
         time = t0 + dt;
         // error control
         double errmax = 0.0;
         for (int i = 0; i < n; i++) errmax = fmax(errmax, fabs(err[i]));
         errmax *= iTOL;
         if ((errmax < 1.0) || finish) {
            dtNext = (errmax > ERRCON) ? SAFETY*dt*pow(errmax, PGROW) : GROW*dt;
            converged = true;
            break;
         }
         dtNext = SAFETY*dt*pow(errmax, PSHRNK);
         dt = (dt > 0.0) ? fmax(dtNext, SHRNK*dt) : fmin(dtNext, SHRNK*dt);
      }
      if (!converged) return false;
      const double Remainder = DelT - time;
      if (finish || (Remainder < 1e-12*DelT)) return true;
      if (dtNext*1.5 > Remainder) { dt = Remainder; finish = true; }
      else { dt = dtNext; finish = false; }
   }
   return false;
}

// ---- CFL -----------------------------------------------------------------------------------
HTR_DEV double spectralRadius(const htr_fluid &f, const Pt &p) {
   const double dx = ld<double>(f.dcsi_d, p), dy = ld<double>(f.deta_d, p), dz = ld<double>(f.dzet_d, p);
   const double a = ld<double>(f.SoS, p), rho = ld<double>(f.rho, p);
   double r = (fabs(ldc<double>(f.velocity, p, 0)) + a)*dx;
   r = fmax(r, (fabs(ldc<double>(f.velocity, p, 1)) + a)*dy);
   r = fmax(r, (fabs(ldc<double>(f.velocity, p, 2)) + a)*dz);
   {
      const double nu = ld<double>(f.mu, p)/rho;
      double v = nu*dx*dx;
      v = fmax(v, nu*dy*dy);
      v = fmax(v, nu*dz*dz);
      v *= 4;
      r = fmax(r, v);
   }
   {
      double Yi[nSpec];
      HTR_UNROLL
      for (int i = 0; i < nSpec; i++) Yi[i] = ldc<double>(f.MassFracs, p, i);
      const double cp = mix::GetHeatCapacity(ld<double>(f.temperature, p), Yi);
      const double DifT = ld<double>(f.lam, p)/(cp*rho);
      double v = DifT*dx*dx;
      v = fmax(v, DifT*dy*dy);
      v = fmax(v, DifT*dz*dz);
      v *= 4;
      r = fmax(r, v);
   }
   {
      double v = 0.0;
      HTR_UNROLL
      for (int s = 0; s < nSpec; s++) {
         const double D = ldc<double>(f.Di, p, s);
         v = fmax(v, D*dx*dx);
         v = fmax(v, D*dy*dy);
         v = fmax(v, D*dz*dz);
      }
      v *= 4;
      r = fmax(r, v);
   }
   return r;
}

} // namespace htr
