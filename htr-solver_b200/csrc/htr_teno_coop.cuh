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
// Lane-cooperative characteristic TENO flux: EIGHT LANES PER FACE instead of one thread per face.
//
// The per-face form (htr_euler.cuh: tenoFaceFlux) keeps the six stencil cells and their 12 split characteristic flux
// vectors in one thread: 255 registers, one or two CTAs per SM, and a chain of ~2500 dependent fp64 operations per
// face -- pure TENO ran at 0.52 G cell-updates/s (2 % of the roofline, profiles/r2_v18_bench_tg128_teno.json), and the
// register cap does not help (128 / 80 registers: 4.02 / 4.64 ms against 4.14 ms, profiles/r2_v19_*).  Here the work
// of one face is spread over a group of eight lanes:
//   1. lane k < 6 loads stencil cell k (Conserved, normal velocity, pressure, speed of sound); the three maximum
//      wave speeds of the stencil are warp-shuffle reductions over the group;
//   2. the Roe state is formed by every lane of the group (L|R conserved states by shuffle from lanes 2 and 3);
//   3. lane k projects ITS cell onto the characteristic fields (the two L*q products and the Lax-Friedrichs split of
//      src/prometeo_rhs.inl:625-700) and leaves the 2 nEq values in the group's shared-memory scratch;
//   4. the 2 nEq one-sided reconstructions (plus and minus side of every field) are dealt to the eight lanes round-robin;
//      the upwind side is a run-time flag (tenoReconstructRT, htr_stencil.cuh) so that lanes reconstructing plus and minus
//      fields do not diverge;
//   5. every lane maps the reconstructed characteristic flux back (K*q) and applies the positivity limiter of the
//      multi-species case, so the whole group holds the face flux and lane e adds component e to Conserved_t.
// Every value is produced by the same device functions, with the same operands in the same order, as in tenoFaceFlux:
// the strict build is bit-identical with the per-face kernels (tests/test_parity_gpu.py::test_coop_teno_*).
// MEASURED (B200, profiles/r2_v21_*): correct and bit-identical, but 2-5x SLOWER than the per-face kernels (128^3 ConstPropMix
// TENO: 8.35 vs 4.14 ms per RHS; air256_front: 29.6 vs 10.6 ms), so it is an option (htr_set_option "coop" = 1), not the default.
// The per-face kernels issue on ~17 % of the cycles, which looked like a latency problem that more, lighter warps would fix;
// but spreading a face over lanes multiplies the instructions: the Roe state (a Newton solve for a multi-component gas), the
// back-projection, the limiter and all L|R loads are repeated by eight lanes, the skew-symmetric flux of a smooth face of a
// deferred Hybrid cell is evaluated eight times, 2 nEq = 10 reconstructions fill two rounds of eight lanes (18 fill three),
// and six of eight lanes own a stencil cell: ~2.5x the warp-instructions per face at the same issue rate.
// A group walks a RUN of cells along the flux direction (the plus face of a cell is the minus face of the next), so a
// face is evaluated once; the deferred cells of the Hybrid scheme are runs of one cell taken from the compacted list.
#pragma once

namespace htr {
namespace coop {

constexpr int GL = 8;                        // lanes per face
constexpr int GROUPS = 16;                   // groups per CTA (128 threads)
constexpr int SMD = 14*nEq;                  // doubles of scratch per group: Fp | Fm of the six cells (12 nEq), 2 nEq reconstructed values

HTR_DEV double shflD(const unsigned mask, const double v, const int src) { return __shfl_sync(mask, v, src); }

template<int dir, int OP>
HTR_DEV void tenoFaceFluxCoop(double *Flux, const htr_fluid &f, const Pt &p, const int t, const htr_rect &B,
                              const double RK0, const double RK1, const double lim_f,
                              double *sm, const int g, const unsigned gmask) {
   const int l0 = (threadIdx.x & 31) & ~(GL - 1);          // first lane of the group
   // 1. own stencil cell (lanes 6, 7 shadow cell 5: they take part in the reductions and in step 4)
   const int k = g < 6 ? g : 5;
   const int off = (k == 0) ? offM2(t) : (k == 1) ? offM1(t) : (k == 2) ? 0 : (k == 3) ? offP1(t) : (k == 4) ? offP2(t) : offP3(t);
   const Pt sk = shift<dir>(p, off, B);
   CellState c;
   loadCell<dir>(c, f, sk);
   const double a = ld<double>(f.SoS, sk);
   double Lam1 = fabs(c.un - a), Lam = fabs(c.un), LamN = fabs(c.un + a);
   HTR_UNROLL
   for (int d = 1; d < GL; d <<= 1) {
      Lam1 = fmax(Lam1, __shfl_xor_sync(gmask, Lam1, d));
      Lam  = fmax(Lam,  __shfl_xor_sync(gmask, Lam, d));
      LamN = fmax(LamN, __shfl_xor_sync(gmask, LamN, d));
   }

   // 2. Roe state of the face (cells 2 | 3)
   double U2[nEq], U3[nEq];
   HTR_UNROLL
   for (int i = 0; i < nEq; i++) { U2[i] = shflD(gmask, c.U[i], l0 + 2); U3[i] = shflD(gmask, c.U[i], l0 + 3); }
   const double p2 = shflD(gmask, c.p, l0 + 2), p3 = shflD(gmask, c.p, l0 + 3);
   const Pt s3 = shift<dir>(p, offP1(t), B);
   RoeState R;
   {
      double YL[nSpec], YR[nSpec], uL[3], uR[3];
      HTR_UNROLL
      for (int i = 0; i < nSpec; i++) { YL[i] = ldc<double>(f.MassFracs, p, i); YR[i] = ldc<double>(f.MassFracs, s3, i); }
      HTR_UNROLL
      for (int i = 0; i < 3; i++) { uL[i] = ldc<double>(f.velocity, p, i); uR[i] = ldc<double>(f.velocity, s3, i); }
      roeAverage(R, U2, U3, YL, YR, p2, p3, uL, uR, ld<double>(f.rho, p), ld<double>(f.rho, s3));
   }

   // 3. split characteristic fluxes of the own cell -> scratch [cell][plus | minus][field]
   {
      double Fp[nEq], Fm[nEq];
      splitFlux<dir>(Fp, Fm, R, c, Lam1, Lam, LamN);
      if (g < 6) {
         HTR_UNROLL
         for (int i = 0; i < nEq; i++) { sm[(2*g)*nEq + i] = Fp[i]; sm[(2*g + 1)*nEq + i] = Fm[i]; }
      }
   }
   __syncwarp(gmask);

   // 4. the 2 nEq reconstructions, round-robin over the lanes
   double *res = sm + 12*nEq;
   for (int r = g; r < 2*nEq; r += GL) {
      const bool plus = r < nEq;
      const double *v = sm + r;                            // plus field i: [k][0][i];  minus field i: [k][1][i] = + nEq + i = r
      res[r] = tenoReconstructRT<OP>(plus, v[0], v[2*nEq], v[4*nEq], v[6*nEq], v[8*nEq], v[10*nEq], t);
   }
   __syncwarp(gmask);

   // 5. back to the conserved variables
   double F[nEq];
   HTR_UNROLL
   for (int i = 0; i < nEq; i++) F[i] = -0.5*(res[i] + res[nEq + i]);
   fromCharacteristic<dir>(Flux, F, R);

   if (nSpec > 1) {
      // positivity of the partial densities after this stage (tenoFaceFlux, src/prometeo_rhs.inl:710-757)
      constexpr double eps_l = -1e-8;
      bool limit = false;
      HTR_UNROLL
      for (int i = 0; i < nSpec; i++) {
         const double dRhoi = lim_f*Flux[i];
         if (((RK0*ldc<double>(f.Conserved_old, p, i) + RK1*U2[i] + dRhoi) < eps_l) ||
             ((RK0*ldc<double>(f.Conserved_old, s3, i) + RK1*U3[i] - dRhoi) < eps_l))
            limit = true;
      }
      if (limit) {
         constexpr double eps = 1e-60;
         double theta = 1.0;
         HTR_UNROLL
         for (int i = 0; i < nEq; i++) F[i] = -0.5*(sm[4*nEq + i] + sm[7*nEq + i]);      // Fp of cell 2, Fm of cell 3
         double FLO[nEq];
         fromCharacteristic<dir>(FLO, F, R);
         for (int i = 0; i < nSpec; i++) {
            const double dRhoi = lim_f*Flux[i];
            const double dRhoi_lo = lim_f*FLO[i];
            double theta_p = 1.0;
            const double rhoi_p = RK0*ldc<double>(f.Conserved_old, p, i) + RK1*U2[i];
            const double rhoHO_p = rhoi_p + dRhoi;
            if (rhoHO_p < eps) {
               const double rhoLO = rhoi_p + dRhoi_lo;
               theta_p = (fabs(rhoHO_p - rhoLO) > eps) ? fmax(0.0, (eps - rhoLO)/(rhoHO_p - rhoLO)) : 1.0;
            }
            double theta_pp1 = 1.0;
            const double rhoi_pp1 = RK0*ldc<double>(f.Conserved_old, s3, i) + RK1*U3[i];
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
   __syncwarp(gmask);                                       // the scratch is free for the next face
}

// eulerFace (htr_kernels.cu) for a group: the skew-symmetric flux of a smooth Hybrid face is evaluated by every lane
template<int dir, int SCHEME>
HTR_DEV void eulerFaceCoop(double *Flux, const htr_fluid &f, const Pt &p, const int t, const htr_rect &B, const EulerParams &ep,
                           double *sm, const int g, const unsigned gmask) {
   if (SCHEME == HTR_SCHEME_SKEWSYMMETRIC) { kgFaceFlux<dir>(Flux, f, p, t, B); return; }
   double lim_f = 0.0;
   if (nSpec > 1)
      lim_f = ep.dt*(ld<double>(m_eOf<dir>(f), p) + ld<double>(m_eOf<dir>(f), shift<dir>(p, offP1(t), B)));
   if (SCHEME == HTR_SCHEME_HYBRID) {
      const htr_acc &ss = sensorOf<dir>(f);
      const bool smooth = ld<uint8_t>(ss, shift<dir>(p, offM1(t), B)) && ld<uint8_t>(ss, p) &&
                          ld<uint8_t>(ss, shift<dir>(p, offP1(t), B));
      if (smooth) kgFaceFlux<dir>(Flux, f, p, t, B);
      else        tenoFaceFluxCoop<dir, OP_TENO>(Flux, f, p, t, B, ep.RK0, ep.RK1, lim_f, sm, g, gmask);
   }
   else if (SCHEME == HTR_SCHEME_TENO)    tenoFaceFluxCoop<dir, OP_TENO>(Flux, f, p, t, B, ep.RK0, ep.RK1, lim_f, sm, g, gmask);
   else if (SCHEME == HTR_SCHEME_TENOA)   tenoFaceFluxCoop<dir, OP_TENOA>(Flux, f, p, t, B, ep.RK0, ep.RK1, lim_f, sm, g, gmask);
   else if (SCHEME == HTR_SCHEME_TENOLAD) tenoFaceFluxCoop<dir, OP_TENOLAD>(Flux, f, p, t, B, ep.RK0, ep.RK1, lim_f, sm, g, gmask);
}

// Conserved_t += m_e*(F[c+1/2] - F[c-1/2]) along `dir` for runs of cells, one group of eight lanes per run.
//   LISTED: the runs are the single cells of the compacted list of deferred cells (htr_kernels.cu: compact_mask_kernel);
//   else:   the tile is cut into runs of `runLen` cells along `dir`; runs are numbered with x fastest among the
//           transverse directions so that the groups of a warp read neighbouring cells.
template<int dir, int SCHEME, bool LISTED>
__global__ void __launch_bounds__(GL*GROUPS)
euler_coop_kernel(const htr_fluid f, const htr_rect tile, const htr_rect B, const EulerParams ep,
                  const int *__restrict__ list, const int *__restrict__ count, const int runLen) {
   __shared__ double smAll[GROUPS*SMD];
   const int g = threadIdx.x & (GL - 1), grp = threadIdx.x/GL;
   const unsigned gmask = 0xFFu << ((threadIdx.x & 31) & ~(GL - 1));
   double *sm = smAll + grp*SMD;
   const int64_t n0 = tile.hi[0] - tile.lo[0] + 1, n1 = tile.hi[1] - tile.lo[1] + 1, n2 = tile.hi[2] - tile.lo[2] + 1;
   const int64_t nd = (dir == 0) ? n0 : ((dir == 1) ? n1 : n2);
   const int64_t runsD = LISTED ? 1 : (nd + runLen - 1)/runLen;
   const int64_t total = LISTED ? (int64_t)*count : runsD*(n0*n1*n2/nd);
   for (int64_t r = (int64_t)blockIdx.x*GROUPS + grp; r < total; r += (int64_t)gridDim.x*GROUPS) {
      Pt c;
      int len = 1;
      if (LISTED) {
         const int64_t idx = list[r];
         c.x = tile.lo[0] + idx % n0; c.y = tile.lo[1] + (idx/n0) % n1; c.z = tile.lo[2] + idx/(n0*n1);
      } else {
         // r = (a, run, b): a = fastest transverse index, b = the other one
         int64_t a, b, run;
         if (dir == 0)      { run = r % runsD; a = (r/runsD) % n1; b = r/(runsD*n1);  c.x = tile.lo[0] + run*runLen; c.y = tile.lo[1] + a; c.z = tile.lo[2] + b; }
         else if (dir == 1) { a = r % n0; run = (r/n0) % runsD; b = r/(n0*runsD);     c.x = tile.lo[0] + a; c.y = tile.lo[1] + run*runLen; c.z = tile.lo[2] + b; }
         else               { a = r % n0; b = (r/n0) % n1; run = r/(n0*n1);           c.x = tile.lo[0] + a; c.y = tile.lo[1] + b; c.z = tile.lo[2] + run*runLen; }
         const int64_t left = nd - run*runLen;
         len = (int)(left < runLen ? left : runLen);
      }
      double Fm[nEq], Fp[nEq];
      {
         const int tc = ld<int>(nTypeOf<dir>(f), c);
         const Pt pm = shift<dir>(c, offM1(tc), B);            // owner of the run's first minus face
         eulerFaceCoop<dir, SCHEME>(Fm, f, pm, ld<int>(nTypeOf<dir>(f), pm), B, ep, sm, g, gmask);
      }
      for (int j = 0; j < len; j++) {
         eulerFaceCoop<dir, SCHEME>(Fp, f, c, ld<int>(nTypeOf<dir>(f), c), B, ep, sm, g, gmask);
         const double m = ld<double>(m_eOf<dir>(f), c);
         double *Ct = at<double>(f.Conserved_t, c);
         HTR_UNROLL
         for (int i = 0; i < nEq; i++) {
            if ((i & (GL - 1)) == g) Ct[i] += m*(Fp[i] - Fm[i]);
            Fm[i] = Fp[i];
         }
         ((dir == 0) ? c.x : ((dir == 1) ? c.y : c.z)) += 1;
      }
   }
}

} // namespace coop
} // namespace htr
