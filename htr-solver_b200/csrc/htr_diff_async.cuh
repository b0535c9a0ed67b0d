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
// Diffusion flux, staged with asynchronous copies:   diffusion_async_kernel<2,TJ>  Conserved_t += Dz
//                                                     diffusion_async_kernel<1,TJ>  Conserved_t += Dy; += Dx
// Same tiles, same arithmetic, same deferral rules and the same result bits as fast::diffusion_staged_kernel
// (htr_fast.cuh; reference: src/prometeo_rhs.inl:1058-1215, order of src/prometeo.rg:560-578); what changes is how a
// tile reaches shared memory.  ncu of the cooperative form (profiles/r2_v10_hot_source_lines.txt) put ~40 % of the
// stall samples of the diffusion kernels on its staging loop: a thread loaded one component of one cell at a time
// (5 requests of 10 cache lines each for a 5-component row that is one contiguous span), waited for it and stored it.
// Here
//   * every raw operand (rho, mu, lam, T, Xi, Di, rho*Yi, velocity) goes global -> shared with 8-byte `cp.async`
//     (LDGSTS in the SASS): no register, no wait per value, all copies of a thread in flight together;
//   * AoS fields are copied row-wise: consecutive lanes take consecutive doubles of the row (a 33-cell row of a
//     5-component field is 165 consecutive doubles = 11 cache lines for 6 requests, was 5 x 10) and keep their
//     component order in shared memory (odd strides 3 / 5 are conflict-free for the face loads);
//   * the in-plane velocity derivatives come from the staged velocity box (one cell wider on each side) instead of
//     8 more global loads per staged cell; only the out-of-plane neighbours (6 or 4 values per staged cell), the
//     metrics and the node types are register loads, issued before the copies so they overlap them.
// Addressing: the uniform view of the fields (march::UF, host: makeUniform -- every field x-dense with one common cell
// pitch): a cell is ONE 32-bit offset n = x + y*SY + z*SZ shared by all fields, rows and the out-of-plane neighbours are
// CTA-uniform offsets, and the periodic wrap in x is a CTA-uniform branch (only the tiles on the seam pay for it).  The
// first version kept the per-field byte strides (three 64-bit multiply-adds per copy) and executed 29 % MORE instructions
// than the cooperative kernel for 3 % less time (profiles/r2_v14_*); a layout that is not uniform takes the cooperative
// kernel.
#pragma once
#include "htr_march.cuh"      // march::UF, march::wrap1

namespace htr {
namespace fast {

HTR_DEV void cpAsync8(double *dst, const double *src) {
   asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" :: "r"((unsigned)__cvta_generic_to_shared(dst)), "l"(src) : "memory");
}
HTR_DEV void cpAsyncWaitAll() { asm volatile("cp.async.wait_all;" ::: "memory"); }

template<int JD, int TJ> struct ATile {
   static constexpr bool DOX = (JD == 1);
   static constexpr int NG = DOX ? 7 : 4;
   static constexpr int W = DOX ? 33 : 32, ROWS = TJ + 1, NC = W*ROWS, NT = 32*TJ;
   // velocity box: one more column on each side; the xy kernel also one more row on each side
   static constexpr int VR0 = DOX ? 1 : 0, VW = W + 2, VR = ROWS + 2*VR0, VL = 3*VW;
   // offsets in doubles.  Scalars and derivatives: [var][cell]; Xi, Di, rho*Yi: [cell][SP]; velocity: [cell][3]
   static constexpr int SP = nSpec | 1;               // species stride in shared memory: odd, so the face loads are conflict-free
   static constexpr int O_RHO = 0, O_MU = NC, O_LAM = 2*NC, O_T = 3*NC, O_MW = 4*NC, O_X = 5*NC, O_DI = O_X + SP*NC,
                        O_RY = O_DI + SP*NC, O_G = O_RY + SP*NC, O_V = O_G + NG*NC, TOTAL = O_V + VL*VR;
   static constexpr size_t SMEM = sizeof(double)*TOTAL;
   static_assert(TOTAL >= nEq*NT, "flux exchange reuses the staging buffer");
   static_assert(NC <= 2*NT, "the derivative pass covers the staged box with two cells per thread");
};

// Face between staged cells ip and iq (velocity-box cells vp, vq): the operations of diffFaceStd (htr_fast.cuh), operand
// for operand, on the layout of ATile.
// x/3 of the normal stress: the production build multiplies by the rounded constant 1/3 (1 ulp, a 20-instruction division
// sequence less per face); the strict build divides like the reference
#ifdef HTR_EXACT_POW
#define HTR_THIRD(x) ((x)/3)
#else
#define HTR_THIRD(x) ((x)*(1.0/3.0))
#endif

template<int dir, class T>
HTR_DEV void diffFaceA(double *Flux, const double *s, const int ip, const int iq, const int vp, const int vq, const double m_s,
                       const int gA, const int gB, const int gC, const int gD) {
   constexpr int NC = T::NC;
   const double rho_s = interpStd(s[T::O_RHO + ip], s[T::O_RHO + iq]);
   const double iMixW_s = 1.0/interpStd(s[T::O_MW + ip], s[T::O_MW + iq]);
   const double Tp = s[T::O_T + ip], Tq = s[T::O_T + iq];
   const double T_s = interpStd(Tp, Tq);
   double heatFlux = interpStd(s[T::O_LAM + ip], s[T::O_LAM + iq])*m_s*(Tp - Tq);

   double hi[nSpec];
   mix::GetSpeciesEnthalpies(hi, T_s);

   double ViCorr = 0.0;
   HTR_UNROLL
   for (int i = 0; i < nSpec; i++) {
      const double YiVi = interpStd(s[T::O_DI + ip*T::SP + i], s[T::O_DI + iq*T::SP + i])*
                          m_s*(s[T::O_X + ip*T::SP + i] - s[T::O_X + iq*T::SP + i])*mix::GetSpeciesMolarWeight(i)*iMixW_s;
      Flux[i] = -rho_s*YiVi;
      heatFlux += rho_s*YiVi*hi[i];
      ViCorr += YiVi;
   }
   HTR_UNROLL
   for (int i = 0; i < nSpec; i++) {
      const double rhoYiViCorr = interpStd(s[T::O_RY + ip*T::SP + i], s[T::O_RY + iq*T::SP + i])*ViCorr;
      Flux[i] += rhoYiViCorr;
      heatFlux -= rhoYiViCorr*hi[i];
   }

   double up[3], uq[3], dN[3];
   HTR_UNROLL
   for (int i = 0; i < 3; i++) { up[i] = s[T::O_V + 3*vp + i]; uq[i] = s[T::O_V + 3*vq + i]; dN[i] = m_s*(uq[i] - up[i]); }
   const double mu_s = interpStd(s[T::O_MU + ip], s[T::O_MU + iq]);
   const double A = interpStd(s[T::O_G + gA*NC + ip], s[T::O_G + gA*NC + iq]);
   const double Bv = interpStd(s[T::O_G + gB*NC + ip], s[T::O_G + gB*NC + iq]);
   const double C = interpStd(s[T::O_G + gC*NC + ip], s[T::O_G + gC*NC + iq]);
   const double D = interpStd(s[T::O_G + gD*NC + ip], s[T::O_G + gD*NC + iq]);
   double sigma[3];
   if (dir == 0) {
      sigma[0] = HTR_THIRD(mu_s*(4*dN[0] - 2*A - 2*D));
      sigma[1] = mu_s*(dN[1] + Bv);
      sigma[2] = mu_s*(dN[2] + C);
   } else if (dir == 1) {
      sigma[0] = mu_s*(dN[0] + Bv);
      sigma[1] = HTR_THIRD(mu_s*(4*dN[1] - 2*A - 2*D));
      sigma[2] = mu_s*(dN[2] + C);
   } else {
      sigma[0] = mu_s*(dN[0] + Bv);
      sigma[1] = mu_s*(dN[1] + D);
      sigma[2] = HTR_THIRD(mu_s*(4*dN[2] - 2*A - 2*C));
   }
   HTR_UNROLL
   for (int i = 0; i < 3; i++) Flux[irU + i] = sigma[i];
   double uSigma = 0.0;
   HTR_UNROLL
   for (int i = 0; i < 3; i++) uSigma += interpStd(up[i], uq[i])*sigma[i];
   Flux[irE] = (uSigma - heatFlux);
}

// What the derivative pass needs from global memory for one staged cell: the out-of-plane velocity neighbours, the
// metrics of the differentiated directions and whether their node types are standard.
template<int JD> struct GPre {
   static constexpr int NO = (JD == 1) ? 3 : 2;      // out-of-plane components: (U,V,W) at z-1/z+1 or (V,W) at y-1/y+1
   double om[NO], op[NO], mx, mo, my;                // mo: metric of the out-of-plane direction; my: deta_d (xy kernel only)
   bool bad, live;
};

// n: the cell; nm, np: its out-of-plane neighbours (wrapped)
template<int JD>
HTR_DEV void loadGPre(GPre<JD> &g, const march::UF &u, const int n, const int nm, const int np) {
   constexpr int OD = (JD == 1) ? 2 : 1;             // out-of-plane direction
   constexpr int C0 = 3 - GPre<JD>::NO;
   const double *va = u.vel + (int64_t)nm*3 + C0, *vb = u.vel + (int64_t)np*3 + C0;
   HTR_UNROLL
   for (int i = 0; i < GPre<JD>::NO; i++) { g.om[i] = __ldg(va + i); g.op[i] = __ldg(vb + i); }
   g.mx = __ldg(u.md[0] + n);
   g.mo = __ldg(u.md[OD] + n);
   g.my = (JD == 1) ? __ldg(u.md[1] + n) : 0.0;
   const int tx = __ldg(u.nT[0] + n), ty = __ldg(u.nT[1] + n), tz = (JD == 1) ? __ldg(u.nT[2] + n) : (int)Std_node;
   g.bad = (tx != Std_node) | (ty != Std_node) | (tz != Std_node);      // independent loads (no short circuit)
}

// Derivatives and molar weight of staged cell n from the copied tile (derivative slots as in stageDiffPart:
// JD = 1: dUdX dVdX dUdY dVdY dUdZ dVdZ dWdZ;  JD = 2: dUdX dWdX dVdY dWdY).
template<int JD, class T>
HTR_DEV void derivPass(double *s, const int n, const GPre<JD> &g) {
   constexpr int NC = T::NC;
   const int rr = n/T::W, cc = n - rr*T::W;
   const int v = 3*((rr + T::VR0)*T::VW + cc + 1);
   const double *sV = s + T::O_V;
   double u[3];
   HTR_UNROLL
   for (int i = 0; i < 3; i++) u[i] = sV[v + i];
   constexpr int c1 = (JD == 1) ? 1 : 2;
   s[T::O_G + 0*NC + n] = derivStd(sV[v - 3], u[0], sV[v + 3], g.mx);
   s[T::O_G + 1*NC + n] = derivStd(sV[v - 3 + c1], u[c1], sV[v + 3 + c1], g.mx);
   if (JD == 1) {
      s[T::O_G + 2*NC + n] = derivStd(sV[v - T::VL], u[0], sV[v + T::VL], g.my);
      s[T::O_G + 3*NC + n] = derivStd(sV[v - T::VL + 1], u[1], sV[v + T::VL + 1], g.my);
      HTR_UNROLL
      for (int i = 0; i < 3; i++) s[T::O_G + (4 + i)*NC + n] = derivStd(g.om[i], u[i], g.op[i], g.mo);
   } else {
      s[T::O_G + 2*NC + n] = derivStd(g.om[0], u[1], g.op[0], g.mo);
      s[T::O_G + 3*NC + n] = derivStd(g.om[1], u[2], g.op[1], g.mo);
   }
   double Xi[nSpec];
   HTR_UNROLL
   for (int i = 0; i < nSpec; i++) Xi[i] = s[T::O_X + n*T::SP + i];
   s[T::O_MW + n] = mix::GetMolarWeightFromXi(Xi);
}

template<int JD, int TJ>
__global__ void __launch_bounds__(32*TJ, HTR_D_MINB)
diffusion_async_kernel(const march::UF u, uint8_t *__restrict__ mask, int *__restrict__ flag) {
   using T = ATile<JD, TJ>;
   using march::wrap1;
   constexpr bool DOX = T::DOX;
   constexpr int W = T::W, NC = T::NC, NT = T::NT, ROWS = T::ROWS;
   constexpr int KD = (JD == 1) ? 2 : 1;              // the direction the grid's third index runs along (= out of plane)
   extern __shared__ double smem[];
   double *s = smem;

   const int i = threadIdx.x, j = threadIdx.y;
   const int tid = j*32 + i;
   // tile-relative coordinates: the tile is [0, n-1]
   const int cx0 = DOX ? ((int)blockIdx.x*31 - 1) : (int)blockIdx.x*32;
   const int cJ0 = (int)blockIdx.y*(TJ - 1) - 1;
   const int hiX = u.n[0] - 1, hiJ = u.n[JD] - 1;
   const int cX = cx0 + i, cJ = cJ0 + j;
   const bool inTile = cX <= hiX && cJ <= hiJ;
   const bool upd = inTile && j >= 1 && (!DOX || i >= 1);
   const int xMax = hiX + 1, JMax = hiJ + 1;          // last staged column / row of the box
   const int SJ = (JD == 1) ? u.SY : u.SZ, SK = (JD == 1) ? u.SZ : u.SY;
   // (one plane per CTA: a loop over several planes of the tile made the compiler hoist the index setup out of it and
   //  spill it -- 248 bytes of stack, diffusion y+x 21.3 -> 25.8 ms on B200, profiles/r2_v16_*; prefetching the rows of the
   //  tile of the CTA 444 / 888 / 1776 launches later into L2 changed nothing either: 21.1 / 22.3 / 22.9 ms, r2_v17_*)
   const int cK = blockIdx.z;
   const int kOff = cK*SK;
   // CTA-uniform: does a column of the (velocity) box leave the global box in x?  Only then x is wrapped per element.
   const bool xSeam = (cx0 - 1 < u.blo[0]) || (cx0 + W > u.bhi[0]);
   const int dKm = (wrap1(cK - 1, u.blo[KD], u.bhi[KD]) - cK)*SK, dKp = (wrap1(cK + 1, u.blo[KD], u.bhi[KD]) - cK)*SK;
   auto rowOff = [&](const int J) { return wrap1(J, u.blo[JD], u.bhi[JD])*SJ + kOff; };
   auto colOf = [&](const int x) { return xSeam ? wrap1(x, u.blo[0], u.bhi[0]) : x; };

   // ---- register loads of the derivative pass (cells tid and tid + NT of the staged box), issued first ----
   GPre<JD> g0, g1;
   {
      const int rr = tid/W, cc = tid - rr*W;                      // tid < NT <= NC
      g0.live = (cx0 + cc <= xMax) && (cJ0 + rr <= JMax);
      g0.bad = false;
      if (g0.live) { const int n = colOf(cx0 + cc) + rowOff(cJ0 + rr); loadGPre<JD>(g0, u, n, n + dKm, n + dKp); }
      const int n1 = tid + NT, r1 = n1/W, c1 = n1 - r1*W;
      g1.live = (n1 < NC) && (cx0 + c1 <= xMax) && (cJ0 + r1 <= JMax);
      g1.bad = false;
      if (g1.live) { const int n = colOf(cx0 + c1) + rowOff(cJ0 + r1); loadGPre<JD>(g1, u, n, n + dKm, n + dKp); }
   }

   // ---- asynchronous copies of the raw operands ----
   // scalars: one staged cell per lane
   for (int m = tid; m < NC; m += NT) {
      const int rr = m/W, cc = m - rr*W;
      if (cx0 + cc <= xMax && cJ0 + rr <= JMax) {
         const int n = colOf(cx0 + cc) + rowOff(cJ0 + rr);
         cpAsync8(s + T::O_RHO + m, u.rho + n);
         cpAsync8(s + T::O_MU + m, u.mu + n);
         cpAsync8(s + T::O_LAM + m, u.lam + n);
         cpAsync8(s + T::O_T + m, u.T + n);
      }
   }
   // species fields: a warp takes 32 consecutive doubles of a row of the box
   {
      constexpr int RL = W*nSpec, CH = (RL + 31)/32;
      for (int it = j; it < ROWS*CH; it += TJ) {
         const int rr = it/CH, e = (it - rr*CH)*32 + i;
         const int cc = e/nSpec, k = e - cc*nSpec;
         if (e < RL && cx0 + cc <= xMax && cJ0 + rr <= JMax) {
            const int n = colOf(cx0 + cc) + rowOff(cJ0 + rr);
            const int d = (rr*W + cc)*T::SP + k;
            cpAsync8(s + T::O_X + d, u.X + ((int64_t)n*nSpec + k));
            cpAsync8(s + T::O_DI + d, u.Di + ((int64_t)n*nSpec + k));
            cpAsync8(s + T::O_RY + d, u.C + ((int64_t)n*nEq + k));
         }
      }
   }
   // velocity box (columns cx0-1 .. cx0+W, rows cJ0-VR0 .. ); its corners are never read
   {
      constexpr int VCH = (T::VL + 31)/32;
      for (int it = j; it < T::VR*VCH; it += TJ) {
         const int rr = it/VCH, e = (it - rr*VCH)*32 + i;
         const int cc = e/3, k = e - cc*3;
         const int x = cx0 - 1 + cc, J = cJ0 - T::VR0 + rr;
         const bool extRow = DOX && (rr == 0 || rr == T::VR - 1), extCol = (cc == 0 || cc == T::VW - 1);
         if (e < T::VL && x <= xMax + 1 && J <= JMax + T::VR0 && !(extRow && extCol)) {
            const int n = colOf(x) + rowOff(J);
            cpAsync8(s + T::O_V + rr*T::VL + e, u.vel + ((int64_t)n*3 + k));
         }
      }
   }

   const bool jAct = inTile && (!DOX || i >= 1), xAct = DOX && inTile && j >= 1;
   const int nq = colOf(cX) + rowOff(cJ);             // own cell, wrapped (it may be lo-1)
   bool bad = g0.bad || g1.bad;
   if (JD == 2 && jAct) bad = bad || (__ldg(u.nT[2] + nq) != Std_node);
   // own face metrics: requested here, used after the derivative pass
   const double msJ = jAct ? __ldg(u.ms[JD] + nq) : 0.0;
   const double msX = xAct ? __ldg(u.ms[0] + nq) : 0.0;

   cpAsyncWaitAll();
   __syncthreads();
   if (g0.live) derivPass<JD, T>(s, tid, g0);
   if (g1.live) derivPass<JD, T>(s, tid + NT, g1);
   const bool anyBad = __syncthreads_or(bad);

   const int own = j*W + i;
   const int vOwn = (j + T::VR0)*T::VW + i + 1;
   double FJ[nEq], Fx[nEq];
   HTR_UNROLL
   for (int e = 0; e < nEq; e++) { FJ[e] = 0.0; Fx[e] = 0.0; }
   const bool ok = !anyBad;
   if (ok && jAct) {
      if (JD == 1) diffFaceA<1, T>(FJ, s, own, own + W, vOwn, vOwn + T::VW, msJ, 0, 1, 5, 6);
      else         diffFaceA<2, T>(FJ, s, own, own + W, vOwn, vOwn + T::VW, msJ, 0, 1, 2, 3);
   }
   if (DOX && ok && xAct) diffFaceA<0, T>(Fx, s, own, own + 1, vOwn, vOwn + 1, msX, 3, 2, 4, 6);
   double Fm[nEq];
   if (DOX) {
      HTR_UNROLL
      for (int e = 0; e < nEq; e++) Fm[e] = __shfl_up_sync(0xffffffffu, Fx[e], 1);
   }
   const int nc = cX + cJ*SJ + kOff;                  // own cell, unwrapped (an updated cell lies in the tile)
   double *Ct = u.Ct + (int64_t)nc*nEq;
   __syncthreads();
   double *sF = smem;
   HTR_UNROLL
   for (int e = 0; e < nEq; e++) sF[e*NT + tid] = FJ[e];
   __syncthreads();
   double ct[nEq];
   if (upd) {
      HTR_UNROLL
      for (int e = 0; e < nEq; e++) ct[e] = Ct[e];
      const double mJ = __ldg(u.md[JD] + nc), mx = DOX ? __ldg(u.md[0] + nc) : 0.0;
      if (ok) {
         HTR_UNROLL
         for (int e = 0; e < nEq; e++) {
            double v = ct[e];
            v += mJ*(FJ[e] - sF[e*NT + tid - 32]);
            if (DOX) v += mx*(Fx[e] - Fm[e]);
            ct[e] = v;
         }
      }
      const int cy = (JD == 1) ? cJ : cK, cz = (JD == 1) ? cK : cJ;
      mask[((int64_t)cz*u.n[1] + cy)*u.n[0] + cX] = !ok;
      if (!ok) *flag = 1;
   }
   // rows of Conserved_t leave through shared memory as contiguous spans (a deferred tile rewrites its old values)
   __syncthreads();
   double *sOut = smem + j*(nEq*32);
   constexpr int X0 = DOX ? 1 : 0, NXC = 32 - X0;      // first updated thread column, updated cells per row
   if (upd) {
      HTR_UNROLL
      for (int e = 0; e < nEq; e++) sOut[(i - X0)*nEq + e] = ct[e];
   }
   __syncwarp();
   if (j >= 1 && cJ <= hiJ) {
      const int left = hiX - (cx0 + X0) + 1;
      const int nval = nEq*(int)(left < NXC ? left : NXC);
      double *row = u.Ct + (int64_t)(cx0 + X0 + cJ*SJ + kOff)*nEq;
      HTR_UNROLL
      for (int e = 0; e < nEq; e++) { const int m = e*32 + i; if (m < nval) row[m] = sOut[m]; }
   }
}

} // namespace fast
} // namespace htr
