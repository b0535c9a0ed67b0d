// Diffusion kernels of the fast path with TMA bulk staging (htr_bulk.cuh): same tiles, arithmetic and
// deferral rules as fast::diffusion_staged_kernel (htr_fast.cuh), but the tile rows of every field are
// moved by cp.async.bulk into shared memory in the fields' own [row][x][comp] layout, and the cell-centred
// velocity derivatives are then taken from shared memory.
//
//   diffusion_bulk_kernel<2,TJ>   Conserved_t += Dz          tile = 32 (x) x TJ (z) cells of one y row
//   diffusion_bulk_kernel<1,TJ>   Conserved_t += Dy; += Dx   tile = 32 (x) x TJ (y) cells of one z plane
//
// Shared-memory box: W = 36 columns starting at an even cell xs <= (first thread cell) - 1, rows 0..TJ of
// the state fields, rows -1..TJ+1 of the in-plane velocity (xy kernel: d/dy), and rows 0..TJ of the velocity
// in the two neighbouring planes (xy: z-1, z+1; z kernel: y-1, y+1).
#pragma once
#include "htr_fast.cuh"
#include "htr_bulk.cuh"

namespace htr {
namespace fast {

template<int JD, int TJ> struct DB {
   static constexpr bool DOX = (JD == 1);
   static constexpr int W = 36, ROWS = TJ + 1, NCB = W*ROWS, UROWS = TJ + 3, NT = 32*TJ;
   static constexpr int NG = DOX ? 7 : 4;
   // offsets in doubles
   static constexpr int O_R = 0;                         // rho, mu, lam, T : [4][ROWS][W]
   static constexpr int O_X = O_R + 4*NCB;               // [ROWS][W][nSpec]
   static constexpr int O_D = O_X + nSpec*NCB;           // [ROWS][W][nSpec]
   static constexpr int O_C = O_D + nSpec*NCB;           // [ROWS][W][nEq]   (Conserved, species part used)
   static constexpr int O_U0 = O_C + nEq*NCB;            // [UROWS][W][3]    rows -1..TJ+1
   static constexpr int O_UM = O_U0 + 3*W*UROWS;         // [ROWS][W][3]     plane K-1
   static constexpr int O_UP = O_UM + 3*NCB;             // [ROWS][W][3]     plane K+1
   static constexpr int O_G = O_UP + 3*NCB;              // [NG][ROWS][W]    derivatives
   static constexpr int ND = O_G + NG*NCB;
   static constexpr size_t SMEM = 16 + sizeof(double)*ND;
   static constexpr uint32_t ROW_STATE_BYTES = 8u*W*(4 + 2*nSpec + nEq + 3 + 3 + 3);   // one state row incl. its 3 velocity rows
   static constexpr uint32_t ROW_U_BYTES = 8u*W*3;
};

// face between box cells n (row rr, col cc) and n + dn; same expression order as diffFaceStd / the reference
template<int dir, int JD, int TJ>
HTR_DEV void diffFaceRaw(double *Flux, const double *d, const int n, const int dn, const double m_s,
                         const int gA, const int gB, const int gC, const int gD) {
   using L = DB<JD, TJ>;
   constexpr int W = L::W, NCB = L::NCB;
   const int q = n + dn;
   const double rho_s = interpStd(d[L::O_R + n], d[L::O_R + q]);
   double Xp[nSpec], Xq[nSpec];
   HTR_UNROLL
   for (int i = 0; i < nSpec; i++) { Xp[i] = d[L::O_X + nSpec*n + i]; Xq[i] = d[L::O_X + nSpec*q + i]; }
   const double iMixW_s = 1.0/interpStd(mix::GetMolarWeightFromXi(Xp), mix::GetMolarWeightFromXi(Xq));
   const double Tp = d[L::O_R + 3*NCB + n], Tq = d[L::O_R + 3*NCB + q];
   const double T_s = interpStd(Tp, Tq);
   double heatFlux = interpStd(d[L::O_R + 2*NCB + n], d[L::O_R + 2*NCB + q])*m_s*(Tp - Tq);

   double hi[nSpec];
   mix::GetSpeciesEnthalpies(hi, T_s);

   double ViCorr = 0.0;
   HTR_UNROLL
   for (int i = 0; i < nSpec; i++) {
      const double YiVi = interpStd(d[L::O_D + nSpec*n + i], d[L::O_D + nSpec*q + i])*
                          m_s*(Xp[i] - Xq[i])*mix::GetSpeciesMolarWeight(i)*iMixW_s;
      Flux[i] = -rho_s*YiVi;
      heatFlux += rho_s*YiVi*hi[i];
      ViCorr += YiVi;
   }
   HTR_UNROLL
   for (int i = 0; i < nSpec; i++) {
      const double rhoYiViCorr = interpStd(d[L::O_C + nEq*n + i], d[L::O_C + nEq*q + i])*ViCorr;
      Flux[i] += rhoYiViCorr;
      heatFlux -= rhoYiViCorr*hi[i];
   }

   const int un = 3*(n + W), uq_ = 3*(q + W);          // velocity rows are shifted by one (row -1 first)
   double up[3], uq[3], dN[3];
   HTR_UNROLL
   for (int i = 0; i < 3; i++) { up[i] = d[L::O_U0 + un + i]; uq[i] = d[L::O_U0 + uq_ + i]; dN[i] = m_s*(uq[i] - up[i]); }
   const double mu_s = interpStd(d[L::O_R + NCB + n], d[L::O_R + NCB + q]);
   const double A = interpStd(d[L::O_G + gA*NCB + n], d[L::O_G + gA*NCB + q]);
   const double Bv = interpStd(d[L::O_G + gB*NCB + n], d[L::O_G + gB*NCB + q]);
   const double C = interpStd(d[L::O_G + gC*NCB + n], d[L::O_G + gC*NCB + q]);
   const double D = interpStd(d[L::O_G + gD*NCB + n], d[L::O_G + gD*NCB + q]);
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
}

template<int JD, int TJ>
__global__ void __launch_bounds__(32*TJ, HTR_DB_MINB)
diffusion_bulk_kernel(const rfluid f, const rbox tile, const rbox B, uint8_t *__restrict__ mask, int *__restrict__ flag,
                      const int bulkOK) {
   using L = DB<JD, TJ>;
   constexpr bool DOX = L::DOX;
   constexpr int W = L::W, NCB = L::NCB, NT = L::NT, ROWS = L::ROWS;
   constexpr int KD = (JD == 1) ? 2 : 1;
   extern __shared__ __align__(16) unsigned char smraw[];
   uint64_t *bar = reinterpret_cast<uint64_t*>(smraw);
   double *d = reinterpret_cast<double*>(smraw + 16);
   static_assert(L::ND >= nEq*NT, "flux exchange reuses the staging buffer");

   const int i = threadIdx.x, j = threadIdx.y;
   const int tid = j*32 + i;
   const int cx0 = DOX ? (tile.lo[0] - 1 + (int)blockIdx.x*31) : (tile.lo[0] + (int)blockIdx.x*32);
   const int cJ0 = tile.lo[JD] - 1 + (int)blockIdx.y*(TJ - 1);
   const int cK = tile.lo[KD] + blockIdx.z;
   const int xs = (cx0 - 1) - (((cx0 - 1) + ((bulkOK >> 1) & 1)) & 1);
   const int co = (int)((cx0 - 1) - xs);
   auto mk = [&](int x, int J, int K) { RPt p; p.x = x; (JD == 1 ? p.y : p.z) = J; (JD == 1 ? p.z : p.y) = K; return p; };
   const RPt c = mk(cx0 + i, cJ0 + j, cK);
   const int cJ = cJ0 + j;

   const int xlo = (tile.lo[0] - 3 > B.lo[0]) ? tile.lo[0] - 3 : B.lo[0];
   const int xhi = (tile.hi[0] + 3 < B.hi[0]) ? tile.hi[0] + 3 : B.hi[0];
   const bool useBulk = (bulkOK & 1) && xs >= xlo && xs + W - 1 <= xhi;
   const int Km = wrapCoord(cK - 1, B.lo[KD], B.hi[KD]), Kp = wrapCoord(cK + 1, B.lo[KD], B.hi[KD]);

   if (tid == 0) bulk::mbarInit(bar, 1);
   __syncthreads();

   auto rowValid = [&](int rr) { return cJ0 + rr <= tile.hi[JD] + 1; };              // state rows 0..TJ
   auto urowValid = [&](int ru) { return cJ0 + ru - 1 <= tile.hi[JD] + 2; };         // velocity rows -1..TJ+1 (index ru = row+1)
   if (useBulk) {
      if (j == 0) {
         if (i == 0) {
            uint32_t bytes = 0;
            for (int rr = 0; rr < ROWS; rr++) {
               if (rowValid(rr)) bytes += L::ROW_STATE_BYTES - L::ROW_U_BYTES;
               if (urowValid(rr + 1)) bytes += L::ROW_U_BYTES;
            }
            if (DOX) { if (urowValid(0)) bytes += L::ROW_U_BYTES; if (urowValid(TJ + 2)) bytes += L::ROW_U_BYTES; }
            bulk::mbarExpectTx(bar, bytes);
         }
         __syncwarp();
         for (int item = i; item < 10*ROWS + (DOX ? 2 : 0); item += 32) {
            if (item >= 10*ROWS) {                            // the two extra in-plane velocity rows (xy kernel)
               const int ru = (item == 10*ROWS) ? 0 : TJ + 2;
               if (!urowValid(ru)) continue;
               const RPt q = mk(xs, wrapCoord(cJ0 + ru - 1, B.lo[JD], B.hi[JD]), cK);
               bulk::copyRow(d + L::O_U0 + 3*W*ru, at<const double>(f.velocity, q), 24u*W, bar);
               continue;
            }
            const int rr = item/10, fld = item - 10*rr;
            if (fld == 7 ? !urowValid(rr + 1) : !rowValid(rr)) continue;
            const int J = wrapCoord(cJ0 + rr, B.lo[JD], B.hi[JD]);
            const RPt q = mk(xs, J, cK);
            switch (fld) {
            case 0: bulk::copyRow(d + L::O_R + rr*W, at<const double>(f.rho, q), 8u*W, bar); break;
            case 1: bulk::copyRow(d + L::O_R + NCB + rr*W, at<const double>(f.mu, q), 8u*W, bar); break;
            case 2: bulk::copyRow(d + L::O_R + 2*NCB + rr*W, at<const double>(f.lam, q), 8u*W, bar); break;
            case 3: bulk::copyRow(d + L::O_R + 3*NCB + rr*W, at<const double>(f.temperature, q), 8u*W, bar); break;
            case 4: bulk::copyRow(d + L::O_X + nSpec*rr*W, at<const double>(f.MolarFracs, q), 8u*nSpec*W, bar); break;
            case 5: bulk::copyRow(d + L::O_D + nSpec*rr*W, at<const double>(f.Di, q), 8u*nSpec*W, bar); break;
            case 6: bulk::copyRow(d + L::O_C + nEq*rr*W, at<const double>(f.Conserved, q), 8u*nEq*W, bar); break;
            case 7: bulk::copyRow(d + L::O_U0 + 3*W*(rr + 1), at<const double>(f.velocity, q), 24u*W, bar); break;
            case 8: bulk::copyRow(d + L::O_UM + 3*rr*W, at<const double>(f.velocity, mk(xs, J, Km)), 24u*W, bar); break;
            default: bulk::copyRow(d + L::O_UP + 3*rr*W, at<const double>(f.velocity, mk(xs, J, Kp)), 24u*W, bar); break;
            }
         }
      }
   } else {
      // cooperative loads (edge tiles, unaligned layouts): one box cell per thread and pass
      for (int n = tid; n < W*L::UROWS; n += NT) {
         const int ru = n/W, cc = n - ru*W;
         const int rr = ru - 1;
         const int x = xs + cc, J = cJ0 + rr;
         if (x < tile.lo[0] - 3 || x > tile.hi[0] + 3 || J < tile.lo[JD] - 3 || J > tile.hi[JD] + 3) continue;
         const int xw = wrapCoord(x, B.lo[0], B.hi[0]), Jw = wrapCoord(J, B.lo[JD], B.hi[JD]);
         const RPt q = mk(xw, Jw, cK);
         HTR_UNROLL
         for (int k = 0; k < 3; k++) d[L::O_U0 + 3*n + k] = ldc<double>(f.velocity, q, k);
         if (rr < 0 || rr > TJ) continue;
         const int m = rr*W + cc;
         d[L::O_R + m] = ld<double>(f.rho, q);
         d[L::O_R + NCB + m] = ld<double>(f.mu, q);
         d[L::O_R + 2*NCB + m] = ld<double>(f.lam, q);
         d[L::O_R + 3*NCB + m] = ld<double>(f.temperature, q);
         HTR_UNROLL
         for (int k = 0; k < nSpec; k++) {
            d[L::O_X + nSpec*m + k] = ldc<double>(f.MolarFracs, q, k);
            d[L::O_D + nSpec*m + k] = ldc<double>(f.Di, q, k);
            d[L::O_C + nEq*m + k] = ldc<double>(f.Conserved, q, k);
         }
         const RPt qm = mk(xw, Jw, Km), qp = mk(xw, Jw, Kp);
         HTR_UNROLL
         for (int k = 0; k < 3; k++) { d[L::O_UM + 3*m + k] = ldc<double>(f.velocity, qm, k); d[L::O_UP + 3*m + k] = ldc<double>(f.velocity, qp, k); }
      }
   }

   // while the rows fly: metrics and node types of the cells whose derivatives this thread will take
   constexpr int PER = (NCB + NT - 1)/NT;
   double md[PER][3];
   bool bad = false, mine[PER];
   HTR_UNROLL
   for (int k = 0; k < PER; k++) {
      const int n = tid + k*NT;
      mine[k] = false;
      md[k][0] = md[k][1] = md[k][2] = 0.0;
      if (n < NCB) {
         const int rr = n/W, cc = n - rr*W;
         const int x = xs + cc, J = cJ0 + rr;
         if (cc >= 1 && cc <= W - 2 && x >= tile.lo[0] - 1 && x <= tile.hi[0] + 1 && J <= tile.hi[JD] + 1) {
            mine[k] = true;
            const RPt q = mk(wrapCoord(x, B.lo[0], B.hi[0]), wrapCoord(J, B.lo[JD], B.hi[JD]), cK);
            md[k][0] = ld<double>(f.dcsi_d, q);
            md[k][1] = ld<double>(f.deta_d, q);
            bad = bad || (ld<int>(f.nType_x, q) != Std_node) || (ld<int>(f.nType_y, q) != Std_node);
            if (DOX) { md[k][2] = ld<double>(f.dzet_d, q); bad = bad || (ld<int>(f.nType_z, q) != Std_node); }
         }
      }
   }
   const bool inTile = c.x <= tile.hi[0] && cJ <= tile.hi[JD];
   const bool jAct = inTile && (!DOX || i >= 1), xAct = DOX && inTile && j >= 1;
   const RPt qo = mk(wrapCoord(c.x, B.lo[0], B.hi[0]), wrapCoord(cJ, B.lo[JD], B.hi[JD]), cK);   // own cell, wrapped
   if (JD == 2 && jAct) bad = bad || (ld<int>(f.nType_z, qo) != Std_node);
   double msJ = 0.0, msx = 0.0;
   if (jAct) msJ = ld<double>(m_sOf<JD>(f), qo);
   if (xAct) msx = ld<double>(f.dcsi_s, qo);

   if (useBulk) bulk::mbarWait(bar, 0);
   else __syncthreads();

   // cell-centred velocity derivatives from shared memory
   HTR_UNROLL
   for (int k = 0; k < PER; k++) {
      const int n = tid + k*NT;
      if (!mine[k]) continue;
      const int nu = 3*(n + W);                                // this cell in the in-plane velocity box
      const double *U0 = d + L::O_U0, *UM = d + L::O_UM + 3*n, *UP = d + L::O_UP + 3*n;
      double *G = d + L::O_G + n;
      if (DOX) {
         G[0*NCB] = derivStd(U0[nu - 3 + 0], U0[nu + 0], U0[nu + 3 + 0], md[k][0]);             // dUdX
         G[1*NCB] = derivStd(U0[nu - 3 + 1], U0[nu + 1], U0[nu + 3 + 1], md[k][0]);             // dVdX
         G[2*NCB] = derivStd(U0[nu - 3*W + 0], U0[nu + 0], U0[nu + 3*W + 0], md[k][1]);         // dUdY
         G[3*NCB] = derivStd(U0[nu - 3*W + 1], U0[nu + 1], U0[nu + 3*W + 1], md[k][1]);         // dVdY
         HTR_UNROLL
         for (int a = 0; a < 3; a++) G[(4 + a)*NCB] = derivStd(UM[a], U0[nu + a], UP[a], md[k][2]);   // dUdZ dVdZ dWdZ
      } else {
         G[0*NCB] = derivStd(U0[nu - 3 + 0], U0[nu + 0], U0[nu + 3 + 0], md[k][0]);             // dUdX
         G[1*NCB] = derivStd(U0[nu - 3 + 2], U0[nu + 2], U0[nu + 3 + 2], md[k][0]);             // dWdX
         G[2*NCB] = derivStd(UM[1], U0[nu + 1], UP[1], md[k][1]);                               // dVdY
         G[3*NCB] = derivStd(UM[2], U0[nu + 2], UP[2], md[k][1]);                               // dWdY
      }
   }
   const bool anyBad = __syncthreads_or(bad);

   const int own = j*W + i + 1 + co;
   double FJ[nEq], Fx[nEq];
   HTR_UNROLL
   for (int e = 0; e < nEq; e++) { FJ[e] = 0.0; Fx[e] = 0.0; }
   const bool ok = !anyBad;
   if (ok && jAct) {
      if (JD == 1) diffFaceRaw<1, JD, TJ>(FJ, d, own, W, msJ, 0, 1, 5, 6);
      else         diffFaceRaw<2, JD, TJ>(FJ, d, own, W, msJ, 0, 1, 2, 3);
   }
   if (DOX && ok && xAct) diffFaceRaw<0, JD, TJ>(Fx, d, own, 1, msx, 3, 2, 4, 6);
   double Fm[nEq];
   if (DOX) {
      HTR_UNROLL
      for (int e = 0; e < nEq; e++) Fm[e] = __shfl_up_sync(0xffffffffu, Fx[e], 1);
   }
   __syncthreads();
   double *sF = d;
   HTR_UNROLL
   for (int e = 0; e < nEq; e++) sF[e*NT + tid] = FJ[e];
   __syncthreads();
   if (inTile && j >= 1 && (!DOX || i >= 1)) {
      if (ok) {
         double *Ct = at<double>(f.Conserved_t, c);
         const double mJ = ld<double>(m_dOf<JD>(f), c);
         const double mx = DOX ? ld<double>(f.dcsi_d, c) : 0.0;
         HTR_UNROLL
         for (int e = 0; e < nEq; e++) {
            double v = Ct[e];
            v += mJ*(FJ[e] - sF[e*NT + tid - 32]);
            if (DOX) v += mx*(Fx[e] - Fm[e]);
            Ct[e] = v;
         }
      }
      mask[maskIndex(c, tile)] = !ok;
      if (!ok) *flag = 1;
   }
}

} // namespace fast
} // namespace htr
