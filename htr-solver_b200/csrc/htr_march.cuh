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

#ifndef HTR_MZ_MINB
#define HTR_MZ_MINB 2
#endif
#ifndef HTR_MP_MINB
#define HTR_MP_MINB 2
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

// KG flux of the face between points c2 and c3 of the six-point stencil c0..c5 (Std_node); arithmetic of
// fast::kgStdFlux / kgFaceFlux (src/prometeo_rhs.inl:42-135)
template<int dir>
HTR_DEV void kgFlux6(double *Flux, const KGPoint &c0, const KGPoint &c1, const KGPoint &c2, const KGPoint &c3,
                     const KGPoint &c4, const KGPoint &c5) {
   const KGPoint *c[6] = {&c0, &c1, &c2, &c3, &c4, &c5};
   double fl[nEq + 1];
   HTR_UNROLL
   for (int i = 0; i < nEq + 1; i++) fl[i] = 0.0;
   HTR_UNROLL
   for (int l = 0; l < 3; l++) {
      double acc[nEq + 1];
      HTR_UNROLL
      for (int i = 0; i < nEq + 1; i++) acc[i] = 0.0;
      HTR_UNROLL
      for (int m = 0; m <= l; m++) {
         double S[nEq + 1];
         kgPair(S, *c[2 - m], *c[2 - m + l + 1]);
         HTR_UNROLL
         for (int i = 0; i < nEq + 1; i++) acc[i] += S[i];
      }
      const double cl = c_KennedyCoeff[Std_node][l];
      HTR_UNROLL
      for (int i = 0; i < nEq + 1; i++) fl[i] += cl*acc[i];
   }
   HTR_UNROLL
   for (int i = 0; i < nEq; i++) Flux[i] = 0.25*fl[i];
   Flux[irU + dir] -= fl[nEq];
}

// =============================================================================================
// Conserved_t = 0 (+ chemistry source) + Euler-z        block = 32 (x) x MZ_TY (y) columns
// =============================================================================================
constexpr int MZ_TY = 4, MZ_NT = 32*MZ_TY;

struct ZRaw { double rho, p, rhoE, u[3], Y[nSpec]; };

HTR_DEV void loadZRaw(ZRaw &r, const UF &f, const int n) {
   r.rho = __ldg(f.rho + n);
   r.p = __ldg(f.p + n);
   r.rhoE = __ldg(f.C + (int64_t)n*nEq + irE);
   HTR_UNROLL
   for (int c = 0; c < 3; c++) r.u[c] = __ldg(f.vel + (int64_t)n*3 + c);
   HTR_UNROLL
   for (int s = 0; s < nSpec; s++) r.Y[s] = __ldg(f.Y + (int64_t)n*nSpec + s);
}
HTR_DEV void toPoint(KGPoint &c, const ZRaw &r) {
   c.rho = r.rho; c.p = r.p;
   c.H = (r.rhoE + r.p)/r.rho;
   HTR_UNROLL
   for (int i = 0; i < 3; i++) c.u[i] = r.u[i];
   c.un = r.u[2];
   HTR_UNROLL
   for (int s = 0; s < nSpec; s++) c.Y[s] = r.Y[s];
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

   // ring slot of plane q: (q - (z0-3)) mod 6; the face z+1/2 uses planes z-2 .. z+3
   KGPoint c[6];
   HTR_UNROLL
   for (int k = 0; k < 6; k++) { ZRaw r; loadZRaw(r, f, HTR_ZOFF(z0 - 3 + k)); toPoint(c[k], r); }
   // rolling sensor / node-type state of the faces
   bool s0 = true, s1 = true;                       // sensor at z-1 and z (before the step's shift: z-2, z-1)
   if (HYBRID) { s0 = __ldg(f.sens[2] + HTR_ZOFF(z0 - 2)) != 0; s1 = __ldg(f.sens[2] + HTR_ZOFF(z0 - 1)) != 0; }
   double Fm[nEq];
   bool haveM;
   {  // face z0 - 1/2: stencil planes z0-3 .. z0+2 = slots 0..5
      bool sm = true;
      if (HYBRID) { const bool s2 = __ldg(f.sens[2] + HTR_ZOFF(z0)) != 0; sm = s0 && s1 && s2; s0 = s1; s1 = s2; }
      haveM = sm && (__ldg(f.nT[2] + HTR_ZOFF(z0 - 1)) == Std_node);
      kgFlux6<2>(Fm, c[0], c[1], c[2], c[3], c[4], c[5]);
   }
   ZRaw nx;                                          // plane z+3 of the step being prepared (one step ahead)
   loadZRaw(nx, f, HTR_ZOFF(z0 + 3));

   for (int zb = z0; zb <= z1; zb += 6) {
      HTR_UNROLL
      for (int s = 0; s < 6; s++) {
         const int z = zb + s;
         if (z <= z1) {                             // block-uniform
            toPoint(c[s], nx);                      // plane z+3 replaces plane z-3
            if (z < z1) loadZRaw(nx, f, HTR_ZOFF(z + 4));
            const int n = HTR_ZOFF(z);
            double Tc = 0.0;
            if (CHEM) Tc = __ldg(f.T + n);
            const double mz = __ldg(f.me[2] + n);
            bool sm = true;
            if (HYBRID) { const bool s2 = __ldg(f.sens[2] + HTR_ZOFF(z + 1)) != 0; sm = s0 && s1 && s2; s0 = s1; s1 = s2; }
            const bool haveP = sm && (__ldg(f.nT[2] + n) == Std_node);
            double Fp[nEq];
            kgFlux6<2>(Fp, c[(s + 1)%6], c[(s + 2)%6], c[(s + 3)%6], c[(s + 4)%6], c[(s + 5)%6], c[s]);
            double base[nEq];
            HTR_UNROLL
            for (int e = 0; e < nEq; e++) base[e] = 0.0;
            if (CHEM) {
               const KGPoint &o = c[(s + 3)%6];      // own plane z
               double w[nSpec];
               mix::GetProductionRates(w, o.rho, o.p, Tc, o.Y);
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
      }
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
// shared-memory map (doubles)
constexpr int OFF_V = 0;                               // [3][NCE*3] velocity ring, AoS
constexpr int OFF_R1 = 3*NCE*3;
constexpr int E_RHO = 0, E_P = NCE, E_H = 2*NCE, E_Y = 3*NCE, E_END = NCE*(3 + nSpec);
constexpr int OFF_SF = E_END;                          // [nEq][PNT] face exchange (beyond the E data, inside the D data)
constexpr int OFF_SO = OFF_SF + nEq*PNT;               // [8][32*nEq] output rows
constexpr int D_RHO = 0, D_MU = NCD, D_LAM = 2*NCD, D_T = 3*NCD, D_X = 4*NCD, D_DI = D_X + NCD*nSpec,
              D_RY = D_DI + NCD*nSpec, D_G = D_RY + NCD*nSpec, D_END = D_G + NG*NCD;
constexpr int R1_DBL = (OFF_SO + 8*32*nEq > D_END) ? OFF_SO + 8*32*nEq : D_END;
constexpr int NBYTES = 2*NCE + 2*PNT;                  // sensX, sensY on the stencil box; have-flags; deferred flags
constexpr size_t PM_SMEM = sizeof(double)*(OFF_R1 + R1_DBL) + ((NBYTES + 15)/16)*16;

// operands of one cell of a diffusion face
struct DOps { double rho, mu, lam, T, u[3], X[nSpec], Di[nSpec], rY[nSpec], g[4]; };

// own cell of plane n (element offset) straight from global memory; u and g are filled by the caller
HTR_DEV void loadDOps(DOps &o, const UF &f, const int n) {
   o.rho = __ldg(f.rho + n); o.mu = __ldg(f.mu + n); o.lam = __ldg(f.lam + n); o.T = __ldg(f.T + n);
   HTR_UNROLL
   for (int s = 0; s < nSpec; s++) {
      o.X[s] = __ldg(f.X + (int64_t)n*nSpec + s);
      o.Di[s] = __ldg(f.Di + (int64_t)n*nSpec + s);
      o.rY[s] = __ldg(f.C + (int64_t)n*nEq + s);
   }
}
// staged cell `idx` of the diffusion box; gA..gD = slots of the staged derivatives the face of this direction uses
HTR_DEV void ldsDOps(DOps &o, const double *__restrict__ d, const double *__restrict__ v, const int idx, const int vidx,
                     const int gA, const int gB, const int gC, const int gD) {
   o.rho = d[D_RHO + idx]; o.mu = d[D_MU + idx]; o.lam = d[D_LAM + idx]; o.T = d[D_T + idx];
   HTR_UNROLL
   for (int c = 0; c < 3; c++) o.u[c] = v[vidx*3 + c];
   HTR_UNROLL
   for (int s = 0; s < nSpec; s++) {
      o.X[s] = d[D_X + idx*nSpec + s];
      o.Di[s] = d[D_DI + idx*nSpec + s];
      o.rY[s] = d[D_RY + idx*nSpec + s];
   }
   o.g[0] = d[D_G + gA*NCD + idx]; o.g[1] = d[D_G + gB*NCD + idx];
   o.g[2] = d[D_G + gC*NCD + idx]; o.g[3] = d[D_G + gD*NCD + idx];
}

// Diffusion flux of the face between cells P and Q = P + 1 along dir (Std_node).  Expression for expression
// fast::diffFaceStd / diffusionFaceFlux (src/prometeo_rhs.inl:1058-1215); g[] holds the transverse derivatives in
// the order   dir 0: dVdY dUdY dUdZ dWdZ    dir 1: dUdX dVdX dVdZ dWdZ    dir 2: dUdX dWdX dVdY dWdY
template<int dir>
HTR_DEV void diffFace(double *Flux, const DOps &P, const DOps &Q, const double m_s) {
   const double rho_s = interpStd(P.rho, Q.rho);
   const double iMixW_s = 1.0/interpStd(mix::GetMolarWeightFromXi(P.X), mix::GetMolarWeightFromXi(Q.X));
   const double Tp = P.T, Tq = Q.T;
   const double T_s = interpStd(Tp, Tq);
   double heatFlux = interpStd(P.lam, Q.lam)*m_s*(Tp - Tq);

   double hi[nSpec];
   mix::GetSpeciesEnthalpies(hi, T_s);

   double ViCorr = 0.0;
   HTR_UNROLL
   for (int i = 0; i < nSpec; i++) {
      const double YiVi = interpStd(P.Di[i], Q.Di[i])*m_s*(P.X[i] - Q.X[i])*mix::GetSpeciesMolarWeight(i)*iMixW_s;
      Flux[i] = -rho_s*YiVi;
      heatFlux += rho_s*YiVi*hi[i];
      ViCorr += YiVi;
   }
   HTR_UNROLL
   for (int i = 0; i < nSpec; i++) {
      const double rhoYiViCorr = interpStd(P.rY[i], Q.rY[i])*ViCorr;
      Flux[i] += rhoYiViCorr;
      heatFlux -= rhoYiViCorr*hi[i];
   }

   double dN[3];
   HTR_UNROLL
   for (int i = 0; i < 3; i++) dN[i] = m_s*(Q.u[i] - P.u[i]);
   const double mu_s = interpStd(P.mu, Q.mu);
   const double A = interpStd(P.g[0], Q.g[0]);
   const double Bv = interpStd(P.g[1], Q.g[1]);
   const double Cv = interpStd(P.g[2], Q.g[2]);
   const double Dv = interpStd(P.g[3], Q.g[3]);
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
   for (int i = 0; i < 3; i++) uSigma += interpStd(P.u[i], Q.u[i])*sigma[i];
   Flux[irE] = (uSigma - heatFlux);
}

// x/y derivatives of the velocity at stencil-box cell v of one staged velocity plane, as the z face uses them:
// g = dUdX dWdX dVdY dWdY
HTR_DEV void zFaceDerivs(double *g, const double *__restrict__ v, const int vi, const double mx, const double my) {
   g[0] = derivStd(v[(vi - 1)*3 + 0], v[vi*3 + 0], v[(vi + 1)*3 + 0], mx);
   g[1] = derivStd(v[(vi - 1)*3 + 2], v[vi*3 + 2], v[(vi + 1)*3 + 2], mx);
   g[2] = derivStd(v[(vi - PW)*3 + 1], v[vi*3 + 1], v[(vi + PW)*3 + 1], my);
   g[3] = derivStd(v[(vi - PW)*3 + 2], v[vi*3 + 2], v[(vi + PW)*3 + 2], my);
}

template<bool HYBRID, bool DIFF>
__global__ void __launch_bounds__(PNT, HTR_MP_MINB)
plane_march_kernel(const UF f, uint8_t *__restrict__ mask, int *__restrict__ flag, const int kch) {
   extern __shared__ double smem[];
   double *sV = smem + OFF_V;
   double *R1 = smem + OFF_R1;
   double *sF = R1 + OFF_SF;
   uint8_t *sSx = reinterpret_cast<uint8_t*>(R1 + R1_DBL);
   uint8_t *sSy = sSx + NCE;
   uint8_t *sHave = sSy + NCE;
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
   if (DIFF) {
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
   }
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
   // any node type of plane z on the diffusion box that the fast arithmetic does not cover?
   auto badNodes = [&](const int zo) -> bool {
      bool bad = false;
      if (DIFF) {
         if (dV[0]) {
            const int n = dOfs[0] + zo;
            bad = (__ldg(f.nT[0] + n) != Std_node) || (__ldg(f.nT[1] + n) != Std_node) || (__ldg(f.nT[2] + n) != Std_node);
         }
         if (partB == 0) {
            const int n = dOfs[1] + zo;
            bad = bad || (__ldg(f.nT[0] + n) != Std_node) || (__ldg(f.nT[1] + n) != Std_node) || (__ldg(f.nT[2] + n) != Std_node);
         }
      } else if (faceAct) {
         const int n = nOwn + zo;
         bad = (__ldg(f.nT[0] + n) != Std_node) || (__ldg(f.nT[1] + n) != Std_node);
      }
      return bad;
   };

   // ---- prologue: velocity planes z0-1, z0; node types of z0-1, z0; diffusion z face z0-1/2 ------------
   // ring slot of plane z: (z - z0 + 1) mod 3
   bool badP = false, badC = false;
   double FzM[nEq], gC[4];                               // diffusion z flux of face p-1/2; z-face derivatives of plane p
   if (DIFF) {
      stageV(sV, HTR_ZP(z0 - 1));
      stageV(sV + NCE*3, HTR_ZP(z0));
      badP = __syncthreads_or(badNodes(HTR_ZP(z0 - 1)));
      badC = __syncthreads_or(badNodes(HTR_ZP(z0)));
      DOps P, Q;
      const int nm = nOwn + HTR_ZP(z0 - 1), nc = nOwn + HTR_ZP(z0);
      loadDOps(P, f, nm); loadDOps(Q, f, nc);
      HTR_UNROLL
      for (int c = 0; c < 3; c++) { P.u[c] = sV[ownE*3 + c]; Q.u[c] = sV[NCE*3 + ownE*3 + c]; }
      zFaceDerivs(P.g, sV, ownE, __ldg(f.md[0] + nm), __ldg(f.md[1] + nm));
      zFaceDerivs(gC, sV + NCE*3, ownE, __ldg(f.md[0] + nc), __ldg(f.md[1] + nc));
      HTR_UNROLL
      for (int k = 0; k < 4; k++) Q.g[k] = gC[k];
      diffFace<2>(FzM, P, Q, __ldg(f.ms[2] + nm));
   }

   for (int p = z0; p <= z1; p++) {
      const int k3 = (p - z0)%3;
      double *vm = sV + k3*(NCE*3);                                 // plane p-1
      double *vc = sV + ((k3 + 1)%3)*(NCE*3);                       // plane p
      double *vn = sV + ((k3 + 2)%3)*(NCE*3);                       // plane p+1
      const int zo = HTR_ZP(p), zn = HTR_ZP(p + 1);
      const int nO = nOwn + zo;

      // ---- (1) stage the E data of plane p, the velocity of plane p+1 (DIFF) or p, the sensors ----------
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
      if (DIFF) stageV(vn, zn); else stageV(vc, zo);
      // barrier B1, which also gathers the node-type check of plane p+1 (DIFF) or p
      const bool badN = __syncthreads_or(badNodes(DIFF ? zn : zo)) != 0;
      if (!DIFF) badC = false;
      const bool bad3 = badP || badC || badN;                                      // CTA-uniform: this plane's cells are deferred
      const bool badF = badC || badN;                                              // the faces of this step are not computed

      // ---- (2) Euler y and x faces ----------------------------------------------------------------------
      double Fy[nEq], Fx[nEq];
      bool haveY = false, haveX = false;
      if (faceAct && !badF) {
         haveY = !HYBRID || (sSy[ownE - PW] && sSy[ownE] && sSy[ownE + PW]);
         haveX = !HYBRID || (sSx[ownE - 1] && sSx[ownE] && sSx[ownE + 1]);
         auto pt = [&](KGPoint &c, const int idx, const int dir) {
            c.rho = R1[E_RHO + idx]; c.p = R1[E_P + idx]; c.H = R1[E_H + idx];
            HTR_UNROLL
            for (int q = 0; q < 3; q++) c.u[q] = vc[idx*3 + q];
            HTR_UNROLL
            for (int s = 0; s < nSpec; s++) c.Y[s] = R1[E_Y + idx*nSpec + s];
            c.un = c.u[dir];
         };
         if (haveY) {
            KGPoint c[6];
            HTR_UNROLL
            for (int k = 0; k < 6; k++) pt(c[k], ownE + (k - 2)*PW, 1);
            kgFlux6<1>(Fy, c[0], c[1], c[2], c[3], c[4], c[5]);
         }
         if (haveX) {
            KGPoint c[6];
            HTR_UNROLL
            for (int k = 0; k < 6; k++) pt(c[k], ownE + (k - 2), 0);
            kgFlux6<0>(Fx, c[0], c[1], c[2], c[3], c[4], c[5]);
         }
      }
      if (!haveY) { HTR_UNROLL for (int e = 0; e < nEq; e++) Fy[e] = 0.0; }
      if (!haveX) { HTR_UNROLL for (int e = 0; e < nEq; e++) Fx[e] = 0.0; }
      HTR_UNROLL
      for (int e = 0; e < nEq; e++) sF[e*PNT + tid] = Fy[e];
      sHave[tid] = haveY;
      double ct[nEq];
      if (upd) {
         HTR_UNROLL
         for (int e = 0; e < nEq; e++) ct[e] = f.Ct[(int64_t)nO*nEq + e];
      }
      __syncthreads();                                                             // B2
      bool okE = false;
      {
         double FxM[nEq];
         HTR_UNROLL
         for (int e = 0; e < nEq; e++) FxM[e] = __shfl_up_sync(0xffffffffu, Fx[e], 1);
         const bool haveXm = __shfl_up_sync(0xffffffffu, (int)haveX, 1) != 0;
         if (upd) {
            okE = haveX && haveXm && haveY && sHave[tid - 32];
            const double my = __ldg(f.me[1] + nO), mx = __ldg(f.me[0] + nO);
            HTR_UNROLL
            for (int e = 0; e < nEq; e++) {
               double v = ct[e];
               v += my*(Fy[e] - sF[e*PNT + tid - 32]);
               v += mx*(Fx[e] - FxM[e]);
               ct[e] = v;
            }
         }
      }

      if (DIFF) {
         // upper operands of the z face p+1/2 (own cell of plane p+1): in flight during the staging below
         const int nN = nOwn + zn;
         __syncthreads();                                                          // B3: sF has been read
         // ---- (3) stage the diffusion operands of plane p and the velocity derivatives --------------------
         if (!badF) {
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
         }
         __syncthreads();                                                          // B4
         // ---- (4) diffusion faces: z (p+1/2), y, x ---------------------------------------------------------
         double Dz[nEq], Dy[nEq], Dx[nEq];
         double gN[4];
         HTR_UNROLL
         for (int e = 0; e < nEq; e++) { Dz[e] = 0.0; Dy[e] = 0.0; Dx[e] = 0.0; }
         if (faceAct && !badF) {       // the z face p+1/2 is the next step's minus face: it only depends on planes p, p+1
            {
               DOps P, Qz;
               loadDOps(Qz, f, nN);          // upper operands of the z face p+1/2: own cell of plane p+1, straight from global
               ldsDOps(P, R1, vc, ownD, ownE, 0, 0, 0, 0);
               HTR_UNROLL
               for (int k = 0; k < 4; k++) P.g[k] = gC[k];
               HTR_UNROLL
               for (int c = 0; c < 3; c++) Qz.u[c] = vn[ownE*3 + c];
               zFaceDerivs(gN, vn, ownE, __ldg(f.md[0] + nN), __ldg(f.md[1] + nN));
               HTR_UNROLL
               for (int k = 0; k < 4; k++) Qz.g[k] = gN[k];
               diffFace<2>(Dz, P, Qz, __ldg(f.ms[2] + nO));
            }
            {
               DOps P, Q;
               ldsDOps(P, R1, vc, ownD, ownE, 0, 1, 5, 6);
               ldsDOps(Q, R1, vc, ownD + DW, ownE + PW, 0, 1, 5, 6);
               diffFace<1>(Dy, P, Q, __ldg(f.ms[1] + nO));
            }
            {
               DOps P, Q;
               ldsDOps(P, R1, vc, ownD, ownE, 3, 2, 4, 6);
               ldsDOps(Q, R1, vc, ownD + 1, ownE + 1, 3, 2, 4, 6);
               diffFace<0>(Dx, P, Q, __ldg(f.ms[0] + nO));
            }
         } else {
            HTR_UNROLL
            for (int k = 0; k < 4; k++) gN[k] = 0.0;
            if (faceAct) zFaceDerivs(gN, vn, ownE, __ldg(f.md[0] + nN), __ldg(f.md[1] + nN));      // keeps the carried derivatives valid across a deferred plane
         }
         __syncthreads();                                                          // B5: the D data has been read
         HTR_UNROLL
         for (int e = 0; e < nEq; e++) sF[e*PNT + tid] = Dy[e];
         __syncthreads();                                                          // B6
         {
            double DxM[nEq];
            HTR_UNROLL
            for (int e = 0; e < nEq; e++) DxM[e] = __shfl_up_sync(0xffffffffu, Dx[e], 1);
            if (upd) {
               const double mz = __ldg(f.md[2] + nO), my = __ldg(f.md[1] + nO), mx = __ldg(f.md[0] + nO);
               HTR_UNROLL
               for (int e = 0; e < nEq; e++) {
                  double v = ct[e];
                  v += mz*(Dz[e] - FzM[e]);
                  v += my*(Dy[e] - sF[e*PNT + tid - 32]);
                  v += mx*(Dx[e] - DxM[e]);
                  ct[e] = v;
               }
            }
         }
         HTR_UNROLL
         for (int e = 0; e < nEq; e++) FzM[e] = Dz[e];
         HTR_UNROLL
         for (int k = 0; k < 4; k++) gC[k] = gN[k];
      }

      // ---- (5) Conserved_t rows leave as contiguous spans; deferred cells keep their value -----------------
      const bool deferred = bad3 || !okE;
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
      badP = DIFF ? badC : false; badC = badN;
   }
   #undef HTR_ZP
}

} // namespace march
} // namespace htr
