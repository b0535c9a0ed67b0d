// TEST INFRASTRUCTURE ONLY (checker, never the product path).
//
// Boundary part of the reference oracle: the NSCBC right-hand-side tasks (src/prometeo_rhs.inl:1285-1930,
// cpu_base_impl loops src/prometeo_rhs.cc:529-1138) and the boundary-cell primitive updates of src/prometeo_bc.cc:80-270,
// driven over raw host arrays.  Like ref_driver.cc this file copies nothing: it #includes the reference headers where
// they lie and restates only the accessor set-up and the loop over the boundary box.  Compiled as its own translation
// unit (prometeo_metric.inl has no include guard) and linked into oracle/_ref/libhtr_ref_<Mix>.so.
#include "legion.h"
using namespace Legion;
#include <math.h>
#include <stdio.h>
#include <string.h>
#include <assert.h>

// prometeo_metric.inl (no include guard) comes with both prometeo_rhs.inl and prometeo_variables.hpp, so the file is
// compiled twice: plain = the NSCBC right-hand-side tasks, -DHTR_BC_PART_SET = the boundary-cell primitive updates.
#ifdef HTR_BC_PART_SET
#include "prometeo_bc.hpp"
#else
#include "prometeo_rhs.hpp"
#include "prometeo_rhs.inl"
#include "prometeo_bc_types.h"
#endif

#include "htr_oracle_api.h"

extern Mix *htr_ref_g_mix;
static inline const Mix &mix() { assert(htr_ref_g_mix); return *htr_ref_g_mix; }

namespace {
struct Acc {
   coord_t nx, ny, nz;
   Rect<3> bounds;
   explicit Acc(const htr_oracle_fields *f) : nx(f->nx), ny(f->ny), nz(f->nz),
      bounds(Point<3>(0, 0, 0), Point<3>(f->nx - 1, f->ny - 1, f->nz - 1)) {}
   template<typename FT> AccessorRO<FT, 3> ro(const void *p) const {
      const coord_t s = sizeof(FT);
      return AccessorRO<FT, 3>((void*)p, s, s*nx, s*nx*ny);
   }
   template<typename FT> AccessorRW<FT, 3> rw(void *p) const {
      const coord_t s = sizeof(FT);
      return AccessorRW<FT, 3>(p, s, s*nx, s*nx*ny);
   }
   template<typename FT> AccessorWO<FT, 3> wo(void *p) const {
      const coord_t s = sizeof(FT);
      return AccessorWO<FT, 3>(p, s, s*nx, s*nx*ny);
   }
};

template<direction dir> struct Sel;
template<> struct Sel<Xdir> {
   static const int32_t *nType(const htr_oracle_fields *f) { return f->nType_x; }
   static const double *m_d(const htr_oracle_fields *f) { return f->dcsi_d; }
};
template<> struct Sel<Ydir> {
   static const int32_t *nType(const htr_oracle_fields *f) { return f->nType_y; }
   static const double *m_d(const htr_oracle_fields *f) { return f->deta_d; }
};
template<> struct Sel<Zdir> {
   static const int32_t *nType(const htr_oracle_fields *f) { return f->nType_z; }
   static const double *m_d(const htr_oracle_fields *f) { return f->dzet_d; }
};

#define HTR_BC_LOOP(lo, hi) \
   for (int k = (int)lo[2]; k <= (int)hi[2]; k++) \
      for (int j = (int)lo[1]; j <= (int)hi[1]; j++) \
         for (int i = (int)lo[0]; i <= (int)hi[0]; i++)

#ifndef HTR_BC_PART_SET
// src/prometeo_rhs.cc:529-574 (minus side), :606-651 (plus side)
template<direction dir, class Task>
void nscbc_inflow(const htr_oracle_fields *f, const htr_oracle_bc_fields *b, const int64_t *lo, const int64_t *hi) {
   Acc a(f);
   const AccessorRO<VecNSp, 3> MassFracs = a.ro<VecNSp>(f->MassFracs);
   const AccessorRO<double, 3> pressure = a.ro<double>(f->pressure);
   const AccessorRO<  Vec3, 3> velocity = a.ro<Vec3>(f->velocity);
   const AccessorRO<   int, 3> nType = a.ro<int>(Sel<dir>::nType(f));
   const AccessorRO<double, 3> m_d = a.ro<double>(Sel<dir>::m_d(f));
   const AccessorRO<double, 3> rho = a.ro<double>(f->rho);
   const AccessorRO<double, 3> SoS = a.ro<double>(f->SoS);
   const AccessorRO<double, 3> temperature = a.ro<double>(f->temperature);
   const AccessorRO<  Vec3, 3> dudt = a.ro<Vec3>(b->dudtBoundary);
   const AccessorRO<double, 3> dTdt = a.ro<double>(b->dTdtBoundary);
   const AccessorRW<VecNEq, 3> Conserved_t = a.rw<VecNEq>(f->Conserved_t);
   HTR_BC_LOOP(lo, hi) {
      const Point<3> p(i, j, k);
      Task::addLODIfluxes(Conserved_t[p], MassFracs, pressure, velocity, SoS[p], rho[p], temperature[p],
                          dudt[p], dTdt[p], p, nType[p], m_d[p], mix());
   }
}

// src/prometeo_rhs.cc:683-738 (minus side), :794-849 (plus side)
template<direction dir, class Task>
void nscbc_outflow(const htr_oracle_fields *f, const int64_t *lo, const int64_t *hi, double MaxMach, double LengthScale, double PInf) {
   Acc a(f);
   const AccessorRO<  Vec3, 3> velocity = a.ro<Vec3>(f->velocity);
   const AccessorRO<   int, 3> tN = a.ro<int>(Sel<dir>::nType(f)), t1 = a.ro<int>(Sel<getT1(dir)>::nType(f)), t2 = a.ro<int>(Sel<getT2(dir)>::nType(f));
   const AccessorRO<double, 3> mN = a.ro<double>(Sel<dir>::m_d(f)), m1 = a.ro<double>(Sel<getT1(dir)>::m_d(f)), m2 = a.ro<double>(Sel<getT2(dir)>::m_d(f));
   const AccessorRO<double, 3> rho = a.ro<double>(f->rho);
   const AccessorRO<double, 3> mu = a.ro<double>(f->mu);
   const AccessorRO<VecNSp, 3> MassFracs = a.ro<VecNSp>(f->MassFracs);
   const AccessorRO<double, 3> pressure = a.ro<double>(f->pressure);
   const AccessorRO<double, 3> SoS = a.ro<double>(f->SoS);
   const AccessorRO<double, 3> temperature = a.ro<double>(f->temperature);
   const AccessorRO<VecNEq, 3> Conserved = a.ro<VecNEq>(f->Conserved);
   const AccessorRW<VecNEq, 3> Conserved_t = a.rw<VecNEq>(f->Conserved_t);
   HTR_BC_LOOP(lo, hi) {
      const Point<3> p(i, j, k);
      Task::addLODIfluxes(Conserved_t[p], tN, t1, t2, mN, m1, m2, MassFracs, rho, mu, pressure, velocity,
                          SoS[p], temperature[p], Conserved[p], p, a.bounds, MaxMach, LengthScale, PInf, mix());
   }
}

// src/prometeo_rhs.cc:905-953 (minus side), :985-1033 (plus side)
template<direction dir, class Task>
void nscbc_farfield(const htr_oracle_fields *f, const htr_oracle_bc_fields *b, const int64_t *lo, const int64_t *hi,
                    double MaxMach, double LengthScale, double PInf) {
   Acc a(f);
   const AccessorRO<   int, 3> nType = a.ro<int>(Sel<dir>::nType(f));
   const AccessorRO<double, 3> m_d = a.ro<double>(Sel<dir>::m_d(f));
   const AccessorRO<double, 3> rho = a.ro<double>(f->rho);
   const AccessorRO<VecNSp, 3> MassFracs = a.ro<VecNSp>(f->MassFracs);
   const AccessorRO<double, 3> pressure = a.ro<double>(f->pressure);
   const AccessorRO<double, 3> temperature = a.ro<double>(f->temperature);
   const AccessorRO<  Vec3, 3> velocity = a.ro<Vec3>(f->velocity);
   const AccessorRO<double, 3> SoS = a.ro<double>(f->SoS);
   const AccessorRO<VecNEq, 3> Conserved = a.ro<VecNEq>(f->Conserved);
   const AccessorRO<VecNSp, 3> Xi_profile = a.ro<VecNSp>(b->MolarFracs_profile);
   const AccessorRO<double, 3> T_profile = a.ro<double>(b->temperature_profile);
   const AccessorRO<  Vec3, 3> v_profile = a.ro<Vec3>(b->velocity_profile);
   const AccessorRW<VecNEq, 3> Conserved_t = a.rw<VecNEq>(f->Conserved_t);
   HTR_BC_LOOP(lo, hi) {
      const Point<3> p(i, j, k);
      Task::addLODIfluxes(Conserved_t[p], rho, MassFracs, pressure, velocity, nType[p], m_d[p], SoS[p], temperature[p],
                          Conserved[p], T_profile[p], v_profile[p], Xi_profile[p], PInf, MaxMach, LengthScale, p, mix());
   }
}
#endif
} // namespace

extern "C" {
// params: iShock, MolarFracs[nSpec], pressure0, temperature0, velocity0[3], pressure1, temperature1, velocity1[3]
// (struct IncomingShockParams, src/prometeo_bc_types.h:46-59)
static IncomingShockParams shockParams(const double *q) {
   IncomingShockParams P;
   P.iShock = (int)q[0];
   for (int s = 0; s < nSpec; s++) P.MolarFracs[s] = q[1 + s];
   const double *r = q + 1 + nSpec;
   P.pressure0 = r[0]; P.temperature0 = r[1]; for (int i = 0; i < 3; i++) P.velocity0[i] = r[2 + i];
   P.pressure1 = r[5]; P.temperature1 = r[6]; for (int i = 0; i < 3; i++) P.velocity1[i] = r[7 + i];
   return P;
}

#ifndef HTR_BC_PART_SET
#define HTR_DIRSIDE(FUNC, MINUS, PLUS, ...) \
   if (side == 0) { \
      if (dir == 0) FUNC<Xdir, MINUS<Xdir> >(__VA_ARGS__); else if (dir == 1) FUNC<Ydir, MINUS<Ydir> >(__VA_ARGS__); else FUNC<Zdir, MINUS<Zdir> >(__VA_ARGS__); \
   } else { \
      if (dir == 0) FUNC<Xdir, PLUS<Xdir> >(__VA_ARGS__); else if (dir == 1) FUNC<Ydir, PLUS<Ydir> >(__VA_ARGS__); else FUNC<Zdir, PLUS<Zdir> >(__VA_ARGS__); \
   }

void htr_oracle_nscbc_inflow(int dir, int side, const htr_oracle_fields *f, const htr_oracle_bc_fields *b,
                             const int64_t *lo, const int64_t *hi) {
   HTR_DIRSIDE(nscbc_inflow, UpdateUsingFluxNSCBCInflowMinusSideTask, UpdateUsingFluxNSCBCInflowPlusSideTask, f, b, lo, hi)
}
void htr_oracle_nscbc_outflow(int dir, int side, const htr_oracle_fields *f, const int64_t *lo, const int64_t *hi,
                              double MaxMach, double LengthScale, double PInf) {
   HTR_DIRSIDE(nscbc_outflow, UpdateUsingFluxNSCBCOutflowMinusSideTask, UpdateUsingFluxNSCBCOutflowPlusSideTask, f, lo, hi, MaxMach, LengthScale, PInf)
}
void htr_oracle_nscbc_farfield(int dir, int side, const htr_oracle_fields *f, const htr_oracle_bc_fields *b,
                               const int64_t *lo, const int64_t *hi, double MaxMach, double LengthScale, double PInf) {
   HTR_DIRSIDE(nscbc_farfield, UpdateUsingFluxNSCBCFarFieldMinusSideTask, UpdateUsingFluxNSCBCFarFieldPlusSideTask, f, b, lo, hi, MaxMach, LengthScale, PInf)
}

// src/prometeo_rhs.cc:1076-1138: the y-plus boundary, upstream state below iShock, downstream state from iShock on
void htr_oracle_incoming_shock(const htr_oracle_fields *f, const int64_t *lo, const int64_t *hi,
                               double MaxMach, double LengthScale, const double *params) {
   Acc a(f);
   const IncomingShockParams P = shockParams(params);
   const AccessorRO<   int, 3> nType = a.ro<int>(f->nType_y);
   const AccessorRO<double, 3> m_d = a.ro<double>(f->deta_d);
   const AccessorRO<double, 3> rho = a.ro<double>(f->rho);
   const AccessorRO<VecNSp, 3> MassFracs = a.ro<VecNSp>(f->MassFracs);
   const AccessorRO<double, 3> pressure = a.ro<double>(f->pressure);
   const AccessorRO<double, 3> temperature = a.ro<double>(f->temperature);
   const AccessorRO<  Vec3, 3> velocity = a.ro<Vec3>(f->velocity);
   const AccessorRO<double, 3> SoS = a.ro<double>(f->SoS);
   const AccessorRO<VecNEq, 3> Conserved = a.ro<VecNEq>(f->Conserved);
   const AccessorRW<VecNEq, 3> Conserved_t = a.rw<VecNEq>(f->Conserved_t);
   const VecNSp MolarFracs(P.MolarFracs);
   const Vec3 velocity0(P.velocity0), velocity1(P.velocity1);
   HTR_BC_LOOP(lo, hi) {
      const Point<3> p(i, j, k);
      if (i < P.iShock)
         UpdateUsingFluxNSCBCFarFieldPlusSideTask<Ydir>::addLODIfluxes(Conserved_t[p], rho, MassFracs, pressure, velocity,
            nType[p], m_d[p], SoS[p], temperature[p], Conserved[p], P.temperature0, velocity0, MolarFracs,
            P.pressure0, MaxMach, LengthScale, p, mix());
      else
         UpdateUsingFluxNSCBCFarFieldPlusSideTask<Ydir>::addLODIfluxes(Conserved_t[p], rho, MassFracs, pressure, velocity,
            nType[p], m_d[p], SoS[p], temperature[p], Conserved[p], P.temperature1, velocity1, MolarFracs,
            P.pressure1, MaxMach, LengthScale, p, mix());
   }
}

#else
// src/prometeo_bc.cc:80-135 (SetNSCBC_InflowBCTask<dir>::cpu_base_impl)
void htr_oracle_set_nscbc_inflow_bc(int dir, const htr_oracle_fields *f, const htr_oracle_bc_fields *b,
                                    const int64_t *lo, const int64_t *hi, double Pbc) {
   Acc a(f);
   const AccessorRO<VecNEq, 3> Conserved = a.ro<VecNEq>(f->Conserved);
   const AccessorRO<double, 3> SoS = a.ro<double>(f->SoS);
   const AccessorRO<VecNSp, 3> Xi_profile = a.ro<VecNSp>(b->MolarFracs_profile);
   const AccessorRO<double, 3> T_profile = a.ro<double>(b->temperature_profile);
   const AccessorRO<  Vec3, 3> v_profile = a.ro<Vec3>(b->velocity_profile);
   const AccessorWO<double, 3> pressure = a.wo<double>(f->pressure);
   const AccessorWO<double, 3> temperature = a.wo<double>(f->temperature);
   const AccessorWO<VecNSp, 3> MolarFracs = a.wo<VecNSp>(f->MolarFracs);
   const AccessorWO<  Vec3, 3> velocity = a.wo<Vec3>(f->velocity);
   HTR_BC_LOOP(lo, hi) {
      const Point<3> p(i, j, k);
      MolarFracs[p] = Xi_profile[p];
      temperature[p] = T_profile[p];
      velocity[p] = v_profile[p];
      if (fabs(v_profile[p][dir]) >= SoS[p]) pressure[p] = Pbc;
      else SetNSCBC_InflowBCTask<Xdir>::setInflowPressure(Conserved, Xi_profile, T_profile, pressure, p, mix());
   }
}

// src/prometeo_bc.cc:160-196 (SetNSCBC_OutflowBCTask::cpu_base_impl)
void htr_oracle_set_nscbc_outflow_bc(const htr_oracle_fields *f, const int64_t *lo, const int64_t *hi) {
   Acc a(f);
   const AccessorRO<VecNEq, 3> Conserved = a.ro<VecNEq>(f->Conserved);
   const AccessorRW<double, 3> temperature = a.rw<double>(f->temperature);
   const AccessorWO<double, 3> pressure = a.wo<double>(f->pressure);
   const AccessorWO<VecNSp, 3> MolarFracs = a.wo<VecNSp>(f->MolarFracs);
   const AccessorWO<  Vec3, 3> velocity = a.wo<Vec3>(f->velocity);
   HTR_BC_LOOP(lo, hi) {
      const Point<3> p(i, j, k);
      UpdatePrimitiveFromConservedTask::UpdatePrimitive(Conserved, temperature, pressure, MolarFracs, velocity, p, mix());
   }
}

// src/prometeo_bc.cc:203-264 (SetIncomingShockBCTask::cpu_base_impl)
void htr_oracle_set_incoming_shock_bc(const htr_oracle_fields *f, const int64_t *lo, const int64_t *hi, const double *params) {
   Acc a(f);
   const IncomingShockParams P = shockParams(params);
   const AccessorRO<VecNEq, 3> Conserved = a.ro<VecNEq>(f->Conserved);
   const AccessorRW<double, 3> temperature = a.rw<double>(f->temperature);
   const AccessorWO<double, 3> pressure = a.wo<double>(f->pressure);
   const AccessorWO<VecNSp, 3> MolarFracs = a.wo<VecNSp>(f->MolarFracs);
   const AccessorWO<  Vec3, 3> velocity = a.wo<Vec3>(f->velocity);
   HTR_BC_LOOP(lo, hi) {
      const Point<3> p(i, j, k);
      if ((i < P.iShock - 1) or (i > P.iShock + 1)) {
         UpdatePrimitiveFromConservedTask::UpdatePrimitive(Conserved, temperature, pressure, MolarFracs, velocity, p, mix());
      } else if (i == P.iShock - 1) {
         MolarFracs[p]  = VecNSp(P.MolarFracs);
         velocity[p]    = 0.75*Vec3(P.velocity0) + 0.25*Vec3(P.velocity1);
         temperature[p] = 0.75*P.temperature0    + 0.25*P.temperature1;
         pressure[p]    = 0.75*P.pressure0       + 0.25*P.pressure1;
      } else if (i == P.iShock) {
         MolarFracs[p]  = VecNSp(P.MolarFracs);
         velocity[p]    = 0.25*Vec3(P.velocity0) + 0.75*Vec3(P.velocity1);
         temperature[p] = 0.25*P.temperature0    + 0.75*P.temperature1;
         pressure[p]    = 0.25*P.pressure0       + 0.75*P.pressure1;
      } else {
         MolarFracs[p]  = VecNSp(P.MolarFracs);
         velocity[p]    = Vec3(P.velocity1);
         temperature[p] = P.temperature1;
         pressure[p]    = P.pressure1;
      }
   }
}
#endif
} // extern "C"
