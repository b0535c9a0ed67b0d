// TEST INFRASTRUCTURE ONLY (checker, never the product path).
//
// Drives the REFERENCE's own CPU numerics over raw host arrays.  Nothing from
// /root/reference is copied: this file #includes the reference headers where
// they lie (-I /root/reference/src ...) against oracle/stub/{legion.h,
// config_schema.h}; the loops below restate only the trivial cpu_base_impl
// scaffolding (accessor set-up + `omp parallel for` over the tile), citing the
// task body each one follows.  Output: oracle/_ref/libhtr_ref_<Mix>.so
// (git-ignored).  See oracle/Makefile.
#include "legion.h"
using namespace Legion;
#include <math.h>
#include <stdio.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

// prometeo_metric.inl has no include guard and is pulled in by both
// prometeo_rhs.inl and prometeo_sensor.inl, so (like the reference, which builds
// prometeo_rhs.cc and prometeo_sensor.cc separately) this file is compiled twice:
// plain (rhs + mixture), -DHTR_PART_SENSOR and -DHTR_PART_VARS (variables, chem, cfl).
#if defined(HTR_PART_METRIC)
#include "prometeo_metric.hpp"
#elif defined(HTR_PART_SENSOR)
#include "prometeo_sensor.hpp"
#include "prometeo_sensor.inl"
#elif defined(HTR_PART_VARS)
#include "prometeo_variables.hpp"
#include "prometeo_chem.hpp"
#include "prometeo_cfl.hpp"
#else
#include "prometeo_rhs.hpp"
#include "prometeo_rhs.inl"
#endif

#include "htr_oracle_api.h"

#define HTR_STR2(x) #x
#define HTR_STR(x) HTR_STR2(x)

#if defined(HTR_PART_METRIC)
// src/prometeo_metric.cc:36-190 (InitializeMetricTask::cpu_base_impl): the nine metric fields of a box from its cell-centre
// coordinates, with the reference's own reconstructCoordEuler / ComputeDiffusionMetrics (src/prometeo_metric.hpp:113-160);
// the box is the whole (periodic or bounded, per nType) domain.  centerCoordinates: [z][y][x][3] doubles.
extern "C" void htr_oracle_initialize_metric(const htr_oracle_fields *f, const double *centerCoordinates, const double *width) {
   const coord_t nx = f->nx, ny = f->ny, nz = f->nz;
   const Rect<3> B(Point<3>(0, 0, 0), Point<3>(nx - 1, ny - 1, nz - 1));
   const coord_t sv = sizeof(Vec3), sd = sizeof(double), si = sizeof(int);
   const AccessorRO<Vec3, 3> cc((void*)centerCoordinates, sv, sv*nx, sv*nx*ny);
   const AccessorRO<int, 3> tx((void*)f->nType_x, si, si*nx, si*nx*ny), ty((void*)f->nType_y, si, si*nx, si*nx*ny), tz((void*)f->nType_z, si, si*nx, si*nx*ny);
   #define HTR_WO(name) const AccessorWO<double, 3> name((void*)f->name, sd, sd*nx, sd*nx*ny)
   HTR_WO(dcsi_e); HTR_WO(deta_e); HTR_WO(dzet_e); HTR_WO(dcsi_d); HTR_WO(deta_d); HTR_WO(dzet_d); HTR_WO(dcsi_s); HTR_WO(deta_s); HTR_WO(dzet_s);
   #undef HTR_WO
   typedef InitializeMetricTask M;
   const coord_t xsize = getSize<Xdir>(B), ysize = getSize<Ydir>(B), zsize = getSize<Zdir>(B);
   for (int k = 0; k < nz; k++)
      for (int j = 0; j < ny; j++) {
         double xM_e, xP_e;
         {
            const Point<3> p(0, j, k);
            const Point<3> pm1 = warpPeriodic<Xdir, Minus>(B, p, xsize, offM1(tx[p]));
            xM_e = M::reconstructCoordEuler<Xdir>(cc, pm1, width[0], tx[pm1], xsize, B);
            xM_e = unwarpCoordinate<Xdir>(xM_e, width[0], -1, p, B);
         }
         for (int i = 0; i < nx; i++) {
            const Point<3> p(i, j, k);
            xP_e = M::reconstructCoordEuler<Xdir>(cc, p, width[0], tx[p], xsize, B);
            dcsi_e[p] = 1.0/(xP_e - xM_e);
            M::ComputeDiffusionMetrics<Xdir>(dcsi_d, dcsi_s, cc, p, width[0], tx[p], xsize, B);
            xM_e = xP_e;
         }
      }
   for (int k = 0; k < nz; k++)
      for (int i = 0; i < nx; i++) {
         double yM_e, yP_e;
         {
            const Point<3> p(i, 0, k);
            const Point<3> pm1 = warpPeriodic<Ydir, Minus>(B, p, ysize, offM1(ty[p]));
            yM_e = M::reconstructCoordEuler<Ydir>(cc, pm1, width[1], ty[pm1], ysize, B);
            yM_e = unwarpCoordinate<Ydir>(yM_e, width[1], -1, p, B);
         }
         for (int j = 0; j < ny; j++) {
            const Point<3> p(i, j, k);
            yP_e = M::reconstructCoordEuler<Ydir>(cc, p, width[1], ty[p], ysize, B);
            deta_e[p] = 1.0/(yP_e - yM_e);
            M::ComputeDiffusionMetrics<Ydir>(deta_d, deta_s, cc, p, width[1], ty[p], ysize, B);
            yM_e = yP_e;
         }
      }
   for (int j = 0; j < ny; j++)
      for (int i = 0; i < nx; i++) {
         double zM_e, zP_e;
         {
            const Point<3> p(i, j, 0);
            const Point<3> pm1 = warpPeriodic<Zdir, Minus>(B, p, zsize, offM1(tz[p]));
            zM_e = M::reconstructCoordEuler<Zdir>(cc, pm1, width[2], tz[pm1], zsize, B);
            zM_e = unwarpCoordinate<Zdir>(zM_e, width[2], -1, p, B);
         }
         for (int k = 0; k < nz; k++) {
            const Point<3> p(i, j, k);
            zP_e = M::reconstructCoordEuler<Zdir>(cc, p, width[2], tz[p], zsize, B);
            dzet_e[p] = 1.0/(zP_e - zM_e);
            M::ComputeDiffusionMetrics<Zdir>(dzet_d, dzet_s, cc, p, width[2], tz[p], zsize, B);
            zM_e = zP_e;
         }
      }
}
#define HTR_METRIC_ONLY 1
#endif

#ifndef HTR_METRIC_ONLY
#if defined(HTR_PART_SENSOR) || defined(HTR_PART_VARS)
extern Mix *htr_ref_g_mix;
#else
static char mix_storage[sizeof(Mix)] __attribute__((aligned(16)));
Mix *htr_ref_g_mix = NULL;
#endif
#define g_mix htr_ref_g_mix
static inline const Mix &mix() { assert(g_mix); return *g_mix; }

namespace {
struct Acc {
   const htr_oracle_fields *f;
   coord_t nx, ny, nz;
   Rect<3> bounds;
   explicit Acc(const htr_oracle_fields *f_) : f(f_), nx(f_->nx), ny(f_->ny), nz(f_->nz),
      bounds(Point<3>(0, 0, 0), Point<3>(f_->nx - 1, f_->ny - 1, f_->nz - 1)) {}
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

template<direction dir> struct DirSel;
template<> struct DirSel<Xdir> {
   static const int32_t *nType(const htr_oracle_fields *f) { return f->nType_x; }
   static const double  *m_e(const htr_oracle_fields *f) { return f->dcsi_e; }
   static const double  *m_d(const htr_oracle_fields *f) { return f->dcsi_d; }
   static const double  *m_s(const htr_oracle_fields *f) { return f->dcsi_s; }
   static uint8_t *sensor(const htr_oracle_fields *f) { return f->shockSensorX; }
};
template<> struct DirSel<Ydir> {
   static const int32_t *nType(const htr_oracle_fields *f) { return f->nType_y; }
   static const double  *m_e(const htr_oracle_fields *f) { return f->deta_e; }
   static const double  *m_d(const htr_oracle_fields *f) { return f->deta_d; }
   static const double  *m_s(const htr_oracle_fields *f) { return f->deta_s; }
   static uint8_t *sensor(const htr_oracle_fields *f) { return f->shockSensorY; }
};
template<> struct DirSel<Zdir> {
   static const int32_t *nType(const htr_oracle_fields *f) { return f->nType_z; }
   static const double  *m_e(const htr_oracle_fields *f) { return f->dzet_e; }
   static const double  *m_d(const htr_oracle_fields *f) { return f->dzet_d; }
   static const double  *m_s(const htr_oracle_fields *f) { return f->dzet_s; }
   static uint8_t *sensor(const htr_oracle_fields *f) { return f->shockSensorZ; }
};

#if !defined(HTR_PART_SENSOR) && !defined(HTR_PART_VARS)
// The reference task classes keep updateRHSSpan public; derive to reach them.
// Loop structure: src/prometeo_rhs.cc:69-117 (Hybrid), :188-222 (TENO family),
// :341-376 (SkewSymmetric): `omp parallel for collapse(3)` over crossPlane<dir>.
template<direction dir>
void euler_flux(int scheme, const htr_oracle_fields *f, double RK0, double RK1, double deltaTime) {
   Acc a(f);
   const AccessorRW<VecNEq, 3> Conserved_t = a.rw<VecNEq>(f->Conserved_t);
   const AccessorRO<double, 3> m_e   = a.ro<double>(DirSel<dir>::m_e(f));
   const AccessorRO<   int, 3> nType = a.ro<int>(DirSel<dir>::nType(f));
   const AccessorRO<  bool, 3> sens  = a.ro<bool>(DirSel<dir>::sensor(f));
   const AccessorRO<VecNEq, 3> Conserved = a.ro<VecNEq>(f->Conserved);
#if (nSpec > 1)
   const AccessorRO<VecNEq, 3> Conserved_old = a.ro<VecNEq>(f->Conserved_old);
#endif
   const AccessorRO<double, 3> rho = a.ro<double>(f->rho);
   const AccessorRO<double, 3> SoS = a.ro<double>(f->SoS);
   const AccessorRO<VecNSp, 3> MassFracs = a.ro<VecNSp>(f->MassFracs);
   const AccessorRO<  Vec3, 3> velocity = a.ro<Vec3>(f->velocity);
   const AccessorRO<double, 3> pressure = a.ro<double>(f->pressure);
   const Rect<3> r = a.bounds;
   const Rect<3> cPlane = crossPlane<dir>(r);
   const coord_t span = getSize<dir>(r);
   (void)RK0; (void)RK1; (void)deltaTime;
#ifdef _OPENMP
   #pragma omp parallel for collapse(3)
#endif
   for (int k = cPlane.lo.z; k <= cPlane.hi.z; k++)
      for (int j = cPlane.lo.y; j <= cPlane.hi.y; j++)
         for (int i = cPlane.lo.x; i <= cPlane.hi.x; i++) {
            double KGSum[6*(nEq+1)];
            const int x = i - cPlane.lo.x, y = j - cPlane.lo.y, z = k - cPlane.lo.z;
            switch (scheme) {
            case HTR_SCHEME_HYBRID:
               UpdateUsingHybridEulerFluxTask<dir>::updateRHSSpan(KGSum,
                  Conserved_t, m_e, nType, sens, Conserved,
#if (nSpec > 1)
                  Conserved_old,
#endif
                  rho, SoS, MassFracs, velocity, pressure,
#if (nSpec > 1)
                  RK0, RK1, deltaTime,
#endif
                  0, span, x, y, z, r, r, mix());
               break;
#define HTR_TENO_CASE(ID, OP) \
            case ID: \
               UpdateUsingTENOEulerFluxTask<dir, OP>::updateRHSSpan( \
                  Conserved_t, m_e, nType, Conserved, HTR_OLD \
                  rho, SoS, MassFracs, velocity, pressure, HTR_RK \
                  0, span, x, y, z, r, r, mix()); \
               break;
#if (nSpec > 1)
#define HTR_OLD Conserved_old,
#define HTR_RK RK0, RK1, deltaTime,
#else
#define HTR_OLD
#define HTR_RK
#endif
            HTR_TENO_CASE(HTR_SCHEME_TENO, TENO_Op<>)
            HTR_TENO_CASE(HTR_SCHEME_TENOA, TENOA_Op<>)
            HTR_TENO_CASE(HTR_SCHEME_TENOLAD, TENOLAD_Op<>)
#undef HTR_TENO_CASE
#undef HTR_OLD
#undef HTR_RK
            case HTR_SCHEME_SKEWSYMMETRIC:
               UpdateUsingSkewSymmetricEulerFluxTask<dir>::updateRHSSpan(KGSum,
                  Conserved_t, m_e, nType, Conserved,
                  rho, MassFracs, velocity, pressure,
                  0, span, x, y, z, r, r);
               break;
            }
         }
}

// src/prometeo_rhs.cc:410-467
template<direction dir>
void diffusion_flux(const htr_oracle_fields *f) {
   Acc a(f);
   const AccessorRW<VecNEq, 3> Conserved_t = a.rw<VecNEq>(f->Conserved_t);
   const AccessorRO<double, 3> m_s  = a.ro<double>(DirSel<dir>::m_s(f));
   const AccessorRO<double, 3> m_d  = a.ro<double>(DirSel<dir>::m_d(f));
   const AccessorRO<double, 3> m_d1 = a.ro<double>(DirSel<getT1(dir)>::m_d(f));
   const AccessorRO<double, 3> m_d2 = a.ro<double>(DirSel<getT2(dir)>::m_d(f));
   const AccessorRO<int, 3> nType  = a.ro<int>(DirSel<dir>::nType(f));
   const AccessorRO<int, 3> nType1 = a.ro<int>(DirSel<getT1(dir)>::nType(f));
   const AccessorRO<int, 3> nType2 = a.ro<int>(DirSel<getT2(dir)>::nType(f));
   const AccessorRO<double, 3> rho = a.ro<double>(f->rho);
   const AccessorRO<double, 3> mu  = a.ro<double>(f->mu);
   const AccessorRO<double, 3> lam = a.ro<double>(f->lam);
   const AccessorRO<VecNSp, 3> Di  = a.ro<VecNSp>(f->Di);
   const AccessorRO<double, 3> temperature = a.ro<double>(f->temperature);
   const AccessorRO<  Vec3, 3> velocity = a.ro<Vec3>(f->velocity);
   const AccessorRO<VecNSp, 3> Xi = a.ro<VecNSp>(f->MolarFracs);
   const AccessorRO<VecNEq, 3> Conserved = a.ro<VecNEq>(f->Conserved);
   const Rect<3> r = a.bounds;
   const Rect<3> cPlane = crossPlane<dir>(r);
   const coord_t span = getSize<dir>(r);
#ifdef _OPENMP
   #pragma omp parallel for collapse(3)
#endif
   for (int k = cPlane.lo.z; k <= cPlane.hi.z; k++)
      for (int j = cPlane.lo.y; j <= cPlane.hi.y; j++)
         for (int i = cPlane.lo.x; i <= cPlane.hi.x; i++)
            UpdateUsingDiffusionFluxTask<dir>::updateRHSSpan(Conserved_t, m_s, m_d, m_d1, m_d2,
               nType, nType1, nType2, rho, mu, lam, Di, temperature, velocity, Xi, Conserved,
               0, span, i - cPlane.lo.x, j - cPlane.lo.y, k - cPlane.lo.z, r, r, mix());
}

#elif defined(HTR_PART_SENSOR)
// src/prometeo_sensor.cc:83-138
template<direction dir>
void shock_sensor(const htr_oracle_fields *f) {
   Acc a(f);
   const AccessorRO<VecNEq, 3> Conserved = a.ro<VecNEq>(f->Conserved);
   const AccessorRO<double, 3> Ducros = a.ro<double>(f->DucrosSensor);
   const AccessorRO<int, 3> nType = a.ro<int>(DirSel<dir>::nType(f));
   const AccessorWO<bool, 3> out = a.wo<bool>(DirSel<dir>::sensor(f));
   const Rect<3> B = a.bounds;
   const coord_t size = getSize<dir>(B);
#ifdef _OPENMP
   #pragma omp parallel for collapse(3)
#endif
   for (int k = 0; k < a.nz; k++)
      for (int j = 0; j < a.ny; j++)
         for (int i = 0; i < a.nx; i++) {
            const Point<3> p(i, j, k);
            const int t = nType[p];
            const Point<3> pM2 = warpPeriodic<dir, Minus>(B, p, size, offM2(t));
            const Point<3> pM1 = warpPeriodic<dir, Minus>(B, p, size, offM1(t));
            const Point<3> pP1 = warpPeriodic<dir, Plus >(B, p, size, offP1(t));
            const Point<3> pP2 = warpPeriodic<dir, Plus >(B, p, size, offP2(t));
            const Point<3> pP3 = warpPeriodic<dir, Plus >(B, p, size, offP3(t));
            const double Phi = std::max(std::max(std::max(std::max(std::max(
                                 Ducros[pM2], Ducros[pM1]), Ducros[p]), Ducros[pP1]), Ducros[pP2]), Ducros[pP3]);
            bool sensor = true;
            for (int h = 0; h < nSpec; h++)
               sensor = sensor && TENOsensor::TENOA(Conserved[pM2][h], Conserved[pM1][h], Conserved[p][h],
                                                    Conserved[pP1][h], Conserved[pP2][h], Conserved[pP3][h], t, Phi);
            out[p] = sensor;
         }
}
#endif
} // namespace

extern "C" {
#if !defined(HTR_PART_SENSOR) && !defined(HTR_PART_VARS)

int htr_oracle_nspec(void) { return nSpec; }
int htr_oracle_neq(void) { return nEq; }
const char *htr_oracle_kind(void) { return "reference"; }
const char *htr_oracle_mixture(void) { return HTR_STR(EOS); }
int htr_oracle_sizeof_mix(void) { return (int)sizeof(Mix); }
int htr_oracle_max_threads(void) {
#ifdef _OPENMP
   return omp_get_max_threads();
#else
   return 1;
#endif
}

void htr_oracle_set_threads(int n) {
#ifdef _OPENMP
   if (n > 0) omp_set_num_threads(n);       // overrides an inherited OMP_NUM_THREADS=1 (torchrun sets it)
#else
   (void)n;
#endif
}

int htr_oracle_init_constprop(double gasConstant, double gamma, double prandtl,
                              int visc_model, double p0, double p1, double p2) {
#ifdef HTR_EOS_IS_CONSTPROP
   static Config config;
   memset(&config, 0, sizeof(config));
   config.Flow.mixture.type = MixtureModel_ConstPropMix;
   config.Flow.mixture.u.ConstPropMix.gasConstant = gasConstant;
   config.Flow.mixture.u.ConstPropMix.gamma = gamma;
   config.Flow.mixture.u.ConstPropMix.prandtl = prandtl;
   config.Flow.mixture.u.ConstPropMix.viscosityModel.type = visc_model;
   if (visc_model == ViscosityModel_Constant)
      config.Flow.mixture.u.ConstPropMix.viscosityModel.u.Constant.Visc = p0;
   else if (visc_model == ViscosityModel_PowerLaw) {
      config.Flow.mixture.u.ConstPropMix.viscosityModel.u.PowerLaw.ViscRef = p0;
      config.Flow.mixture.u.ConstPropMix.viscosityModel.u.PowerLaw.TempRef = p1;
   } else {
      config.Flow.mixture.u.ConstPropMix.viscosityModel.u.Sutherland.ViscRef = p0;
      config.Flow.mixture.u.ConstPropMix.viscosityModel.u.Sutherland.TempRef = p1;
      config.Flow.mixture.u.ConstPropMix.viscosityModel.u.Sutherland.SRef = p2;
   }
   g_mix = new (mix_storage) Mix(config);
   return 0;
#else
   (void)gasConstant; (void)gamma; (void)prandtl; (void)visc_model; (void)p0; (void)p1; (void)p2;
   return -1;
#endif
}

int htr_oracle_init_multicomp(double LRef, double PRef, double TRef,
                              int n, const char *names, const double *molarFracs) {
#ifdef HTR_EOS_IS_CONSTPROP
   (void)LRef; (void)PRef; (void)TRef; (void)n; (void)names; (void)molarFracs;
   return -1;
#else
   static Config config;
   memset(&config, 0, sizeof(config));
   config.Flow.mixture.type = HTR_MIXTURE_MODEL;
   MultiCompRef &m = config.Flow.mixture.u.HTR_MIXTURE_MEMBER;
   m.LRef = LRef; m.PRef = PRef; m.TRef = TRef;
   m.XiRef.Species.length = n;
   for (int i = 0; i < n; i++) {
      strncpy(m.XiRef.Species.values[i].Name, names + 10*i, 10);
      m.XiRef.Species.values[i].Name[9] = 0;
      m.XiRef.Species.values[i].MolarFrac = molarFracs[i];
   }
   g_mix = new (mix_storage) Mix(config);
   return 0;
#endif
}

void htr_oracle_mix_eval(double P, double T, const double *Yi_, double *out) {
   const Mix &m = mix();
   VecNSp Yi; for (int i = 0; i < nSpec; i++) Yi[i] = Yi_[i];
   const double MixW = m.GetMolarWeightFromYi(Yi);
   VecNSp Xi; m.GetMolarFractions(Xi, MixW, Yi);
   const double rho = m.GetRho(P, T, MixW);
   const double gamma = m.GetGamma(T, MixW, Yi);
   const double e = m.GetInternalEnergy(T, Yi);
   for (int i = 0; i < 16 + 4*nSpec; i++) out[i] = 0.0;
   out[0] = MixW; out[1] = rho; out[2] = m.GetHeatCapacity(T, Yi); out[3] = m.GetEnthalpy(T, Yi);
   out[4] = e; out[5] = m.GetViscosity(T, Xi); out[6] = m.GetHeatConductivity(T, Xi);
   out[7] = gamma; out[8] = m.GetSpeedOfSound(T, gamma, MixW); out[9] = m.Getdpde(rho, gamma);
   out[10] = m.GetMolarWeightFromXi(Xi);
   out[11] = m.GetTFromInternalEnergy(e, T + 0.05*T, Yi);
   VecNSp Di; m.GetDiffusivity(Di, P, T, MixW, Xi);
   VecNSp w;  m.GetProductionRates(w, rho, P, T, Yi);
   VecNSp dp; m.Getdpdrhoi(dp, gamma, T, Yi);
   for (int i = 0; i < nSpec; i++) {
      out[16 + i] = Xi[i]; out[16 + nSpec + i] = Di[i];
      out[16 + 2*nSpec + i] = w[i]; out[16 + 3*nSpec + i] = dp[i];
   }
}
// Raw bytes of the reference's `struct Mix` (the POD the reference ships in every task's Args,
// src/prometeo_rhs.hpp:183) and the size of its leading constant species/reaction tables.
int htr_oracle_mix_bytes(void *out) { memcpy(out, &mix(), sizeof(Mix)); return (int)sizeof(Mix); }
int htr_oracle_mix_table_bytes(void) {
#ifdef HTR_EOS_IS_CONSTPROP
   return 0;
#else
   return (int)((const char*)&mix().PRef - (const char*)&mix());
#endif
}
double htr_oracle_species_enthalpy(int i, double T) { return mix().GetSpeciesEnthalpy(i, T); }
double htr_oracle_species_W(int i) { return mix().GetSpeciesMolarWeight(i); }

void htr_oracle_euler_flux(int scheme, int dir, const htr_oracle_fields *f,
                           double RK0, double RK1, double deltaTime) {
   if (dir == 0) euler_flux<Xdir>(scheme, f, RK0, RK1, deltaTime);
   else if (dir == 1) euler_flux<Ydir>(scheme, f, RK0, RK1, deltaTime);
   else euler_flux<Zdir>(scheme, f, RK0, RK1, deltaTime);
}
void htr_oracle_diffusion_flux(int dir, const htr_oracle_fields *f) {
   if (dir == 0) diffusion_flux<Xdir>(f);
   else if (dir == 1) diffusion_flux<Ydir>(f);
   else diffusion_flux<Zdir>(f);
}
// unitTests/mathUtilsTest/CPPmathUtilsTest.cc:6-29
void htr_oracle_lu3_solve(const double *A, const double *b, double *x) {
   MyMatrix<double, 3, 3> M;
   for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) M(i, j) = A[3*i + j];
   LUdec<3> LU; LU.ludcmp(M);
   MyArray<double, 3> v; for (int i = 0; i < 3; i++) v[i] = b[i];
   LU.lubksb(v);
   for (int i = 0; i < 3; i++) x[i] = v[i];
}
// unitTests/mathUtilsTest/CPPmathUtilsTest.cc:31-55
namespace { class RobertsonLike : public Rosenbrock<3> {
   inline void rhs(MyArray<double, 3> &r, const MyArray<double, 3> &x) {
      r[0] = -0.013* x[0] - 1000.0*x[0]*x[2];
      r[1] = -2500.0*x[1]*x[2];
      r[2] = -0.013* x[0] - 1000.0*x[0]*x[2] - 2500.0*x[1]*x[2];
   };
}; }
void htr_oracle_rosenbrock3(double *x, double dtTry, double DelT) {
   RobertsonLike p; MyArray<double, 3> v; for (int i = 0; i < 3; i++) v[i] = x[i];
   p.solve(v, dtTry, DelT);
   for (int i = 0; i < 3; i++) x[i] = v[i];
}
#elif defined(HTR_PART_VARS)
// src/prometeo_chem.cc:78-112
void htr_oracle_add_chemistry_sources(const htr_oracle_fields *f) {
   Acc a(f);
   const AccessorRO<double, 3> rho = a.ro<double>(f->rho), P = a.ro<double>(f->pressure), T = a.ro<double>(f->temperature);
   const AccessorRO<VecNSp, 3> Y = a.ro<VecNSp>(f->MassFracs);
   const AccessorRW<VecNEq, 3> Ct = a.rw<VecNEq>(f->Conserved_t);
#ifdef _OPENMP
   #pragma omp parallel for collapse(3)
#endif
   for (int k = 0; k < a.nz; k++)
      for (int j = 0; j < a.ny; j++)
         for (int i = 0; i < a.nx; i++) {
            const Point<3> p(i, j, k);
            VecNSp w; mix().GetProductionRates(w, rho[p], P[p], T[p], Y[p]);
            for (int s = 0; s < nSpec; s++) Ct[p][s] += w[s];
         }
}

// src/prometeo_chem.cc:36-71
void htr_oracle_update_chemistry(const htr_oracle_fields *f, double dt) {
   Acc a(f);
   const AccessorRO<VecNEq, 3> Ct = a.ro<VecNEq>(f->Conserved_t);
   const AccessorRW<VecNEq, 3> C = a.rw<VecNEq>(f->Conserved), Cto = a.rw<VecNEq>(f->Conserved_t_old);
   const AccessorRW<double, 3> T = a.rw<double>(f->temperature);
#ifdef _OPENMP
   #pragma omp parallel for collapse(3) schedule(dynamic, 16)
#endif
   for (int k = 0; k < a.nz; k++)
      for (int j = 0; j < a.ny; j++)
         for (int i = 0; i < a.nx; i++) {
            const Point<3> p(i, j, k);
            Cto[p] = Ct[p];
            ImplicitSolver s = ImplicitSolver(Cto[p], T[p], mix());
            s.solve(C[p], dt, dt);
         }
}

// The same task with the number of rhs() evaluations of every cell reported: (2 nEq + 1) per integration step plus 3
// per step attempt, i.e. a signature of the adaptive step / retry sequence (tests compare it with the GPU kernel's).
// rhs() below restates ImplicitSolver::rhs (src/prometeo_chem.hpp:66-87), which is private there.
namespace { class CountingSolver : public Rosenbrock<nEq> {
public:
   CountingSolver(const VecNEq &S_, double &T_, const Mix &m_) : calls(0), S(S_), temperature(T_), m(m_) {}
   int calls;
private:
   inline void rhs(VecNEq &r, const VecNEq &x) {
      calls++;
      VecNSp rhoYi;
      for (int i = 0; i < nSpec; i++) rhoYi[i] = x[i];
      const double rho = m.GetRhoFromRhoYi(rhoYi);
      VecNSp Yi; m.GetYi(Yi, rho, rhoYi);
      m.ClipYi(Yi);
      const double MixW = m.GetMolarWeightFromYi(Yi);
      const double rhoInv = 1.0/rho;
      Vec3 velocity;
      for (int i = 0; i < 3; i++) velocity[i] = x[i + irU]*rhoInv;
      const double kineticEnergy = 0.5*velocity.mod2();
      const double InternalEnergy = x[irE]*rhoInv - kineticEnergy;
      temperature = m.GetTFromInternalEnergy(InternalEnergy, temperature, Yi);
      const double pressure = m.GetPFromRhoAndT(rho, MixW, temperature);
      VecNSp w; m.GetProductionRates(w, rho, pressure, temperature, Yi);
      for (int i = 0; i < nSpec; i++) r[i] = w[i] + S[i];
      for (int i = nSpec; i < nEq; i++) r[i] = S[i];
   }
   const VecNEq &S; double &temperature; const Mix &m;
}; }
void htr_oracle_update_chemistry_counted(const htr_oracle_fields *f, double dt, int *rhs_calls) {
   Acc a(f);
   const AccessorRO<VecNEq, 3> Ct = a.ro<VecNEq>(f->Conserved_t);
   const AccessorRW<VecNEq, 3> C = a.rw<VecNEq>(f->Conserved), Cto = a.rw<VecNEq>(f->Conserved_t_old);
   const AccessorRW<double, 3> T = a.rw<double>(f->temperature);
#ifdef _OPENMP
   #pragma omp parallel for collapse(3) schedule(dynamic, 16)
#endif
   for (int k = 0; k < a.nz; k++)
      for (int j = 0; j < a.ny; j++)
         for (int i = 0; i < a.nx; i++) {
            const Point<3> p(i, j, k);
            Cto[p] = Ct[p];
            CountingSolver s(Cto[p], T[p], mix());
            s.solve(C[p], dt, dt);
            rhs_calls[(k*a.ny + j)*a.nx + i] = s.calls;
         }
}

// src/prometeo_variables.cc:138-173
void htr_oracle_update_primitive_from_conserved(const htr_oracle_fields *f) {
   Acc a(f);
   const AccessorRO<VecNEq, 3> C = a.ro<VecNEq>(f->Conserved);
   const AccessorRW<double, 3> T = a.rw<double>(f->temperature);
   const AccessorWO<double, 3> P = a.wo<double>(f->pressure);
   const AccessorWO<VecNSp, 3> X = a.wo<VecNSp>(f->MolarFracs);
   const AccessorWO<Vec3, 3> v = a.wo<Vec3>(f->velocity);
#ifdef _OPENMP
   #pragma omp parallel for collapse(3)
#endif
   for (int k = 0; k < a.nz; k++)
      for (int j = 0; j < a.ny; j++)
         for (int i = 0; i < a.nx; i++)
            UpdatePrimitiveFromConservedTask::UpdatePrimitive(C, T, P, X, v, Point<3>(i, j, k), mix());
}

// src/prometeo_variables.cc:36-86
void htr_oracle_update_properties_from_primitive(const htr_oracle_fields *f) {
   Acc a(f);
   const AccessorRO<double, 3> P = a.ro<double>(f->pressure), T = a.ro<double>(f->temperature);
   const AccessorRO<VecNSp, 3> X = a.ro<VecNSp>(f->MolarFracs);
   const AccessorRO<Vec3, 3> v = a.ro<Vec3>(f->velocity);
   const AccessorWO<VecNSp, 3> Y = a.wo<VecNSp>(f->MassFracs), Di = a.wo<VecNSp>(f->Di);
   const AccessorWO<double, 3> rho = a.wo<double>(f->rho), mu = a.wo<double>(f->mu),
                               lam = a.wo<double>(f->lam), SoS = a.wo<double>(f->SoS);
#ifdef _OPENMP
   #pragma omp parallel for collapse(3)
#endif
   for (int k = 0; k < a.nz; k++)
      for (int j = 0; j < a.ny; j++)
         for (int i = 0; i < a.nx; i++)
            UpdatePropertiesFromPrimitiveTask::UpdateProperties(P, T, X, v, Y, rho, mu, lam, Di, SoS, Point<3>(i, j, k), mix());
}

// src/prometeo_variables.cc:92-132
void htr_oracle_update_conserved_from_primitive(const htr_oracle_fields *f) {
   Acc a(f);
   const AccessorRO<VecNSp, 3> Y = a.ro<VecNSp>(f->MassFracs);
   const AccessorRO<double, 3> T = a.ro<double>(f->temperature), rho = a.ro<double>(f->rho);
   const AccessorRO<Vec3, 3> v = a.ro<Vec3>(f->velocity);
   const AccessorWO<VecNEq, 3> C = a.wo<VecNEq>(f->Conserved);
#ifdef _OPENMP
   #pragma omp parallel for collapse(3)
#endif
   for (int k = 0; k < a.nz; k++)
      for (int j = 0; j < a.ny; j++)
         for (int i = 0; i < a.nx; i++)
            UpdateConservedFromPrimitiveTask::UpdateConserved(Y, T, v, rho, C, Point<3>(i, j, k), mix());
}

// src/prometeo_cfl.cc:36-93 (max over the tile; the sum-reduction is order independent)
double htr_oracle_max_spectral_radius(const htr_oracle_fields *f) {
   Acc a(f);
   const AccessorRO<double, 3> dx = a.ro<double>(f->dcsi_d), dy = a.ro<double>(f->deta_d), dz = a.ro<double>(f->dzet_d);
   const AccessorRO<double, 3> T = a.ro<double>(f->temperature), rho = a.ro<double>(f->rho), mu = a.ro<double>(f->mu),
                               lam = a.ro<double>(f->lam), SoS = a.ro<double>(f->SoS);
   const AccessorRO<VecNSp, 3> Y = a.ro<VecNSp>(f->MassFracs), Di = a.ro<VecNSp>(f->Di);
   const AccessorRO<Vec3, 3> v = a.ro<Vec3>(f->velocity);
   double r = 0.0;
#ifdef _OPENMP
   #pragma omp parallel for collapse(3) reduction(max:r)
#endif
   for (int k = 0; k < a.nz; k++)
      for (int j = 0; j < a.ny; j++)
         for (int i = 0; i < a.nx; i++)
            r = std::max(r, CalculateMaxSpectralRadiusTask::CalculateMaxSpectralRadius(
                               dx, dy, dz, T, Y, v, rho, mu, lam, Di, SoS, Point<3>(i, j, k), mix()));
   return r;
}

#else
void htr_oracle_shock_sensor(int dir, const htr_oracle_fields *f) {
   if (dir == 0) shock_sensor<Xdir>(f);
   else if (dir == 1) shock_sensor<Ydir>(f);
   else shock_sensor<Zdir>(f);
}

// src/prometeo_sensor.cc:37-79
void htr_oracle_ducros_sensor(const htr_oracle_fields *f, double vorticityScale) {
   Acc a(f);
   const AccessorRO<Vec3, 3> velocity = a.ro<Vec3>(f->velocity);
   const AccessorRO<int, 3> tx = a.ro<int>(f->nType_x), ty = a.ro<int>(f->nType_y), tz = a.ro<int>(f->nType_z);
   const AccessorRO<double, 3> mx = a.ro<double>(f->dcsi_d), my = a.ro<double>(f->deta_d), mz = a.ro<double>(f->dzet_d);
   const AccessorWO<double, 3> out = a.wo<double>(f->DucrosSensor);
   const double eps = std::max(vorticityScale*vorticityScale, 1e-6);
#ifdef _OPENMP
   #pragma omp parallel for collapse(3)
#endif
   for (int k = 0; k < a.nz; k++)
      for (int j = 0; j < a.ny; j++)
         for (int i = 0; i < a.nx; i++) {
            const Point<3> p(i, j, k);
            out[p] = UpdateDucrosSensorTask::DucrosSensor(velocity, tx, ty, tz, mx, my, mz, p, a.bounds, eps);
         }
}

#endif
} // extern "C"
#endif // HTR_METRIC_ONLY
