// TEST INFRASTRUCTURE ONLY.  Executes the Legion adapter (htr-solver_b200/csrc/legion_adapter.cc) outside Legion:
// builds, for every task the adapter implements, the std::vector<PhysicalRegion> the reference's region requirements
// produce -- region ORDER and the FIELDS of each region exactly as the reference's own task bodies open them
// (src/prometeo_rhs.cc:43-71,163-190,322-342,416-446; src/prometeo_sensor.cc:43-59,89-100; src/prometeo_chem.cc:42-52,
// 84-94; src/prometeo_variables.cc:42-60,98-113,144-157; src/prometeo_cfl.cc:42-63) -- fills the task's Args and calls
// T::gpu_base_impl.  The regions wrap DEVICE pointers (oracle/stub/legion.h: a field that is not part of a region, a
// ptr() outside the instance, or the Rect of a sparse ghost domain abort, like Legion would), so a wrong region index or
// field id in the adapter fails here instead of in a production run.  tests/test_adapter_gpu.py compares every task with
// the direct C-ABI call, bit for bit.
#include "legion.h"
using namespace Legion;
#include "prometeo_rhs.hpp"
#include "prometeo_sensor.hpp"
#include "prometeo_chem.hpp"
#include "prometeo_variables.hpp"
#include "prometeo_cfl.hpp"
#include "prometeo_bc.hpp"
#include "htr_b200.h"
#include <initializer_list>
#include <utility>

namespace {
typedef std::pair<FieldID, const htr_acc*> FA;

PhysicalRegion region(const htr_rect &b, bool dense, std::initializer_list<FA> fields) {
   PhysicalRegion r;
   r.is = IndexSpace(Rect<3>(Point<3>(b.lo[0], b.lo[1], b.lo[2]), Point<3>(b.hi[0], b.hi[1], b.hi[2])), dense);
   for (const FA &f : fields) {
      StubField s;
      s.fid = f.first; s.base = f.second->base;
      for (int i = 0; i < 3; i++) s.s[i] = f.second->stride[i];
      r.fields.push_back(s);
   }
   return r;
}
Rect<3> toLegion(const htr_rect &b) { return Rect<3>(Point<3>(b.lo[0], b.lo[1], b.lo[2]), Point<3>(b.hi[0], b.hi[1], b.hi[2])); }

struct DirF { FieldID nType, m_e, m_d, m_s, sensor; const htr_acc *aNType, *aMe, *aMd, *aMs, *aSensor; };
DirF dirFields(const htr_fluid &f, int d) {
   if (d == 0) return {FID_nType_x, FID_dcsi_e, FID_dcsi_d, FID_dcsi_s, FID_shockSensorX, &f.nType_x, &f.dcsi_e, &f.dcsi_d, &f.dcsi_s, &f.shockSensorX};
   if (d == 1) return {FID_nType_y, FID_deta_e, FID_deta_d, FID_deta_s, FID_shockSensorY, &f.nType_y, &f.deta_e, &f.deta_d, &f.deta_s, &f.shockSensorY};
   return {FID_nType_z, FID_dzet_e, FID_dzet_d, FID_dzet_s, FID_shockSensorZ, &f.nType_z, &f.dzet_e, &f.dzet_d, &f.dzet_s, &f.shockSensorZ};
}

template<class T> void fillMix(T &args, const void *mix) { memcpy((void*)&args.mix, mix, sizeof(Mix)); }
// Args holds a Mix (no default constructor): zeroed raw storage, like the byte buffer Regent hands to the task
template<class A> struct RawArgs {
   alignas(16) unsigned char b[sizeof(A)];
   RawArgs() { memset(b, 0, sizeof(b)); }
   A &get() { return *reinterpret_cast<A*>(b); }
};

// region 0 of the TENO / Hybrid tasks (EulerGhost); MassFracs is part of it for Hybrid only
PhysicalRegion eulerGhost(const htr_fluid &f, const htr_rect &alloc, bool withY) {
   PhysicalRegion r = region(alloc, false, {{FID_Conserved, &f.Conserved}, {FID_SoS, &f.SoS}, {FID_rho, &f.rho},
                                            {FID_velocity, &f.velocity}, {FID_pressure, &f.pressure}});
#if (nSpec > 1)
   { PhysicalRegion o = region(alloc, false, {{FID_Conserved_old, &f.Conserved_old}}); r.fields.push_back(o.fields[0]); }
#endif
   if (withY) { PhysicalRegion o = region(alloc, false, {{FID_MassFracs, &f.MassFracs}}); r.fields.push_back(o.fields[0]); }
   return r;
}

template<direction dir>
int runDir(int task, const htr_fluid &f, const htr_rect &tile, const htr_rect &alloc, const htr_rect &B, const void *mix,
           double s0, double s1, double s2) {
   const int d = (dir == Xdir) ? 0 : (dir == Ydir ? 1 : 2);
   const DirF D = dirFields(f, d);
   std::vector<PhysicalRegion> R;
   std::vector<Future> fut;
   Runtime rt;
   switch (task) {
   case 0: {   // UpdateUsingHybridEulerFlux, src/prometeo_rhs.cc:43-71
      R.push_back(eulerGhost(f, alloc, true));
      R.push_back(region(alloc, false, {{D.sensor, D.aSensor}, {D.m_e, D.aMe}}));
      R.push_back(region(alloc, false, {{D.nType, D.aNType}}));
      R.push_back(region(tile, true, {{FID_Conserved_t, &f.Conserved_t}}));
RawArgs<      typename UpdateUsingHybridEulerFluxTask<dir>::Args> raw; auto &a = raw.get();
      a.RK_coeffs[0] = s0; a.RK_coeffs[1] = s1; a.Fluid_bounds = toLegion(B); fillMix(a, mix);
      Future ft; ft.v = s2; fut.push_back(ft);
      UpdateUsingHybridEulerFluxTask<dir>::gpu_base_impl(a, R, fut, 0, &rt);
      return 0; }
   case 1: case 2: case 3: {   // UpdateUsingTENOEulerFlux<dir, Op>, src/prometeo_rhs.cc:163-190
      R.push_back(eulerGhost(f, alloc, false));
      R.push_back(region(alloc, false, {{FID_MassFracs, &f.MassFracs}, {D.m_e, D.aMe}}));
      R.push_back(region(alloc, false, {{D.nType, D.aNType}}));
      R.push_back(region(tile, true, {{FID_Conserved_t, &f.Conserved_t}}));
      Future ft; ft.v = s2; fut.push_back(ft);
      #define HTR_RUN_TENO(Op) {RawArgs< typename UpdateUsingTENOEulerFluxTask<dir, Op>::Args> raw; auto &a = raw.get(); \
         a.RK_coeffs[0] = s0; a.RK_coeffs[1] = s1; a.Fluid_bounds = toLegion(B); fillMix(a, mix); \
         UpdateUsingTENOEulerFluxTask<dir, Op>::gpu_base_impl(a, R, fut, 0, &rt); }
      if (task == 1) HTR_RUN_TENO(TENO_Op<>) else if (task == 2) HTR_RUN_TENO(TENOA_Op<>) else HTR_RUN_TENO(TENOLAD_Op<>)
      #undef HTR_RUN_TENO
      return 0; }
   case 4: {   // UpdateUsingSkewSymmetricEulerFlux, src/prometeo_rhs.cc:322-342
      R.push_back(region(alloc, false, {{FID_Conserved, &f.Conserved}, {FID_rho, &f.rho}, {FID_MassFracs, &f.MassFracs},
                                        {FID_velocity, &f.velocity}, {FID_pressure, &f.pressure}}));
      R.push_back(region(alloc, false, {{D.nType, D.aNType}}));
      R.push_back(region(alloc, false, {{D.m_e, D.aMe}}));
      R.push_back(region(tile, true, {{FID_Conserved_t, &f.Conserved_t}}));
RawArgs<      typename UpdateUsingSkewSymmetricEulerFluxTask<dir>::Args> raw; auto &a = raw.get();
      a.Fluid_bounds = toLegion(B); fillMix(a, mix);
      UpdateUsingSkewSymmetricEulerFluxTask<dir>::gpu_base_impl(a, R, fut, 0, &rt);
      return 0; }
   case 5: {   // UpdateUsingDiffusionFlux, src/prometeo_rhs.cc:416-446 (five regions: Fluid appears read-only and read-write)
      const DirF D1 = dirFields(f, d == 0 ? 1 : 0), D2 = dirFields(f, d == 2 ? 1 : 2);
      R.push_back(region(alloc, false, {{FID_Conserved, &f.Conserved}, {FID_temperature, &f.temperature}, {FID_MolarFracs, &f.MolarFracs},
                                        {FID_rho, &f.rho}, {FID_mu, &f.mu}, {FID_lam, &f.lam}, {FID_Di, &f.Di},
                                        {D1.nType, D1.aNType}, {D2.nType, D2.aNType}, {D1.m_d, D1.aMd}, {D2.m_d, D2.aMd}}));
      R.push_back(region(alloc, false, {{FID_velocity, &f.velocity}}));
      R.push_back(region(alloc, false, {{D.nType, D.aNType}, {D.m_s, D.aMs}}));
      R.push_back(region(tile, true, {{D.m_d, D.aMd}}));
      R.push_back(region(tile, true, {{FID_Conserved_t, &f.Conserved_t}}));
RawArgs<      typename UpdateUsingDiffusionFluxTask<dir>::Args> raw; auto &a = raw.get();
      a.Fluid_bounds = toLegion(B); fillMix(a, mix);
      UpdateUsingDiffusionFluxTask<dir>::gpu_base_impl(a, R, fut, 0, &rt);
      return 0; }
   case 7: {   // UpdateShockSensor, src/prometeo_sensor.cc:89-100
      R.push_back(region(alloc, false, {{FID_Conserved, &f.Conserved}, {FID_DucrosSensor, &f.DucrosSensor}}));
      R.push_back(region(alloc, false, {{D.nType, D.aNType}}));
      R.push_back(region(tile, true, {{D.sensor, D.aSensor}}));
RawArgs<      typename UpdateShockSensorTask<dir>::Args> raw; auto &a = raw.get();
      a.Fluid_bounds = toLegion(B);
      UpdateShockSensorTask<dir>::gpu_base_impl(a, R, fut, 0, &rt);
      return 0; }
   }
   return -1;
}
} // namespace

extern "C" {

// task: 0 Hybrid 1 TENO 2 TENOA 3 TENOLAD 4 SkewSymmetric Euler flux, 5 diffusion flux, 6 Ducros sensor, 7 shock sensor,
// 8 AddChemistrySources, 9 UpdateChemistry, 10 UpdatePropertiesFromPrimitive, 11 UpdateConservedFromPrimitive,
// 12 UpdatePrimitiveFromConserved, 13 CalculateMaxSpectralRadius.
// alloc = rectangle of the allocation (tile + ghosts): the bounding box of the (sparse) ghost regions.
// s0, s1, s2: RK_coeffs[0], RK_coeffs[1], deltaTime | vorticityScale | Integrator_deltaTime.  d_result: device double (task 13).
int htr_adapter_run(int task, int dir, const htr_fluid *f, const htr_rect *tile, const htr_rect *alloc, const htr_rect *bounds,
                    const void *mix, double s0, double s1, double s2, void *stream, double *d_result) {
   Realm::Cuda::stub_task_stream() = stream;
   DeferredValue<double>::stub_storage() = d_result;
   if (task <= 5 || task == 7) {
      if (dir == 0) return runDir<Xdir>(task, *f, *tile, *alloc, *bounds, mix, s0, s1, s2);
      if (dir == 1) return runDir<Ydir>(task, *f, *tile, *alloc, *bounds, mix, s0, s1, s2);
      return runDir<Zdir>(task, *f, *tile, *alloc, *bounds, mix, s0, s1, s2);
   }
   std::vector<PhysicalRegion> R;
   std::vector<Future> fut;
   Runtime rt;
   switch (task) {
   case 6: {   // UpdateDucrosSensor, src/prometeo_sensor.cc:43-59
      R.push_back(region(*alloc, false, {{FID_velocity, &f->velocity}}));
      R.push_back(region(*tile, true, {{FID_nType_x, &f->nType_x}, {FID_nType_y, &f->nType_y}, {FID_nType_z, &f->nType_z},
                                       {FID_dcsi_d, &f->dcsi_d}, {FID_deta_d, &f->deta_d}, {FID_dzet_d, &f->dzet_d}}));
      R.push_back(region(*tile, true, {{FID_DucrosSensor, &f->DucrosSensor}}));
RawArgs<      UpdateDucrosSensorTask::Args> raw; auto &a = raw.get();
      a.Fluid_bounds = toLegion(*bounds); a.vorticityScale = s0;
      UpdateDucrosSensorTask::gpu_base_impl(a, R, fut, 0, &rt);
      return 0; }
   case 8: {   // AddChemistrySources, src/prometeo_chem.cc:84-94
      R.push_back(region(*tile, true, {{FID_rho, &f->rho}, {FID_pressure, &f->pressure}, {FID_temperature, &f->temperature}, {FID_MassFracs, &f->MassFracs}}));
      R.push_back(region(*tile, true, {{FID_Conserved_t, &f->Conserved_t}}));
RawArgs<      AddChemistrySourcesTask::Args> raw; auto &a = raw.get(); fillMix(a, mix);
      AddChemistrySourcesTask::gpu_base_impl(a, R, fut, 0, &rt);
      return 0; }
   case 9: {   // UpdateChemistry, src/prometeo_chem.cc:42-57 (Integrator_deltaTime is a future)
      R.push_back(region(*tile, true, {{FID_Conserved_t, &f->Conserved_t}}));
      R.push_back(region(*tile, true, {{FID_Conserved, &f->Conserved}, {FID_Conserved_t_old, &f->Conserved_t_old}, {FID_temperature, &f->temperature}}));
      Future ft; ft.v = s0; fut.push_back(ft);
RawArgs<      UpdateChemistryTask::Args> raw; auto &a = raw.get(); fillMix(a, mix);
      UpdateChemistryTask::gpu_base_impl(a, R, fut, 0, &rt);
      return 0; }
   case 10: {  // UpdatePropertiesFromPrimitive, src/prometeo_variables.cc:42-60
      R.push_back(region(*tile, true, {{FID_pressure, &f->pressure}, {FID_temperature, &f->temperature}, {FID_MolarFracs, &f->MolarFracs}, {FID_velocity, &f->velocity}}));
      R.push_back(region(*tile, true, {{FID_MassFracs, &f->MassFracs}, {FID_rho, &f->rho}, {FID_mu, &f->mu}, {FID_lam, &f->lam}, {FID_Di, &f->Di}, {FID_SoS, &f->SoS}}));
RawArgs<      UpdatePropertiesFromPrimitiveTask::Args> raw; auto &a = raw.get(); fillMix(a, mix);
      UpdatePropertiesFromPrimitiveTask::gpu_base_impl(a, R, fut, 0, &rt);
      return 0; }
   case 11: {  // UpdateConservedFromPrimitive, src/prometeo_variables.cc:98-131: the launch domain is a union of rectangles
      R.push_back(region(*tile, true, {{FID_MassFracs, &f->MassFracs}, {FID_temperature, &f->temperature}, {FID_velocity, &f->velocity}, {FID_rho, &f->rho}}));
      PhysicalRegion w = region(*tile, false, {{FID_Conserved, &f->Conserved}});
      // three pieces: the low z plane, the high z plane, and the low-y face of the planes in between (BC-like shells)
      Rect<3> t = toLegion(*tile);
      Rect<3> a1 = t; a1.hi.z = t.lo.z;
      Rect<3> a2 = t; a2.lo.z = t.hi.z;
      w.is.pieces.push_back(a1);
      if (t.hi.z > t.lo.z) w.is.pieces.push_back(a2);
      if (t.hi.z > t.lo.z + 1) { Rect<3> a3 = t; a3.lo.z = t.lo.z + 1; a3.hi.z = t.hi.z - 1; a3.hi.y = t.lo.y; w.is.pieces.push_back(a3); }
      R.push_back(w);
RawArgs<      UpdateConservedFromPrimitiveTask::Args> raw; auto &a = raw.get(); fillMix(a, mix);
      UpdateConservedFromPrimitiveTask::gpu_base_impl(a, R, fut, 0, &rt);
      return 0; }
   case 12: {  // UpdatePrimitiveFromConserved, src/prometeo_variables.cc:144-157
      R.push_back(region(*tile, true, {{FID_Conserved, &f->Conserved}}));
      R.push_back(region(*tile, true, {{FID_temperature, &f->temperature}, {FID_pressure, &f->pressure}, {FID_MolarFracs, &f->MolarFracs}, {FID_velocity, &f->velocity}}));
RawArgs<      UpdatePrimitiveFromConservedTask::Args> raw; auto &a = raw.get(); fillMix(a, mix);
      UpdatePrimitiveFromConservedTask::gpu_base_impl(a, R, fut, 0, &rt);
      return 0; }
   case 13: {  // CalculateMaxSpectralRadius, src/prometeo_cfl.cc:42-63
      R.push_back(region(*tile, true, {{FID_dcsi_d, &f->dcsi_d}, {FID_deta_d, &f->deta_d}, {FID_dzet_d, &f->dzet_d}, {FID_MassFracs, &f->MassFracs},
                                       {FID_temperature, &f->temperature}, {FID_velocity, &f->velocity}, {FID_rho, &f->rho}, {FID_mu, &f->mu},
                                       {FID_lam, &f->lam}, {FID_Di, &f->Di}, {FID_SoS, &f->SoS}}));
RawArgs<      CalculateMaxSpectralRadiusTask::Args> raw; auto &a = raw.get(); fillMix(a, mix);
      CalculateMaxSpectralRadiusTask::gpu_base_impl(a, R, fut, 0, &rt);
      return 0; }
   }
   return -1;
}

// Boundary tasks.  task: 0 UpdateUsingFluxNSCBCInflow, 1 ...Outflow, 2 ...FarField (side 0 minus / 1 plus), 3 UpdateUsingFluxIncomingShock
// (y-plus), 4 SetNSCBC_InflowBC, 5 SetNSCBC_OutflowBC, 6 SetIncomingShockBC.  bc = the plane of boundary cells (the written region);
// regions and their fields as the reference's bodies open them (src/prometeo_rhs.cc:538-556,692-716,914-934,1085-1100;
// src/prometeo_bc.cc:89-110,167-180,210-223).  params: the flattened IncomingShockParams of include/htr_b200.h.
} // extern "C"
namespace {
void unflattenShock(IncomingShockParams &P, const double *q) {
   P.iShock = (int)q[0];
   for (int s = 0; s < nSpec; s++) P.MolarFracs[s] = q[1 + s];
   const double *r = q + 1 + nSpec;
   P.pressure0 = r[0]; P.temperature0 = r[1]; for (int i = 0; i < 3; i++) P.velocity0[i] = r[2 + i];
   P.pressure1 = r[5]; P.temperature1 = r[6]; for (int i = 0; i < 3; i++) P.velocity1[i] = r[7 + i];
}
template<direction dir>
int runBC(int task, int side, const htr_fluid &f, const htr_bc_fields &b, const htr_rect &bc, const htr_rect &alloc, const htr_rect &B,
          const void *mix, double MaxMach, double LengthScale, double P, const double *params) {
   const int d = (dir == Xdir) ? 0 : (dir == Ydir ? 1 : 2);
   const DirF D = dirFields(f, d), D1 = dirFields(f, d == 0 ? 1 : 0), D2 = dirFields(f, d == 2 ? 1 : 2);
   std::vector<PhysicalRegion> R;
   std::vector<Future> fut;
   Runtime rt;
   Future ft; ft.v = MaxMach;
   #define HTR_RUN_SIDE(MINUS, PLUS, FILL) \
      if (side == 0) { RawArgs<typename MINUS<dir>::Args> raw; auto &a = raw.get(); fillMix(a, mix); FILL; MINUS<dir>::gpu_base_impl(a, R, fut, 0, &rt); } \
      else           { RawArgs<typename PLUS<dir>::Args> raw; auto &a = raw.get(); fillMix(a, mix); FILL; PLUS<dir>::gpu_base_impl(a, R, fut, 0, &rt); }
   switch (task) {
   case 0:
      R.push_back(region(alloc, false, {{FID_MassFracs, &f.MassFracs}, {FID_pressure, &f.pressure}, {FID_velocity, &f.velocity}}));
      R.push_back(region(bc, true, {{D.nType, D.aNType}, {D.m_d, D.aMd}, {FID_rho, &f.rho}, {FID_SoS, &f.SoS}, {FID_temperature, &f.temperature},
                                    {FID_dudtBoundary, &b.dudtBoundary}, {FID_dTdtBoundary, &b.dTdtBoundary}}));
      R.push_back(region(bc, true, {{FID_Conserved_t, &f.Conserved_t}}));
      HTR_RUN_SIDE(UpdateUsingFluxNSCBCInflowMinusSideTask, UpdateUsingFluxNSCBCInflowPlusSideTask, (void)0)
      return 0;
   case 1:
      R.push_back(region(alloc, false, {{FID_velocity, &f.velocity}}));
      R.push_back(region(alloc, false, {{D.nType, D.aNType}, {D1.nType, D1.aNType}, {D2.nType, D2.aNType}, {D.m_d, D.aMd}, {D1.m_d, D1.aMd}, {D2.m_d, D2.aMd},
                                        {FID_rho, &f.rho}, {FID_mu, &f.mu}, {FID_MassFracs, &f.MassFracs}, {FID_pressure, &f.pressure}}));
      R.push_back(region(bc, true, {{FID_SoS, &f.SoS}, {FID_temperature, &f.temperature}, {FID_Conserved, &f.Conserved}}));
      R.push_back(region(bc, true, {{FID_Conserved_t, &f.Conserved_t}}));
      fut.push_back(ft);
      HTR_RUN_SIDE(UpdateUsingFluxNSCBCOutflowMinusSideTask, UpdateUsingFluxNSCBCOutflowPlusSideTask,
                   a.Fluid_bounds = toLegion(B); a.LengthScale = LengthScale; a.PInf = P)
      return 0;
   case 2:
      R.push_back(region(alloc, false, {{FID_rho, &f.rho}, {FID_MassFracs, &f.MassFracs}, {FID_pressure, &f.pressure},
                                        {FID_temperature, &f.temperature}, {FID_velocity, &f.velocity}}));
      R.push_back(region(bc, true, {{D.nType, D.aNType}, {D.m_d, D.aMd}, {FID_SoS, &f.SoS}, {FID_Conserved, &f.Conserved},
                                    {FID_MolarFracs_profile, &b.MolarFracs_profile}, {FID_temperature_profile, &b.temperature_profile},
                                    {FID_velocity_profile, &b.velocity_profile}}));
      R.push_back(region(bc, true, {{FID_Conserved_t, &f.Conserved_t}}));
      fut.push_back(ft);
      HTR_RUN_SIDE(UpdateUsingFluxNSCBCFarFieldMinusSideTask, UpdateUsingFluxNSCBCFarFieldPlusSideTask, a.LengthScale = LengthScale; a.PInf = P)
      return 0;
   case 4: {
      R.push_back(region(bc, true, {{FID_Conserved, &f.Conserved}, {FID_SoS, &f.SoS}, {FID_MolarFracs_profile, &b.MolarFracs_profile},
                                    {FID_temperature_profile, &b.temperature_profile}, {FID_velocity_profile, &b.velocity_profile}}));
      R.push_back(region(bc, true, {{FID_pressure, &f.pressure}, {FID_temperature, &f.temperature}, {FID_MolarFracs, &f.MolarFracs}, {FID_velocity, &f.velocity}}));
      RawArgs<typename SetNSCBC_InflowBCTask<dir>::Args> raw; auto &a = raw.get(); fillMix(a, mix); a.Pbc = P;
      SetNSCBC_InflowBCTask<dir>::gpu_base_impl(a, R, fut, 0, &rt);
      return 0; }
   }
   #undef HTR_RUN_SIDE
   (void)params;
   return -1;
}
} // namespace

extern "C" {
int htr_adapter_run_bc(int task, int dir, int side, const htr_fluid *f, const htr_bc_fields *b, const htr_rect *bc, const htr_rect *alloc,
                       const htr_rect *bounds, const void *mix, double MaxMach, double LengthScale, double P, const double *params, void *stream) {
   Realm::Cuda::stub_task_stream() = stream;
   if (task <= 2 || task == 4) {
      if (dir == 0) return runBC<Xdir>(task, side, *f, *b, *bc, *alloc, *bounds, mix, MaxMach, LengthScale, P, params);
      if (dir == 1) return runBC<Ydir>(task, side, *f, *b, *bc, *alloc, *bounds, mix, MaxMach, LengthScale, P, params);
      return runBC<Zdir>(task, side, *f, *b, *bc, *alloc, *bounds, mix, MaxMach, LengthScale, P, params);
   }
   std::vector<PhysicalRegion> R;
   std::vector<Future> fut;
   Runtime rt;
   switch (task) {
   case 3: {
      R.push_back(region(*alloc, false, {{FID_rho, &f->rho}, {FID_MassFracs, &f->MassFracs}, {FID_pressure, &f->pressure},
                                         {FID_temperature, &f->temperature}, {FID_velocity, &f->velocity}}));
      R.push_back(region(*bc, true, {{FID_nType_y, &f->nType_y}, {FID_deta_d, &f->deta_d}, {FID_SoS, &f->SoS}, {FID_Conserved, &f->Conserved}}));
      R.push_back(region(*bc, true, {{FID_Conserved_t, &f->Conserved_t}}));
      Future ft; ft.v = MaxMach; fut.push_back(ft);
      RawArgs<UpdateUsingFluxIncomingShockTask<Ydir>::Args> raw; auto &a = raw.get(); fillMix(a, mix);
      a.LengthScale = LengthScale; unflattenShock(a.params, params);
      UpdateUsingFluxIncomingShockTask<Ydir>::gpu_base_impl(a, R, fut, 0, &rt);
      return 0; }
   case 5: case 6: {
      R.push_back(region(*bc, true, {{FID_Conserved, &f->Conserved}}));
      R.push_back(region(*bc, true, {{FID_temperature, &f->temperature}, {FID_pressure, &f->pressure}, {FID_MolarFracs, &f->MolarFracs}, {FID_velocity, &f->velocity}}));
      if (task == 5) {
         RawArgs<SetNSCBC_OutflowBCTask::Args> raw; auto &a = raw.get(); fillMix(a, mix);
         SetNSCBC_OutflowBCTask::gpu_base_impl(a, R, fut, 0, &rt);
      } else {
         RawArgs<SetIncomingShockBCTask::Args> raw; auto &a = raw.get(); fillMix(a, mix); unflattenShock(a.params, params);
         SetIncomingShockBCTask::gpu_base_impl(a, R, fut, 0, &rt);
      }
      return 0; }
   }
   return -1;
}

int htr_adapter_sizeof_mix(void) { return (int)sizeof(Mix); }

} // extern "C"
