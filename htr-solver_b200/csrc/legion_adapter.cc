// Legion adapter: the reference-side binding of the C ABI in include/htr_b200.h.
//
// It is compiled INSIDE the reference tree in place of the flux / sensor / chemistry / variables /
// cfl task bodies of src/prometeo_{rhs,sensor,chem,variables,cfl}.cu: it defines the same
// `T::gpu_base_impl(args, regions, futures, ctx, runtime)` members that the reference's own
// registration (TaskHelper::register_hybrid_variants<T>, src/Utils/task_helper.hpp:186-211, called
// from register_{rhs,sensor,chem,variables,cfl}_tasks) binds to the TOC_PROC variant of each task
// id -- task ids, Args layouts, region order and the registrar stay untouched.  Each body only
// unpacks its PhysicalRegions into affine accessors and calls one htr_* function; no numerics here.
// Region / field order per task: src/prometeo_rhs.cu:108-131,:247-269,:409-425,:540-566,
// src/prometeo_sensor.cu, src/prometeo_chem.cu, src/prometeo_variables.cu, src/prometeo_cfl.cu:92-110.
// Legion itself is not available here: the adapter is compiled against oracle/stub/legion.h (`make -C oracle adapter`) and
// EXECUTED against that stub by tests/test_adapter_gpu.py (oracle/adapter_driver.cc builds the PhysicalRegions of every
// task the way the reference's region requirements list them, with sparse ghost regions); it has never run inside Legion.
#include "legion.h"
using namespace Legion;
#include "prometeo_rhs.hpp"
#include "prometeo_sensor.hpp"
#include "prometeo_chem.hpp"
#include "prometeo_variables.hpp"
#include "prometeo_cfl.hpp"
#include "prometeo_bc.hpp"
#include "htr_b200.h"

namespace {
typedef const std::vector<PhysicalRegion> Regions;

// Every field is addressed through the DENSE rectangle of the written region (the task's execution domain): the ghost
// regions are unions of shifted / wrapped pieces (src/prometeo_partitioner.rg:1239+), their domains are sparse and must
// never be converted to a Rect (Legion aborts) -- the reference indexes ghost accessors by point for the same reason.
// All fields of a tile live in one instance that contains the tile, so ptr(tile) is valid for each of them.
struct Unpacker {
   Regions &regions; Rect<3> tile;
   Unpacker(Regions &r, int written, Context ctx, Runtime *runtime) : regions(r),
      tile(runtime->get_index_space_domain(ctx, r[written].get_logical_region().get_index_space())) {}
   // base pointer of global point (0,0,0) + byte strides out of Legion's affine accessor
   template<typename ACC, typename FT> htr_acc get(int reg, FieldID fid) const {
      const ACC a(regions[reg], fid);
      size_t s[3];
      FT *p = a.ptr(tile, s);                  // pointer to tile.lo, strides in elements of FT
      htr_acc o;
      for (int i = 0; i < 3; i++) o.stride[i] = (int64_t)(s[i]*sizeof(FT));
      o.base = (char*)p - (tile.lo.x*o.stride[0] + tile.lo.y*o.stride[1] + tile.lo.z*o.stride[2]);
      return o;
   }
   template<typename FT> htr_acc ro(int reg, FieldID fid) const { return get<AccessorRO<FT, 3>, FT>(reg, fid); }
   template<typename FT> htr_acc rw(int reg, FieldID fid) const { return get<AccessorRW<FT, 3>, FT>(reg, fid); }
   template<typename FT> htr_acc wo(int reg, FieldID fid) const { return get<AccessorWO<FT, 3>, FT>(reg, fid); }
};

htr_rect toRect(const Rect<3> &r) { return htr_rect{{r.lo.x, r.lo.y, r.lo.z}, {r.hi.x, r.hi.y, r.hi.z}}; }
htr_fluid emptyFluid() { htr_fluid f; memset(&f, 0, sizeof(f)); return f; }
void ok(int rc) { assert(rc == HTR_OK); (void)rc; }
// The stream the task's work must be ordered on.  With Realm's CUDA runtime hijack every launch on the default stream is
// redirected to the task's stream (what the reference relies on, src/prometeo_rhs.cu:146); without it the task stream has
// to be asked for.
void *taskStream() {
#ifdef REALM_USE_CUDART_HIJACK
   return 0;
#else
   return (void*)Realm::Cuda::get_task_cuda_stream();
#endif
}

void loadMix(const Mix &m) {         // args.mix is byte-identical to the library's parameter block; once per DEVICE
   if (!htr_mixture_loaded()) ok(htr_set_mixture(&m, sizeof(Mix)));
}

// direction-dependent members of htr_fluid and field ids (src/prometeo_rhs.cc:1100-1200 FID_nType etc.)
struct DirFields { htr_acc htr_fluid::*nType, htr_fluid::*m_e, htr_fluid::*m_d, htr_fluid::*m_s, htr_fluid::*sensor;
                   FieldID fNType, fMe, fMd, fMs, fSensor; };
const DirFields DIR[3] = {
   {&htr_fluid::nType_x, &htr_fluid::dcsi_e, &htr_fluid::dcsi_d, &htr_fluid::dcsi_s, &htr_fluid::shockSensorX,
    FID_nType_x, FID_dcsi_e, FID_dcsi_d, FID_dcsi_s, FID_shockSensorX},
   {&htr_fluid::nType_y, &htr_fluid::deta_e, &htr_fluid::deta_d, &htr_fluid::deta_s, &htr_fluid::shockSensorY,
    FID_nType_y, FID_deta_e, FID_deta_d, FID_deta_s, FID_shockSensorY},
   {&htr_fluid::nType_z, &htr_fluid::dzet_e, &htr_fluid::dzet_d, &htr_fluid::dzet_s, &htr_fluid::shockSensorZ,
    FID_nType_z, FID_dzet_e, FID_dzet_d, FID_dzet_s, FID_shockSensorZ}};
constexpr int dirIndex(direction d) { return d == Xdir ? 0 : (d == Ydir ? 1 : 2); }

// fields every Euler task reads from regions[0] (EulerGhost)
// (MassFracs comes with regions[0] for the Hybrid and SkewSymmetric tasks, with regions[1] for the TENO tasks:
// src/prometeo_rhs.cc:53 vs :177 -- asking regions[0] for it there is a Legion runtime error)
void eulerGhost(htr_fluid &f, const Unpacker &u, bool teno, int massFracsRegion = 0) {
   f.Conserved = u.ro<VecNEq>(0, FID_Conserved);
   f.rho       = u.ro<double>(0, FID_rho);
   f.MassFracs = u.ro<VecNSp>(massFracsRegion, FID_MassFracs);
   f.velocity  = u.ro<Vec3>(0, FID_velocity);
   f.pressure  = u.ro<double>(0, FID_pressure);
   if (teno) {
      f.SoS = u.ro<double>(0, FID_SoS);
#if (nSpec > 1)
      f.Conserved_old = u.ro<VecNEq>(0, FID_Conserved_old);
#endif
   }
}
} // namespace

// ---- Euler fluxes ------------------------------------------------------------------------------
template<direction dir>
void UpdateUsingHybridEulerFluxTask<dir>::gpu_base_impl(const Args &args, Regions &regions,
                      const std::vector<Future> &futures, Context ctx, Runtime *runtime) {
   assert(regions.size() == 4);
   const Unpacker u(regions, 3, ctx, runtime);
   const DirFields &D = DIR[dirIndex(dir)];
   htr_fluid f = emptyFluid();
   eulerGhost(f, u, true);
   f.*D.sensor = u.ro<bool>(1, D.fSensor);
   f.*D.m_e    = u.ro<double>(1, D.fMe);
   f.*D.nType  = u.ro<int>(2, D.fNType);
   f.Conserved_t = u.rw<VecNEq>(3, FID_Conserved_t);
   loadMix(args.mix);
   const htr_rect tile = toRect(u.tile), B = toRect(args.Fluid_bounds);
   double dt = 0.0;
#if (nSpec > 1)
   dt = futures[0].get_result<double>();
#endif
   (void)futures;
   ok(htr_euler_flux(HTR_SCHEME_HYBRID, dirIndex(dir), &f, &tile, &B, args.RK_coeffs[0], args.RK_coeffs[1], dt, taskStream()));
}

template<class Op> struct SchemeOf;
template<> struct SchemeOf<TENO_Op<> >    { enum { id = HTR_SCHEME_TENO }; };
template<> struct SchemeOf<TENOA_Op<> >   { enum { id = HTR_SCHEME_TENOA }; };
template<> struct SchemeOf<TENOLAD_Op<> > { enum { id = HTR_SCHEME_TENOLAD }; };

template<direction dir, class Op>
void UpdateUsingTENOEulerFluxTask<dir, Op>::gpu_base_impl(const Args &args, Regions &regions,
                      const std::vector<Future> &futures, Context ctx, Runtime *runtime) {
   assert(regions.size() == 4);
   const Unpacker u(regions, 3, ctx, runtime);
   const DirFields &D = DIR[dirIndex(dir)];
   htr_fluid f = emptyFluid();
   eulerGhost(f, u, true, 1);
   f.*D.m_e   = u.ro<double>(1, D.fMe);
   f.*D.nType = u.ro<int>(2, D.fNType);
   f.Conserved_t = u.rw<VecNEq>(3, FID_Conserved_t);
   loadMix(args.mix);
   const htr_rect tile = toRect(u.tile), B = toRect(args.Fluid_bounds);
   double dt = 0.0;
#if (nSpec > 1)
   dt = futures[0].get_result<double>();
#endif
   (void)futures;
   ok(htr_euler_flux(SchemeOf<Op>::id, dirIndex(dir), &f, &tile, &B, args.RK_coeffs[0], args.RK_coeffs[1], dt, taskStream()));
}

template<direction dir>
void UpdateUsingSkewSymmetricEulerFluxTask<dir>::gpu_base_impl(const Args &args, Regions &regions,
                      const std::vector<Future> &futures, Context ctx, Runtime *runtime) {
   assert(regions.size() == 4);
   (void)futures;
   const Unpacker u(regions, 3, ctx, runtime);
   const DirFields &D = DIR[dirIndex(dir)];
   htr_fluid f = emptyFluid();
   eulerGhost(f, u, false);
   f.*D.nType = u.ro<int>(1, D.fNType);
   f.*D.m_e   = u.ro<double>(2, D.fMe);
   f.Conserved_t = u.rw<VecNEq>(3, FID_Conserved_t);
   loadMix(args.mix);
   const htr_rect tile = toRect(u.tile), B = toRect(args.Fluid_bounds);
   ok(htr_euler_flux(HTR_SCHEME_SKEWSYMMETRIC, dirIndex(dir), &f, &tile, &B, 1.0, 0.0, 0.0, taskStream()));
}

// ---- diffusion flux (5 regions, src/prometeo_rhs.cu:540-566) -------------------------------------
template<direction dir>
void UpdateUsingDiffusionFluxTask<dir>::gpu_base_impl(const Args &args, Regions &regions,
                      const std::vector<Future> &futures, Context ctx, Runtime *runtime) {
   assert(regions.size() == 5);
   (void)futures;
   const Unpacker u(regions, 4, ctx, runtime);
   const int d = dirIndex(dir), d1 = dirIndex(getT1(dir)), d2 = dirIndex(getT2(dir));
   htr_fluid f = emptyFluid();
   f.Conserved   = u.ro<VecNEq>(0, FID_Conserved);
   f.temperature = u.ro<double>(0, FID_temperature);
   f.MolarFracs  = u.ro<VecNSp>(0, FID_MolarFracs);
   f.rho = u.ro<double>(0, FID_rho); f.mu = u.ro<double>(0, FID_mu); f.lam = u.ro<double>(0, FID_lam);
   f.Di  = u.ro<VecNSp>(0, FID_Di);
   f.*DIR[d1].nType = u.ro<int>(0, DIR[d1].fNType);  f.*DIR[d2].nType = u.ro<int>(0, DIR[d2].fNType);
   f.*DIR[d1].m_d   = u.ro<double>(0, DIR[d1].fMd);  f.*DIR[d2].m_d   = u.ro<double>(0, DIR[d2].fMd);
   f.velocity = u.ro<Vec3>(1, FID_velocity);
   f.*DIR[d].nType = u.ro<int>(2, DIR[d].fNType);
   f.*DIR[d].m_s   = u.ro<double>(2, DIR[d].fMs);
   f.*DIR[d].m_d   = u.ro<double>(3, DIR[d].fMd);
   f.Conserved_t = u.rw<VecNEq>(4, FID_Conserved_t);
   loadMix(args.mix);
   const htr_rect tile = toRect(u.tile), B = toRect(args.Fluid_bounds);
   ok(htr_diffusion_flux(d, &f, &tile, &B, taskStream()));
}

// ---- sensors -------------------------------------------------------------------------------------
void UpdateDucrosSensorTask::gpu_base_impl(const Args &args, Regions &regions,
                      const std::vector<Future> &futures, Context ctx, Runtime *runtime) {
   assert(regions.size() == 3);
   (void)futures;
   const Unpacker u(regions, 2, ctx, runtime);
   htr_fluid f = emptyFluid();
   f.velocity = u.ro<Vec3>(0, FID_velocity);
   for (int d = 0; d < 3; d++) { f.*DIR[d].nType = u.ro<int>(1, DIR[d].fNType); f.*DIR[d].m_d = u.ro<double>(1, DIR[d].fMd); }
   f.DucrosSensor = u.wo<double>(2, FID_DucrosSensor);
   const htr_rect tile = toRect(u.tile), B = toRect(args.Fluid_bounds);
   ok(htr_ducros_sensor(&f, &tile, &B, args.vorticityScale, taskStream()));
}

template<direction dir>
void UpdateShockSensorTask<dir>::gpu_base_impl(const Args &args, Regions &regions,
                      const std::vector<Future> &futures, Context ctx, Runtime *runtime) {
   assert(regions.size() == 3);
   (void)futures;
   const Unpacker u(regions, 2, ctx, runtime);
   const DirFields &D = DIR[dirIndex(dir)];
   htr_fluid f = emptyFluid();
   f.Conserved    = u.ro<VecNEq>(0, FID_Conserved);
   f.DucrosSensor = u.ro<double>(0, FID_DucrosSensor);
   f.*D.nType  = u.ro<int>(1, D.fNType);
   f.*D.sensor = u.wo<bool>(2, D.fSensor);
   const htr_rect tile = toRect(u.tile), B = toRect(args.Fluid_bounds);
   ok(htr_shock_sensor(dirIndex(dir), &f, &tile, &B, taskStream()));
}

// ---- chemistry -----------------------------------------------------------------------------------
void AddChemistrySourcesTask::gpu_base_impl(const Args &args, Regions &regions,
                      const std::vector<Future> &futures, Context ctx, Runtime *runtime) {
   assert(regions.size() == 2);
   (void)futures;
   const Unpacker u(regions, 1, ctx, runtime);
   htr_fluid f = emptyFluid();
   f.rho = u.ro<double>(0, FID_rho); f.pressure = u.ro<double>(0, FID_pressure);
   f.temperature = u.ro<double>(0, FID_temperature); f.MassFracs = u.ro<VecNSp>(0, FID_MassFracs);
   f.Conserved_t = u.rw<VecNEq>(1, FID_Conserved_t);
   loadMix(args.mix);
   const htr_rect tile = toRect(u.tile);
   ok(htr_add_chemistry_sources(&f, &tile, taskStream()));
}

void UpdateChemistryTask::gpu_base_impl(const Args &args, Regions &regions,
                      const std::vector<Future> &futures, Context ctx, Runtime *runtime) {
   assert(regions.size() == 2);
   const Unpacker u(regions, 1, ctx, runtime);
   htr_fluid f = emptyFluid();
   f.Conserved_t     = u.ro<VecNEq>(0, FID_Conserved_t);
   f.Conserved       = u.rw<VecNEq>(1, FID_Conserved);
   f.Conserved_t_old = u.rw<VecNEq>(1, FID_Conserved_t_old);
   f.temperature     = u.rw<double>(1, FID_temperature);
   loadMix(args.mix);
   const htr_rect tile = toRect(u.tile);
   ok(htr_update_chemistry(&f, &tile, futures[0].get_result<double>(), taskStream()));
}

// ---- variables -----------------------------------------------------------------------------------
void UpdatePropertiesFromPrimitiveTask::gpu_base_impl(const Args &args, Regions &regions,
                      const std::vector<Future> &futures, Context ctx, Runtime *runtime) {
   assert(regions.size() == 2);
   (void)futures;
   const Unpacker u(regions, 1, ctx, runtime);
   htr_fluid f = emptyFluid();
   f.pressure = u.ro<double>(0, FID_pressure); f.temperature = u.ro<double>(0, FID_temperature);
   f.MolarFracs = u.ro<VecNSp>(0, FID_MolarFracs); f.velocity = u.ro<Vec3>(0, FID_velocity);
   f.MassFracs = u.wo<VecNSp>(1, FID_MassFracs); f.rho = u.wo<double>(1, FID_rho); f.mu = u.wo<double>(1, FID_mu);
   f.lam = u.wo<double>(1, FID_lam); f.Di = u.wo<VecNSp>(1, FID_Di); f.SoS = u.wo<double>(1, FID_SoS);
   loadMix(args.mix);
   const htr_rect tile = toRect(u.tile);
   ok(htr_update_properties_from_primitive(&f, &tile, taskStream()));
}

void UpdatePrimitiveFromConservedTask::gpu_base_impl(const Args &args, Regions &regions,
                      const std::vector<Future> &futures, Context ctx, Runtime *runtime) {
   assert(regions.size() == 2);
   (void)futures;
   const Unpacker u(regions, 1, ctx, runtime);
   htr_fluid f = emptyFluid();
   f.Conserved = u.ro<VecNEq>(0, FID_Conserved);
   f.temperature = u.rw<double>(1, FID_temperature); f.pressure = u.wo<double>(1, FID_pressure);
   f.MolarFracs = u.wo<VecNSp>(1, FID_MolarFracs); f.velocity = u.wo<Vec3>(1, FID_velocity);
   loadMix(args.mix);
   const htr_rect tile = toRect(u.tile);
   ok(htr_update_primitive_from_conserved(&f, &tile, taskStream()));
}
// The launch domain of UpdateConservedFromPrimitive is a union of rectangles (boundary-condition cells,
// src/prometeo_variables.cc:113-131, src/prometeo_variables.cu:184): one call per rectangle, each addressed on its own.
void UpdateConservedFromPrimitiveTask::gpu_base_impl(const Args &args, Regions &regions,
                      const std::vector<Future> &futures, Context ctx, Runtime *runtime) {
   assert(regions.size() == 2);
   (void)futures;
   loadMix(args.mix);
   const AccessorRO<VecNSp, 3> aY(regions[0], FID_MassFracs);
   const AccessorRO<double, 3> aT(regions[0], FID_temperature);
   const AccessorRO<  Vec3, 3> aU(regions[0], FID_velocity);
   const AccessorRO<double, 3> aR(regions[0], FID_rho);
   const AccessorWO<VecNEq, 3> aC(regions[1], FID_Conserved);
   const Domain dom = runtime->get_index_space_domain(ctx, regions[1].get_logical_region().get_index_space());
   for (RectInDomainIterator<3> Rit(dom); Rit(); Rit++) {
      const Rect<3> r = *Rit;
      htr_fluid f = emptyFluid();
      size_t s[3];
      #define HTR_PIECE(dst, acc, FT) { FT *p = (FT*)acc.ptr(r, s); \
         for (int i = 0; i < 3; i++) f.dst.stride[i] = (int64_t)(s[i]*sizeof(FT)); \
         f.dst.base = (char*)p - (r.lo.x*f.dst.stride[0] + r.lo.y*f.dst.stride[1] + r.lo.z*f.dst.stride[2]); }
      HTR_PIECE(MassFracs, aY, VecNSp) HTR_PIECE(temperature, aT, double) HTR_PIECE(velocity, aU, Vec3)
      HTR_PIECE(rho, aR, double) HTR_PIECE(Conserved, aC, VecNEq)
      #undef HTR_PIECE
      const htr_rect tile = toRect(r);
      ok(htr_update_conserved_from_primitive(&f, &tile, taskStream()));
   }
}

// ---- CFL -----------------------------------------------------------------------------------------
DeferredValue<double> CalculateMaxSpectralRadiusTask::gpu_base_impl(const Args &args, Regions &regions,
                      const std::vector<Future> &futures, Context ctx, Runtime *runtime) {
   assert(regions.size() == 1);
   (void)futures;
   const Unpacker u(regions, 0, ctx, runtime);
   htr_fluid f = emptyFluid();
   f.dcsi_d = u.ro<double>(0, FID_dcsi_d); f.deta_d = u.ro<double>(0, FID_deta_d); f.dzet_d = u.ro<double>(0, FID_dzet_d);
   f.MassFracs = u.ro<VecNSp>(0, FID_MassFracs); f.temperature = u.ro<double>(0, FID_temperature);
   f.velocity = u.ro<Vec3>(0, FID_velocity);
   f.rho = u.ro<double>(0, FID_rho); f.mu = u.ro<double>(0, FID_mu); f.lam = u.ro<double>(0, FID_lam);
   f.Di = u.ro<VecNSp>(0, FID_Di); f.SoS = u.ro<double>(0, FID_SoS);
   loadMix(args.mix);
   const htr_rect tile = toRect(u.tile);
   DeferredValue<double> r(0.0);
   ok(htr_max_spectral_radius(&f, &tile, r.ptr(), taskStream()));
   return r;
}

// ---- non-periodic boundaries: NSCBC right-hand-side tasks (src/prometeo_rhs.cu:625-1560) -------------------------
// Region order as in the reference's bodies: the written region (last) is the plane of boundary cells; the fields read at
// the interior neighbour come with the Fluid / ghost regions in front of it.
namespace {
htr_bc_fields emptyBC() { htr_bc_fields b; memset(&b, 0, sizeof(b)); return b; }
void flattenShock(const IncomingShockParams &P, double *q) {      // include/htr_b200.h: params of htr_nscbc_incoming_shock
   q[0] = P.iShock;
   for (int s = 0; s < nSpec; s++) q[1 + s] = P.MolarFracs[s];
   double *r = q + 1 + nSpec;
   r[0] = P.pressure0; r[1] = P.temperature0; for (int i = 0; i < 3; i++) r[2 + i] = P.velocity0[i];
   r[5] = P.pressure1; r[6] = P.temperature1; for (int i = 0; i < 3; i++) r[7 + i] = P.velocity1[i];
}
template<direction dir, class Args>
void nscbcInflow(int side, const Args &args, Regions &regions, Context ctx, Runtime *runtime) {
   assert(regions.size() == 3);
   const Unpacker u(regions, 2, ctx, runtime);
   const DirFields &D = DIR[dirIndex(dir)];
   htr_fluid f = emptyFluid();
   htr_bc_fields b = emptyBC();
   f.MassFracs = u.ro<VecNSp>(0, FID_MassFracs); f.pressure = u.ro<double>(0, FID_pressure); f.velocity = u.ro<Vec3>(0, FID_velocity);
   f.*D.nType = u.ro<int>(1, D.fNType); f.*D.m_d = u.ro<double>(1, D.fMd);
   f.rho = u.ro<double>(1, FID_rho); f.SoS = u.ro<double>(1, FID_SoS); f.temperature = u.ro<double>(1, FID_temperature);
   b.dudtBoundary = u.ro<Vec3>(1, FID_dudtBoundary); b.dTdtBoundary = u.ro<double>(1, FID_dTdtBoundary);
   f.Conserved_t = u.rw<VecNEq>(2, FID_Conserved_t);
   loadMix(args.mix);
   const htr_rect bc = toRect(u.tile);
   ok(htr_nscbc_inflow(dirIndex(dir), side, &f, &b, &bc, taskStream()));
}
template<direction dir, class Args>
void nscbcOutflow(int side, const Args &args, Regions &regions, const std::vector<Future> &futures, Context ctx, Runtime *runtime) {
   assert(regions.size() == 4 && futures.size() == 1);
   const Unpacker u(regions, 3, ctx, runtime);
   htr_fluid f = emptyFluid();
   f.velocity = u.ro<Vec3>(0, FID_velocity);
   for (int d = 0; d < 3; d++) { f.*DIR[d].nType = u.ro<int>(1, DIR[d].fNType); f.*DIR[d].m_d = u.ro<double>(1, DIR[d].fMd); }
   f.rho = u.ro<double>(1, FID_rho); f.mu = u.ro<double>(1, FID_mu);
   f.MassFracs = u.ro<VecNSp>(1, FID_MassFracs); f.pressure = u.ro<double>(1, FID_pressure);
   f.SoS = u.ro<double>(2, FID_SoS); f.temperature = u.ro<double>(2, FID_temperature); f.Conserved = u.ro<VecNEq>(2, FID_Conserved);
   f.Conserved_t = u.rw<VecNEq>(3, FID_Conserved_t);
   loadMix(args.mix);
   const htr_rect bc = toRect(u.tile), B = toRect(args.Fluid_bounds);
   ok(htr_nscbc_outflow(dirIndex(dir), side, &f, &bc, &B, futures[0].get_result<double>(), args.LengthScale, args.PInf, taskStream()));
}
// fields of the far-field and incoming-shock tasks (src/prometeo_rhs.cu:1225-1245)
template<direction dir>
void farFieldFluid(htr_fluid &f, const Unpacker &u) {
   const DirFields &D = DIR[dirIndex(dir)];
   f.*D.nType = u.ro<int>(1, D.fNType); f.*D.m_d = u.ro<double>(1, D.fMd);
   f.rho = u.ro<double>(0, FID_rho); f.MassFracs = u.ro<VecNSp>(0, FID_MassFracs); f.pressure = u.ro<double>(0, FID_pressure);
   f.temperature = u.ro<double>(0, FID_temperature); f.velocity = u.ro<Vec3>(0, FID_velocity);
   f.SoS = u.ro<double>(1, FID_SoS); f.Conserved = u.ro<VecNEq>(1, FID_Conserved);
   f.Conserved_t = u.rw<VecNEq>(2, FID_Conserved_t);
}
template<direction dir, class Args>
void nscbcFarField(int side, const Args &args, Regions &regions, const std::vector<Future> &futures, Context ctx, Runtime *runtime) {
   assert(regions.size() == 3 && futures.size() == 1);
   const Unpacker u(regions, 2, ctx, runtime);
   htr_fluid f = emptyFluid();
   htr_bc_fields b = emptyBC();
   farFieldFluid<dir>(f, u);
   b.MolarFracs_profile = u.ro<VecNSp>(1, FID_MolarFracs_profile);
   b.temperature_profile = u.ro<double>(1, FID_temperature_profile);
   b.velocity_profile = u.ro<Vec3>(1, FID_velocity_profile);
   loadMix(args.mix);
   const htr_rect bc = toRect(u.tile);
   ok(htr_nscbc_farfield(dirIndex(dir), side, &f, &b, &bc, futures[0].get_result<double>(), args.LengthScale, args.PInf, taskStream()));
}
} // namespace

#define HTR_NSCBC_BODY(TASK, CALL) \
   template<direction dir> void TASK<dir>::gpu_base_impl(const Args &args, Regions &regions, \
                         const std::vector<Future> &futures, Context ctx, Runtime *runtime) { (void)futures; CALL; }
HTR_NSCBC_BODY(UpdateUsingFluxNSCBCInflowMinusSideTask,    nscbcInflow<dir>(HTR_SIDE_MINUS, args, regions, ctx, runtime))
HTR_NSCBC_BODY(UpdateUsingFluxNSCBCInflowPlusSideTask,     nscbcInflow<dir>(HTR_SIDE_PLUS, args, regions, ctx, runtime))
HTR_NSCBC_BODY(UpdateUsingFluxNSCBCOutflowMinusSideTask,   nscbcOutflow<dir>(HTR_SIDE_MINUS, args, regions, futures, ctx, runtime))
HTR_NSCBC_BODY(UpdateUsingFluxNSCBCOutflowPlusSideTask,    nscbcOutflow<dir>(HTR_SIDE_PLUS, args, regions, futures, ctx, runtime))
HTR_NSCBC_BODY(UpdateUsingFluxNSCBCFarFieldMinusSideTask,  nscbcFarField<dir>(HTR_SIDE_MINUS, args, regions, futures, ctx, runtime))
HTR_NSCBC_BODY(UpdateUsingFluxNSCBCFarFieldPlusSideTask,   nscbcFarField<dir>(HTR_SIDE_PLUS, args, regions, futures, ctx, runtime))

template<direction dir>
void UpdateUsingFluxIncomingShockTask<dir>::gpu_base_impl(const Args &args, Regions &regions,
                      const std::vector<Future> &futures, Context ctx, Runtime *runtime) {
   assert(regions.size() == 3 && futures.size() == 1);
   static_assert(dir == Ydir, "the reference only has the y-plus incoming-shock boundary (src/prometeo_rhs.cc:1064-1074)");
   const Unpacker u(regions, 2, ctx, runtime);
   htr_fluid f = emptyFluid();
   farFieldFluid<dir>(f, u);
   loadMix(args.mix);
   double q[1 + nSpec + 10];
   flattenShock(args.params, q);
   const htr_rect bc = toRect(u.tile);
   ok(htr_nscbc_incoming_shock(&f, &bc, futures[0].get_result<double>(), args.LengthScale, q, taskStream()));
}

// ---- boundary-cell primitive updates (src/prometeo_bc.cu:120-410) ----------------------------------------------------
template<direction dir>
void SetNSCBC_InflowBCTask<dir>::gpu_base_impl(const Args &args, Regions &regions,
                      const std::vector<Future> &futures, Context ctx, Runtime *runtime) {
   assert(regions.size() == 2);
   (void)futures;
   const Unpacker u(regions, 1, ctx, runtime);
   htr_fluid f = emptyFluid();
   htr_bc_fields b = emptyBC();
   f.Conserved = u.ro<VecNEq>(0, FID_Conserved); f.SoS = u.ro<double>(0, FID_SoS);
   b.MolarFracs_profile = u.ro<VecNSp>(0, FID_MolarFracs_profile);
   b.temperature_profile = u.ro<double>(0, FID_temperature_profile);
   b.velocity_profile = u.ro<Vec3>(0, FID_velocity_profile);
   f.pressure = u.wo<double>(1, FID_pressure); f.temperature = u.wo<double>(1, FID_temperature);
   f.MolarFracs = u.wo<VecNSp>(1, FID_MolarFracs); f.velocity = u.wo<Vec3>(1, FID_velocity);
   loadMix(args.mix);
   const htr_rect bc = toRect(u.tile);
   ok(htr_set_nscbc_inflow_bc(dirIndex(dir), &f, &b, &bc, args.Pbc, taskStream()));
}

namespace {
void outflowPrimitives(htr_fluid &f, const Unpacker &u) {
   f.Conserved = u.ro<VecNEq>(0, FID_Conserved);
   f.temperature = u.rw<double>(1, FID_temperature);
   f.pressure = u.wo<double>(1, FID_pressure); f.MolarFracs = u.wo<VecNSp>(1, FID_MolarFracs); f.velocity = u.wo<Vec3>(1, FID_velocity);
}
}
void SetNSCBC_OutflowBCTask::gpu_base_impl(const Args &args, Regions &regions,
                      const std::vector<Future> &futures, Context ctx, Runtime *runtime) {
   assert(regions.size() == 2);
   (void)futures;
   const Unpacker u(regions, 1, ctx, runtime);
   htr_fluid f = emptyFluid();
   outflowPrimitives(f, u);
   loadMix(args.mix);
   const htr_rect bc = toRect(u.tile);
   ok(htr_set_nscbc_outflow_bc(&f, &bc, taskStream()));
}

void SetIncomingShockBCTask::gpu_base_impl(const Args &args, Regions &regions,
                      const std::vector<Future> &futures, Context ctx, Runtime *runtime) {
   assert(regions.size() == 2);
   (void)futures;
   const Unpacker u(regions, 1, ctx, runtime);
   htr_fluid f = emptyFluid();
   outflowPrimitives(f, u);
   loadMix(args.mix);
   double q[1 + nSpec + 10];
   flattenShock(args.params, q);
   const htr_rect bc = toRect(u.tile);
   ok(htr_set_incoming_shock_bc(&f, &bc, q, taskStream()));
}

// ---- explicit instantiations, as at the end of the reference's .cu files ------------------------------
#define HTR_INST3(T) template void T<Xdir>::gpu_base_impl(const Args&, Regions&, const std::vector<Future>&, Context, Runtime*); \
                     template void T<Ydir>::gpu_base_impl(const Args&, Regions&, const std::vector<Future>&, Context, Runtime*); \
                     template void T<Zdir>::gpu_base_impl(const Args&, Regions&, const std::vector<Future>&, Context, Runtime*);
HTR_INST3(UpdateUsingHybridEulerFluxTask)
HTR_INST3(UpdateUsingSkewSymmetricEulerFluxTask)
HTR_INST3(UpdateUsingDiffusionFluxTask)
HTR_INST3(UpdateShockSensorTask)
#define HTR_INST_TENO(Op) \
   template void UpdateUsingTENOEulerFluxTask<Xdir, Op>::gpu_base_impl(const Args&, Regions&, const std::vector<Future>&, Context, Runtime*); \
   template void UpdateUsingTENOEulerFluxTask<Ydir, Op>::gpu_base_impl(const Args&, Regions&, const std::vector<Future>&, Context, Runtime*); \
   template void UpdateUsingTENOEulerFluxTask<Zdir, Op>::gpu_base_impl(const Args&, Regions&, const std::vector<Future>&, Context, Runtime*);
HTR_INST_TENO(TENO_Op<>)
HTR_INST_TENO(TENOA_Op<>)
HTR_INST_TENO(TENOLAD_Op<>)
HTR_INST3(UpdateUsingFluxNSCBCInflowMinusSideTask)
HTR_INST3(UpdateUsingFluxNSCBCInflowPlusSideTask)
HTR_INST3(UpdateUsingFluxNSCBCOutflowMinusSideTask)
HTR_INST3(UpdateUsingFluxNSCBCOutflowPlusSideTask)
HTR_INST3(UpdateUsingFluxNSCBCFarFieldMinusSideTask)
HTR_INST3(UpdateUsingFluxNSCBCFarFieldPlusSideTask)
template void UpdateUsingFluxIncomingShockTask<Ydir>::gpu_base_impl(const Args&, Regions&, const std::vector<Future>&, Context, Runtime*);
HTR_INST3(SetNSCBC_InflowBCTask)
