/* htr_b200.h -- C ABI of the B200-native HTR right-hand-side engine.
 *
 * One shared library per mixture (libhtr_b200_<Mix>.so), like the reference builds one
 * executable per -DEOS=<Mix> (src/prometeo_const.h:37-41, src/Makefile:68-70).
 * Every entry point is the Legion-free body of one reference leaf task: it takes what the
 * task's gpu_base_impl extracts from its PhysicalRegions -- one affine accessor
 * (base pointer + byte strides, Realm::AffineAccessor semantics) per field, the
 * execution rectangle (index space of the written region) and args.Fluid_bounds --
 * and launches hand-written sm_100a kernels on the caller's stream.  All pointers in
 * htr_fluid are DEVICE pointers.  No torch types, no Legion types.
 *
 * The Legion side (register_{rhs,sensor,chem,variables,cfl}_tasks, src/prometeo_*.h:39)
 * is htr-solver_b200/csrc/legion_adapter.cc, which only unpacks regions into these calls
 * (INTEGRATION.md).
 *
 * Return value: 0 on success, a negative HTR_E* code on a usage error; CUDA launch
 * failures are returned as 1000+cudaError_t.  The reference aborts (assert) on the
 * same conditions (src/prometeo_rhs.cc:43-44).
 */
#ifndef HTR_B200_H
#define HTR_B200_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define HTR_OK 0
#define HTR_EINVAL (-1)   /* bad argument (direction, scheme, empty accessor that the task reads) */
#define HTR_ENOMIX (-2)   /* mixture not loaded: call htr_set_mixture first (LoadMixtureTask, src/prometeo_mixture.cu:35-47) */
#define HTR_EMIXTYPE (-3) /* constructor for another mixture than the one this library was built for */
#define HTR_ETOOBIG (-4)  /* htr_rhs_no_alloc: the tile has more cells than the library's static scratch holds */

/* Affine accessor: address(p) = base + p.x*stride[0] + p.y*stride[1] + p.z*stride[2], p in GLOBAL
 * grid coordinates, strides in BYTES (replaces FieldAccessor<...,Realm::AffineAccessor<T,3,coord_t>>,
 * src/Utils/task_helper.hpp:99-102).  Vector fields are MyArray<double,n>: n contiguous doubles
 * per cell (src/prometeo_types.h:321-323). */
typedef struct { void *base; int64_t stride[3]; } htr_acc;

/* Inclusive rectangle, Legion Rect<3> semantics. */
typedef struct { int64_t lo[3]; int64_t hi[3]; } htr_rect;

/* The Fluid_columns fields the hot path touches (src/prometeo_types.h:49-113), same names. */
typedef struct {
   htr_acc nType_x, nType_y, nType_z;            /* int32 */
   htr_acc dcsi_e, deta_e, dzet_e;               /* double */
   htr_acc dcsi_d, deta_d, dzet_d;
   htr_acc dcsi_s, deta_s, dzet_s;
   htr_acc rho, mu, lam, Di, SoS;                /* Di: nSpec doubles */
   htr_acc pressure, temperature, MassFracs, MolarFracs, velocity;
   htr_acc Conserved, Conserved_old, Conserved_t, Conserved_t_old;   /* nEq doubles */
   htr_acc shockSensorX, shockSensorY, shockSensorZ;                 /* C++ bool, 1 byte */
   htr_acc DucrosSensor;
} htr_fluid;

/* Euler schemes = the reference's task families (src/prometeo_types.h:184-283 TID_UpdateUsing*EulerFlux*). */
enum { HTR_SCHEME_HYBRID = 0, HTR_SCHEME_TENO = 1, HTR_SCHEME_TENOA = 2,
       HTR_SCHEME_TENOLAD = 3, HTR_SCHEME_SKEWSYMMETRIC = 4 };
enum { HTR_XDIR = 0, HTR_YDIR = 1, HTR_ZDIR = 2 };

/* ---- library identity ---- */
int  htr_nspec(void);
int  htr_neq(void);
const char *htr_mixture_name(void);
int  htr_sizeof_mix(void);        /* == sizeof(reference struct Mix) for this mixture */
const char *htr_build_info(void); /* arch, flags */

/* Library switches (process-wide; A/B testing and measured alternatives, same results in every setting -- bit for bit in the
 * strict build).  Defaults in brackets:
 *   "fast"     [1] htr_rhs uses the staged kernels for the Hybrid and SkewSymmetric schemes; 0 = the per-task kernels only
 *   "march"    [1] bit 0: Euler-z by the register-ring march; bit 1: the fused plane march; bit 2: plane march for Euler y, x only
 *   "dasync"   [1] diffusion tiles staged by cp.async row copies (0 = cooperative loads)
 *   "diff3"    [0] diffusion z, y, x in one kernel        "list" [1] deferred cells from a compacted list (0 = block scan)
 *   "coop"     [0] lane-cooperative TENO (eight lanes per face)
 *   "bulk" / "tensor" [1] TMA row / tensor-map copies in the Euler-xy kernel
 *   "pipeline" [1] z-slab copy/compute overlap in htr_host_rhs; "slab" [0 = 14 planes] its slab thickness in tiles of 7 planes
 * Unknown names return HTR_EINVAL.  (No counterpart in the reference: its variants are compile-time.) */
int  htr_set_option(const char *name, int value);

/* ---- mixture (replaces Mix::Mix(const Config&) + LoadMixtureTask) ----
 * The constructors fill `mix_out` (htr_sizeof_mix() bytes) with a POD that is byte-identical
 * to the reference's struct Mix, so an adapter may instead pass args.mix straight to
 * htr_set_mixture.  ConstPropMix: src/Mixtures/ConstPropMix.inl:31-52 (visc_model 0 Constant{p0=Visc},
 * 1 PowerLaw{p0=ViscRef,p1=TempRef}, 2 Sutherland{p0=ViscRef,p1=TempRef,p2=SRef}).
 * Multi-component: src/Mixtures/AirMix.hpp:319-332 + MultiComponent.inl:438-488; names = n x 10 chars. */
int htr_mix_init_constprop(void *mix_out, double gasConstant, double gamma, double prandtl,
                           int visc_model, double p0, double p1, double p2);
int htr_mix_init_multicomp(void *mix_out, double LRef, double PRef, double TRef,
                           int n, const char *names, const double *molarFracs);
int htr_set_mixture(const void *mix, size_t bytes);   /* copies into __constant__ memory of the current device */

/* ---- leaf tasks (device pointers; asynchronous on `stream`, a cudaStream_t) ----
 * tile   = index space of the region the task writes (src/prometeo_rhs.cc:70-71)
 * bounds = args.Fluid_bounds, the global box stencil points wrap on
 *          (src/Utils/PointDomain_helper.hpp:156-182) */
/* UpdateUsing{Hybrid,TENO,TENOA,TENOLAD,SkewSymmetric}EulerFlux{X,Y,Z}Task (src/prometeo_rhs.cc:37-376) */
int htr_euler_flux(int scheme, int dir, const htr_fluid *f, const htr_rect *tile, const htr_rect *bounds,
                   double RK_coeffs0, double RK_coeffs1, double deltaTime, void *stream);
/* UpdateUsingDiffusionFlux{X,Y,Z}Task (src/prometeo_rhs.cc:410-467) */
int htr_diffusion_flux(int dir, const htr_fluid *f, const htr_rect *tile, const htr_rect *bounds, void *stream);
/* UpdateDucrosSensorTask (src/prometeo_sensor.cc:37-79) */
int htr_ducros_sensor(const htr_fluid *f, const htr_rect *tile, const htr_rect *bounds,
                      double vorticityScale, void *stream);
/* UpdateShockSensorTask<dir> (src/prometeo_sensor.cc:83-138) */
int htr_shock_sensor(int dir, const htr_fluid *f, const htr_rect *tile, const htr_rect *bounds, void *stream);
/* AddChemistrySourcesTask (src/prometeo_chem.cc:78-112) */
int htr_add_chemistry_sources(const htr_fluid *f, const htr_rect *tile, void *stream);
/* UpdateChemistryTask (src/prometeo_chem.cc:36-71) */
int htr_update_chemistry(const htr_fluid *f, const htr_rect *tile, double Integrator_deltaTime, void *stream);
/* Cells whose implicit integration hit the step/iteration caps since the last call (the reference
 * asserts there, src/Utils/math_utils.hpp:233,240).  Synchronises the device and resets the counter. */
int htr_chemistry_failures(void);
/* diagnostic (parity tests): htr_update_chemistry then also writes the number of right-hand-side evaluations of every
 * cell of its tile (dense int32, x fastest) to the DEVICE buffer d_counts; NULL switches it off. */
int htr_chemistry_rhs_counts(int *d_counts);
/* the same after synchronising `stream`, the stream the update was launched on (any kind of stream) */
int htr_chemistry_failures_on(void *stream);
/* 1 if htr_set_mixture has been called on the CURRENT device (all library state is per device: one process may drive
 * several GPUs, src/prometeo_mapper.cc:231-262) */
int htr_mixture_loaded(void);

/* fp64 pipe probe (measurement only): runs independent DFMA chains for about `seconds` on the current device and
 * returns the sustained (tflops[0], last batch) and burst (tflops[1], best batch) DFMA rate in TFLOP/s: the
 * denominator of the fp64 roofline of the chemistry / TENO kernels. */
int htr_probe_fp64(double seconds, double *tflops, void *stream);
/* UpdatePrimitiveFromConservedTask / UpdatePropertiesFromPrimitiveTask / UpdateConservedFromPrimitiveTask
 * (src/prometeo_variables.cc:138-173, :36-86, :92-132) */
int htr_update_primitive_from_conserved(const htr_fluid *f, const htr_rect *tile, void *stream);
int htr_update_properties_from_primitive(const htr_fluid *f, const htr_rect *tile, void *stream);
/* UpdatePrimitiveFromConserved for ghost copies of a neighbour block's cells whose `temperature` was exchanged
 * together with `Conserved`: pressure, velocity and MolarFracs from the converged temperature without iterating again
 * (a second Newton pass from the converged value moves T within its 1e-8 tolerance, so ghost and owner would differ).
 * The reference has no such task: Legion copies all fields of the ghost region (src/prometeo_partitioner.rg). */
int htr_update_ghost_primitive_from_conserved(const htr_fluid *f, const htr_rect *tile, void *stream);
int htr_update_conserved_from_primitive(const htr_fluid *f, const htr_rect *tile, void *stream);
/* Runge-Kutta glue (src/prometeo_rk.rg; the reference has Regent tasks only, no C++ task to swap, so an
 * adapter registers these as new leaf tasks - INTEGRATION.md section 6).
 *   htr_rk_initialize_temporaries       InitializeTemporaries, :46-56      Conserved_old = Conserved
 *   htr_rk_initialize_time_derivatives  mkInitializeTimeDerivatives, :58-86  Conserved_t = 0, or -Conserved_t_old
 *                                       with implicit chemistry
 *   htr_rk_update_vars                  mkUpdateVars(STAGE), :88-111; stage = 1..3, coefficients RK_C
 *                                       (src/prometeo_const.rg:54-58); dt is halved with implicit chemistry
 *   htr_rk_stage                        UpdateVars(STAGE) + UpdatePrimitiveFromConserved +
 *                                       UpdatePropertiesFromPrimitive in one pass over the tile (the three
 *                                       launches of src/prometeo.rg:1427, :452, :470 when the tile has no
 *                                       boundary nodes); same operations in the same order as the three calls in sequence
 *                                       (bit-identical in the strict build; FMA contraction may differ by an ulp
 *                                       in the production build) */
int htr_rk_initialize_temporaries(const htr_fluid *f, const htr_rect *tile, void *stream);
int htr_rk_initialize_time_derivatives(const htr_fluid *f, const htr_rect *tile, int implicitChemistry, void *stream);
int htr_rk_update_vars(int stage, double Integrator_deltaTime, int implicitChemistry, const htr_fluid *f,
                       const htr_rect *tile, void *stream);
int htr_rk_stage(int stage, double Integrator_deltaTime, int implicitChemistry, const htr_fluid *f,
                 const htr_rect *tile, void *stream);
/* CalculateMaxSpectralRadiusTask (src/prometeo_cfl.cc:36-93): the tile maximum is written to
 * *d_result (device double, the DeferredValue of the GPU variant, src/prometeo_cfl.hpp:181-189). */
int htr_max_spectral_radius(const htr_fluid *f, const htr_rect *tile, double *d_result, void *stream);

/* ---- fused stage (B200 fast path; same arithmetic, one pass per direction group) ----
 * One UpdateDerivatives pass (src/prometeo.rg:502-578): Conserved_t = 0 (+ chemistry source)
 * + Euler z,y,x + diffusion z,y,x, in the reference's accumulation order. */
int htr_rhs(int scheme, int diffusion, int chemistry, const htr_fluid *f, const htr_rect *tile,
            const htr_rect *bounds, double RK_coeffs0, double RK_coeffs1, double deltaTime, void *stream);

/* htr_rhs for callers that must not allocate (CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).
 * NO COUNTERPART IN THE REFERENCE.  Same arguments, same kernels and same
 * result as htr_rhs; the scratch of the staged path (deferral flags, two mask bytes per cell, the list of deferred cells:
 * 6 bytes per cell) lives in a file-scope `static __device__` array of 12 MiB that every GPU loading the library holds until
 * exit, instead of the cudaMalloc'ed per-stream buffers htr_rhs grows on demand.  Calls no cudaMalloc / cudaFree / allocating
 * routine.  Limit: tiles of at most htr_rhs_no_alloc_max_cells() = 128^3 = 2 097 152 cells; a larger tile is rejected with
 * HTR_ETOOBIG (and a line on stderr) before anything is launched.  All calls on one device share the array: calls that
 * overlap on the GPU (different streams, no ordering between them) race; calls ordered on one stream are safe.  The array
 * keeps the last call's contents; the entry point clears what must start at zero itself (flags, list counter). */
int htr_rhs_no_alloc(int scheme, int diffusion, int chemistry, const htr_fluid *f, const htr_rect *tile,
                     const htr_rect *bounds, double RK_coeffs0, double RK_coeffs1, double deltaTime, void *stream);
int64_t htr_rhs_no_alloc_max_cells(void);

/* The same pass restricted to some of its staged kernels (timing / profiling only: a partial pass leaves
 * Conserved_t incomplete).  parts = bit mask: 1 Conserved_t = 0 + chem + Euler z; 2 += Euler y, x; 4 += diffusion z;
 * 8 += diffusion y, x.  15 = htr_rhs.  Hybrid and SkewSymmetric schemes with the "fast" option on. */
int htr_rhs_parts(int parts, int scheme, int diffusion, int chemistry, const htr_fluid *f, const htr_rect *tile,
                  const htr_rect *bounds, double RK_coeffs0, double RK_coeffs1, double deltaTime, void *stream);
/* One RK3 time step of a block without boundary nodes (MainLoopBody, src/prometeo.rg:1348-1470):
 * UpdateShockSensors (Hybrid only); with implicitChemistry the operator-split half step (:1360-1395:
 * UpdateDerivatives(RK_coeffs = (0, 1), deltaTime/2) without chemistry sources, UpdateChemistry(deltaTime),
 * primitive/property recovery); InitializeTemporaries; then for STAGE = 1..3 htr_rhs with RK_coeffs = RK_C[STAGE][1..2]
 * and deltaTime*RK_C[STAGE][3] (halved, and Conserved_t offset by -Conserved_t_old, with implicitChemistry) followed
 * by htr_rk_stage.  `chemistry` adds the explicit source (ignored with implicitChemistry, like the reference).
 * The halo of `f` (periodic wrap inside `bounds`, or ghost planes filled by the caller) must be valid for every
 * stage: with tile == bounds (one periodic block) nothing else is needed. */
int htr_time_step(int scheme, int diffusion, int chemistry, int implicitChemistry, double vorticityScale,
                  const htr_fluid *f, const htr_rect *tile, const htr_rect *bounds, double Integrator_deltaTime,
                  void *stream);

/* ---- host-buffer stage (the call a host-resident driver binds) ----
 * h_* are HOST pointers (pinned for full copy bandwidth), dense [z][y][x][comp] over `tile`.
 * Copies Conserved (and Conserved_old for the multispecies limiter; NULL = same as Conserved) and the
 * previous temperature to the device workspace `f`, recovers primitives and properties
 * (src/prometeo.rg:450-471), optionally refreshes the shock sensors (src/prometeo.rg:1356), evaluates
 * UpdateDerivatives and copies Conserved_t (and the converged temperature) back.  Asynchronous on
 * `stream`; synchronise it before reading h_Conserved_t. */
int htr_host_rhs(int scheme, int diffusion, int chemistry, int update_sensors, double vorticityScale,
                 const htr_fluid *f, const htr_rect *tile, const htr_rect *bounds,
                 const double *h_Conserved, const double *h_Conserved_old, double *h_temperature,
                 double *h_Conserved_t, double RK_coeffs0, double RK_coeffs1, double deltaTime, void *stream);

/* ---- non-periodic boundaries: NSCBC right-hand-side tasks and boundary-cell primitive updates ----
 * Boundary fields of struct Fluid_columns (src/prometeo_types.h:49-113) read by these tasks; same accessor convention as
 * htr_fluid.  An accessor a task does not read may be {NULL}. */
typedef struct {
   htr_acc dudtBoundary, dTdtBoundary;                                  /* Vec3, double */
   htr_acc velocity_profile, temperature_profile, MolarFracs_profile;   /* Vec3, double, nSpec doubles */
} htr_bc_fields;
#define HTR_SIDE_MINUS 0   /* the x/y/zNeg boundary (…MinusSideTask) */
#define HTR_SIDE_PLUS  1   /* the x/y/zPos boundary (…PlusSideTask) */

/* UpdateUsingFluxNSCBCInflow{Minus,Plus}SideTask<dir> (src/prometeo_rhs.hpp:479-582, addLODIfluxes src/prometeo_rhs.inl:1288-1412,
 * kernels src/prometeo_rhs.cu:630-845): Conserved_t of the boundary cells `bc` -- zeroed for a supersonic inlet, species
 * LODI fluxes otherwise. */
int htr_nscbc_inflow(int dir, int side, const htr_fluid *f, const htr_bc_fields *b, const htr_rect *bc, void *stream);
/* UpdateUsingFluxNSCBCOutflow{Minus,Plus}SideTask<dir> (src/prometeo_rhs.hpp:584-720, src/prometeo_rhs.inl:1418-1638,
 * src/prometeo_rhs.cu:863-1170).  MaxMach is the value of the reference's future (CalculateMaxMachNumber). */
int htr_nscbc_outflow(int dir, int side, const htr_fluid *f, const htr_rect *bc, const htr_rect *bounds,
                      double MaxMach, double LengthScale, double PInf, void *stream);
/* UpdateUsingFluxNSCBCFarField{Minus,Plus}SideTask<dir> (src/prometeo_rhs.hpp:722-838, src/prometeo_rhs.inl:1644-1925,
 * src/prometeo_rhs.cu:1180-1465): far-field state from the *_profile fields. */
int htr_nscbc_farfield(int dir, int side, const htr_fluid *f, const htr_bc_fields *b, const htr_rect *bc,
                       double MaxMach, double LengthScale, double PInf, void *stream);
/* UpdateUsingFluxIncomingShockTask<Ydir> (src/prometeo_rhs.hpp:840-875, src/prometeo_rhs.cc:1076-1138, src/prometeo_rhs.cu:1475-1560):
 * the y-plus boundary with the upstream state for x < iShock and the downstream state from iShock on.
 * params (HOST doubles; struct IncomingShockParams, src/prometeo_bc_types.h:46-59, flattened):
 * iShock, MolarFracs[nSpec], pressure0, temperature0, velocity0[3], pressure1, temperature1, velocity1[3]. */
int htr_nscbc_incoming_shock(const htr_fluid *f, const htr_rect *bc, double MaxMach, double LengthScale,
                             const double *params, void *stream);
/* SetNSCBC_InflowBCTask<dir> (src/prometeo_bc.hpp:113-152, src/prometeo_bc.cc:80-135, src/prometeo_bc.cu:120-218): primitives
 * of the inflow cells from the profiles; pressure = Pbc where the profile is supersonic, else from the transported density. */
int htr_set_nscbc_inflow_bc(int dir, const htr_fluid *f, const htr_bc_fields *b, const htr_rect *bc, double Pbc, void *stream);
/* SetNSCBC_OutflowBCTask (src/prometeo_bc.cc:160-196, src/prometeo_bc.cu:228-290): UpdatePrimitive on the boundary cells. */
int htr_set_nscbc_outflow_bc(const htr_fluid *f, const htr_rect *bc, void *stream);
/* SetIncomingShockBCTask (src/prometeo_bc.cc:203-264, src/prometeo_bc.cu:296-410): outflow treatment except for the three
 * cells iShock-1 .. iShock+1, where the shock is injected; params as for htr_nscbc_incoming_shock. */
int htr_set_incoming_shock_bc(const htr_fluid *f, const htr_rect *bc, const double *params, void *stream);

/* ---- multi-GPU: block decomposition and ghost-plane exchange (mirrors src/prometeo_partitioner.rg:436-446 tiling,
 * :1107-1352 ghost partitions; Legion/Realm moves the ghosts implicitly in the reference) ----
 * One process per GPU.  The global box is cut into tiles = [tx, ty, tz] equal blocks, rank = cx + tx*(cy + ty*cz), periodic
 * neighbours.  htr_comm_get_unique_id on rank 0 returns 128 bytes that the caller distributes to all ranks (MPI, a file,
 * torch.distributed ...); htr_comm_init is collective.  With one rank no NCCL is needed (all wraps are local copies). */
typedef struct htr_comm_s htr_comm;
int htr_comm_get_unique_id(void *id128);
int htr_comm_init(htr_comm **comm, const void *id128, int rank, int nranks, int tx, int ty, int tz);
int htr_comm_free(htr_comm *comm);
int htr_comm_rank_of_tile(const htr_comm *comm, int cx, int cy, int cz);
/* Fills the `halo` ghost planes per side of `nfields` fields of this rank's block (interior = `interior`, global
 * coordinates; cellBytes[k] = bytes per cell of field k) from the neighbouring blocks: three phases x -> y -> z whose
 * slabs include the ghosts of the earlier phases (edges and corners arrive without diagonal messages), per phase one
 * pack kernel, ncclSend/ncclRecv to the +/- neighbours inside one group, one unpack kernel; all on `stream`. */
int htr_halo_exchange(htr_comm *comm, int nfields, const htr_acc *fields, const int *cellBytes, const htr_rect *interior,
                      int halo, void *stream);
/* in-place MAX over all ranks of `count` device doubles (the spectral radius of the CFL condition, src/prometeo.rg:1161-1165) */
int htr_allreduce_max(htr_comm *comm, double *d_values, int count, void *stream);

#ifdef __cplusplus
}
#endif
#endif
