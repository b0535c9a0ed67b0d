/* TEST INFRASTRUCTURE ONLY (checker, never the product path).
 *
 * C ABI shared by the two CPU oracles of the HTR per-cell RHS hot path:
 *   oracle/_ref/libhtr_ref_<Mix>.so   -- the reference's own CPU numerics
 *       (/root/reference/src/prometeo_*.{hpp,inl}, Mixtures/*, Utils/*) compiled
 *       IN PLACE against oracle/stub/{legion.h,config_schema.h}; built by
 *       oracle/Makefile only where /root/reference exists.  kind = "reference".
 *   oracle/libhtr_port_<Mix>.so       -- oracle/port.cc, our CPU restatement
 *       of the same algorithm.  kind = "port".
 *
 * Data model: ONE periodic tile that is the whole domain (the reference wraps
 * stencil points by modulo on the global bounds,
 * src/Utils/PointDomain_helper.hpp:156-187), no halos, C layout [z][y][x][comp]
 * exactly as a single-tile Legion instance (src/prometeo_sensor.cc:65).
 * nSpec / nEq are compile-time (-DEOS=<Mix>) as in src/prometeo_const.h:37-41.
 */
#ifndef HTR_ORACLE_API_H
#define HTR_ORACLE_API_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

/* Field table; member order mirrors struct Fluid_columns (src/prometeo_types.h:49-113). */
typedef struct {
   int64_t nx, ny, nz;
   int32_t *nType_x, *nType_y, *nType_z;
   double  *dcsi_e, *deta_e, *dzet_e;
   double  *dcsi_d, *deta_d, *dzet_d;
   double  *dcsi_s, *deta_s, *dzet_s;
   double  *rho, *mu, *lam, *Di, *SoS;
   double  *pressure, *temperature, *MassFracs, *MolarFracs, *velocity;
   double  *Conserved, *Conserved_old, *Conserved_t, *Conserved_t_old;
   uint8_t *shockSensorX, *shockSensorY, *shockSensorZ;   /* C++ bool, 1 byte */
   double  *DucrosSensor;
} htr_oracle_fields;

/* Boundary fields used by the NSCBC tasks (struct Fluid_columns, src/prometeo_types.h:49-113); dense over the same box. */
typedef struct {
   double *dudtBoundary, *dTdtBoundary;                                /* Vec3, double */
   double *velocity_profile, *temperature_profile, *MolarFracs_profile; /* Vec3, double, VecNSp */
} htr_oracle_bc_fields;

/* Euler schemes (src/config_schema.lua EulerSchemes; task families src/prometeo_rhs.hpp). */
enum { HTR_SCHEME_HYBRID = 0, HTR_SCHEME_TENO = 1, HTR_SCHEME_TENOA = 2,
       HTR_SCHEME_TENOLAD = 3, HTR_SCHEME_SKEWSYMMETRIC = 4 };

int  htr_oracle_nspec(void);
int  htr_oracle_neq(void);
const char *htr_oracle_kind(void);      /* "reference" or "port" */
const char *htr_oracle_mixture(void);   /* "ConstPropMix", "AirMix", ... */
int  htr_oracle_sizeof_mix(void);

/* Mixture construction (replaces Mix::Mix(const Config&)):
 * ConstPropMix: src/Mixtures/ConstPropMix.inl:31-52; visc_model 0 Constant{p0=Visc},
 *   1 PowerLaw{p0=ViscRef,p1=TempRef}, 2 Sutherland{p0=ViscRef,p1=TempRef,p2=SRef}.
 * Multicomponent: src/Mixtures/AirMix.hpp:319-332 -> StoreReferenceQuantities
 *   (src/Mixtures/MultiComponent.inl:438-488); names is n x 10 chars. */
int htr_oracle_init_constprop(double gasConstant, double gamma, double prandtl,
                              int visc_model, double p0, double p1, double p2);
int htr_oracle_init_multicomp(double LRef, double PRef, double TRef,
                              int n, const char *names, const double *molarFracs);

/* Point-wise mixture wrappers used to pin the oracle on unitTests/mixTest goldens
 * (src/prometeo_mixture.cc:57-220).  out has 16 + 4*nSpec doubles:
 *  [0]MixW(Yi) [1]rho [2]cp [3]h [4]e [5]mu [6]lam [7]gamma [8]SoS [9]dpde [10]MixW(Xi)
 *  [11]T(e) [16..]Xi(Yi) [16+n..]Di [16+2n..]w_i [16+3n..]dpdrhoi ; inputs are
 *  non-dimensional P,T and mass fractions. */
void htr_oracle_mix_eval(double P, double T, const double *Yi, double *out);
int htr_oracle_mix_bytes(void *out);       /* copies sizeof(Mix) bytes, returns the size */
int htr_oracle_mix_table_bytes(void);      /* offset of Mix::PRef (0 for ConstPropMix) */
double htr_oracle_species_enthalpy(int i, double T);
double htr_oracle_species_W(int i);

/* Tasks.  dir 0/1/2 = X/Y/Z.  Each one is the reference task's cpu_base_impl loop
 * over the whole tile (file:line in the .cc driver). */
void htr_oracle_euler_flux(int scheme, int dir, const htr_oracle_fields *f,
                           double RK_coeffs0, double RK_coeffs1, double deltaTime);
void htr_oracle_diffusion_flux(int dir, const htr_oracle_fields *f);
void htr_oracle_ducros_sensor(const htr_oracle_fields *f, double vorticityScale);
void htr_oracle_shock_sensor(int dir, const htr_oracle_fields *f);
void htr_oracle_add_chemistry_sources(const htr_oracle_fields *f);
void htr_oracle_update_chemistry(const htr_oracle_fields *f, double Integrator_deltaTime);
/* NSCBC right-hand-side tasks (src/prometeo_rhs.inl:1285-1930) and boundary-cell primitive updates (src/prometeo_bc.cc:80-264)
 * over the box lo..hi of boundary cells.  side 0 = minus (left) boundary, 1 = plus.  Incoming shock: the y-plus boundary;
 * params = iShock, MolarFracs[nSpec], pressure0, temperature0, velocity0[3], pressure1, temperature1, velocity1[3]. */
void htr_oracle_nscbc_inflow(int dir, int side, const htr_oracle_fields *f, const htr_oracle_bc_fields *b,
                             const int64_t *lo, const int64_t *hi);
void htr_oracle_nscbc_outflow(int dir, int side, const htr_oracle_fields *f, const int64_t *lo, const int64_t *hi,
                              double MaxMach, double LengthScale, double PInf);
void htr_oracle_nscbc_farfield(int dir, int side, const htr_oracle_fields *f, const htr_oracle_bc_fields *b,
                               const int64_t *lo, const int64_t *hi, double MaxMach, double LengthScale, double PInf);
void htr_oracle_incoming_shock(const htr_oracle_fields *f, const int64_t *lo, const int64_t *hi,
                               double MaxMach, double LengthScale, const double *params);
void htr_oracle_set_nscbc_inflow_bc(int dir, const htr_oracle_fields *f, const htr_oracle_bc_fields *b,
                                    const int64_t *lo, const int64_t *hi, double Pbc);
void htr_oracle_set_nscbc_outflow_bc(const htr_oracle_fields *f, const int64_t *lo, const int64_t *hi);
void htr_oracle_set_incoming_shock_bc(const htr_oracle_fields *f, const int64_t *lo, const int64_t *hi, const double *params);
void htr_oracle_update_primitive_from_conserved(const htr_oracle_fields *f);
void htr_oracle_update_properties_from_primitive(const htr_oracle_fields *f);
void htr_oracle_update_conserved_from_primitive(const htr_oracle_fields *f);
double htr_oracle_max_spectral_radius(const htr_oracle_fields *f);

/* Known-answer helpers for Utils/math_utils.hpp (unitTests/mathUtilsTest). */
void htr_oracle_lu3_solve(const double *A9_rowmajor, const double *b3, double *x3);
void htr_oracle_rosenbrock3(double *x3, double dt, double dtTry);

/* Number of OpenMP threads the library was built to use (1 when serial). */
int htr_oracle_max_threads(void);
void htr_oracle_update_chemistry_counted(const htr_oracle_fields *f, double dt, int *rhs_calls);   /* + rhs() evaluations per cell */
/* InitializeMetricTask (src/prometeo_metric.cc:36-190): all nine metric fields from cell-centre coordinates [z][y][x][3] */
void htr_oracle_initialize_metric(const htr_oracle_fields *f, const double *centerCoordinates, const double *width);
void htr_oracle_set_threads(int n);      /* OpenMP threads of the timed build (bench.py: all host cores, also under torchrun) */
#ifdef __cplusplus
}
#endif
#endif
