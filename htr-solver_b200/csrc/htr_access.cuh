// Device-side view of the C-ABI accessors (include/htr_b200.h): global-coordinate affine
// addressing with byte strides, and the periodic wrap of stencil points on Fluid_bounds
// (semantics of src/Utils/PointDomain_helper.hpp:156-187; the modulo is only evaluated for
// points that actually leave the box).
#pragma once
#include "../../include/htr_b200.h"
#include "htr_config.h"

namespace htr {

struct Pt { int64_t x, y, z; };

template<typename T>
__device__ __forceinline__ T *at(const htr_acc &a, const Pt &p) {
   return reinterpret_cast<T*>(static_cast<char*>(a.base) + p.x*a.stride[0] + p.y*a.stride[1] + p.z*a.stride[2]);
}
template<typename T>
__device__ __forceinline__ T ld(const htr_acc &a, const Pt &p) { return __ldg(at<const T>(a, p)); }
template<typename T>
__device__ __forceinline__ T ldc(const htr_acc &a, const Pt &p, int c) { return __ldg(at<const T>(a, p) + c); }

// p shifted by `off` cells along `dir`, wrapped into [lo, hi] of the global box
template<int dir>
__device__ __forceinline__ Pt shift(const Pt &p, int off, const htr_rect &B) {
   Pt q = p;
   int64_t &c = (dir == 0) ? q.x : ((dir == 1) ? q.y : q.z);
   c += off;
   const int64_t size = B.hi[dir] - B.lo[dir] + 1;
   if (c > B.hi[dir] || c < B.lo[dir])      // rare path: full modulo (also right for boxes thinner than the stencil)
      c = ((c - B.lo[dir]) % size + size) % size + B.lo[dir];
   return q;
}

// ---- tile-relative 32-bit addressing (staged kernels) -----------------------------------------
// The staged kernels address cells relative to tile.lo with 32-bit coordinates and strides: one
// IMAD.WIDE per term instead of a 64x64-bit multiply per term per field (the per-field byte strides made
// 9-16 % of their instructions).  Built on the host by makeRelative() (htr_kernels.cu); a layout whose
// strides do not fit 31 bits takes the per-task kernels.
struct racc { char *base; int s[3]; };           // base = address of cell tile.lo
struct rbox { int lo[3], hi[3]; };               // inclusive, relative to tile.lo, clamped to +-2^30
struct RPt { int x, y, z; };
struct rfluid {                                  // same members, same order as htr_fluid
   racc nType_x, nType_y, nType_z;
   racc dcsi_e, deta_e, dzet_e;
   racc dcsi_d, deta_d, dzet_d;
   racc dcsi_s, deta_s, dzet_s;
   racc rho, mu, lam, Di, SoS;
   racc pressure, temperature, MassFracs, MolarFracs, velocity;
   racc Conserved, Conserved_old, Conserved_t, Conserved_t_old;
   racc shockSensorX, shockSensorY, shockSensorZ;
   racc DucrosSensor;
};
static_assert(sizeof(rfluid)/sizeof(racc) == sizeof(htr_fluid)/sizeof(htr_acc), "rfluid mirrors htr_fluid");

template<typename T>
__device__ __forceinline__ T *at(const racc &a, const RPt &p) {
   return reinterpret_cast<T*>(a.base + (int64_t)p.x*a.s[0] + (int64_t)p.y*a.s[1] + (int64_t)p.z*a.s[2]);
}
template<typename T>
__device__ __forceinline__ T ld(const racc &a, const RPt &p) { return __ldg(at<const T>(a, p)); }
template<typename T>
__device__ __forceinline__ T ldc(const racc &a, const RPt &p, int c) { return __ldg(at<const T>(a, p) + c); }

template<int dir> __device__ __forceinline__ const racc &m_dOf(const rfluid &f)   { return dir == 0 ? f.dcsi_d : (dir == 1 ? f.deta_d : f.dzet_d); }
template<int dir> __device__ __forceinline__ const racc &m_sOf(const rfluid &f)   { return dir == 0 ? f.dcsi_s : (dir == 1 ? f.deta_s : f.dzet_s); }

template<int dir> __device__ __forceinline__ const htr_acc &nTypeOf(const htr_fluid &f) { return dir == 0 ? f.nType_x : (dir == 1 ? f.nType_y : f.nType_z); }
template<int dir> __device__ __forceinline__ const htr_acc &m_eOf(const htr_fluid &f)   { return dir == 0 ? f.dcsi_e : (dir == 1 ? f.deta_e : f.dzet_e); }
template<int dir> __device__ __forceinline__ const htr_acc &m_dOf(const htr_fluid &f)   { return dir == 0 ? f.dcsi_d : (dir == 1 ? f.deta_d : f.dzet_d); }
template<int dir> __device__ __forceinline__ const htr_acc &m_sOf(const htr_fluid &f)   { return dir == 0 ? f.dcsi_s : (dir == 1 ? f.deta_s : f.dzet_s); }
template<int dir> __device__ __forceinline__ const htr_acc &sensorOf(const htr_fluid &f){ return dir == 0 ? f.shockSensorX : (dir == 1 ? f.shockSensorY : f.shockSensorZ); }

__device__ __forceinline__ double mod2(const double *v) { return v[0]*v[0] + v[1]*v[1] + v[2]*v[2]; }

// transverse directions and velocity-component indices (src/Utils/PointDomain_helper.hpp:89-109)
constexpr int T1(int dir) { return dir == 0 ? 1 : 0; }
constexpr int T2(int dir) { return dir == 2 ? 1 : 2; }

} // namespace htr
