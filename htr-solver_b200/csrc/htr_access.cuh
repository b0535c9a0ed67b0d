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
