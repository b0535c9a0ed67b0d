// TEST INFRASTRUCTURE ONLY -- minimal stand-in for Legion's "legion.h".
//
// The reference's hot-path numerics (/root/reference/src/prometeo_*.{hpp,inl},
// Mixtures/*, Utils/*) are header-style C++ that only *name* a handful of
// Legion types.  This stub provides exactly those names so that the reference
// sources can be compiled IN PLACE (never copied) into oracle/_ref/ and driven
// over raw host arrays by oracle/ref_driver.cc.  It contains no HTR code.
// Surface list: SURVEY.md section 8(c).
#ifndef HTR_ORACLE_STUB_LEGION_H
#define HTR_ORACLE_STUB_LEGION_H

#include <cstdint>
#include <cstddef>
#include <cstring>
#include <vector>
#include <cassert>
#include <cmath>
#include <algorithm>
#include <cstdio>
#include <cstdlib>

enum legion_privilege_mode_stub { READ_ONLY = 1, READ_WRITE = 7, WRITE_ONLY = 0x10000002, WRITE_DISCARD = 0x10000002, REDUCE = 8 };

namespace Realm {
   template<typename FT, int N, typename T> struct AffineAccessor {};
   namespace Cuda {
      // realm/cuda/cuda_access.h: the stream of the running GPU task (used when the CUDA runtime hijack is off)
      inline void *&stub_task_stream() { static void *s = 0; return s; }
      inline void *get_task_cuda_stream() { return stub_task_stream(); }
   }
}

namespace Legion {
   typedef long long coord_t;
   typedef unsigned int FieldID;
   typedef unsigned int LayoutConstraintID;
   typedef unsigned int TaskID;
   typedef void* Context;

   template<int N, typename T = coord_t> struct Point;
   template<typename T> struct Point<1, T> {
      T x;
      Point() : x(0) {}
      Point(T x_) : x(x_) {}
      operator T() const { return x; }
      T operator[](int) const { return x; }
   };
   template<typename T> struct Point<3, T> {
      T x, y, z;
      Point() : x(0), y(0), z(0) {}
      Point(T x_, T y_, T z_) : x(x_), y(y_), z(z_) {}
      T operator[](int i) const { return i == 0 ? x : (i == 1 ? y : z); }
   };
   template<typename T> inline Point<3, T> operator+(const Point<3, T> &a, const Point<3, T> &b) { return Point<3, T>(a.x + b.x, a.y + b.y, a.z + b.z); }
   template<typename T> inline Point<3, T> operator-(const Point<3, T> &a, const Point<3, T> &b) { return Point<3, T>(a.x - b.x, a.y - b.y, a.z - b.z); }
   template<typename T> inline bool operator==(const Point<3, T> &a, const Point<3, T> &b) { return a.x == b.x && a.y == b.y && a.z == b.z; }
   template<int N, typename T = coord_t> struct Rect {
      Point<N, T> lo, hi;
      Rect() {}
      Rect(const Point<N, T> &l, const Point<N, T> &h) : lo(l), hi(h) {}
      size_t volume() const;
   };
   template<> inline size_t Rect<3, coord_t>::volume() const {
      return size_t(hi.x - lo.x + 1)*size_t(hi.y - lo.y + 1)*size_t(hi.z - lo.z + 1);
   }
   template<> inline size_t Rect<1, coord_t>::volume() const { return size_t(hi.x - lo.x + 1); }

   struct Domain {
      Rect<3> r3; Rect<1> r1; bool dense;
      std::vector<Rect<3> > pieces;
      Domain() : dense(true) {}
      Domain(const Rect<3> &r, bool d = true) : r3(r), dense(d) {}
      operator Rect<3>() const {
         if (!dense) { fprintf(stderr, "stub legion.h: Domain -> Rect<3> conversion of a SPARSE index space (Legion aborts here)\n"); abort(); }
         return r3;
      }
      operator Rect<1>() const { return r1; }
   };

   // An index space is a bounding rectangle plus a flag: the reference's ghost partitions are unions of shifted /
   // wrapped pieces (src/prometeo_partitioner.rg:1239+), i.e. SPARSE -- converting such a domain to a Rect aborts in
   // Legion, and so it does here (tests/test_adapter_gpu.py drives the adapter with sparse ghost regions).
   struct IndexSpace {
      Rect<3> bounds; bool dense;
      std::vector<Rect<3> > pieces;          // the rectangles of a sparse space (RectInDomainIterator walks them)
      IndexSpace() : dense(true) {}
      IndexSpace(const Rect<3> &b, bool d = true) : bounds(b), dense(d) {}
   };
   struct FieldSpace { unsigned id; };
   struct LogicalRegion {
      unsigned tree_id; unsigned long long index_space_id; unsigned is_tid; unsigned type_tag; unsigned field_space_id;
      const IndexSpace *stub_is;
      LogicalRegion() : tree_id(0), index_space_id(0), is_tid(0), type_tag(0), field_space_id(0), stub_is(0) {}
      IndexSpace get_index_space() const { return stub_is ? *stub_is : IndexSpace(); }
   };
   // One field of a physical instance: base addresses global point (0,0,0), byte strides (Realm::AffineAccessor).
   struct StubField { FieldID fid; void *base; coord_t s[3]; };
   struct PhysicalRegion {
      IndexSpace is;
      std::vector<StubField> fields;       // exactly the fields this region requirement names
      LogicalRegion get_logical_region() const { LogicalRegion l; l.stub_is = &is; return l; }
      const StubField *find(FieldID fid) const {
         for (size_t i = 0; i < fields.size(); i++) if (fields[i].fid == fid) return &fields[i];
         return 0;
      }
   };
   struct Future {
      double v;
      template<typename T> T get_result() const { return T(v); }
   };
   struct Task {
      void *args; size_t arglen; void *local_args; size_t local_arglen;
      std::vector<Future> futures;
   };
   struct Processor { enum Kind { NO_KIND, TOC_PROC, LOC_PROC, UTIL_PROC, IO_PROC, PROC_GROUP, PROC_SET, OMP_PROC, PY_PROC }; };
   struct Memory { enum Kind { NO_MEMKIND, GLOBAL_MEM, SYSTEM_MEM, REGDMA_MEM, SOCKET_MEM, Z_COPY_MEM, GPU_FB_MEM }; };
   struct ProcessorConstraint { ProcessorConstraint(Processor::Kind) {} };
   struct OrderingConstraint { template<typename V> OrderingConstraint(const V&, bool) {} };
   struct LayoutConstraintRegistrar {
      LayoutConstraintRegistrar() {}
      template<typename A, typename B> LayoutConstraintRegistrar(A, B) {}
      template<typename C> LayoutConstraintRegistrar& add_constraint(const C&) { return *this; }
   };
   struct TaskVariantRegistrar {
      TaskVariantRegistrar(TaskID, const char*) {}
      template<typename C> TaskVariantRegistrar& add_constraint(const C&) { return *this; }
      TaskVariantRegistrar& add_layout_constraint_set(unsigned, LayoutConstraintID) { return *this; }
      void set_leaf(bool = true) {}
   };
   struct Runtime {
      Domain get_index_space_domain(Context, const IndexSpace &is) { Domain d(is.bounds, is.dense); d.pieces = is.pieces; return d; }
      template<void (*F)(const Task*, const std::vector<PhysicalRegion>&, Context, Runtime*)>
      static void preregister_task_variant(const TaskVariantRegistrar&, const char*) {}
      template<typename R, R (*F)(const Task*, const std::vector<PhysicalRegion>&, Context, Runtime*)>
      static void preregister_task_variant(const TaskVariantRegistrar&, const char*) {}
      static LayoutConstraintID preregister_layout(const LayoutConstraintRegistrar&) { return 0; }
   };
   typedef int ReductionOpID;
   // Legion's built-in sum reduction: the primary template covers the scalar types, src/prometeo_redop.inl specialises
   // apply / fold for Vec3, VecNSp, ...
   template<typename T> class SumReduction {
   public:
      typedef T LHS;
      typedef T RHS;
      template<bool EXCLUSIVE> static void apply(LHS &lhs, RHS rhs);
      template<bool EXCLUSIVE> static void fold(RHS &rhs1, RHS rhs2);
   };
   template<typename T> template<bool EXCLUSIVE> inline void SumReduction<T>::apply(LHS &lhs, RHS rhs) { lhs += rhs; }
   template<typename T> template<bool EXCLUSIVE> inline void SumReduction<T>::fold(RHS &rhs1, RHS rhs2) { rhs1 += rhs2; }
   // GPU variants return a DeferredValue whose storage is device-visible; the stub keeps a caller-settable pointer
   template<typename T> struct DeferredValue {
      T v; T *p;
      static T *&stub_storage() { static T *s = 0; return s; }       // test harness: device buffer for the result
      DeferredValue(T x = T()) : v(x), p(stub_storage()) {}
      T* ptr() { return p ? p : &v; }
   };
   template<typename T, int N> struct DeferredBuffer {};
   template<int N> struct RectInDomainIterator {
      std::vector<Rect<3> > rects; size_t k;
      RectInDomainIterator(const Domain &d) : rects(d.pieces), k(0) { if (rects.empty()) rects.push_back(d.r3); }
      bool operator()() const { return k < rects.size(); }
      const Rect<3>& operator*() const { return rects[k]; }
      RectInDomainIterator& operator++(int) { k++; return *this; }
   };

   // Accessor over a raw strided host array (byte strides, like Realm::AffineAccessor).
   template<int PRIV, typename FT, int N, typename T = coord_t, typename A = Realm::AffineAccessor<FT, N, T> >
   struct FieldAccessor;
   template<int PRIV, typename FT, typename T, typename A>
   struct FieldAccessor<PRIV, FT, 3, T, A> {
      char *base; coord_t sx, sy, sz;   // base addresses global point (0,0,0)
      FieldAccessor() : base(0), sx(0), sy(0), sz(0) {}
      Rect<3> bounds; bool bounded;
      // the accessor of a field the region requirement does not name is a Legion runtime error
      FieldAccessor(const PhysicalRegion &r, FieldID fid) : base(0), sx(0), sy(0), sz(0), bounds(r.is.bounds), bounded(true) {
         const StubField *f = r.find(fid);
         if (!f) { fprintf(stderr, "stub legion.h: field %u is not part of this physical region\n", fid); abort(); }
         base = (char*)f->base; sx = f->s[0]; sy = f->s[1]; sz = f->s[2];
      }
      FieldAccessor(void *b, coord_t sx_, coord_t sy_, coord_t sz_) : base((char*)b), sx(sx_), sy(sy_), sz(sz_), bounded(false) {}
      FT& operator[](const Point<3, T> &p) const { return *(FT*)(base + p.x*sx + p.y*sy + p.z*sz); }
      // Legion: pointer to r.lo and strides in units of field_size; r must lie inside the instance
      FT* ptr(const Rect<3, T> &r, size_t strides[3], size_t field_size = sizeof(FT)) const {
         if (bounded && (r.lo.x < bounds.lo.x || r.lo.y < bounds.lo.y || r.lo.z < bounds.lo.z ||
                         r.hi.x > bounds.hi.x || r.hi.y > bounds.hi.y || r.hi.z > bounds.hi.z)) {
            fprintf(stderr, "stub legion.h: ptr(rect) outside the bounds of the instance\n"); abort();
         }
         strides[0] = sx/field_size; strides[1] = sy/field_size; strides[2] = sz/field_size;
         return (FT*)(base + r.lo.x*sx + r.lo.y*sy + r.lo.z*sz);
      }
   };
   template<int PRIV, typename FT, typename T, typename A>
   struct FieldAccessor<PRIV, FT, 1, T, A> {
      char *base; coord_t sx;
      FieldAccessor() : base(0), sx(0) {}
      FieldAccessor(const PhysicalRegion&, FieldID) : base(0), sx(0) {}
      FieldAccessor(void *b, coord_t sx_) : base((char*)b), sx(sx_) {}
      FT& operator[](const Point<1, T> &p) const { return *(FT*)(base + p.x*sx); }
      FT& operator[](coord_t i) const { return *(FT*)(base + i*sx); }
   };
   // reduction accessor: acc[i] <<= v folds v with the reduction operator (src/prometeo_bc.hpp:85-88 is the only user here)
   template<typename REDOP, bool EXCLUSIVE>
   struct StubReducer {
      typename REDOP::LHS *lhs;
      void operator<<=(const typename REDOP::RHS &v) const { REDOP::template apply<EXCLUSIVE>(*lhs, v); }
   };
   template<typename REDOP, bool EXCLUSIVE, int N, typename T = coord_t, typename A = void>
   struct ReductionAccessor {
      char *base; coord_t sx;
      ReductionAccessor() : base(0), sx(0) {}
      ReductionAccessor(const PhysicalRegion&, FieldID, ReductionOpID) : base(0), sx(0) {}
      StubReducer<REDOP, EXCLUSIVE> operator[](coord_t i) const { return StubReducer<REDOP, EXCLUSIVE>{(typename REDOP::LHS*)(base + i*sx)}; }
   };
   // bit casts used by the compare-and-swap loops of the reduction operators (src/prometeo_redop.inl)
   namespace TypePunning {
      template<typename T> class Pointer {
         void *p;
      public:
         Pointer(void *ptr) : p(ptr) {}
         operator T*() const { return (T*)p; }
         T operator[](size_t i) const { T v; memcpy(&v, (char*)p + i*sizeof(T), sizeof(T)); return v; }
      };
      template<typename T1, typename T2> class Alias {
         union { T1 a; T2 b; } u;
      public:
         Alias() { memset(&u, 0, sizeof(u)); }
         void load(const Pointer<T1> &ptr) { u.a = ptr[0]; }
         T1 as_one() const { return u.a; }
         T2 as_two() const { return u.b; }
         Alias& operator=(T2 v) { u.b = v; return *this; }
      };
   }
}
#endif
