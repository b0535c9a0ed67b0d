// Ghost-plane exchange of a block-decomposed box behind the C ABI (include/htr_b200.h: htr_comm_*, htr_halo_exchange).
//
// What prometeo_partitioner declares and Legion/Realm executes in the reference (src/prometeo_partitioner.rg:436-446
// tiling, :1107-1352 ghost partitions): the global box is cut into tiles = [tx, ty, tz] equal blocks, one per GPU, every
// block holds `halo` ghost planes per side in the same allocation.  Here the copies are explicit: per direction one pack
// kernel gathers the two face slabs of ALL fields into two contiguous messages, NCCL moves them to the +/- neighbours
// (ncclGroupStart / ncclSend / ncclRecv / ncclGroupEnd over NVLink), one unpack kernel scatters the received slabs into
// the ghost planes.  The three phases run x -> y -> z and the slabs of a phase span the ghosts received in the earlier
// phases, so edges and corners (read by the transverse velocity gradients of the diffusion flux,
// src/prometeo_partitioner.rg:1287-1301) arrive without diagonal messages.  A direction with one tile wraps onto the block
// itself with a single copy kernel and no message.
//
// NCCL is bound at run time (dlopen of libnccl.so.2: inside a PyTorch process that is the copy torch already loaded), so
// the library has no link-time dependency on it and single-GPU users never load it.
#pragma once
#include <dlfcn.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <string.h>
#include "../../include/htr_b200.h"

namespace htr {
namespace halo {

// ---- the few NCCL entry points used, resolved with dlsym (types per nccl.h 2.x) ------------------------------------
typedef struct ncclComm *ncclComm_t;
typedef struct { char internal[128]; } ncclUniqueId;
enum { ncclSuccess = 0 };
enum { ncclChar = 0, ncclUint8 = 1, ncclFloat64 = 8 };       // ncclDataType_t
enum { ncclMax = 2 };                                        // ncclRedOp_t
struct Nccl {
   void *lib;
   int (*GetUniqueId)(ncclUniqueId*);
   int (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int);
   int (*CommDestroy)(ncclComm_t);
   int (*GroupStart)();
   int (*GroupEnd)();
   int (*Send)(const void*, size_t, int, int, ncclComm_t, cudaStream_t);
   int (*Recv)(void*, size_t, int, int, ncclComm_t, cudaStream_t);
   int (*AllReduce)(const void*, void*, size_t, int, int, ncclComm_t, cudaStream_t);
};
static Nccl *nccl() {
   static Nccl n = {};
   static bool tried = false;
   if (!tried) {
      tried = true;
      n.lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
      if (!n.lib) n.lib = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
      if (n.lib) {
         #define HTR_SYM(f) *(void**)(&n.f) = dlsym(n.lib, "nccl" #f)
         HTR_SYM(GetUniqueId); HTR_SYM(CommInitRank); HTR_SYM(CommDestroy); HTR_SYM(GroupStart); HTR_SYM(GroupEnd);
         HTR_SYM(Send); HTR_SYM(Recv); HTR_SYM(AllReduce);
         #undef HTR_SYM
         if (!n.GetUniqueId || !n.CommInitRank || !n.CommDestroy || !n.GroupStart || !n.GroupEnd || !n.Send || !n.Recv || !n.AllReduce) n.lib = nullptr;
      }
   }
   return n.lib ? &n : nullptr;
}

struct Comm {
   int rank, nranks, tiles[3], tc[3];
   ncclComm_t nc;
   char *buf;            // [send_lo | send_hi | recv_lo | recv_hi], each `cap` bytes
   size_t cap;
   int device;
};

constexpr int MAXF = 16;
struct Field { char *base; int64_t s[3]; int cellBytes; int64_t off; };   // off: byte offset of this field inside a message
struct Job {
   Field f[MAXF]; int nf;
   int64_t lo[3], n[3];          // slab origin (global coordinates) and extents
   int64_t cells;
};

// mode 0: block -> message (pack), 1: message -> block (unpack), 2: block -> block shifted by `shift` cells along `dir`
template<int MODE>
__global__ void __launch_bounds__(256) slab_kernel(const Job j, char *__restrict__ msg, const int dir, const int64_t shift) {
   const int64_t total = j.cells;
   for (int64_t c = (int64_t)blockIdx.x*blockDim.x + threadIdx.x; c < total; c += (int64_t)gridDim.x*blockDim.x) {
      const int64_t x = c % j.n[0], y = (c/j.n[0]) % j.n[1], z = c/(j.n[0]*j.n[1]);
      int64_t p[3] = {j.lo[0] + x, j.lo[1] + y, j.lo[2] + z};
      for (int k = 0; k < j.nf; k++) {
         const Field &f = j.f[k];
         char *cell = f.base + p[0]*f.s[0] + p[1]*f.s[1] + p[2]*f.s[2];
         char *other;
         if (MODE == 2) { int64_t q[3] = {p[0], p[1], p[2]}; q[dir] += shift; other = f.base + q[0]*f.s[0] + q[1]*f.s[1] + q[2]*f.s[2]; }
         else other = msg + f.off + c*f.cellBytes;
         char *dst = (MODE == 0) ? other : ((MODE == 1) ? cell : other);
         const char *src = (MODE == 0) ? cell : ((MODE == 1) ? other : cell);
         if ((f.cellBytes & 7) == 0 && (((uintptr_t)dst | (uintptr_t)src) & 7) == 0) {
            for (int b = 0; b < f.cellBytes; b += 8) *reinterpret_cast<double*>(dst + b) = *reinterpret_cast<const double*>(src + b);
         } else {
            for (int b = 0; b < f.cellBytes; b++) dst[b] = src[b];
         }
      }
   }
}

static inline unsigned gridFor(int64_t cells) {
   const int64_t g = (cells + 255)/256;
   return (unsigned)(g < 1 ? 1 : (g > 148*16 ? 148*16 : g));
}

} // namespace halo
} // namespace htr

extern "C" {

int htr_comm_get_unique_id(void *id128) {
   using namespace htr::halo;
   Nccl *n = nccl();
   if (!n || !id128) return HTR_EINVAL;
   ncclUniqueId id;
   if (n->GetUniqueId(&id) != ncclSuccess) return 2000;
   memcpy(id128, &id, sizeof(id));
   return HTR_OK;
}

int htr_comm_init(htr_comm **out, const void *id128, int rank, int nranks, int tx, int ty, int tz) {
   using namespace htr::halo;
   if (!out || nranks < 1 || rank < 0 || rank >= nranks || tx < 1 || ty < 1 || tz < 1 || tx*ty*tz != nranks) return HTR_EINVAL;
   Comm *c = new Comm();
   c->rank = rank; c->nranks = nranks; c->tiles[0] = tx; c->tiles[1] = ty; c->tiles[2] = tz;
   c->tc[0] = rank % tx; c->tc[1] = (rank/tx) % ty; c->tc[2] = rank/(tx*ty);        // htr-solver_b200/halo.py: tile_coords
   c->nc = nullptr; c->buf = nullptr; c->cap = 0;
   cudaGetDevice(&c->device);
   if (nranks > 1) {
      Nccl *n = nccl();
      if (!n || !id128) { delete c; return HTR_EINVAL; }
      ncclUniqueId id; memcpy(&id, id128, sizeof(id));
      const int rc = n->CommInitRank(&c->nc, nranks, id, rank);
      if (rc != ncclSuccess) { delete c; return 2000 + rc; }
   }
   *out = reinterpret_cast<htr_comm*>(c);
   return HTR_OK;
}

int htr_comm_free(htr_comm *comm) {
   using namespace htr::halo;
   Comm *c = reinterpret_cast<Comm*>(comm);
   if (!c) return HTR_OK;
   if (c->buf) cudaFree(c->buf);
   if (c->nc && nccl()) nccl()->CommDestroy(c->nc);
   delete c;
   return HTR_OK;
}

int htr_comm_rank_of_tile(const htr_comm *comm, int cx, int cy, int cz) {
   const htr::halo::Comm *c = reinterpret_cast<const htr::halo::Comm*>(comm);
   if (!c) return -1;
   const int *t = c->tiles;
   auto m = [](int a, int n) { return ((a % n) + n) % n; };
   return m(cx, t[0]) + t[0]*(m(cy, t[1]) + t[1]*m(cz, t[2]));
}

int htr_halo_exchange(htr_comm *comm, int nfields, const htr_acc *fields, const int *cellBytes, const htr_rect *interior,
                      int halo, void *stream) {
   using namespace htr::halo;
   Comm *c = reinterpret_cast<Comm*>(comm);
   if (!c || !fields || !cellBytes || !interior || nfields < 1 || nfields > MAXF || halo < 1) return HTR_EINVAL;
   cudaStream_t st = (cudaStream_t)stream;
   const int64_t lo[3] = {interior->lo[0], interior->lo[1], interior->lo[2]};
   const int64_t n[3] = {interior->hi[0] - lo[0] + 1, interior->hi[1] - lo[1] + 1, interior->hi[2] - lo[2] + 1};
   const int64_t h = halo;
   for (int d = 0; d < 3; d++) if (n[d] < h) return HTR_EINVAL;
   Job base;
   base.nf = nfields;
   for (int d = 0; d < 3; d++) {
      // slab of phase d: dimensions below d span interior + ghosts (already exchanged), dimensions above d the interior only
      Job jb = base;
      for (int e = 0; e < 3; e++) { jb.lo[e] = (e < d) ? lo[e] - h : lo[e]; jb.n[e] = (e < d) ? n[e] + 2*h : n[e]; }
      jb.n[d] = h;
      jb.cells = jb.n[0]*jb.n[1]*jb.n[2];
      int64_t off = 0;                                  // 16-byte aligned field sections inside a message
      for (int k = 0; k < nfields; k++) {
         jb.f[k].base = (char*)fields[k].base;
         for (int e = 0; e < 3; e++) jb.f[k].s[e] = fields[k].stride[e];
         jb.f[k].cellBytes = cellBytes[k];
         jb.f[k].off = off;
         off += ((int64_t)cellBytes[k]*jb.cells + 15)/16*16;
      }
      const size_t msgBytes = (size_t)off;
      const unsigned grid = gridFor(jb.cells);
      Job sendLo = jb, sendHi = jb, recvLo = jb, recvHi = jb;
      sendLo.lo[d] = lo[d];                 // first h interior planes  -> minus neighbour's high ghosts
      sendHi.lo[d] = lo[d] + n[d] - h;      // last h interior planes   -> plus neighbour's low ghosts
      recvLo.lo[d] = lo[d] - h;
      recvHi.lo[d] = lo[d] + n[d];
      if (c->tiles[d] == 1) {
         // periodic wrap onto the block itself: two shifted copies, no message
         slab_kernel<2><<<grid, 256, 0, st>>>(sendLo, nullptr, d, n[d]);       // low interior planes  -> high ghosts
         slab_kernel<2><<<grid, 256, 0, st>>>(sendHi, nullptr, d, -n[d]);      // high interior planes -> low ghosts
         const cudaError_t e = cudaGetLastError();
         if (e != cudaSuccess) return 1000 + (int)e;
         continue;
      }
      Nccl *nl = nccl();
      if (!nl || !c->nc) return HTR_EINVAL;
      if (c->cap < msgBytes) {
         if (c->buf) { cudaStreamSynchronize(st); cudaFree(c->buf); c->buf = nullptr; }
         if (cudaMalloc((void**)&c->buf, 4*msgBytes) != cudaSuccess) { c->cap = 0; return 1000 + (int)cudaErrorMemoryAllocation; }
         c->cap = msgBytes;
      }
      char *outLo = c->buf, *outHi = c->buf + c->cap, *inLo = c->buf + 2*c->cap, *inHi = c->buf + 3*c->cap;
      slab_kernel<0><<<grid, 256, 0, st>>>(sendLo, outLo, d, 0);
      slab_kernel<0><<<grid, 256, 0, st>>>(sendHi, outHi, d, 0);
      int tcm[3] = {c->tc[0], c->tc[1], c->tc[2]}, tcp[3] = {c->tc[0], c->tc[1], c->tc[2]};
      tcm[d] -= 1; tcp[d] += 1;
      const int minus = htr_comm_rank_of_tile(comm, tcm[0], tcm[1], tcm[2]), plus = htr_comm_rank_of_tile(comm, tcp[0], tcp[1], tcp[2]);
      // with two tiles both neighbours are the same rank: messages between a pair are matched in order, so the peer's first
      // message (its low planes) is received into the HIGH ghosts first
      int rc = nl->GroupStart();
      if (rc == ncclSuccess) rc = nl->Send(outLo, msgBytes, ncclChar, minus, c->nc, st);
      if (rc == ncclSuccess) rc = nl->Send(outHi, msgBytes, ncclChar, plus, c->nc, st);
      if (rc == ncclSuccess) rc = nl->Recv(inHi, msgBytes, ncclChar, plus, c->nc, st);
      if (rc == ncclSuccess) rc = nl->Recv(inLo, msgBytes, ncclChar, minus, c->nc, st);
      const int rc2 = nl->GroupEnd();
      if (rc != ncclSuccess || rc2 != ncclSuccess) return 2000 + (rc != ncclSuccess ? rc : rc2);
      slab_kernel<1><<<grid, 256, 0, st>>>(recvLo, inLo, d, 0);
      slab_kernel<1><<<grid, 256, 0, st>>>(recvHi, inHi, d, 0);
      const cudaError_t e = cudaGetLastError();
      if (e != cudaSuccess) return 1000 + (int)e;
   }
   return HTR_OK;
}

int htr_allreduce_max(htr_comm *comm, double *d_values, int count, void *stream) {
   using namespace htr::halo;
   Comm *c = reinterpret_cast<Comm*>(comm);
   if (!c || !d_values || count < 1) return HTR_EINVAL;
   if (c->nranks == 1) return HTR_OK;
   Nccl *nl = nccl();
   if (!nl || !c->nc) return HTR_EINVAL;
   const int rc = nl->AllReduce(d_values, d_values, (size_t)count, ncclFloat64, ncclMax, c->nc, (cudaStream_t)stream);
   return rc == ncclSuccess ? HTR_OK : 2000 + rc;
}

} // extern "C"
