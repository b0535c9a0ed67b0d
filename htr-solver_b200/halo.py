"""Block decomposition and ghost-plane exchange for the RHS hot path.

Mirrors what `prometeo_partitioner` declares and Legion/Realm executes in the reference
(src/prometeo_partitioner.rg:436-446 tiling, :1107-1352 ghost partitions; one instance per tile
holding interior + ghosts, src/prometeo_mapper.cc:1142-1204): the global box is cut into
tiles=[tx,ty,tz] equal blocks, one per GPU; every block keeps HALO = 3 ghost planes per side
(the Euler stencil reach, partitioner.rg:1239-1252).  Ghost data moves with NCCL send/recv
(torch.distributed P2P) in three phases x -> y -> z whose slabs include the ghosts received in the
earlier phases, so edges and corners (needed by the transverse velocity gradients of the
diffusion flux, partitioner.rg:1287-1301) arrive without diagonal messages.  Periodic wrap =
neighbour index modulo the tile count; a direction with one tile wraps onto the block itself
(local copy).

What is exchanged per RHS evaluation: Conserved and temperature only (80 B/cell for AirMix);
primitives and properties of the ghost cells are recomputed locally by the same kernels.  The
exchanged temperature is the owner's converged value, so the ghost Newton solve returns it
unchanged and ghost data are bit-identical with the owner's.
"""
import os
import numpy as np

HALO = 3


def tile_coords(rank, tiles):
    return (rank % tiles[0], (rank // tiles[0]) % tiles[1], rank // (tiles[0] * tiles[1]))


def rank_of(tc, tiles):
    return (tc[0] % tiles[0]) + tiles[0] * ((tc[1] % tiles[1]) + tiles[1] * (tc[2] % tiles[2]))


def neighbours(rank, tiles):
    """[(minus_rank, plus_rank)] per direction, periodic."""
    tc = tile_coords(rank, tiles)
    out = []
    for d in range(3):
        lo = list(tc); hi = list(tc)
        lo[d] -= 1; hi[d] += 1
        out.append((rank_of(lo, tiles), rank_of(hi, tiles)))
    return out


def phase_slices(n, d, h=HALO):
    """Index slices (z, y, x order, into arrays with h ghost planes per side) of phase d:
    returns (send_lo, send_hi, recv_lo, recv_hi).  Dimensions < d span interior+ghosts (already
    exchanged), dimensions > d span the interior only."""
    def span(dd):
        return slice(0, n[dd] + 2 * h) if dd < d else slice(h, h + n[dd])

    def mk(sl_d):
        s = [span(0), span(1), span(2)]
        s[d] = sl_d
        return (s[2], s[1], s[0])          # arrays are [z][y][x]
    send_lo = mk(slice(h, 2 * h))                       # first h interior planes  -> minus neighbour's high ghosts
    send_hi = mk(slice(n[d], n[d] + h))                 # last h interior planes   -> plus neighbour's low ghosts
    recv_lo = mk(slice(0, h))
    recv_hi = mk(slice(n[d] + h, n[d] + 2 * h))
    return send_lo, send_hi, recv_lo, recv_hi


def exchange(tensors, n, rank, tiles, dist=None, h=HALO):
    """Host-logic double of Comm.exchange in plain torch (CPU tests with gloo, tests/test_halo_cpu.py; the GPU path is the
    C++ exchange).  Fills the ghost planes of every tensor in `tensors` ([z][y][x](comp), h ghosts per side).
    Runs on whatever stream / device the tensors live on; `dist` = torch.distributed (None or
    world 1: all wraps are local)."""
    import torch
    nb = neighbours(rank, tiles)
    for d in range(3):
        s_lo, s_hi, r_lo, r_hi = phase_slices(n, d, h)
        if tiles[d] == 1:
            for t in tensors:
                t[r_hi].copy_(t[s_lo])
                t[r_lo].copy_(t[s_hi])
            continue
        minus, plus = nb[d]
        # one packed message per direction and side
        out_lo = torch.cat([t[s_lo].reshape(-1).view(torch.uint8) for t in tensors])
        out_hi = torch.cat([t[s_hi].reshape(-1).view(torch.uint8) for t in tensors])
        in_lo = torch.empty_like(out_hi)    # from the minus neighbour: its high planes
        in_hi = torch.empty_like(out_lo)
        if tiles[d] == 2:
            # both neighbours are the same rank: tag by ordering (send lo first, receive hi-side first)
            ops = [dist.P2POp(dist.isend, out_lo, minus), dist.P2POp(dist.isend, out_hi, plus),
                   dist.P2POp(dist.irecv, in_hi, plus), dist.P2POp(dist.irecv, in_lo, minus)]
        else:
            ops = [dist.P2POp(dist.isend, out_lo, minus), dist.P2POp(dist.isend, out_hi, plus),
                   dist.P2POp(dist.irecv, in_hi, plus), dist.P2POp(dist.irecv, in_lo, minus)]
        for w in dist.batch_isend_irecv(ops):
            w.wait()
        off_lo = off_hi = 0
        for t in tensors:
            v = t[r_lo]
            nb_ = v.numel() * v.element_size()
            v.copy_(in_lo[off_lo:off_lo + nb_].view(t.dtype).view(v.shape)); off_lo += nb_
            v = t[r_hi]
            nb_ = v.numel() * v.element_size()
            v.copy_(in_hi[off_hi:off_hi + nb_].view(t.dtype).view(v.shape)); off_hi += nb_


class Comm:
    """The C++ side of the exchange (include/htr_b200.h: htr_comm_*, htr_halo_exchange; csrc/htr_halo.cuh): pack kernel,
    ncclSend/ncclRecv inside one group, unpack kernel per phase, on the caller's stream.  torch.distributed is only used
    once, to hand rank 0's 128-byte NCCL id to the other ranks."""

    def __init__(self, solver, tiles, rank, world, dist=None):
        import ctypes as C
        import torch
        self.C, self.L, self.torch = C, solver.L, torch
        L = self.L
        L.htr_comm_init.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int]
        L.htr_halo_exchange.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
        L.htr_allreduce_max.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
        L.htr_comm_free.argtypes = [C.c_void_p]
        idbuf = torch.zeros(128, dtype=torch.uint8)
        if world > 1:
            if rank == 0:
                raw = (C.c_uint8 * 128)()
                rc = L.htr_comm_get_unique_id(raw)
                assert rc == 0, f"htr_comm_get_unique_id failed ({rc}): is libnccl.so.2 loadable?"
                idbuf = torch.tensor(list(raw), dtype=torch.uint8)
            idbuf = idbuf.cuda()
            dist.broadcast(idbuf, 0)
            idbuf = idbuf.cpu()
        raw = (C.c_uint8 * 128)(*idbuf.tolist())
        self.h = C.c_void_p()
        rc = L.htr_comm_init(C.byref(self.h), raw, rank, world, tiles[0], tiles[1], tiles[2])
        assert rc == 0, f"htr_comm_init failed ({rc})"

    def exchange(self, fl, names, h=HALO):
        """Ghost planes of the named fields of block `fl` from the neighbouring blocks (current stream)."""
        from .core import Acc
        C = self.C
        st = fl.struct()
        accs = (Acc * len(names))(*[getattr(st, n) for n in names])
        cb = (C.c_int * len(names))(*[fl.t[n][0, 0, 0].numel() * fl.t[n].element_size() for n in names])
        tile = fl.tile()
        rc = self.L.htr_halo_exchange(self.h, len(names), accs, cb, C.byref(tile), h,
                                      C.c_void_p(self.torch.cuda.current_stream().cuda_stream))
        if rc != 0:
            raise RuntimeError(f"htr_halo_exchange failed with code {rc}")

    def allreduce_max(self, t):
        rc = self.L.htr_allreduce_max(self.h, self.C.c_void_p(t.data_ptr()), t.numel(),
                                      self.C.c_void_p(self.torch.cuda.current_stream().cuda_stream))
        if rc != 0:
            raise RuntimeError(f"htr_allreduce_max failed with code {rc}")

    def __del__(self):
        try:
            if self.h:
                self.L.htr_comm_free(self.h)
        except Exception:
            pass


def shell_rects(origin, n, h=HALO):
    """The ghost region as 6 disjoint boxes (global coordinates, inclusive): x slabs over the interior
    y,z; y slabs over x incl. ghosts; z slabs over x,y incl. ghosts."""
    o = origin
    lo = [o[0], o[1], o[2]]; hi = [o[0] + n[0] - 1, o[1] + n[1] - 1, o[2] + n[2] - 1]
    rects = []
    for d in range(3):
        a = [lo[k] - (h if k < d else 0) for k in range(3)]
        b = [hi[k] + (h if k < d else 0) for k in range(3)]
        r1a, r1b = list(a), list(b); r1a[d] = lo[d] - h; r1b[d] = lo[d] - 1
        r2a, r2b = list(a), list(b); r2a[d] = hi[d] + 1; r2b[d] = hi[d] + h
        rects += [(tuple(r1a), tuple(r1b)), (tuple(r2a), tuple(r2b))]
    return rects


# Shell thickness per direction (x, y, z): at least the stencil reach (HALO), and exactly what ONE tile of the staged
# kernels updates in that direction (csrc/htr_fast.cuh: tiles overlap by one cell per flux direction, so a 32 x 8 tile
# updates 31 x 7 cells and the 8-plane z tiles 7 planes): a shell is then one tile thick in every kernel.
SHELL_W = (31, 7, 7)


def boundary_rects(origin, n, w=HALO):
    """Interior cells within w (scalar or per-direction) of the block boundary as 6 disjoint boxes + the
    inner box that needs no ghost data.  Returns (inner, [shell boxes]); boxes may be empty (lo > hi)."""
    o = origin
    w = (w, w, w) if np.isscalar(w) else tuple(w)
    assert min(w) >= HALO
    lo = [o[0], o[1], o[2]]; hi = [o[0] + n[0] - 1, o[1] + n[1] - 1, o[2] + n[2] - 1]
    inner = (tuple(l + wd for l, wd in zip(lo, w)), tuple(u - wd for u, wd in zip(hi, w)))
    shells = []
    a, b = list(lo), list(hi)
    for d in (2, 1, 0):            # peel z slabs (full x,y), then y slabs, then x slabs
        s1a, s1b = list(a), list(b); s1b[d] = min(lo[d] + w[d] - 1, hi[d])
        s2a, s2b = list(a), list(b); s2a[d] = max(hi[d] - w[d] + 1, s1b[d] + 1)
        shells += [(tuple(s1a), tuple(s1b)), (tuple(s2a), tuple(s2b))]
        a[d] = lo[d] + w[d]; b[d] = hi[d] - w[d]
    return inner, shells


def overlap_default(tiles, env=None):
    """Exchange policy of HaloExchanger.rhs: True = ghost exchange on a side stream under the inner box, shells afterwards;
    False = exchange first, then the whole block as one htr_rhs call.  The overlap pays from two remote directions on (see
    HaloExchanger.__init__ for the measurement); HTR_OVERLAP=0/1 in the environment forces either form."""
    env = os.environ.get("HTR_OVERLAP") if env is None else env
    if env not in (None, ""):
        return env != "0"
    return sum(1 for t in tiles if t > 1) >= 2


class HaloExchanger:
    """Per-rank driver of one block: ghost exchange overlapped with the interior RHS."""

    def __init__(self, solver, fl, tiles, tc, rank, world):
        import torch
        import torch.distributed as dist
        from .core import Rect
        self.torch, self.dist, self.Rect = torch, dist, Rect
        self.s, self.fl, self.tiles, self.tc, self.rank, self.world = solver, fl, tuple(tiles), tuple(tc), rank, world
        self.n = (fl.nx, fl.ny, fl.nz)
        assert fl.halos() == (HALO, HALO, HALO)
        self.comm = torch.cuda.Stream()
        self.nccl = Comm(solver, self.tiles, rank, world, dist)
        self._cfl = torch.zeros(1, dtype=torch.float64, device="cuda")
        self.ghost_boxes = [Rect.make(a, b) for a, b in shell_rects(fl.origin, self.n)]
        w = tuple(min(SHELL_W[d], max(HALO, self.n[d] // 4)) for d in range(3))
        inner, shells = boundary_rects(fl.origin, self.n, w)
        self.inner = Rect.make(*inner)
        self.shells = [Rect.make(a, b) for a, b in shells if all(x <= y for x, y in zip(a, b))]
        self.nowrap = Rect.make((-2 ** 40,) * 3, (2 ** 40,) * 3)
        # Overlap the exchange with the inner box only when it is long enough to pay for the box decomposition.  Measured on
        # B200, 512^3 per GPU (profiles/r2_v28_*, 2 GPUs): the block as inner box + 6 shells costs 64.2 ms against 60.9 ms as
        # one htr_rhs call (84 launches on boxes one tile thick), the exchange + ghost recovery it hides 2.0 ms per remote
        # direction -- so with ONE remote direction the step is faster with the exchange up front and the whole block in one
        # call; from two remote directions on the overlap wins.  HTR_OVERLAP=0/1 forces either form.
        self.overlap = overlap_default(self.tiles)

    # ghost planes of the state, then ghost primitives/properties (on the current stream)
    def exchange_state(self):
        fl = self.fl
        self.nccl.exchange(fl, ["Conserved", "temperature"])
        for box in self.ghost_boxes:      # the owner's bits: T is the owner's converged value, no second Newton pass
            self.s.UpdateGhostPrimitiveFromConserved(fl, tile=box)
            self.s.UpdatePropertiesFromPrimitive(fl, tile=box)

    def exchange_old(self):
        """Ghost planes of Conserved_old.  The multi-species positivity limiter of the TENO fluxes reads it at both
        cells of a face (src/prometeo_rhs.inl, partial densities after the stage), hence in the first ghost plane:
        call this whenever Conserved_old was set on the interior only.  initialize_temporaries() needs no message."""
        self.nccl.exchange(self.fl, ["Conserved_old"])

    def initialize_temporaries(self):
        """InitializeTemporaries (src/prometeo_rk.rg:46-56) over interior AND ghosts; the ghost planes of Conserved must
        be current (exchange_state / sensors just ran), so Conserved_old needs no exchange of its own."""
        self.fl.t["Conserved_old"].copy_(self.fl.t["Conserved"])

    def time_step(self, scheme, Integrator_deltaTime, diffusion=True, chemistry=False, vorticityScale=1.0):
        """One explicit RK3 step of the block (MainLoopBody, src/prometeo.rg:1348-1470; the multi-GPU form of
        Solver.TimeStep): sensors with their exchanges, InitializeTemporaries, then per stage the overlapped
        exchange + RHS and the fused UpdateVars + primitive/property recovery of the interior."""
        from .core import RK_C
        fl, s = self.fl, self.s
        if scheme == "Hybrid":
            self.sensors(vorticityScale)
        else:
            self.exchange_state()
        self.initialize_temporaries()
        self._ghosts_current = True      # the ghosts hold the state stage 1 differentiates: its rhs() needs no second exchange
        for stage in (1, 2, 3):
            c1, c2, c3 = RK_C[stage - 1]
            self.rhs(scheme, diffusion, chemistry, (c1, c2), Integrator_deltaTime*c3)
            s.UpdateVarsAndRecover(fl, stage, Integrator_deltaTime, False, tile=fl.tile())

    def sensors(self, vorticityScale):
        """UpdateShockSensors (src/prometeo_sensor.rg:90-124) with the two ghost exchanges it implies."""
        fl, s = self.fl, self.s
        self.exchange_state()
        s.UpdateDucrosSensor(fl, vorticityScale, bounds=self.nowrap)
        self.nccl.exchange(fl, ["DucrosSensor"])
        for d in range(3):
            s.UpdateShockSensor(fl, d, bounds=self.nowrap)
        self.nccl.exchange(fl, ["shockSensorX", "shockSensorY", "shockSensorZ"])

    def _fluxes(self, box, scheme, diffusion, rk, dt):
        fl, s = self.fl, self.s
        for d in (2, 1, 0):
            s.UpdateUsingEulerFlux(fl, scheme, d, rk, dt, tile=box, bounds=self.nowrap)
        if diffusion:
            for d in (2, 1, 0):
                s.UpdateUsingDiffusionFlux(fl, d, tile=box, bounds=self.nowrap)

    def rhs(self, scheme, diffusion, chemistry, rk, dt):
        """One UpdateDerivatives pass of the block.  The ghost exchange runs on a side stream while the
        cells that need no ghost data are updated; the boundary shells follow.  Every cell belongs to
        exactly one box and each box runs the reference's order (chem, Euler z,y,x, diffusion z,y,x)."""
        torch, fl, s = self.torch, self.fl, self.s
        main = torch.cuda.current_stream()
        fresh, self._ghosts_current = getattr(self, "_ghosts_current", False), False
        if not self.overlap:     # exchange up front, then the whole block as one htr_rhs call
            if not fresh:
                self.exchange_state()
            s.UpdateDerivatives(fl, scheme, diffusion, chemistry, rk, dt, tile=fl.tile(), bounds=self.nowrap)
            return
        if not fresh:            # time_step: sensors() has just exchanged this very state for stage 1
            self.comm.wait_stream(main)
            with torch.cuda.stream(self.comm):
                self.exchange_state()
        # each box is one htr_rhs call (0 + chem + Euler z,y,x + diffusion z,y,x over the box)
        s.UpdateDerivatives(fl, scheme, diffusion, chemistry, rk, dt, tile=self.inner, bounds=self.nowrap)
        main.wait_stream(self.comm)
        for box in self.shells:
            s.UpdateDerivatives(fl, scheme, diffusion, chemistry, rk, dt, tile=box, bounds=self.nowrap)

    def max_spectral_radius(self):
        """CalculateMaxSpectralRadius of the block followed by the MAX over all ranks (src/prometeo.rg:1161-1165): the
        same time step on every rank."""
        t = self.s.CalculateMaxSpectralRadius(self.fl, sync=False)
        self.nccl.allreduce_max(t)
        return float(t.item())

    def launches_per_step(self, diffusion, chemistry, scheme="Hybrid"):
        from .core import rhs_launches
        boxes = (1 + len(self.shells)) if self.overlap else 1
        return rhs_launches(scheme, diffusion, chemistry) * boxes + 2 * len(self.ghost_boxes)
