"""Multi-GPU layer: one process per GPU (torch.distributed, NCCL over NVLink on
the B200 box, gloo in CPU tests), each owning one shard of the particle set.

What shards how:
  * likelihood_weight / incorporate / resimulation_mh / log_likelihood_at are
    per-particle: rank g owns particles [g*P_local, (g+1)*P_local), no
    communication (the reference fans these out with traces.map,
    trace_set.py:287-294,339-340).  Philox streams are addressed by GLOBAL
    particle index, so results do not depend on the GPU count.
  * resample (trace_set.py:202-235) is the one exchange step:
      all_reduce(max) -> fixed-point local CDF vs the global max ->
      all_gather(mass, sum exp) -> every rank finds the children whose
      threshold falls in ITS mass interval (exact integer arithmetic, so all
      ranks agree bit-for-bit with the single-GPU result) -> packed ancestor
      rows travel with all_to_all -> unpack into child slots ->
      lw := logsumexp - log(P_total).
  * subsampled_mh (subsampled_mh.py:76-171): observations (local sections)
    shard, particles are replicated; per minibatch round each rank reduces its
    local (sum dllh, sum dllh^2) per particle, one all_reduce(sum) of 2*P
    doubles, then the identical sequential-test decision on every rank.

The collective plumbing is written against a small `Shard` protocol so the
same code runs over the CUDA library (CudaShard) and over a CPU stand-in in
the gloo tests.
"""
import math

import numpy as np

MASK32 = 0xFFFFFFFF


# ---- Philox4x32-10 on python ints (host side needs the ONE shared uniform) ----
def philox4x32_10(c, k):
    c0, c1, c2, c3 = c
    k0, k1 = k
    for _ in range(10):
        p0 = 0xD2511F53 * c0
        p1 = 0xCD9E8D57 * c2
        n0 = ((p1 >> 32) ^ c1 ^ k0) & MASK32
        n2 = ((p0 >> 32) ^ c3 ^ k1) & MASK32
        c0, c1, c2, c3 = n0, p1 & MASK32, n2, p0 & MASK32
        k0 = (k0 + 0x9E3779B9) & MASK32
        k1 = (k1 + 0xBB67AE85) & MASK32
    return c0, c1, c2, c3


def global_uniform(seed, step):
    """draw 0 of the stream (seed, GLOBAL_PARTICLE, step) -- csrc/philox.cuh"""
    b = philox4x32_10((MASK32, MASK32, step & MASK32, 0), (seed & MASK32, (seed >> 32) & MASK32))
    x = ((b[0] << 32) | b[1]) >> 11
    return min((x + 0.5) / 9007199254740992.0, 0.99999999999999988898)    # strictly below 1, as philox.cuh


def frac_bits_for(P_total):
    lg = 0
    while (1 << lg) < P_total:
        lg += 1
    return min(62 - lg, 52)


def systematic_child_range(offset, mass, total, n_new, u):
    """[k_lo, k_hi): children whose threshold floor(((k<<32)+U0)*total / (n_new<<32))
    lies in [offset, offset+mass) -- exact integer arithmetic, mirrors
    ancestor_search_kernel (csrc/k_weights.cu)."""
    U0 = min(int(u * 4294967296.0), MASK32)
    den = n_new << 32

    def first_k_with_threshold_at_least(v):
        need = -(-(v * den) // total)          # ceil(v*den/total)
        k = -(-(need - U0) // (1 << 32))       # ceil((need-U0)/2^32)
        return min(max(k, 0), n_new)

    return first_k_with_threshold_at_least(offset), first_k_with_threshold_at_least(offset + mass)


def lse_from_parts(global_max, sumexps, any_nan=False, global_min=None):
    """utils.logsumexp (lite/utils.py:70-91) from per-shard partial sums: NaN if any weight is NaN or if
    +inf and -inf are both present; the max itself when it is infinite."""
    if any_nan:
        return float("nan")
    if math.isinf(global_max):
        if global_min is not None and global_min == -global_max:
            return float("nan")
        return global_max
    return global_max + math.log(float(np.sum(sumexps)))


class CudaShard(object):
    """Shard protocol over the C ABI; tensors live on this rank's GPU."""

    def __init__(self, trace_set):
        import torch
        self.ts = trace_set
        self.torch = torch
        self.device = torch.device("cuda", trace_set.device)
        self.P = trace_set.n_particles

    # scalars
    def weight_max(self):
        import ctypes
        from . import _lib
        m = ctypes.c_double()
        _lib.call("vk_shard_weight_max", self.ts._h, ctypes.byref(m))
        return m.value

    def weight_stats(self):
        """(max ignoring NaN, min ignoring NaN, any NaN) of this shard's log weights"""
        import ctypes
        from . import _lib
        v = (ctypes.c_double * 3)()
        _lib.call("vk_shard_weight_stats", self.ts._h, v)
        return v[0], v[1], bool(v[2])

    def build_cdf(self, global_max, frac_bits):
        import ctypes
        from . import _lib
        mass, se = ctypes.c_uint64(), ctypes.c_double()
        _lib.call("vk_shard_build_cdf", self.ts._h, float(global_max), int(frac_bits), ctypes.byref(mass),
                  ctypes.byref(se))
        return mass.value, se.value

    def ancestors(self, offset, total, n_new, k_begin, k_count, mode, step):
        from . import _lib
        out = self.torch.empty(max(k_count, 1), dtype=self.torch.int32, device=self.device)
        _lib.call("vk_shard_ancestors", self.ts._h, offset, total, n_new, k_begin, k_count, mode, step, None,
                  out.data_ptr())
        return out[:k_count]

    def row_bytes(self):
        import ctypes
        from . import _lib
        n = ctypes.c_int64()
        _lib.call("vk_state_row_bytes", self.ts._h, ctypes.byref(n))
        return n.value

    def _order(self):
        # tensors handed to the library may have been produced by torch / NCCL work that is still
        # in flight on torch's current stream; the library runs on the handle's stream
        self.torch.cuda.current_stream().synchronize()

    def pack(self, particles):
        from . import _lib
        self._order()
        n = particles.numel()
        out = self.torch.empty((n, self.row_bytes()), dtype=self.torch.uint8, device=self.device)
        if n:
            _lib.call("vk_pack_state", self.ts._h, particles.contiguous().data_ptr(), n, out.data_ptr())
        return out

    def unpack(self, slots, packed):
        from . import _lib
        self._order()
        n = slots.numel()
        if n:
            _lib.call("vk_unpack_state", self.ts._h, slots.contiguous().data_ptr(), n, packed.contiguous().data_ptr())

    def fill_log_weights(self, v):
        from . import _lib
        _lib.call("vk_fill_log_weights", self.ts._h, float(v))

    def incorporate(self):
        return self.ts.incorporate(return_increments=False)

    def tensor(self, values, dtype):
        return self.torch.tensor(values, dtype=dtype, device=self.device)


def _all_to_all_rows(torch, dist, rows, send_counts, group=None, recv_counts=None):
    """rows: (n, width) tensor whose first axis is grouped by destination rank with
    send_counts[r] rows for rank r.  Returns the rows received, grouped by source.
    recv_counts, when the caller can derive them (systematic resampling), saves the
    count exchange and its host round trip."""
    world = dist.get_world_size(group)
    if dist.get_backend(group) == "nccl":
        if recv_counts is None:
            sc = torch.tensor(send_counts, dtype=torch.int64, device=rows.device)
            rc = torch.empty(world, dtype=torch.int64, device=rows.device)
            dist.all_to_all_single(rc, sc, group=group)
            recv_counts = rc.tolist()
        out = torch.empty((sum(recv_counts),) + tuple(rows.shape[1:]), dtype=rows.dtype, device=rows.device)
        dist.all_to_all_single(out, rows.contiguous(), output_split_sizes=list(recv_counts),
                               input_split_sizes=list(send_counts), group=group)
        return out, list(recv_counts)
    # gloo (CPU tests): gather everything, keep what is addressed to this rank
    rank = dist.get_rank(group)
    gathered = [None] * world
    dist.all_gather_object(gathered, (list(send_counts), rows.cpu().numpy()), group=group)
    parts, got = [], []
    for src in range(world):
        counts, data = gathered[src]
        lo = sum(counts[:rank])
        parts.append(data[lo:lo + counts[rank]])
        got.append(counts[rank])
    if recv_counts is not None:
        assert list(recv_counts) == got, "derived receive counts disagree with the senders"
    out = torch.from_numpy(np.concatenate(parts, axis=0)) if parts else rows[:0]
    return out.to(rows.device), got


def _overlap(lo, hi, a, b):
    return max(0, min(hi, b) - max(lo, a))


def sharded_resample(shard, seed, step, mode="systematic", group=None):
    """Global resample of P_total = world * P_local particles.  Returns the global
    logsumexp.  `step` must be the same on every rank (the handles advance their
    RNG step counters in lock step)."""
    import torch
    import torch.distributed as dist
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    P_local = shard.P
    P_total = world * P_local
    sys_mode = mode == "systematic"
    # 1. global max; min and NaN flag travel with it for the edge cases of utils.logsumexp
    gmax, gmin, any_nan = _global_weight_stats(shard, torch, dist, group)
    # 2. fixed-point masses relative to the global max
    mass, sumexp = shard.build_cdf(gmax, frac_bits_for(P_total))
    mine = shard.tensor([float(mass >> 32), float(mass & MASK32), sumexp], torch.float64)   # exact: 2 x 32-bit halves
    allv = torch.empty(world * 3, dtype=torch.float64, device=mine.device)
    dist.all_gather_into_tensor(allv, mine, group=group)
    allh = allv.cpu().numpy().reshape(world, 3)                  # one host round trip for all scalars
    masses = [(int(allh[r, 0]) << 32) | int(allh[r, 1]) for r in range(world)]
    lse = lse_from_parts(gmax, [float(allh[r, 2]) for r in range(world)], any_nan, gmin)
    offsets = [sum(masses[:r]) for r in range(world)]
    offset, total = offsets[rank], sum(masses)
    if sys_mode:
        # 3. every rank derives EVERY rank's contiguous child range from the gathered masses, so the
        #    all-to-all split sizes need no exchange; children arrive ordered by k, i.e. in slot order
        u = global_uniform(seed, step)
        ranges = [systematic_child_range(offsets[r], masses[r], total, P_total, u) for r in range(world)]
        k_lo, k_hi = ranges[rank]
        anc = shard.ancestors(offset, total, P_total, k_lo, k_hi - k_lo, 1, step)
        send_counts = [_overlap(k_lo, k_hi, d * P_local, (d + 1) * P_local) for d in range(world)]
        recv_counts = [_overlap(ranges[r][0], ranges[r][1], rank * P_local, (rank + 1) * P_local)
                       for r in range(world)]
        assert sum(recv_counts) == P_local, "every child slot must be filled exactly once"
        packed = shard.pack(anc.to(torch.int32))
        recv_rows, _ = _all_to_all_rows(torch, dist, packed, send_counts, group, recv_counts)
        slots = torch.arange(P_local, dtype=torch.int32, device=recv_rows.device)
        shard.unpack(slots, recv_rows)
    else:
        anc_all = shard.ancestors(offset, total, P_total, 0, P_total, 0, step)
        ks = torch.nonzero(anc_all >= 0).flatten()
        anc = anc_all[ks]
        # 4. route: child k lives on rank k // P_local, slot k % P_local (ks ascending => grouped by destination)
        dest = torch.div(ks, P_local, rounding_mode="floor")
        send_counts = torch.bincount(dest, minlength=world).tolist() if ks.numel() else [0] * world
        packed = shard.pack(anc.to(torch.int32))
        slots = (ks - dest * P_local).to(torch.int32).reshape(-1, 1)
        recv_rows, rc = _all_to_all_rows(torch, dist, packed, send_counts, group)
        recv_slots, _ = _all_to_all_rows(torch, dist, slots, send_counts, group, rc)
        assert recv_slots.shape[0] == P_local, "every child slot must be filled exactly once"
        shard.unpack(recv_slots.flatten().contiguous(), recv_rows)
    # 5. even weights (utils.log_domain_even_out), then incorporate (trace_set.py:206-207)
    shard.fill_log_weights(lse - math.log(P_total))
    shard.incorporate()
    return lse


def _global_weight_stats(shard, torch, dist, group):
    """(max, min, any NaN) over all ranks; one all_reduce(MAX) of (max, -min, nan flag)"""
    mx, mn, nan = shard.weight_stats()
    v = shard.tensor([mx, -mn, 1.0 if nan else 0.0], torch.float64)
    dist.all_reduce(v, op=dist.ReduceOp.MAX, group=group)
    v = v.tolist()
    return float(v[0]), -float(v[1]), bool(v[2] > 0.0)


class ShardedTraceSet(object):
    """The particle set of one job, sharded over the ranks of a process group.
    Wraps this rank's B200TraceSet; per-particle operations are forwarded
    unchanged, resample and subsampled_mh add the collectives."""

    def __init__(self, model, particles_per_rank, seed=1, device=None, group=None, replicate_particles=False):
        """replicate_particles=True: every rank holds the SAME particles (global indices
        0..P-1) -- the layout of observation-sharded subsampled_mh."""
        import torch
        import torch.distributed as dist
        from .trace_set import B200TraceSet
        self.group = group
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)
        if device is None:
            device = torch.cuda.current_device()
        self.seed = int(seed)
        self.replicated = bool(replicate_particles)
        self.local = B200TraceSet(model, particles_per_rank, seed=seed, device=device,
                                  particle_base=0 if self.replicated else self.rank * particles_per_rank)
        # one dedicated stream for torch ops, NCCL collectives (they order against torch's current
        # stream) and the library's kernels of this particle set
        self.stream = torch.cuda.Stream(device=device)
        self.local.use_stream(self.stream.cuda_stream)
        self.shard = CudaShard(self.local)
        self.n_particles = particles_per_rank if self.replicated else self.world * particles_per_rank

    def __getattr__(self, name):       # per-particle operations: no communication
        return getattr(self.local, name)

    def enable_peer_memory(self):
        """Switch `resample` to the NVLink peer-memory path (csrc/p2p_resample.cuh): CUDA IPC handles of the
        mailbox, the ancestor array and the state buffers are exchanged ONCE here (host side, any backend of
        the process group); afterwards a resample is a stream-ordered kernel sequence with no collective and
        no host synchronisation.  Ranks must be the GPUs of one node with peer access."""
        import ctypes
        import torch.distributed as dist
        from . import _lib
        buf = (ctypes.c_ubyte * (64 * 32))()
        n = ctypes.c_int32()
        _lib.call("vk_p2p_export", self.local._h, buf, 32, ctypes.byref(n))
        mine = bytes(buf[:64 * n.value])
        gathered = [None] * self.world
        dist.all_gather_object(gathered, mine, group=self.group)
        if any(len(g) != len(mine) for g in gathered):
            from .exception import VentureException
            raise VentureException("evaluation", "peer-memory resample: ranks hold different models")
        blob = b"".join(gathered)
        arr = (ctypes.c_ubyte * len(blob)).from_buffer_copy(blob)
        _lib.call("vk_p2p_open", self.local._h, self.world, self.rank, arr, n.value)
        dist.barrier(group=self.group)
        self.peer_memory = True

    def peer_status(self):
        """(timed_out, global logsumexp before the last resample); synchronises this rank's stream."""
        import ctypes
        from . import _lib
        e, lse = ctypes.c_int32(), ctypes.c_double()
        _lib.call("vk_p2p_status", self.local._h, ctypes.byref(e), ctypes.byref(lse))
        return bool(e.value), lse.value

    def resample(self, P=None, mode="multinomial"):
        from .exception import VentureException
        if P is not None and int(P) != self.n_particles:
            raise VentureException("invalid_argument", "resample: a sharded particle set keeps its particle count")
        if self.replicated:
            raise VentureException("invalid_argument", "resample: this set's particles are replicated, not sharded")
        if getattr(self, "peer_memory", False):
            from . import _lib
            _lib.call("vk_p2p_resample", self.local._h,
                      _lib.RESAMPLE_SYSTEMATIC if mode == "systematic" else _lib.RESAMPLE_MULTINOMIAL)
            return None          # stream-ordered; peer_status() returns the LSE when it is wanted
        import torch
        step = self.local.rng_step
        self.local.rng_step = step + 1
        with torch.cuda.stream(self.stream):
            return sharded_resample(self.shard, self.seed, step,
                                    "systematic" if mode == "systematic" else "multinomial", self.group)

    def logsumexp(self):
        import torch
        import torch.distributed as dist
        gmax, gmin, any_nan = _global_weight_stats(self.shard, torch, dist, self.group)
        _, se = self.shard.build_cdf(gmax, frac_bits_for(self.n_particles))
        s = self.shard.tensor([se], torch.float64)
        dist.all_reduce(s, op=dist.ReduceOp.SUM, group=self.group)
        return lse_from_parts(gmax, [float(s.item())], any_nan, gmin)

    # -- observation-sharded subsampled MH (particles replicated on every rank) --
    def observe_shard(self, X_local, y_local, global_offset, n_total):
        from . import _lib
        self.local.observe(X_local, y_local)
        _lib.call("vk_blr_set_observation_shard", self.local._h, int(global_offset), int(n_total))
        self._n_total = int(n_total)

    def subsampled_mh(self, scope, block, Nbatch, k0, epsilon, useDeltaKernels, deltaKernelArgs, updateValues,
                      transitions):
        import ctypes
        import torch
        import torch.distributed as dist
        from . import _lib
        if getattr(self, "peer_memory", False) and self.replicated:
            # the partial sums travel through peer mailboxes INSIDE the library's persistent kernel
            # (csrc/blr_subsampled_persistent.cuh): no collective, no host round trip
            return self.local.subsampled_mh(scope, block, Nbatch, k0, epsilon, useDeltaKernels, deltaKernelArgs,
                                            updateValues, transitions)
        h = self.local._h
        s, b = self.local._scope_block(scope, block)
        P = self.local.n_particles
        Nt, Nbatch = self._n_total, int(Nbatch)
        delta_mode = 2 if useDeltaKernels == 2 else 0
        _lib.call("vk_ssmh_begin", h)
        n_steps = int(transitions) * (self.local.numBlocksInScope() if b == _lib.BLOCK_EACH else 1)
        for t in range(n_steps):
            _lib.call("vk_ssmh_propose", h, b, t, delta_mode, float(deltaKernelArgs or 0.0))
            for q0 in range(0, Nt, Nbatch):
                q1 = min(q0 + Nbatch, Nt)
                p = ctypes.c_void_p()
                _lib.call("vk_ssmh_round", h, q0, q1, ctypes.byref(p))
                sums = _wrap_device(torch, p.value, P * 2, self.local.device)
                with torch.cuda.stream(self.stream):      # the library runs on this stream too (use_stream)
                    dist.all_reduce(sums, op=dist.ReduceOp.SUM, group=self.group)
                self.stream.synchronize()
                n_active = ctypes.c_int32()
                _lib.call("vk_ssmh_decide", h, ctypes.c_void_p(p.value), q1, int(k0), float(epsilon),
                          ctypes.byref(n_active))
                if n_active.value == 0:
                    break
            _lib.call("vk_ssmh_commit", h)
        stats = np.empty((P, 3))
        _lib.call("vk_ssmh_stats", h, n_steps, _lib.ptr(stats))
        return stats


class _DevArray(object):
    def __init__(self, ptr, n):
        self.__cuda_array_interface__ = {"shape": (n,), "typestr": "<f8", "data": (ptr, False), "version": 2}


def _wrap_device(torch, ptr, n_doubles, device):
    """zero-copy float64 view of library-owned device memory"""
    return torch.as_tensor(_DevArray(ptr, n_doubles), device=torch.device("cuda", device))
