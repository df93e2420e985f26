"""One `PolynomialBatch::from_values` commit sharded over the GPUs of a box, one process per GPU.

plonky2 has no multi-device path (fri/oracle.rs is single-process, rayon over columns); this is the
partitioning BASELINE.json's north star lays out, on top of the `bsx_shard_*` entry points of
include/bsx_b200.h:

  columns -> ranks   iNTT + coset LDE of this rank's columns, no communication;
  leaf rows -> ranks  rank g owns leaves [g * chunk, (g + 1) * chunk) of every column.  Because the
                     device transforms leave every column in Merkle-leaf order, that is one contiguous
                     range per column;
  exchange           mode "p2p": the last NTT pass stores each LDE word into the owner's row slab over
                     NVLink (slab addresses exchanged once as CUDA IPC handles), then one barrier;
                     mode "nccl": the pass fills a destination-major staging buffer and the host runs one
                     all-to-all over `torch.distributed`;
  trees              every rank hashes its rows and builds its 2^cap_height / G cap subtrees; the caps
                     are all-gathered (512 bytes at cap_height 4).

`torch.distributed` is plumbing only: rendezvous, the handle/cap all-gathers, the barrier and (mode
"nccl") the all-to-all.  All arithmetic is in libbsx_b200.so.
"""
import ctypes as C

import numpy as np

from .binding import BSX_DEVICE, BSX_HOST, BsxError, MerkleTree, Timing, _np, _ptr, lib

IPC_BYTES = 64


def _log2_strict(n):
    if n <= 0 or n & (n - 1):
        raise BsxError("Not a power of two: %d" % n)
    return n.bit_length() - 1


class ShardPlan:
    """Who owns what.  Pure host arithmetic (covered by the CPU tests); must agree with api_shard.cu."""

    def __init__(self, log_n, n_cols, rate_bits, cap_height, world):
        self.log_n, self.n_cols, self.rate_bits, self.cap_height, self.world = log_n, n_cols, rate_bits, cap_height, world
        self.log_world = _log2_strict(world)
        self.log_nb = log_n + rate_bits
        if self.log_world > self.log_nb:
            raise BsxError("more ranks than leaves")
        if not 0 <= cap_height <= self.log_nb:
            raise BsxError("cap_height=%d should be at most log2(leaves.len())=%d" % (cap_height, self.log_nb))
        self.chunk = 1 << (self.log_nb - self.log_world)  # leaf rows per rank

    def column_range(self, rank):
        """Balanced contiguous column ranges: the first (C mod G) ranks take one extra column."""
        q, r = divmod(self.n_cols, self.world)
        c0 = rank * q + min(rank, r)
        return c0, c0 + q + (1 if rank < r else 0)

    def owner(self, leaf):
        return leaf >> (self.log_nb - self.log_world)

    def local_leaf(self, leaf):
        return leaf & (self.chunk - 1)

    @property
    def local_cap_height(self):
        return max(self.cap_height - self.log_world, 0)

    @property
    def reduce_levels(self):
        """two_to_one levels left above the per-rank roots when there are more ranks than cap entries."""
        return max(self.log_world - self.cap_height, 0)

    def staging_block_words(self, rank):
        """Words of each of the G destination blocks in rank's staging buffer (all equal)."""
        c0, c1 = self.column_range(rank)
        return (c1 - c0) * self.chunk

    def slab_offset_words(self, src_rank):
        """Where src_rank's block lands in a destination's row slab."""
        return self.column_range(src_rank)[0] * self.chunk

    def staging_index(self, rank, col, leaf):
        """Word index in rank's staging buffer of LDE word (col, leaf), col in rank's column range.
        The contract between the transform kernel's store and the all-to-all."""
        c0, c1 = self.column_range(rank)
        assert c0 <= col < c1
        return self.owner(leaf) * (c1 - c0) * self.chunk + (col - c0) * self.chunk + self.local_leaf(leaf)


class _DevArray:
    """A raw device pointer seen as a 1-D int64 CUDA array (for torch.as_tensor; zero copy)."""

    def __init__(self, ptr, words):
        self.__cuda_array_interface__ = {"shape": (int(words),), "typestr": "<i8", "data": (int(ptr), False),
                                         "version": 2}


def device_tensor(ptr, words, device):
    import torch
    if words == 0 or not ptr:
        return torch.empty(0, dtype=torch.int64, device="cuda:%d" % device)
    return torch.as_tensor(_DevArray(ptr, words), device="cuda:%d" % device)


def exchange_staged(send, recv, plan, rank, group=None):
    """The all-to-all of mode "nccl" on any pair of 1-D int64 torch tensors (CUDA for NCCL, CPU for the
    gloo tests): block d of `send` -> rank d, landing at slab_offset_words(rank) in its `recv`."""
    import torch.distributed as dist
    n_in = [plan.staging_block_words(rank)] * plan.world
    n_out = [plan.staging_block_words(s) for s in range(plan.world)]
    assert send.numel() == sum(n_in) and recv.numel() == plan.n_cols * plan.chunk
    if dist.get_backend(group) == "nccl":
        dist.all_to_all_single(recv, send, output_split_sizes=n_out, input_split_sizes=n_in, group=group)
        return
    # gloo: paired point-to-point transfers, same layout
    ops, off_in = [], 0
    for peer in range(plan.world):
        blk = send[off_in:off_in + n_in[peer]]
        off_in += n_in[peer]
        dst = recv[plan.slab_offset_words(peer):plan.slab_offset_words(peer) + n_out[peer]]
        if peer == rank:
            dst.copy_(blk)
        else:
            if n_in[peer]:
                ops.append(dist.P2POp(dist.isend, blk.contiguous(), peer, group=group))
            if n_out[peer]:
                ops.append(dist.P2POp(dist.irecv, dst, peer, group=group))
    if ops:
        for w in dist.batch_isend_irecv(ops):
            w.wait()


def assemble_cap(local_caps, plan, ctx=None):
    """Cap of the whole batch from the per-rank caps in rank order (merkle_tree.rs: the cap is the layer of
    2^cap_height nodes, left to right).  More ranks than cap entries: hash the rank roots up."""
    cap = np.concatenate([np.asarray(c, dtype=np.uint64).reshape(-1, 4) for c in local_caps], axis=0)
    if plan.reduce_levels:
        if ctx is None:
            raise BsxError("cap_height < log2(world) needs a Context to hash the rank roots")
        out = np.zeros((cap.shape[0] >> plan.reduce_levels, 4), dtype=np.uint64)
        ctx.check(lib().bsx_merkle_reduce(ctx._h, _ptr(_np(cap)), cap.shape[0], plan.reduce_levels, _ptr(out)))
        cap = out
    assert cap.shape[0] == 1 << plan.cap_height
    return cap


def top_siblings(rank_caps, plan, rank, ctx):
    """Merkle siblings above `rank`'s local root when there are more ranks than cap entries (bottom-up,
    `plan.reduce_levels` digests): what the owner appends to its local proof to make it the whole-tree proof."""
    level = np.concatenate([np.asarray(c, dtype=np.uint64).reshape(-1, 4) for c in rank_caps], axis=0)
    sib, idx = [], rank
    for _ in range(plan.reduce_levels):
        sib.append(level[idx ^ 1].copy())
        out = np.zeros((level.shape[0] >> 1, 4), dtype=np.uint64)
        ctx.check(lib().bsx_merkle_reduce(ctx._h, _ptr(_np(level)), level.shape[0], 1, _ptr(out)))
        level, idx = out, idx >> 1
    return np.array(sib, dtype=np.uint64).reshape(-1, 4)


class ShardedPolynomialBatch:
    """This rank's part of one committed batch: its coefficient columns, its leaf rows and cap subtrees,
    and the cap of the whole batch.  `merkle_tree.get / prove` answer for leaves this rank owns
    (`plan.owner(leaf) == rank`, local index `plan.local_leaf(leaf)`); with cap_height >= log2(world) the
    proof is word for word the single-GPU proof."""

    def __init__(self, ctx, plan, rank, handle):
        self.ctx, self.plan, self.rank, self._h = ctx, plan, rank, handle
        self.merkle_tree = None
        self.cap = None
        self.timing = None
        self.group, self.mode = None, "p2p"
        self.rank_caps = None  # every rank's local cap (set by `run`; needed for proofs when world > 2^cap_height)

    # ---- construction pieces (also used one by one by the single-process tests)
    @classmethod
    def create(cls, ctx, plan, rank):
        h = C.c_void_p()
        ctx.check(lib().bsx_shard_create(ctx._h, plan.log_n, plan.n_cols, plan.rate_bits, plan.cap_height, rank,
                                         plan.world, C.byref(h)))
        return cls(ctx, plan, rank, h)

    @property
    def slab_ptr(self):
        return lib().bsx_shard_slab_device(self._h) or 0

    @property
    def staging_ptr(self):
        return lib().bsx_shard_staging_device(self._h) or 0

    def ipc_handle(self):
        buf = (C.c_uint8 * IPC_BYTES)()
        self.ctx.check(lib().bsx_shard_slab_ipc(self._h, buf))
        return bytes(buf)

    def set_peer_ipc(self, peer, handle):
        buf = (C.c_uint8 * IPC_BYTES).from_buffer_copy(handle)
        self.ctx.check(lib().bsx_shard_set_peer_ipc(self._h, peer, buf))

    def set_peer_ptr(self, peer, ptr):
        self.ctx.check(lib().bsx_shard_set_peer_ptr(self._h, peer, C.c_void_p(int(ptr))))

    def transform(self, values_local, is_coeffs=False, use_staging=False, device_ptr=None, timing=False):
        c0, c1 = self.plan.column_range(self.rank)
        if device_ptr is not None:
            src, mem = C.c_void_p(int(device_ptr)), BSX_DEVICE
        else:
            values_local = _np(values_local)
            if values_local.shape != (c1 - c0, 1 << self.plan.log_n):
                raise BsxError("rank %d expects columns [%d, %d) as a (%d, %d) array" %
                               (self.rank, c0, c1, c1 - c0, 1 << self.plan.log_n))
            src, mem = _ptr(values_local), BSX_HOST
        t = Timing() if timing else None
        self.ctx.check(lib().bsx_shard_transform(self._h, src, mem, 1 if is_coeffs else 0, 1 if use_staging else 0,
                                                 C.byref(t) if t is not None else None))
        self.timing = t

    def commit(self):
        h = C.c_void_p()
        self.ctx.check(lib().bsx_shard_commit(self._h, C.byref(h)))
        self.merkle_tree = MerkleTree(self.ctx, _handle=h)
        return self.merkle_tree.cap

    @property
    def polynomials(self):
        c0, c1 = self.plan.column_range(self.rank)
        out = np.zeros((c1 - c0, 1 << self.plan.log_n), dtype=np.uint64)
        self.ctx.check(lib().bsx_shard_coeffs(self._h, _ptr(out), BSX_HOST))
        return out

    # ---- the whole thing, one process per GPU
    @classmethod
    def setup(cls, ctx, log_n, n_cols, rate_bits, cap_height, group=None, mode="p2p"):
        """Collective: allocate this rank's row slab and (mode "p2p") open every peer's slab through CUDA
        IPC.  The returned object can commit any number of batches of this shape with `run`."""
        import torch
        import torch.distributed as dist
        if mode not in ("p2p", "p2p_host", "nccl"):
            raise BsxError("mode must be 'p2p' (device-side barrier), 'p2p_host' (host barrier) or 'nccl'")
        rank, world = dist.get_rank(group), dist.get_world_size(group)
        plan = ShardPlan(log_n, n_cols, rate_bits, cap_height, world)
        self = cls.create(ctx, plan, rank)
        self.group, self.mode = group, mode
        if mode in ("p2p", "p2p_host"):
            dev = "cuda:%d" % ctx.device
            mine = torch.tensor(list(self.ipc_handle()), dtype=torch.uint8, device=dev)
            allh = torch.empty(world * IPC_BYTES, dtype=torch.uint8, device=dev)
            dist.all_gather_into_tensor(allh, mine, group=group)
            allh = allh.cpu().numpy().tobytes()
            for peer in range(world):
                if peer != rank:
                    self.set_peer_ipc(peer, allh[peer * IPC_BYTES:(peer + 1) * IPC_BYTES])
        return self

    def run(self, values_local, device_ptr=None, is_coeffs=False, timing=False):
        """Collective: commit one batch.  Every rank passes ITS columns (`plan.column_range(rank)`); on
        return `cap` is the cap of the whole batch and `merkle_tree` this rank's rows and subtrees.  The cap
        all-gather at the end doubles as the barrier that makes the slabs safe to overwrite by the next run."""
        import torch
        import torch.distributed as dist
        plan, ctx, group = self.plan, self.ctx, self.group
        dev = "cuda:%d" % ctx.device
        if self.merkle_tree is not None:
            self.merkle_tree.close()
            self.merkle_tree = None
        if self.mode == "p2p":
            # the whole commit inside the library: peer stores, device-side flag barrier, hashing, caps exchanged through
            # the mailboxes behind the slabs -- no NCCL call and one host synchronisation per commit
            c0, c1 = plan.column_range(self.rank)
            if device_ptr is not None:
                src, mem = C.c_void_p(int(device_ptr)), BSX_DEVICE
            else:
                values_local = _np(values_local)
                if values_local.shape != (c1 - c0, 1 << plan.log_n):
                    raise BsxError("rank %d expects columns [%d, %d)" % (self.rank, c0, c1))
                src, mem = _ptr(values_local), BSX_HOST
            caps = np.zeros((plan.world, 1 << plan.local_cap_height, 4), dtype=np.uint64)
            h = C.c_void_p()
            t = Timing() if timing else None
            f = lib().bsx_shard_run_p2p
            f.restype, f.argtypes = C.c_int32, [C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.POINTER(C.c_void_p),
                                                C.c_void_p, C.POINTER(Timing)]
            ctx.check(f(self._h, src, mem, 1 if is_coeffs else 0, C.byref(h), _ptr(caps), C.byref(t) if t is not None else None))
            self.timing = t
            self.merkle_tree = MerkleTree(ctx, _handle=h)
            self.rank_caps = list(caps)
            self.cap = assemble_cap(self.rank_caps, plan, ctx)
            return self
        if self.mode == "p2p_host":
            self.transform(values_local, is_coeffs, use_staging=False, device_ptr=device_ptr, timing=timing)
            dist.barrier(group=group)  # every rank's stores into every slab are complete
        else:
            self._run_nccl_chunked(values_local, device_ptr, is_coeffs)
        local = torch.from_numpy(self.commit().astype(np.int64)).to(dev).reshape(-1)
        caps = torch.empty(plan.world * local.numel(), dtype=torch.int64, device=dev)
        dist.all_gather_into_tensor(caps, local, group=group)
        caps = caps.cpu().numpy().astype(np.uint64).reshape(plan.world, -1, 4)
        self.rank_caps = list(caps)
        self.cap = assemble_cap(self.rank_caps, plan, ctx)
        return self

    def _run_nccl_chunked(self, values_local, device_ptr, is_coeffs):
        """Mode "nccl": the local columns are transformed in a few chunks into the destination-major staging buffer, and
        the all-to-all of chunk k (grouped NCCL send/recv between views of the staging buffer and of the row slabs) runs
        on NCCL's stream while chunk k + 1 is being transformed on the library's."""
        import torch
        import torch.distributed as dist
        plan, ctx, group = self.plan, self.ctx, self.group
        c0, c1 = plan.column_range(self.rank)
        n_local, n = c1 - c0, 1 << plan.log_n
        if device_ptr is not None:
            base, mem = int(device_ptr), BSX_DEVICE
        else:
            values_local = _np(values_local)
            if values_local.shape != (n_local, n):
                raise BsxError("rank %d expects columns [%d, %d)" % (self.rank, c0, c1))
            base, mem = values_local.ctypes.data, BSX_HOST
        # every rank must post the same number of exchanges: the count comes from the smallest column range
        nl_min = plan.n_cols // plan.world
        n_chunks = 1 if nl_min < 6 else (2 if nl_min < 12 else 3)
        bounds = [(k * n_local) // n_chunks for k in range(n_chunks + 1)]
        f = lib().bsx_shard_transform_cols
        f.restype, f.argtypes = C.c_int32, [C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_int32]
        lib_stream = torch.cuda.ExternalStream(ctx.stream, device=ctx.device)
        works, keep = [], []
        with torch.cuda.stream(lib_stream):
            for k in range(n_chunks):
                cb, cc = bounds[k], bounds[k + 1] - bounds[k]
                ctx.check(f(self._h, C.c_void_p(base + 8 * cb * n), mem, 1 if is_coeffs else 0, 1, cb, cc, 0))
                send = device_tensor(self.staging_ptr, plan.world * plan.staging_block_words(self.rank), ctx.device)
                recv = device_tensor(self.slab_ptr, plan.n_cols * plan.chunk, ctx.device)
                ins, outs = [], []
                for peer in range(plan.world):
                    ins.append(send[peer * n_local * plan.chunk + cb * plan.chunk:][:cc * plan.chunk])
                    p0, p1 = plan.column_range(peer)
                    pb = [(j * (p1 - p0)) // n_chunks for j in range(n_chunks + 1)]
                    outs.append(recv[(p0 + pb[k]) * plan.chunk:][:(pb[k + 1] - pb[k]) * plan.chunk])
                keep.append((ins, outs))
                works.append(dist.all_to_all(outs, ins, group=group, async_op=True))
            for w in works:
                w.wait()
        torch.cuda.synchronize(ctx.device)
        self.timing = Timing()  # phases interleave with the exchange: no per-phase split in this mode

    # ---- openings: served by the owning rank (row l lives on rank l >> log2(chunk))
    @property
    def proof_len(self):
        return self.plan.log_nb - self.plan.cap_height

    def get(self, leaf):
        """`MerkleTree::get(leaf)` for a leaf this rank owns (global index)."""
        if self.plan.owner(leaf) != self.rank:
            raise BsxError("leaf %d lives on rank %d" % (leaf, self.plan.owner(leaf)))
        return self.merkle_tree.get(self.plan.local_leaf(leaf))

    def prove(self, leaf):
        """`MerkleTree::prove(leaf)` against the cap of the WHOLE batch, for a leaf this rank owns."""
        if self.plan.owner(leaf) != self.rank:
            raise BsxError("leaf %d lives on rank %d" % (leaf, self.plan.owner(leaf)))
        sib = np.asarray(self.merkle_tree.prove(self.plan.local_leaf(leaf)), dtype=np.uint64).reshape(-1, 4)
        if self.plan.reduce_levels:
            if self.rank_caps is None:
                raise BsxError("rank caps unknown: commit with run() or set rank_caps")
            sib = np.concatenate([sib, top_siblings(self.rank_caps, self.plan, self.rank, self.ctx)], axis=0)
        return sib

    def local_openings(self, leaves):
        """Rows and proofs of the listed (global) leaves this rank owns, zeros elsewhere: summing the arrays of
        all ranks gives every opening exactly once."""
        leaves = [int(x) for x in leaves]
        width = self.merkle_tree.shape()[1]
        rows = np.zeros((len(leaves), width), dtype=np.uint64)
        proofs = np.zeros((len(leaves), self.proof_len, 4), dtype=np.uint64)
        for k, leaf in enumerate(leaves):
            if self.plan.owner(leaf) == self.rank:
                rows[k] = self.get(leaf)
                proofs[k] = self.prove(leaf)
        return rows, proofs

    def open(self, leaves):
        """Collective: every rank gets the rows and Merkle proofs of `leaves` (the same list on every rank), each
        answered by its owner; one all-reduce of a few KB (the FRI query phase opens 28 rows per tree)."""
        import torch
        import torch.distributed as dist
        rows, proofs = self.local_openings(leaves)
        dev = "cuda:%d" % self.ctx.device
        flat = np.concatenate([rows.reshape(-1), proofs.reshape(-1)]).view(np.int64)
        t = torch.from_numpy(flat.copy()).to(dev)
        dist.all_reduce(t, op=dist.ReduceOp.SUM, group=self.group)  # exactly one non-zero contribution per word
        flat = t.cpu().numpy().view(np.uint64)
        return flat[:rows.size].reshape(rows.shape), flat[rows.size:].reshape(proofs.shape)

    @classmethod
    def from_values(cls, ctx, values_local, log_n, n_cols, rate_bits, cap_height, group=None, mode="p2p",
                    device_ptr=None, is_coeffs=False, timing=False):
        """`PolynomialBatch::from_values` over `group` (default: the world group, NCCL): setup + run."""
        return cls.setup(ctx, log_n, n_cols, rate_bits, cap_height, group, mode).run(values_local, device_ptr,
                                                                                     is_coeffs, timing)

    def close(self):
        if self.merkle_tree is not None:
            self.merkle_tree.close()
            self.merkle_tree = None
        if self._h is not None:
            lib().bsx_shard_free(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


# ---------------------------------------------------------------------------------------------
# prove_openings over row-sharded oracles
def final_poly_part(sbatches, instance, alpha):
    """This rank's part of the final polynomial (2^degree_bits, 2): the polynomials whose coefficient columns it
    holds, at their alpha powers in the whole instance.  `instance` uses GLOBAL column indices."""
    ctx = sbatches[0].ctx
    n = 1 << sbatches[0].plan.log_n
    hs = (C.c_void_p * len(sbatches))(*[b._h for b in sbatches])
    a = _np(alpha)
    out = np.zeros((n, 2), dtype=np.uint64)
    ctx.check(lib().bsx_shard_final_poly_part(hs, len(sbatches), C.byref(instance.struct), _ptr(a), _ptr(out), BSX_HOST))
    return out


def sum_parts(ctx, parts):
    """Field sum of the ranks' parts."""
    parts = _np(np.stack([np.asarray(p, dtype=np.uint64).reshape(-1) for p in parts]))
    out = np.zeros(parts.shape[1], dtype=np.uint64)
    ctx.check(lib().bsx_field_sum(ctx._h, _ptr(parts), BSX_HOST, parts.shape[0], parts.shape[1], _ptr(out), BSX_HOST))
    return out.reshape(-1, 2)


def fri_from_final_poly(ctx, final_poly, challenger, fri_params):
    """Commit phase, proof of work and FRI-tree openings from the final polynomial -> (FRI-only words, indices)."""
    from .fri import _params_struct
    p = _params_struct(fri_params)
    n_words = C.c_size_t()
    ctx.check(lib().bsx_fri_proof_words(None, 0, C.byref(p), C.byref(n_words)))
    proof = np.zeros(n_words.value, dtype=np.uint64)
    idx = np.zeros(fri_params.config.num_query_rounds, dtype=np.uint64)
    fp = _np(final_poly)
    ctx.check(lib().bsx_fri_proof_from_final_poly(ctx._h, None, 0, _ptr(fp), BSX_HOST, C.byref(challenger._s), C.byref(p),
                                                  _ptr(proof), _ptr(idx)))
    return proof, idx


def sharded_prove_openings(instance, sbatches, challenger, fri_params, gather=None, opener=None):
    """`PolynomialBatch::prove_openings` for oracles committed with ShardedPolynomialBatch.  Collective: every rank
    calls it with the same instance (GLOBAL column indices) and an identical challenger; every rank returns the same
    proof, word for word the single-GPU one.  Only 16 bytes x 2^degree_bits per rank (the parts of the final
    polynomial) and the 28 opened rows with their paths cross the ranks; the FRI rounds themselves are not sharded
    (< 2 % of the work, serialised by Fiat-Shamir) and run redundantly on every rank, which keeps the transcripts
    in step without a broadcast.
    gather(part) -> list of all ranks' parts and opener(sbatch, leaves) -> (rows, proofs) default to
    torch.distributed collectives; the single-process tests pass their own."""
    from .serialization import splice_initial_openings
    ctx = sbatches[0].ctx
    alpha = challenger.get_extension_challenge()
    part = final_poly_part(sbatches, instance, alpha)
    if gather is None:
        import torch
        import torch.distributed as dist
        group = sbatches[0].group
        world = dist.get_world_size(group)
        dev = "cuda:%d" % ctx.device
        mine = torch.from_numpy(part.view(np.int64).reshape(-1).copy()).to(dev)
        allp = torch.empty(world * mine.numel(), dtype=torch.int64, device=dev)
        dist.all_gather_into_tensor(allp, mine, group=group)
        parts = list(allp.cpu().numpy().view(np.uint64).reshape(world, -1))
    else:
        parts = gather(part)
    final_poly = sum_parts(ctx, parts)
    fri_words, idx = fri_from_final_poly(ctx, final_poly, challenger, fri_params)
    leaves = [int(x) for x in idx]
    openings = [(opener(sb, leaves) if opener is not None else sb.open(leaves)) for sb in sbatches]
    c = fri_params.config
    return splice_initial_openings(fri_words, openings, fri_params.degree_bits, c.rate_bits, c.cap_height,
                                   list(fri_params.reduction_arity_bits), c.num_query_rounds)


# ---------------------------------------------------------------------------------------------
# one process driving G devices through the C ABI (bsx_comm_create / bsx_sharded_*): what the Rust shim would call
class Comm:
    """`bsx_comm`: one context per rank; `devices` may repeat a device (emulated ranks share it)."""

    def __init__(self, devices):
        L = _declare_comm()
        devs = np.ascontiguousarray(devices, dtype=np.int32)
        h = C.c_void_p()
        rc = L.bsx_comm_create(int(devs.size), C.c_void_p(devs.ctypes.data), C.byref(h))
        if rc != 0:
            raise BsxError("bsx_comm_create(%r) failed with %d" % (list(devices), rc))
        self._h, self.size = h, int(devs.size)

    def check(self, rc):
        if rc != 0:
            raise BsxError("%s (code %d)" % (lib().bsx_comm_last_error(self._h).decode(), rc))

    def close(self):
        if self._h is not None:
            lib().bsx_comm_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class CommShardedBatch:
    """`bsx_sharded_batch`: one PolynomialBatch spread over the ranks of a Comm; indices are global."""

    def __init__(self, comm, handle, shape, timing=None):
        self.comm, self._h, self.shape, self.timing = comm, handle, shape, timing

    @classmethod
    def from_values(cls, comm, values, rate_bits, cap_height, is_coeffs=False, timing=False):
        L = _declare_comm()
        v = _np(values)
        n_cols, n = v.shape
        h = C.c_void_p()
        t = Timing() if timing else None
        comm.check(L.bsx_sharded_from_values(comm._h, _ptr(v), _log2_strict(n), int(n_cols), int(rate_bits),
                                             int(cap_height), 1 if is_coeffs else 0, C.byref(h),
                                             C.byref(t) if t is not None else None))
        return cls(comm, h, (int(n_cols), _log2_strict(n), int(rate_bits), int(cap_height)), t)

    @property
    def cap(self):
        out = np.zeros((1 << self.shape[3], 4), dtype=np.uint64)
        self.comm.check(lib().bsx_sharded_cap(self._h, _ptr(out)))
        return out

    def get(self, leaves):
        idx = _np(np.atleast_1d(leaves))
        out = np.zeros((idx.size, self.shape[0]), dtype=np.uint64)
        self.comm.check(lib().bsx_sharded_get(self._h, _ptr(idx), int(idx.size), _ptr(out)))
        return out[0] if np.isscalar(leaves) else out

    def prove(self, leaf):
        out = np.zeros((lib().bsx_sharded_proof_len(self._h), 4), dtype=np.uint64)
        self.comm.check(lib().bsx_sharded_prove(self._h, C.c_uint64(int(leaf)), _ptr(out)))
        return out

    def close(self):
        if self._h is not None:
            lib().bsx_sharded_free(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def comm_prove_openings(comm, instance, oracles, challenger, fri_params):
    """`PolynomialBatch::prove_openings` over CommShardedBatch oracles (bsx_sharded_prove_openings)."""
    from .fri import _params_struct
    from .proof import fri_proof_layout
    L = _declare_comm()
    p = _params_struct(fri_params)
    words = fri_proof_layout(fri_params, [o.shape[0] for o in oracles]).total_words
    proof = np.zeros(words, dtype=np.uint64)
    hs = (C.c_void_p * len(oracles))(*[o._h for o in oracles])
    comm.check(L.bsx_sharded_prove_openings(comm._h, hs, len(oracles), C.byref(instance.struct), C.byref(challenger._s),
                                            C.byref(p), _ptr(proof)))
    return proof


def _declare_comm():
    L = lib()
    if getattr(L, "_bsx_comm_declared", False):
        return L
    vp, i32, u64 = C.c_void_p, C.c_int32, C.c_uint64
    pp = C.POINTER(C.c_void_p)
    sig = {
        "bsx_comm_create": (i32, [i32, vp, pp]), "bsx_comm_destroy": (None, [vp]), "bsx_comm_size": (i32, [vp]),
        "bsx_comm_ctx": (vp, [vp, i32]), "bsx_comm_last_error": (C.c_char_p, [vp]),
        "bsx_sharded_from_values": (i32, [vp, vp, i32, i32, i32, i32, i32, pp, C.POINTER(Timing)]),
        "bsx_sharded_from_values_parts": (i32, [vp, vp, i32, i32, i32, i32, i32, i32, pp, C.POINTER(Timing)]),
        "bsx_sharded_free": (None, [vp]), "bsx_sharded_cap": (i32, [vp, vp]), "bsx_sharded_get": (i32, [vp, vp, i32, vp]),
        "bsx_sharded_prove": (i32, [vp, u64, vp]), "bsx_sharded_proof_len": (i32, [vp]),
        "bsx_sharded_shard": (vp, [vp, i32]), "bsx_sharded_tree": (vp, [vp, i32]),
        "bsx_sharded_prove_openings": (i32, [vp, vp, i32, vp, vp, vp, vp]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(L, name)
        fn.restype, fn.argtypes = res, args
    L._bsx_comm_declared = True
    return L
