"""ctypes binding of include/bsx_b200.h and the plonky2-shaped host classes on top of it."""
import ctypes as C
import os

import numpy as np

BSX_HOST, BSX_DEVICE = 0, 1
_DIR = os.path.dirname(os.path.abspath(__file__))
# BSX_B200_LIB: another build of the same library (A/B builds of a kernel under tools/); never a different backend
_SO = os.environ.get("BSX_B200_LIB") or os.path.join(_DIR, "csrc", "libbsx_b200.so")

# every symbol include/bsx_b200.h declares (checked by tests/test_abi.py against the header text)
EXPORTED_SYMBOLS = [
    "bsx_version", "bsx_device_count", "bsx_launch_count", "bsx_ctx_create", "bsx_ctx_destroy", "bsx_last_error", "bsx_ctx_stream",
    "bsx_ctx_sync", "bsx_ctx_mem_used", "bsx_ctx_trim", "bsx_batch_from_values", "bsx_batch_from_coeffs",
    "bsx_batch_from_values_cols", "bsx_batch_from_coeffs_cols", "bsx_batch_from_quotient_values", "bsx_partial_products_and_zs", "bsx_batch_free", "bsx_batch_cap",
    "bsx_batch_tree", "bsx_batch_get_lde_values", "bsx_batch_coeffs", "bsx_batch_export", "bsx_batch_shape",
    "bsx_batch_eval_ext", "bsx_batch_import", "bsx_batch_get_lde_values_packed", "bsx_batch_lde_device", "bsx_batch_digests_device", "bsx_merkle_new", "bsx_merkle_new_colmajor_device",
    "bsx_merkle_free", "bsx_merkle_cap", "bsx_merkle_get", "bsx_merkle_prove", "bsx_merkle_digests",
    "bsx_merkle_shape", "bsx_poseidon_permute", "bsx_poseidon_gate_witness", "bsx_debug_salt_stream", "bsx_ntt", "bsx_challenger_init", "bsx_challenger_observe",
    "bsx_challenger_get", "bsx_fri_proof_words", "bsx_fri_prove_openings", "bsx_fri_proof", "bsx_fri_final_poly", "bsx_fri_final_poly_part", "bsx_shard_final_poly_part", "bsx_field_sum",
    "bsx_fri_proof_from_final_poly",
    "bsx_merkle_reduce", "bsx_merkle_reduce_noalloc", "bsx_shard_column_range", "bsx_shard_create", "bsx_shard_free", "bsx_shard_shape",
    "bsx_shard_slab_device", "bsx_shard_staging_device", "bsx_shard_coeffs_device", "bsx_shard_slab_ipc",
    "bsx_shard_set_peer_ipc", "bsx_shard_set_peer_ptr", "bsx_shard_transform", "bsx_shard_commit", "bsx_shard_coeffs", "bsx_shard_run_p2p", "bsx_shard_transform_cols",
    "bsx_fri_proof_layout", "bsx_fri_proof_bytes", "bsx_proof_with_public_inputs_bytes",
    "bsx_quotient_values", "bsx_compute_quotient_polys",
    "bsx_comm_create", "bsx_comm_destroy", "bsx_comm_size", "bsx_comm_ctx", "bsx_comm_last_error",
    "bsx_sharded_from_values", "bsx_sharded_from_values_parts", "bsx_sharded_free", "bsx_sharded_cap", "bsx_sharded_get",
    "bsx_sharded_prove", "bsx_sharded_proof_len", "bsx_sharded_shard", "bsx_sharded_tree", "bsx_sharded_prove_openings",
]


class BsxError(RuntimeError):
    """A non-zero status from the C ABI (upstream plonky2 would panic)."""


class Timing(C.Structure):
    _fields_ = [("h2d_ms", C.c_float), ("ifft_ms", C.c_float), ("lde_ms", C.c_float), ("leaf_hash_ms", C.c_float),
                ("tree_ms", C.c_float), ("total_ms", C.c_float)]

    def as_dict(self):
        return {k: float(getattr(self, k)) for k, _ in self._fields_}


def library_path():
    return _SO


_lib = None


def lib():
    """Load libbsx_b200.so.  Raises if it has not been built: there is no fallback path."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(_SO):
        raise BsxError("libbsx_b200.so is not built (run `python -c 'import __graft_entry__ as g; g.build()'` "
                       "or `make -C 0-kno-blobstreamx_b200/csrc`); there is no CPU fallback")
    L = C.CDLL(_SO)
    vp, i32, u64, sz = C.c_void_p, C.c_int32, C.c_uint64, C.c_size_t
    pp = C.POINTER(C.c_void_p)
    sig = {
        "bsx_version": (C.c_char_p, []),
        "bsx_device_count": (i32, []),
        "bsx_launch_count": (u64, []),
        "bsx_ctx_create": (i32, [i32, pp]),
        "bsx_ctx_destroy": (None, [vp]),
        "bsx_last_error": (C.c_char_p, [vp]),
        "bsx_ctx_stream": (vp, [vp]),
        "bsx_ctx_sync": (i32, [vp]),
        "bsx_ctx_mem_used": (i32, [vp, C.POINTER(C.c_uint64)]),
        "bsx_ctx_trim": (i32, [vp]),
        "bsx_batch_from_values": (i32, [vp, vp, i32, i32, i32, i32, i32, i32, vp, pp, C.POINTER(Timing)]),
        "bsx_batch_from_coeffs": (i32, [vp, vp, i32, i32, i32, i32, i32, i32, vp, pp, C.POINTER(Timing)]),
        "bsx_batch_from_values_cols": (i32, [vp, vp, i32, i32, i32, i32, i32, vp, pp, C.POINTER(Timing)]),
        "bsx_batch_from_coeffs_cols": (i32, [vp, vp, i32, i32, i32, i32, i32, vp, pp, C.POINTER(Timing)]),
        "bsx_batch_from_quotient_values": (i32, [vp, vp, i32, i32, i32, i32, i32, i32, i32, i32, vp, pp,
                                                 C.POINTER(Timing)]),
        "bsx_partial_products_and_zs": (i32, [vp, vp, vp, i32, i32, i32, i32, vp, vp, vp, i32, vp, i32]),
        "bsx_batch_free": (None, [vp]),
        "bsx_batch_cap": (i32, [vp, vp]),
        "bsx_batch_tree": (vp, [vp]),
        "bsx_batch_get_lde_values": (i32, [vp, u64, u64, vp]),
        "bsx_batch_get_lde_values_packed": (i32, [vp, u64, u64, u64, vp, i32]),
        "bsx_batch_coeffs": (i32, [vp, vp, i32]),
        "bsx_batch_export": (i32, [vp, vp, vp, vp]),
        "bsx_batch_eval_ext": (i32, [vp, i32, i32, vp, i32, vp]),
        "bsx_batch_import": (i32, [vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, pp]),
        "bsx_batch_shape": (i32, [vp] + [C.POINTER(i32)] * 5),
        "bsx_batch_lde_device": (vp, [vp]),
        "bsx_batch_digests_device": (vp, [vp]),
        "bsx_merkle_new": (i32, [vp, vp, i32, i32, i32, i32, pp]),
        "bsx_merkle_new_colmajor_device": (i32, [vp, vp, sz, i32, i32, i32, pp]),
        "bsx_merkle_free": (None, [vp]),
        "bsx_merkle_cap": (i32, [vp, vp]),
        "bsx_merkle_get": (i32, [vp, vp, i32, vp]),
        "bsx_merkle_prove": (i32, [vp, u64, vp]),
        "bsx_merkle_digests": (i32, [vp, vp]),
        "bsx_merkle_shape": (i32, [vp] + [C.POINTER(i32)] * 3),
        "bsx_poseidon_permute": (i32, [vp, vp, i32, sz]),
        "bsx_poseidon_gate_witness": (i32, [vp, vp, i32, sz]),
        "bsx_debug_salt_stream": (i32, [vp, vp, i32, vp, sz]),
        "bsx_ntt": (i32, [vp, vp, i32, i32, i32, i32, u64]),
        "bsx_challenger_init": (None, [vp]),
        "bsx_challenger_observe": (i32, [vp, vp, vp, sz]),
        "bsx_challenger_get": (i32, [vp, vp, vp, i32]),
        "bsx_fri_proof_words": (i32, [vp, i32, vp, C.POINTER(sz)]),
        "bsx_fri_prove_openings": (i32, [vp, vp, i32, vp, vp, vp, vp]),
        "bsx_fri_proof": (i32, [vp, vp, i32, vp, vp, vp, vp]),
        "bsx_fri_final_poly": (i32, [vp, vp, i32, vp, vp, vp]),
        "bsx_fri_final_poly_part": (i32, [vp, vp, i32, vp, vp, i32, vp, i32]),
        "bsx_shard_final_poly_part": (i32, [vp, i32, vp, vp, vp, i32]),
        "bsx_field_sum": (i32, [vp, vp, i32, i32, sz, vp, i32]),
        "bsx_fri_proof_from_final_poly": (i32, [vp, vp, i32, vp, i32, vp, vp, vp, vp]),
        "bsx_merkle_reduce": (i32, [vp, vp, i32, i32, vp]),
        "bsx_merkle_reduce_noalloc": (i32, [vp, vp, i32, i32, i32, vp, i32]),
        "bsx_shard_column_range": (i32, [i32, i32, i32, C.POINTER(i32), C.POINTER(i32)]),
        "bsx_shard_create": (i32, [vp, i32, i32, i32, i32, i32, i32, pp]),
        "bsx_shard_free": (None, [vp]),
        "bsx_shard_shape": (i32, [vp, C.POINTER(i32), C.POINTER(i32), C.POINTER(u64), C.POINTER(i32)]),
        "bsx_shard_slab_device": (vp, [vp]),
        "bsx_shard_staging_device": (vp, [vp]),
        "bsx_shard_coeffs_device": (vp, [vp]),
        "bsx_shard_slab_ipc": (i32, [vp, vp]),
        "bsx_shard_set_peer_ipc": (i32, [vp, i32, vp]),
        "bsx_shard_set_peer_ptr": (i32, [vp, i32, vp]),
        "bsx_shard_transform": (i32, [vp, vp, i32, i32, i32, C.POINTER(Timing)]),
        "bsx_shard_commit": (i32, [vp, pp]),
        "bsx_shard_coeffs": (i32, [vp, vp, i32]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(L, name)
        fn.restype = res
        fn.argtypes = args
    _lib = L
    return L


def _np(a):
    return np.ascontiguousarray(a, dtype=np.uint64)


def _ptr(a):
    """Host numpy array or raw integer address -> c_void_p."""
    if a is None:
        return None
    if isinstance(a, (int, np.integer)):
        return C.c_void_p(int(a))
    return C.c_void_p(a.ctypes.data)


def _log2_strict(n):
    """plonky2_util::log2_strict: panics unless n is a power of two."""
    if n <= 0 or n & (n - 1):
        raise BsxError("Not a power of two: %d" % n)
    return n.bit_length() - 1


class Context:
    """One per (host thread, device): owns a CUDA stream, twiddle tables and the memory pool."""

    def __init__(self, device=0):
        self._h = None
        L = lib()
        h = C.c_void_p()
        rc = L.bsx_ctx_create(int(device), C.byref(h))
        if rc != 0:
            raise BsxError("bsx_ctx_create(device=%d) failed with %d: no usable CUDA device "
                           "(this library has no CPU fallback)" % (device, rc))
        self._h = h
        self.device = int(device)

    def check(self, rc):
        if rc != 0:
            raise BsxError("%s (code %d)" % (lib().bsx_last_error(self._h).decode(), rc))

    @property
    def stream(self):
        return lib().bsx_ctx_stream(self._h)

    def mem_used(self):
        """Bytes of device memory the library currently holds allocations for (pool `used`, not `reserved`)."""
        v = C.c_uint64()
        self.check(lib().bsx_ctx_mem_used(self._h, C.byref(v)))
        return v.value

    def sync(self):
        self.check(lib().bsx_ctx_sync(self._h))

    def trim(self):
        """Give the large buffers the context keeps for reuse back to the device."""
        self.check(lib().bsx_ctx_trim(self._h))

    def close(self):
        if self._h is not None:
            lib().bsx_ctx_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def partial_products_and_zs(self, wires, sigmas, k_is, betas, gammas, max_degree, out_device_ptr=None,
                                device_ptrs=None, shape=None):
        """plonk/prover.rs `all_wires_permutation_partial_products`, columns ordered as the Zs commitment wants them
        (Zs first, then each challenge's partial products).  wires, sigmas: (n_routed, n) values (or `device_ptrs` =
        (wires_ptr, sigmas_ptr) with `shape`); returns the (n_challenges * ceil(n_routed / max_degree), n) array, or
        writes it to `out_device_ptr`."""
        if device_ptrs is not None:
            n_routed, n = shape
            w, s, mem = C.c_void_p(int(device_ptrs[0])), C.c_void_p(int(device_ptrs[1])), BSX_DEVICE
        else:
            wires, sigmas = _np(wires), _np(sigmas)
            if wires.ndim != 2 or wires.shape != sigmas.shape:
                raise BsxError("wires and sigmas must both be (n_routed, n)")
            n_routed, n = wires.shape
            w, s, mem = _ptr(wires), _ptr(sigmas), BSX_HOST
        k_is, betas, gammas = _np(k_is).reshape(-1), _np(betas).reshape(-1), _np(gammas).reshape(-1)
        if k_is.shape[0] != n_routed or betas.shape != gammas.shape:
            raise BsxError("k_is / betas / gammas do not match the wires")
        n_chunks = -(-n_routed // int(max_degree)) if max_degree > 0 else 0
        out = None
        if out_device_ptr is None:
            out = np.zeros((betas.shape[0] * n_chunks, n), dtype=np.uint64)
        self.check(lib().bsx_partial_products_and_zs(
            self._h, w, s, mem, _log2_strict(n), int(n_routed), int(max_degree), _ptr(k_is), _ptr(betas), _ptr(gammas),
            int(betas.shape[0]), _ptr(out) if out is not None else C.c_void_p(int(out_device_ptr)),
            BSX_HOST if out is not None else BSX_DEVICE))
        return out

    # ---- test hooks
    def poseidon_permute(self, states):
        s = _np(states).copy()
        flat = s.reshape(-1, 12)
        self.check(lib().bsx_poseidon_permute(self._h, _ptr(flat), BSX_HOST, flat.shape[0]))
        return s

    def poseidon_gate_witness(self, rows):
        """gates/poseidon.rs PoseidonGenerator on (n, 135) rows: inputs (wires 0..11) and swap (wire 24) are read, the delta
        wires, the S-box input wires and the outputs are filled in.  Returns the completed copy."""
        r = np.ascontiguousarray(_np(rows)).copy()
        if r.ndim != 2 or r.shape[1] != 135:
            raise BsxError("PoseidonGate rows are (n, 135)")
        self.check(lib().bsx_poseidon_gate_witness(self._h, _ptr(r), BSX_HOST, r.shape[0]))
        return r

    def salt_stream(self, key, n, raw=False):
        """n words of the ChaCha20 stream blinding salts come from, under `key` (8 x u32); raw: no rejection mod p."""
        k = np.ascontiguousarray(key, dtype=np.uint32).reshape(8)
        out = np.zeros(int(n), dtype=np.uint64)
        self.check(lib().bsx_debug_salt_stream(self._h, C.c_void_p(k.ctypes.data), 1 if raw else 0, _ptr(out), out.size))
        return out

    def ntt(self, cols, kind="fft", shift=7):
        """cols: (n_cols, n) column vectors; natural order in and out."""
        a = _np(cols).copy()
        two_d = a.reshape(-1, a.shape[-1])
        k = {"fft": 0, "ifft": 1, "coset_fft": 2, "coset_ifft": 3}[kind]
        self.check(lib().bsx_ntt(self._h, _ptr(two_d), BSX_HOST, _log2_strict(two_d.shape[1]), two_d.shape[0], k,
                                 C.c_uint64(shift)))
        return a


class MerkleTree:
    """plonky2 hash/merkle_tree.rs `MerkleTree<GoldilocksField, PoseidonHash>`, device resident."""

    def __init__(self, ctx, leaves=None, cap_height=0, _handle=None, _owner=None):
        """`MerkleTree::new(leaves, cap_height)`; leaves: (n_leaves, leaf_len) row-major."""
        self.ctx = ctx
        self._owner = _owner  # a PolynomialBatch keeping the handle alive
        if _handle is not None:
            self._h = _handle
            return
        self._h = None
        leaves = _np(leaves)
        if leaves.ndim != 2:
            raise BsxError("leaves must be (n_leaves, leaf_len)")
        log_leaves = _log2_strict(leaves.shape[0])
        h = C.c_void_p()
        ctx.check(lib().bsx_merkle_new(ctx._h, _ptr(leaves), BSX_HOST, log_leaves, leaves.shape[1], int(cap_height),
                                       C.byref(h)))
        self._h = h

    @classmethod
    def from_device_colmajor(cls, ctx, dev_ptr, col_stride, log_leaves, leaf_len, cap_height):
        h = C.c_void_p()
        ctx.check(lib().bsx_merkle_new_colmajor_device(ctx._h, C.c_void_p(int(dev_ptr)), int(col_stride),
                                                       int(log_leaves), int(leaf_len), int(cap_height), C.byref(h)))
        return cls(ctx, _handle=h)

    def shape(self):
        a, b, c = C.c_int32(), C.c_int32(), C.c_int32()
        self.ctx.check(lib().bsx_merkle_shape(self._h, C.byref(a), C.byref(b), C.byref(c)))
        return a.value, b.value, c.value

    @property
    def cap(self):
        _, _, ch = self.shape()
        out = np.zeros((1 << ch, 4), dtype=np.uint64)
        self.ctx.check(lib().bsx_merkle_cap(self._h, _ptr(out)))
        return out

    def get(self, leaf_index):
        """`MerkleTree::get(i)` for one index or a list of indices."""
        _, leaf_len, _ = self.shape()
        idx = _np(np.atleast_1d(leaf_index))
        out = np.zeros((idx.size, leaf_len), dtype=np.uint64)
        self.ctx.check(lib().bsx_merkle_get(self._h, _ptr(idx), idx.size, _ptr(out)))
        return out[0] if np.isscalar(leaf_index) else out

    def prove(self, leaf_index):
        """`MerkleTree::prove(leaf_index).siblings`, bottom-up."""
        log_leaves, _, ch = self.shape()
        out = np.zeros((log_leaves - ch, 4), dtype=np.uint64)
        self.ctx.check(lib().bsx_merkle_prove(self._h, C.c_uint64(int(leaf_index)), _ptr(out)))
        return out

    @property
    def digests(self):
        """`MerkleTree.digests` in plonky2's Vec layout."""
        log_leaves, _, ch = self.shape()
        out = np.zeros((2 * ((1 << log_leaves) - (1 << ch)), 4), dtype=np.uint64)
        self.ctx.check(lib().bsx_merkle_digests(self._h, _ptr(out)))
        return out

    def close(self):
        if self._h is not None and self._owner is None:
            lib().bsx_merkle_free(self._h)
        self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class PolynomialBatch:
    """plonky2 fri/oracle.rs `PolynomialBatch<GoldilocksField, PoseidonGoldilocksConfig, 2>`."""

    def __init__(self, ctx, handle, timing=None):
        self.ctx = ctx
        self._h = handle
        self.timing = timing

    @staticmethod
    def _build(fn, ctx, cols, rate_bits, blinding, cap_height, salt, device_ptr, shape, timing):
        if device_ptr is None and isinstance(cols, (list, tuple)):
            # one separately allocated vector per polynomial: `Vec<PolynomialValues<F>>` as the Rust shim sees it
            vecs = [_np(c).reshape(-1) for c in cols]
            n = vecs[0].shape[0] if vecs else 1
            if any(v.shape[0] != n for v in vecs):
                raise BsxError("polynomials of different lengths")
            ptrs = (C.c_void_p * max(len(vecs), 1))(*[v.ctypes.data for v in vecs])
            fn_cols = lib().bsx_batch_from_coeffs_cols if fn is lib().bsx_batch_from_coeffs else lib().bsx_batch_from_values_cols
            h = C.c_void_p()
            t = Timing() if timing else None
            ctx.check(fn_cols(ctx._h, ptrs, _log2_strict(n), len(vecs), int(rate_bits), int(cap_height),
                              1 if blinding else 0, _ptr(_np(salt)) if salt is not None else None, C.byref(h),
                              C.byref(t) if t is not None else None))
            return PolynomialBatch(ctx, h, t)
        if device_ptr is not None:
            n_cols, n = shape
            src, mem = C.c_void_p(int(device_ptr)), BSX_DEVICE
        else:
            cols = _np(cols)
            if cols.ndim != 2:
                raise BsxError("expected (n_cols, n) column vectors")
            n_cols, n = cols.shape
            src, mem = _ptr(cols), BSX_HOST
        log_n = _log2_strict(n)
        if salt is not None and mem == BSX_HOST:
            salt = _np(salt)
        h = C.c_void_p()
        t = Timing() if timing else None
        ctx.check(fn(ctx._h, src, mem, log_n, n_cols, int(rate_bits), int(cap_height), 1 if blinding else 0,
                     _ptr(salt), C.byref(h), C.byref(t) if t is not None else None))
        return PolynomialBatch(ctx, h, t)

    @classmethod
    def from_values(cls, ctx, values, rate_bits, blinding, cap_height, salt=None, device_ptr=None, shape=None,
                    timing=False):
        """`PolynomialBatch::from_values(values, rate_bits, blinding, cap_height, timing, fft_root_table)`.
        values: (n_cols, n) -- one row per PolynomialValues.  `device_ptr`/`shape` pass a device buffer."""
        return cls._build(lib().bsx_batch_from_values, ctx, values, rate_bits, blinding, cap_height, salt,
                          device_ptr, shape, timing)

    @classmethod
    def from_coeffs(cls, ctx, coeffs, rate_bits, blinding, cap_height, salt=None, device_ptr=None, shape=None,
                    timing=False):
        """`PolynomialBatch::from_coeffs(polynomials, rate_bits, blinding, cap_height, timing, fft_root_table)`."""
        return cls._build(lib().bsx_batch_from_coeffs, ctx, coeffs, rate_bits, blinding, cap_height, salt,
                          device_ptr, shape, timing)

    @classmethod
    def from_quotient_values(cls, ctx, quotient_values, degree_log, quotient_degree_factor, rate_bits, blinding,
                             cap_height, salt=None, device_ptr=None, shape=None, timing=False):
        """plonk/prover.rs, between quotient evaluation and the quotient commitment: every row of `quotient_values`
        (the values of one challenge's quotient on the coset 7 * <omega> of size 2^degree_log * 2^q, natural order) goes
        through `coset_ifft(F::coset_shift())`, `trim_to_len(quotient_degree_factor * 2^degree_log)` and
        `.chunks(2^degree_log)`; the chunks are committed with `PolynomialBatch::from_coeffs`.  Raises where upstream
        panics (coefficients beyond the quotient degree)."""
        if device_ptr is not None:
            n_polys, big = shape
            src, mem = C.c_void_p(int(device_ptr)), BSX_DEVICE
        else:
            vals = _np(quotient_values)
            if vals.ndim != 2:
                raise BsxError("expected (n_polys, 2^(degree_log + q)) value vectors")
            n_polys, big = vals.shape
            src, mem = _ptr(vals), BSX_HOST
        q_bits = _log2_strict(big) - int(degree_log)
        if q_bits < 0:
            raise BsxError("quotient value vectors are shorter than the degree")
        h = C.c_void_p()
        t = Timing() if timing else None
        ctx.check(lib().bsx_batch_from_quotient_values(
            ctx._h, src, mem, int(degree_log), q_bits, int(quotient_degree_factor), int(n_polys), int(rate_bits),
            int(cap_height), 1 if blinding else 0, _ptr(_np(salt)) if salt is not None else None, C.byref(h),
            C.byref(t) if t is not None else None))
        return PolynomialBatch(ctx, h, t)

    @classmethod
    def from_parts(cls, ctx, coeffs, leaves, digests, cap, rate_bits):
        """Restore a batch from `export()`'s parts without recomputing anything (the circuit-artefact reader,
        plonky2 util/serialization `read_polynomial_batch`).  coeffs (n_cols, n) natural order; leaves
        (n << rate_bits, n_cols + salt) in leaf order; digests in plonky2's `digests` layout; cap (2^h, 4)."""
        coeffs, leaves, digests, cap = _np(coeffs), _np(leaves), _np(digests), _np(cap)
        n_cols, n = coeffs.shape
        log_n, cap_height = _log2_strict(n), _log2_strict(cap.shape[0])
        if leaves.shape[0] != n << rate_bits:
            raise BsxError("leaves.len() = %d, expected %d" % (leaves.shape[0], n << rate_bits))
        n_salt = leaves.shape[1] - n_cols
        want = 2 * ((n << rate_bits) - (1 << cap_height))
        if digests.reshape(-1, 4).shape[0] != want:
            raise BsxError("digests.len() = %d, expected %d" % (digests.reshape(-1, 4).shape[0], want))
        h = C.c_void_p()
        ctx.check(lib().bsx_batch_import(ctx._h, _ptr(coeffs), _ptr(leaves), _ptr(digests), _ptr(cap), log_n, n_cols,
                                         n_salt, int(rate_bits), cap_height, C.byref(h)))
        return PolynomialBatch(ctx, h, None)

    def to_bytes(self):
        """plonky2 util/serialization `Write::write_polynomial_batch` (layout in serialization.py)."""
        from .serialization import write_polynomial_batch
        sh, parts = self.shape(), self.export()
        return write_polynomial_batch(parts["coeffs"], parts["leaves"], parts["digests"], parts["cap"],
                                      sh["degree_log"], sh["rate_bits"], sh["n_salt"] != 0)

    @classmethod
    def from_bytes(cls, ctx, data):
        """`Read::read_polynomial_batch`: the inverse of `to_bytes`; malformed input raises BsxError."""
        from .serialization import read_polynomial_batch
        try:
            d = read_polynomial_batch(data)
        except ValueError as e:
            raise BsxError(str(e)) from None
        return cls.from_parts(ctx, d["coeffs"], d["leaves"], d["digests"], d["cap"], d["rate_bits"])

    def shape(self):
        v = [C.c_int32() for _ in range(5)]
        self.ctx.check(lib().bsx_batch_shape(self._h, *[C.byref(x) for x in v]))
        return dict(zip(("degree_log", "n_cols", "rate_bits", "cap_height", "n_salt"), [x.value for x in v]))

    @property
    def merkle_tree(self):
        return MerkleTree(self.ctx, _handle=C.c_void_p(lib().bsx_batch_tree(self._h)), _owner=self)

    @property
    def cap(self):
        sh = self.shape()
        out = np.zeros((1 << sh["cap_height"], 4), dtype=np.uint64)
        self.ctx.check(lib().bsx_batch_cap(self._h, _ptr(out)))
        return out

    @property
    def polynomials(self):
        """Coefficient vectors in natural order, (n_cols, n)."""
        sh = self.shape()
        out = np.zeros((sh["n_cols"], 1 << sh["degree_log"]), dtype=np.uint64)
        self.ctx.check(lib().bsx_batch_coeffs(self._h, _ptr(out), BSX_HOST))
        return out

    def get_lde_values(self, index, step=1):
        """`PolynomialBatch::get_lde_values(index, step)`: leaf reverse_bits(index*step), salt stripped."""
        out = np.zeros(self.shape()["n_cols"], dtype=np.uint64)
        self.ctx.check(lib().bsx_batch_get_lde_values(self._h, C.c_uint64(int(index)), C.c_uint64(int(step)),
                                                      _ptr(out)))
        return out

    def get_lde_values_packed(self, index_start, step=1, width=4):
        """`PolynomialBatch::get_lde_values_packed::<P>(index_start, step)` with P::WIDTH = width: (n_cols, width),
        lane i of every column = get_lde_values(index_start + i, step)."""
        rows = np.zeros((int(width), self.shape()["n_cols"]), dtype=np.uint64)
        self.ctx.check(lib().bsx_batch_get_lde_values_packed(self._h, C.c_uint64(int(index_start)),
                                                             C.c_uint64(int(width)), C.c_uint64(int(step)),
                                                             _ptr(rows), BSX_HOST))
        return np.ascontiguousarray(rows.T)

    def eval_commitment(self, points, col_begin=0, n_cols=None):
        """The `eval_commitment(z, c)` closure of plonk/proof.rs `OpeningSet::new`: every polynomial of the batch
        evaluated at extension points.  points: (k, 2) or (2,); returns (k, n_cols, 2) or (n_cols, 2)."""
        pts = np.ascontiguousarray(points, dtype=np.uint64)
        single = pts.ndim == 1
        pts = pts.reshape(-1, 2)
        if n_cols is None:
            n_cols = self.shape()["n_cols"] - col_begin
        out = np.zeros((pts.shape[0], max(n_cols, 0), 2), dtype=np.uint64)
        self.ctx.check(lib().bsx_batch_eval_ext(self._h, int(col_begin), int(n_cols), _ptr(pts), pts.shape[0],
                                                _ptr(out)))
        return out[0] if single else out

    def export(self, coeffs=True, leaves=True, digests=True):
        sh = self.shape()
        n, nb = 1 << sh["degree_log"], 1 << (sh["degree_log"] + sh["rate_bits"])
        w = sh["n_cols"] + sh["n_salt"]
        c = np.zeros((sh["n_cols"], n), dtype=np.uint64) if coeffs else None
        lv = np.zeros((nb, w), dtype=np.uint64) if leaves else None
        d = np.zeros((2 * (nb - (1 << sh["cap_height"])), 4), dtype=np.uint64) if digests else None
        self.ctx.check(lib().bsx_batch_export(self._h, _ptr(c), _ptr(lv), _ptr(d)))
        return {"coeffs": c, "leaves": lv, "digests": d, "cap": self.cap}

    @property
    def lde_device_ptr(self):
        return lib().bsx_batch_lde_device(self._h)

    def close(self):
        if self._h is not None:
            lib().bsx_batch_free(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
