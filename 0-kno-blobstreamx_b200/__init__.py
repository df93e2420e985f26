"""blobstreamx_b200 -- B200-native plonky2 commitment core (host-side mirror over the C ABI).

The product is `csrc/libbsx_b200.so` (hand-written sm_100a CUDA behind `include/bsx_b200.h`).
This package is the thin host mirror of the plonky2 interface the BlobstreamX prover calls on
its hot path -- `PolynomialBatch::from_values / from_coeffs / get_lde_values`, `MerkleTree::new /
get / prove` and `fri_proof` -- with the same names, argument meaning and error behaviour
(plonky2 panics -> Python exceptions), bound with ctypes.  The reference's host language is Rust,
which this image does not have; INTEGRATION.md carries the Rust shim.

There is NO CPU fallback: importing works anywhere (so the ABI can be checked on a CPU box),
but creating a `Context` without a CUDA device, or without the built library, raises.

The directory name (`0-kno-blobstreamx_b200`) is not a Python identifier; load it with
`__graft_entry__.load_package()` which registers it as `blobstreamx_b200`.
"""
from .binding import (  # noqa: F401
    BSX_DEVICE,
    BSX_HOST,
    BsxError,
    Context,
    MerkleTree,
    PolynomialBatch,
    Timing,
    lib,
    library_path,
    EXPORTED_SYMBOLS,
)
from .fri import (  # noqa: F401
    Challenger,
    FriConfig,
    FriInstanceInfo,
    FriParams,
    final_poly,
    fri_proof,
    fri_proof_struct,
    fri_proof_to_bytes,
    fri_proof_words,
    fri_reduction_arity_bits,
    prove_openings,
)
from .proof import (  # noqa: F401
    FriLayout,
    OpeningSet,
    Proof,
    ProofWithPublicInputs,
    fri_proof_bytes,
    fri_proof_layout,
)
from .plonk import (  # noqa: F401
    GATE_ARITHMETIC,
    GATE_ARITHMETIC_EXTENSION,
    GATE_BASE_SUM,
    GATE_CONSTANT,
    GATE_COSET_INTERPOLATION,
    GATE_EXPONENTIATION,
    GATE_MUL_EXTENSION,
    GATE_NOOP,
    GATE_POSEIDON,
    GATE_POSEIDON_MDS,
    GATE_PUBLIC_INPUT,
    GATE_RANDOM_ACCESS,
    GATE_REDUCING,
    GATE_REDUCING_EXTENSION,
    CircuitDescription,
    Gate,
    SyntheticCircuit,
    circuit_digest,
    compute_quotient_polys,
    hash_no_pad,
    prove,
    quotient_values,
    synthetic_circuit,
)
from .parallel import (  # noqa: F401
    Comm,
    CommShardedBatch,
    comm_prove_openings,
    ShardPlan,
    ShardedPolynomialBatch,
    assemble_cap,
    device_tensor,
    exchange_staged,
    final_poly_part,
    fri_from_final_poly,
    sharded_prove_openings,
    sum_parts,
    top_siblings,
)

P = 0xFFFFFFFF00000001
