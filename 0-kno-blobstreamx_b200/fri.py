"""Host-side FRI configuration mirror (plonky2 fri/mod.rs, fri/reduction_strategies.rs)."""
from dataclasses import dataclass, field
from typing import List


def fri_reduction_arity_bits(degree_bits, rate_bits, cap_height, arity_bits=4, final_poly_bits=5):
    """`FriReductionStrategy::ConstantArityBits(arity_bits, final_poly_bits).reduction_arity_bits(..)`."""
    out = []
    while degree_bits > final_poly_bits and degree_bits + rate_bits - arity_bits >= cap_height:
        out.append(arity_bits)
        assert degree_bits >= arity_bits
        degree_bits -= arity_bits
    return out


@dataclass
class FriConfig:
    """plonky2 `FriConfig`; defaults are `CircuitConfig::standard_recursion_config().fri_config`."""
    rate_bits: int = 3
    cap_height: int = 4
    proof_of_work_bits: int = 16
    arity_bits: int = 4
    final_poly_bits: int = 5
    num_query_rounds: int = 28

    def fri_params(self, degree_bits, hiding=False):
        return FriParams(self, hiding, degree_bits,
                         fri_reduction_arity_bits(degree_bits, self.rate_bits, self.cap_height, self.arity_bits,
                                                  self.final_poly_bits))


@dataclass
class FriParams:
    """plonky2 `FriParams`."""
    config: FriConfig
    hiding: bool
    degree_bits: int
    reduction_arity_bits: List[int] = field(default_factory=list)

    def total_arities(self):
        return sum(self.reduction_arity_bits)

    def lde_bits(self):
        return self.degree_bits + self.config.rate_bits

    def final_poly_bits(self):
        return self.degree_bits - self.total_arities()

    def final_poly_len(self):
        return 1 << self.final_poly_bits()
