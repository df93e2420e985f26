"""CPU oracle (test infrastructure only) -- see oracle/plonky2_oracle.h."""
