"""CPU oracle of the SAM hot path: test infrastructure only (see sam_oracle.py header)."""
