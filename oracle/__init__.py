"""CPU oracle of the DAGR hot path -- TEST INFRASTRUCTURE ONLY (see oracle/evg_oracle.c, oracle/dagr_ops.py)."""
