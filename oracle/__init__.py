"""CPU oracle (test infrastructure only — see fp8_oracle.py / oracle_c.c headers)."""
