"""TEST INFRASTRUCTURE ONLY — parity checkers (see oracle/pinyon_oracle.c, oracle/ref_harness.cc)."""
