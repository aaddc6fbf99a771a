"""CPU oracle for the 3DGS-Avatar hot path.  TEST INFRASTRUCTURE ONLY (see gs_oracle.c)."""
