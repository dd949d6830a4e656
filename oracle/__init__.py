"""CPU oracle (test infrastructure only -- never imported by opticalflow_b200)."""
