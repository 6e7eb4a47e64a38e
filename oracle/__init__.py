"""CPU oracle (test infrastructure only; see oracle/photon_oracle.cpp)."""
