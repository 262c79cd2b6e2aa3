"""CPU checkers (TEST INFRASTRUCTURE ONLY): see oracle/oracle.py and oracle/tamcmc_oracle.c."""
