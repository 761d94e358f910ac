"""TEST INFRASTRUCTURE ONLY: CPU oracle for the RT_bench.py hot path (see rt_oracle.c)."""
