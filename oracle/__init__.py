"""CPU oracle for the Q4_K x Q8_K mul_mat path — TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import
this package.  The product path (ggml-neon-opt_b200/) never does.

  oracle.port  ctypes binding of oracle/q4k_oracle.c  (our plain-C restatement)
  oracle.ref   ctypes binding of oracle/_ref/*.so     (the unmodified reference, built by oracle/Makefile)
"""
