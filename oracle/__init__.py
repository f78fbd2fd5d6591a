"""TEST INFRASTRUCTURE ONLY — CPU oracle for the PLS fit / CV path.

`oracle.port`  : ctypes binding of oracle/liboracle.so, our plain-C restatement of the reference
                 algorithm (pls_oracle.c; every function cites the reference lines it follows).
`oracle.ref`   : ctypes binding of oracle/_ref/libplsref.so, the UNMODIFIED reference compiled from
                 /root/reference by oracle/Makefile (present only where it was built).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / `--impl reference` legs may
import this package.  The product (pls_b200, libplsb.so) never does and has no CPU fallback.
"""
