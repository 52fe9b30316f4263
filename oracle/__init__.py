"""TEST INFRASTRUCTURE ONLY.

CPU checkers for the CUDA pair-HMM path: ``pyoracle`` wraps our plain-C restatement (oracle/pairhmm_oracle.c),
``pyref`` wraps the unmodified reference C library compiled into oracle/_ref/ (only where it was built).
Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this package;
the product package (marginalign_trna_b200) never does.
"""
