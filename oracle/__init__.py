"""TEST INFRASTRUCTURE: CPU oracle of the reference physics step (see oracle/vmas_oracle.h).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import this package.
"""
