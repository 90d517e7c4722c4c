"""CPU oracle for the KMeans hot path -- TEST INFRASTRUCTURE ONLY (see hkm_oracle.c header).

Only tests/, __graft_entry__.smoke() and bench.py (cpu_baseline / --impl reference) import this.
"""
