"""CPU oracle for the dipole-field hot path.  TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
reference legs may import this package; the product path
(dipole-normal-prop_b200/) never does.
"""
