"""CPU oracle for the DiLoCo outer-sync hot path.  TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
reference legs may import this package.  nanodiloco_b200/ never does.
"""
