"""CPU oracle for the vit-colmap hot path.  TEST INFRASTRUCTURE, NOT PRODUCT CODE.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may import this package.  vit_colmap_b200/ never does.
"""
