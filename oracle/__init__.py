"""TEST INFRASTRUCTURE -- CPU oracle for the cut-FEM hot path.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
leg may import this package.  The product (ngsxfem_b200) never does.
"""
