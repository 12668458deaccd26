"""CPU oracle for the hfops operator path — TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py (cpu_baseline / --impl reference) may
import this package; the product (heterofusionrcnn_b200/) never does.
"""
