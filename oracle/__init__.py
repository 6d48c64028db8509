"""CPU oracle for the invert_ng / predict_ng hot path -- TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
may import this package.  The product (ska_sdp_func_python_b200) never does.
"""
