"""CPU oracle for the segment()/link() hot path -- TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may import this package; ultrack_b200 never does.
"""
