"""TEST INFRASTRUCTURE -- CPU oracle for the Smith-Waterman hot path.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
reference legs may import this package.  Nothing under mgl_b200/ does.
"""
