"""CPU oracle for the visnav front-end hot path -- TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may import this package.  The product (visual-slam_b200/, include/) never does.
"""
from .oracle import Oracle, load_oracle, load_ref, build  # noqa: F401
