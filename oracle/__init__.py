"""CPU oracle for the KKT evaluation path -- TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this package;
the product (asset_asrl_b200/) never does.  See oracle/oracle.cpp for what is restated and how it is pinned
("parity unpinned" against the reference binary, which cannot be built in this image).
"""
from .binding import Oracle, computeall, build, hardware_threads, use_variant  # noqa: F401
