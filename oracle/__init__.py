"""CPU oracle of the GHOST face-swap hot path - TEST INFRASTRUCTURE ONLY.

Nothing under ghost_b200/ imports this package.  Only tests/, __graft_entry__.smoke() and the
cpu_baseline / --impl reference legs of bench.py may use it, and only as the checker / the
timed CPU baseline, never as part of the product path.
"""
