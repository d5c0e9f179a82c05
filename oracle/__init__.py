"""oracle/ — TEST INFRASTRUCTURE ONLY.

CPU restatements of the reference's algorithm for the pairwise matching path, used as the
parity checker by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
reference legs.  Nothing in the product package imports this.
"""
