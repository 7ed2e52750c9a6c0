"""oracle/ -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

CPU restatements of the reference algorithms on the hot path (and, when it was
built here, the reference's own compiled Cython core).  Only tests/,
__graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this package.  glowtts_b200/ never does.
"""
