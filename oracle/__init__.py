"""TEST INFRASTRUCTURE — CPU restatements of the reference algorithms (checkers only).

Nothing under oracle/ is imported by the product package; only tests/, bench.py's
cpu_baseline / --impl reference legs and __graft_entry__.smoke() may use it.
"""
