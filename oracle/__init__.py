"""TEST INFRASTRUCTURE ONLY -- CPU restatement of the reference hot path.

Nothing under oracle/ may be imported by the product package
(show-attend-and-tell_b200/).  Only tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline / --impl reference legs use it, and only as the
checker / the timed CPU baseline.
"""
