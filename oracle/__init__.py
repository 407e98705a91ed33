"""Oracle package: TEST INFRASTRUCTURE ONLY (parity unpinned, see raster_oracle.py).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
may import from here.  The product package never does.
"""
