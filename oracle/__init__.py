"""TEST INFRASTRUCTURE ONLY: CPU oracle for the bounce loop (see numpy_engine.py).
Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may import this package."""
