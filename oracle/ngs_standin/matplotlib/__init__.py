"""matplotlib is not installed in this image; heat.py:2 imports pyplot without using it in the path under test.
Only on sys.path inside oracle/make_golden.py."""
