"""Stub for `matplotlib` (absent).  TEST INFRASTRUCTURE ONLY."""
