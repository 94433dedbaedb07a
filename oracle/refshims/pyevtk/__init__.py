"""Stub for `pyevtk` (absent).  TEST INFRASTRUCTURE ONLY."""
