"""Baseline harness: runs the reference implementation for side-by-side numbers."""
