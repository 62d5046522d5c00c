"""Utilities: output (PNG/raw), configuration, profiling hooks, logging."""
