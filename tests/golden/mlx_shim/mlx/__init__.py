"""numpy-backed stand-in package for `mlx` (test infrastructure, see core.py)."""
