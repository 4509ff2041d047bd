"""Utilities: synthetic data, checkpoints, export/inference comparison, timing."""
