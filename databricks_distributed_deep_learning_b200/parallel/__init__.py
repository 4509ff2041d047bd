"""Parallel runtime: process group, symmetric memory, collectives, DDP and FSDP wrappers."""
