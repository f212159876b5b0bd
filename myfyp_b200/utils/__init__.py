"""Utilities: topologies, wait helpers, seeding, fault injection, checkpoints."""
