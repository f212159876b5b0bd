"""Flax binding (reference: learning/frameworks/flax/).  Host-side, like the Keras one: needs ``jax``, ``flax`` and ``optax``,
none of which is part of this image; exercised by ``tests/test_frameworks_optional.py`` only where they are installed."""
