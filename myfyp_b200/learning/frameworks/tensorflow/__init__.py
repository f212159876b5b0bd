"""TensorFlow / Keras binding (reference: learning/frameworks/tensorflow/).  Host-side: a Keras model takes part in a
federation through the wire format (lists of NumPy arrays); the B200 kernels are reached through the PyTorch binding only.
Importing the sub-modules needs ``tensorflow``; it is not part of this image, so these modules are exercised by
``tests/test_frameworks_optional.py`` only where it is installed."""
