"""PyTorch binding under the reference's module paths (the implementation lives in ``learning/torch_learner.py`` / ``p2pfl_model.py``)."""
