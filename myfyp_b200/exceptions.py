"""Framework exceptions (reference: p2pfl/exceptions.py:21-36)."""


class NodeRunningException(Exception):
    """The node is (not) running when the opposite was required."""


class LearnerRunningException(Exception):
    """Learner/model/data changed while learning is in progress."""


class ZeroRoundsException(Exception):
    """set_start_learning called with rounds < 1."""
