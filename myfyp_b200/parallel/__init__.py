"""Data plane: device placement, parameter slabs, peer/IPC mapping, NCCL baseline."""
