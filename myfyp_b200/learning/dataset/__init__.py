from myfyp_b200.learning.dataset.array_dataset import ArrayDataset, ArrayDatasetDict
from myfyp_b200.learning.dataset.p2pfl_dataset import DataExportStrategy, P2PFLDataset, PyTorchExportStrategy, TensorExportStrategy
from myfyp_b200.learning.dataset.partition_strategies import (
    DataPartitionStrategy,
    DirichletPartitionStrategy,
    LabelSkewedPartitionStrategy,
    PercentageBasedNonIIDPartitionStrategy,
    RandomIIDPartitionStrategy,
)

__all__ = [
    "ArrayDataset", "ArrayDatasetDict", "DataExportStrategy", "P2PFLDataset", "PyTorchExportStrategy", "TensorExportStrategy",
    "DataPartitionStrategy", "DirichletPartitionStrategy", "LabelSkewedPartitionStrategy",
    "PercentageBasedNonIIDPartitionStrategy", "RandomIIDPartitionStrategy",
]
