"""Model zoo: the reference's MNIST MLP plus the CNN / ResNet-18 CIFAR models the north star names."""

from myfyp_b200.models.cnn import CNN, cnn_model
from myfyp_b200.models.mlp import MLP, mlp_model
from myfyp_b200.models.resnet import ResNet18, resnet18_model

__all__ = ["MLP", "mlp_model", "CNN", "cnn_model", "ResNet18", "resnet18_model"]
