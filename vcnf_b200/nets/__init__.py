from . import mlp, resnet, made, cnn, varquant
from .mlp import MLP
from .cnn import ConvNet2d
from .resnet import ResidualNet, ConvResidualNet
from .made import MADE
from .varquant import Dequantization, VariationalDequantization
