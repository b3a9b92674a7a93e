"""b2nn -- B200-native (sm_100a) training backend behind the asheptunov/graph-neural-net Java API.

The directory name carries a hyphen, so load the package through `__graft_entry__.load_package()`
(registers it as module `gnn_b200`).  Importing it binds libb2nn.so; there is no CPU fallback.
"""
from . import _capi as capi
from . import dp, functions
from .functions import (HALF_SQUARED_ERROR, HALF_SQUARED_ERROR_PRIME, IDENTITY, IDENTITY_PRIME, LEAKY_RELU,
                        LEAKY_RELU_PRIME, RELU, RELU_PRIME, SIGMOID, SIGMOID_PRIME, TANH, TANH_PRIME,
                        ActivationFunction, ActivationPrime, LossFunction, LossFunctionPrime)
from .mnist import MNISTTrainer
from .nets import GeneralNeuralNet, NeuralNet, SoftmaxCrossEntropyNeuralNet, commUniqueId
from .trainer import Dataset, NeuralNetTrainer, ProgressBar, Sampler

capi.lib()  # fail loudly at import time when the CUDA library has not been built
